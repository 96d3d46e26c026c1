// ConvTranspose3d with kernel = stride = 2 (the U-Net up-sampling, UNet.py:265-269) as ONE tensor-core GEMM launch.
//   y[n, co, 2d+pd, 2h+ph, 2w+pw] = b[co] + sum_ci x[n, ci, d, h, w] * W[ci, co, pd, ph, pw]
// i.e. D[input voxel][(parity class, co)] = A[voxel][ci] . B[ci][(class, co)]: every input voxel produces its 2x2x2
// output voxels at once.  The generic path runs it as the stride-2 data gradient, one launch per parity class: the
// input is read 8 times and every launch writes every other voxel (4.0 ms per 256^3 U-Net block).  Here the input
// tile is loaded once and stays in shared memory while the 8*Cout weight columns stream past it in chunks of 256;
// the layer is bound by the HBM write of the 8x larger output, which the epilogue warps stream out (64 bytes per
// thread and column group) into arbitrary strides -- e.g. the channel slice of the skip-concat buffer.
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected thread), warps 2-9 = epilogue
// (two per TMEM lane quadrant, 128 of the 256 chunk columns each).  Two 256-column accumulators in TMEM.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace rg {

constexpr int UP_THREADS = 320;
constexpr int UP_TH = 16, UP_TW = 8;           // input-voxel tile (one depth plane)
constexpr int UP_A_SUB = 128 * 128;            // one 64-channel chunk of the A tile
constexpr int UP_B_STAGE = 256 * 128;          // 256 weight rows x 64 reduction channels
constexpr int UP_MAX_NB = 6, UP_MAX_COUT = 512;
constexpr int UP_STAGE_OUT = 32 * 64;           // epilogue staging: 32 voxels x 32 channels bf16, two buffers per warp
constexpr int UP_OUT_BYTES = 8 * 2 * UP_STAGE_OUT;
constexpr int UP_SMEM_LIMIT = 227 * 1024;

struct alignas(64) UpParams {
  CUtensorMap a_map;       // 5-D {Cin, W, H, D, N}; box {64, 8, 16, 1, 1}
  CUtensorMap b_map;       // 2-D {Cin, 8*Cout};     box {64, 256}
  CUtensorMap y_maps[8];   // per parity class: 5-D {Cout, W, H, D, N} view of y with doubled strides; box {32, 8, 4, 1, 1}
  int cin, cout, kchunks, nchunks;   // Cin / 64, 8*Cout / 256
  int na, nb;                        // A tile buffers (1 or 2), B stages
  int in_d, in_h, in_w, batch;
  int tiles_h, tiles_w;
  long long total_tiles;
  const float* bias;
  int act; float slope;
};

__device__ __forceinline__ void up_tile_coords(const UpParams& p, long long t, int& n, int& d, int& h0, int& w0) {
  unsigned u = (unsigned)t;
  const unsigned iw = u % (unsigned)p.tiles_w; u /= (unsigned)p.tiles_w;
  const unsigned ih = u % (unsigned)p.tiles_h; u /= (unsigned)p.tiles_h;
  d = (int)(u % (unsigned)p.in_d); n = (int)(u / (unsigned)p.in_d);
  h0 = (int)ih * UP_TH; w0 = (int)iw * UP_TW;
}

__global__ void __launch_bounds__(UP_THREADS, 1) upconv2_kernel(const __grid_constant__ UpParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);
  const int a_bytes = p.kchunks * UP_A_SUB;
  const uint32_t smem_a = base;
  const uint32_t smem_b = base + (uint32_t)(p.na * a_bytes);
  uint8_t* out_stage = base_ptr + (size_t)p.na * a_bytes + (size_t)p.nb * UP_B_STAGE;      // 1024-aligned
  uint8_t* tail = out_stage + UP_OUT_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);   // afull[2] aempty[2] bfull[NB] bempty[NB] tfull[2] tempty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 8 * (8 + 2 * UP_MAX_NB));
  float* bias_s = reinterpret_cast<float*>(tail + 8 * (8 + 2 * UP_MAX_NB) + 64);
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto afull = [&](int s) { return bar0 + 8u * s; };
  auto aempty = [&](int s) { return bar0 + 8u * (2 + s); };
  auto bfull = [&](int s) { return bar0 + 8u * (4 + s); };
  auto bempty = [&](int s) { return bar0 + 8u * (4 + UP_MAX_NB + s); };
  auto tfull = [&](int s) { return bar0 + 8u * (4 + 2 * UP_MAX_NB + s); };
  auto tempty = [&](int s) { return bar0 + 8u * (6 + 2 * UP_MAX_NB + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(afull(s), 1); ptx::mbar_init(aempty(s), 1); }
    for (int s = 0; s < p.nb; ++s) { ptx::mbar_init(bfull(s), 1); ptx::mbar_init(bempty(s), 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(tfull(s), 1); ptx::mbar_init(tempty(s), 256); }
    ptx::fence_mbar_init();
    ptx::prefetch_tmap(&p.a_map);
    ptx::prefetch_tmap(&p.b_map);
  }
  if (warp == 1) ptx::tmem_alloc<512>(ptx::smem_u32(tmem_slot));
  for (int i = threadIdx.x; i < p.cout; i += UP_THREADS) bias_s[i] = p.bias ? p.bias[i] : 0.f;
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer =================
      int b_i = 0; uint32_t b_ph = 0;
      long long i = 0;
      for (long long t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
        int n, d, h0, w0;
        up_tile_coords(p, t, n, d, h0, w0);
        const int buf = (int)(i % p.na);
        const uint32_t aph = (uint32_t)((i / p.na) & 1);
        ptx::mbar_wait_backoff(aempty(buf), aph ^ 1u);
        ptx::mbar_arrive_expect_tx(afull(buf), (uint32_t)a_bytes);
        for (int kc = 0; kc < p.kchunks; ++kc)
          ptx::tma_load_5d(smem_a + (uint32_t)(buf * a_bytes + kc * UP_A_SUB), &p.a_map, afull(buf), kc * 64, w0, h0, d, n);
        for (int nc = 0; nc < p.nchunks; ++nc)
          for (int kc = 0; kc < p.kchunks; ++kc) {
            ptx::mbar_wait_backoff(bempty(b_i), b_ph ^ 1u);
            ptx::mbar_arrive_expect_tx(bfull(b_i), (uint32_t)UP_B_STAGE);
            ptx::tma_load_2d(smem_b + (uint32_t)(b_i * UP_B_STAGE), &p.b_map, bfull(b_i), kc * 64, nc * 256);
            if (++b_i == p.nb) { b_i = 0; b_ph ^= 1u; }
          }
      }
    }
  } else if (warp == 1) {
    if (ptx::elect_one_sync()) {
      // ================= MMA issuer =================
      const uint32_t idesc = ptx::idesc_bf16(128, 256, 0, 0);
      const uint32_t hi = ((1024u >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);      // SBO 1024 B | version | SW128
      int b_i = 0; uint32_t b_ph = 0;
      long long i = 0, j = 0;                    // tile counter, chunk counter (accumulator ring)
      for (long long t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
        const int buf = (int)(i % p.na);
        ptx::mbar_wait(afull(buf), (uint32_t)((i / p.na) & 1));
        ptx::tc_fence_after();
        for (int nc = 0; nc < p.nchunks; ++nc, ++j) {
          const int acc = (int)(j & 1);
          if (j >= 2) ptx::mbar_wait(tempty(acc), (uint32_t)(((j >> 1) - 1) & 1));
          ptx::tc_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
          for (int kc = 0; kc < p.kchunks; ++kc) {
            ptx::mbar_wait(bfull(b_i), b_ph);
            ptx::tc_fence_after();
            const uint32_t a_lo = (((smem_a + (uint32_t)(buf * a_bytes + kc * UP_A_SUB)) >> 4) & 0x3FFFu) | (1u << 16);
            const uint32_t b_lo = (((smem_b + (uint32_t)(b_i * UP_B_STAGE)) >> 4) & 0x3FFFu) | (1u << 16);
            if (kc == 0) {
              ptx::umma_bf16(d_tmem, ((uint64_t)hi << 32) | a_lo, ((uint64_t)hi << 32) | b_lo, idesc, 0u);
#pragma unroll
              for (int k = 1; k < 4; ++k)
                ptx::umma_bf16(d_tmem, ((uint64_t)hi << 32) | (uint64_t)(a_lo + 2 * k), ((uint64_t)hi << 32) | (uint64_t)(b_lo + 2 * k),
                               idesc, 1u);
            } else {
              ptx::umma_bf16_x4(d_tmem, a_lo, hi, b_lo, hi, idesc);
            }
            ptx::umma_commit(bempty(b_i));
            if (++b_i == p.nb) { b_i = 0; b_ph ^= 1u; }
          }
          ptx::umma_commit(tfull(acc));
        }
        ptx::umma_commit(aempty(buf));
      }
    }
    __syncwarp();
  } else {
    // ================= epilogue: thread = TMEM lane = input voxel; warp pair (quad, half) splits the 256 columns ======
    // A warp owns 4 (h) x 8 (w) input voxels.  Per group of 32 columns (one parity class, 32 output channels) it
    // packs 64 bytes per voxel into a 64B-swizzled staging buffer and one lane issues a TMA store through that
    // class's strided view of y: full 64-byte segments, clipped at the tensor edge, no per-lane scattered stores
    // (those saturated the L1 store path: 32 half-used sectors per instruction).
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    const int act = p.act;
    const float slope = p.slope;
    uint8_t* my_stage = out_stage + (warp - 2) * 2 * UP_STAGE_OUT;
    const uint32_t my_stage_u32 = ptx::smem_u32(my_stage);
    const uint32_t row_off = (uint32_t)(lane * 64), sw = (uint32_t)((lane >> 1) & 3);
    long long i = 0, j = 0;
    uint32_t gcount = 0;
    for (long long t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
      int n, d, h0, w0;
      up_tile_coords(p, t, n, d, h0, w0);
      for (int nc = 0; nc < p.nchunks; ++nc, ++j) {
        const int acc = (int)(j & 1);
        ptx::mbar_wait_backoff(tfull(acc), (uint32_t)((j >> 1) & 1));
        ptx::tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * 256 + half * 128);
#pragma unroll 1
        for (int g = 0; g < 4; ++g, ++gcount) {
          const int col = nc * 256 + half * 128 + g * 32;       // global column = class * Cout + co
          const int cls = col / p.cout, co = col - cls * p.cout;
          uint32_t r[32];
          ptx::tmem_ld_32x32(taddr + (uint32_t)(g * 32), r);
          // the staging buffer this group writes was read by the TMA store issued two groups ago
          if (lane == 0) ptx::tma_store_wait_read<1>();
          __syncwarp();
          ptx::tmem_ld_wait();
          float f[32];
#pragma unroll
          for (int c = 0; c < 32; c += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(bias_s + co + c);
            f[c] = __uint_as_float(r[c]) + b4.x; f[c + 1] = __uint_as_float(r[c + 1]) + b4.y;
            f[c + 2] = __uint_as_float(r[c + 2]) + b4.z; f[c + 3] = __uint_as_float(r[c + 3]) + b4.w;
          }
          if (act == RG_ACT_RELU) {
#pragma unroll
            for (int c = 0; c < 32; ++c) f[c] = fmaxf(f[c], 0.f);
          } else if (act == RG_ACT_LEAKY) {
#pragma unroll
            for (int c = 0; c < 32; ++c) f[c] = f[c] > 0.f ? f[c] : f[c] * slope;
          } else if (act != RG_ACT_NONE) {
#pragma unroll
            for (int c = 0; c < 32; ++c) f[c] = act_apply(f[c], act, slope);
          }
          uint8_t* dst = my_stage + (gcount & 1u) * UP_STAGE_OUT + row_off;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            uint32_t q[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const __nv_bfloat162 h2 = __floats2bfloat162_rn(f[8 * e + 2 * k], f[8 * e + 2 * k + 1]);
              q[k] = *reinterpret_cast<const uint32_t*>(&h2);
            }
            *reinterpret_cast<uint4*>(dst + (((uint32_t)e ^ sw) << 4)) = make_uint4(q[0], q[1], q[2], q[3]);
          }
          ptx::fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            ptx::tma_store_5d(&p.y_maps[cls], my_stage_u32 + (gcount & 1u) * UP_STAGE_OUT, co, w0, h0 + quad * 4, d, n);
            ptx::tma_store_commit();
          }
        }
        ptx::tc_fence_before();
        ptx::mbar_arrive(tempty(acc));
      }
    }
    if (lane == 0) ptx::tma_store_wait_all<0>();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc<512>(tmem_base);
  }
}

typedef CUresult (*UpEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static UpEncodeFn up_encode_fn() {
  static UpEncodeFn fn = nullptr;
  if (!fn) {
    void* q = nullptr;
    cudaDriverEntryPointQueryResult r;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &q, cudaEnableDefault, &r) == cudaSuccess && r == cudaDriverEntryPointSuccess)
      fn = (UpEncodeFn)q;
  }
  return fn;
}

static int up_num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

constexpr int UP_TAIL_BYTES = 8 * (8 + 2 * UP_MAX_NB) + 64 + UP_MAX_COUT * 4 + 64;

static int up_geometry(const rg_upconv2_desc* d, UpParams* p, size_t* smem) {
  RG_UNSUPPORTED(d->cin % 64 != 0 || d->cin < 64 || d->cin > 512, "upconv2: Cin must be a multiple of 64, <= 512");
  RG_UNSUPPORTED(d->cout % 32 != 0 || d->cout < 32 || d->cout > UP_MAX_COUT, "upconv2: Cout must be a multiple of 32, <= 512");
  for (int i = 0; i < 3; ++i) RG_UNSUPPORTED(d->in_dhw[i] < 1, "upconv2: empty input");
  for (int i = 0; i < 4; ++i)
    RG_UNSUPPORTED(d->x_strides[i] % 8 != 0 || d->y_strides[i] % 8 != 0, "upconv2: strides must be multiples of 16 bytes");
  memset(p, 0, sizeof(*p));
  p->cin = d->cin; p->cout = d->cout;
  p->kchunks = d->cin / 64; p->nchunks = 8 * d->cout / 256;
  const int a_bytes = p->kchunks * UP_A_SUB;
  const int avail = UP_SMEM_LIMIT - 1024 - UP_TAIL_BYTES - UP_OUT_BYTES;
  p->na = 2;
  int nb = (avail - 2 * a_bytes) / UP_B_STAGE;
  if (nb < 3) { p->na = 1; nb = (avail - a_bytes) / UP_B_STAGE; }
  RG_UNSUPPORTED(nb < 2, "upconv2: tile does not fit shared memory");
  p->nb = nb > UP_MAX_NB ? UP_MAX_NB : nb;
  p->in_d = d->in_dhw[0]; p->in_h = d->in_dhw[1]; p->in_w = d->in_dhw[2]; p->batch = d->n;
  p->tiles_h = (p->in_h + UP_TH - 1) / UP_TH;
  p->tiles_w = (p->in_w + UP_TW - 1) / UP_TW;
  p->total_tiles = (long long)p->batch * p->in_d * p->tiles_h * p->tiles_w;
  RG_UNSUPPORTED(p->total_tiles >= (1LL << 31), "upconv2: problem too large for 32-bit tile indices");
  p->act = d->act; p->slope = d->slope;
  *smem = 1024 + (size_t)p->na * a_bytes + (size_t)p->nb * UP_B_STAGE + UP_OUT_BYTES + UP_TAIL_BYTES;
  return RG_OK;
}

}  // namespace rg

using namespace rg;

extern "C" {

int rg_upconv2_supported(const rg_upconv2_desc* d) {
  if (!d) return 0;
  UpParams p;
  size_t smem;
  return up_geometry(d, &p, &smem) == RG_OK ? 1 : 0;
}

int rg_upconv2_run(const rg_upconv2_desc* d, const void* x, const void* w_packed, const float* bias, void* y, void* stream) {
  RG_CHECK_ARG(d && x && w_packed && y, "rg_upconv2_run: null argument");
  UpParams p;
  size_t smem;
  int rc = up_geometry(d, &p, &smem);
  if (rc) return rc;
  RG_CHECK_ARG(((uintptr_t)x & 15) == 0 && ((uintptr_t)y & 15) == 0 && ((uintptr_t)w_packed & 15) == 0,
               "rg_upconv2_run: tensors must be 16-byte aligned");
  UpEncodeFn enc = up_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled unavailable"); return RG_E_CUDA; }
  {
    cuuint64_t dims[5] = {(cuuint64_t)d->cin, (cuuint64_t)d->in_dhw[2], (cuuint64_t)d->in_dhw[1], (cuuint64_t)d->in_dhw[0], (cuuint64_t)d->n};
    cuuint64_t gs[4] = {(cuuint64_t)(d->x_strides[3] * 2), (cuuint64_t)(d->x_strides[2] * 2), (cuuint64_t)(d->x_strides[1] * 2),
                        (cuuint64_t)(d->x_strides[0] * 2)};
    cuuint32_t bx[5] = {64, UP_TW, UP_TH, 1, 1};
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(&p.a_map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, (void*)x, dims, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("upconv2: cuTensorMapEncodeTiled (activation) failed: %d", (int)r); return RG_E_CUDA; }
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)d->cin, (cuuint64_t)(8 * d->cout)};
    cuuint64_t gs[1] = {(cuuint64_t)d->cin * 2};
    cuuint32_t bx[2] = {64, 256};
    cuuint32_t es[2] = {1, 1};
    CUresult r = enc(&p.b_map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)w_packed, dims, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("upconv2: cuTensorMapEncodeTiled (weights) failed: %d", (int)r); return RG_E_CUDA; }
  }
  for (int cls = 0; cls < 8; ++cls) {
    const int pd = (cls >> 2) & 1, ph = (cls >> 1) & 1, pw = cls & 1;
    const char* base = (const char*)y + 2 * (pd * d->y_strides[1] + ph * d->y_strides[2] + pw * d->y_strides[3]);
    cuuint64_t dims[5] = {(cuuint64_t)d->cout, (cuuint64_t)d->in_dhw[2], (cuuint64_t)d->in_dhw[1], (cuuint64_t)d->in_dhw[0], (cuuint64_t)d->n};
    cuuint64_t gs[4] = {(cuuint64_t)(d->y_strides[3] * 4), (cuuint64_t)(d->y_strides[2] * 4), (cuuint64_t)(d->y_strides[1] * 4),
                        (cuuint64_t)(d->y_strides[0] * 2)};
    cuuint32_t bx[5] = {32, UP_TW, 4, 1, 1};
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(&p.y_maps[cls], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, (void*)base, dims, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("upconv2: cuTensorMapEncodeTiled (output) failed: %d", (int)r); return RG_E_CUDA; }
  }
  p.bias = d->has_bias ? bias : nullptr;
  static bool set = false;
  if (!set) {
    if (cudaFuncSetAttribute(upconv2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, UP_SMEM_LIMIT) != cudaSuccess) {
      set_error("upconv2: cannot raise the dynamic shared memory limit");
      return RG_E_CUDA;
    }
    set = true;
  }
  const int grid = (int)(p.total_tiles < up_num_sms() ? p.total_tiles : up_num_sms());
  upconv2_kernel<<<grid, UP_THREADS, smem, (cudaStream_t)stream>>>(p);
  RG_LAUNCH_CHECK("rg_upconv2_run");
  return RG_OK;
}

}  // extern "C"
