// Halo-resident weight gradient (stride 1) on tcgen05 tensor cores.
//   dW[co][ci][kd,kh,kw] = sum_{n,d,h,w} X[n, d+kd-pd, h+kh-ph, w+kw-pw, ci] * dY[n,d,h,w,co]
// GEMM view per tap pair: D[M = 2 taps x 64 ci][N = BN co] += A[M][K = voxels] * B[N][K]; both operands are MN-major
// (the voxel index is the GEMM K and selects the 128-byte smem row), exactly what TMA delivers from NDHWC tensors.
// The plain wgrad kernel (igemm_sm100.cu) loads a fresh 16 KB X box per (tap, chunk): 32 KB per 128x64x128 MMA block,
// L2-feed bound at ~0.36 PFLOP/s on the 7^3 head.  Here one X plane of the tile WITH its (kh,kw) halo is loaded once
// per 128-voxel chunk and every tap of the pass reads it through a row-shifted UMMA descriptor (start + (kh*WB+kw)
// rows, 8-voxel-group stride WB*128 B; the second tap of a pair sits LBO = (shift_b - shift_a)*128 B further), so a
// chunk costs one ~39 KB slice + the dY tile for up to 16 taps.
// A CTA = (pass, output-channel tile, voxel split); a pass = one kd, one 64-channel ci chunk and up to 512/BN tap
// pairs whose accumulators fill TMEM.  Every CTA STORES its partial sums (16-byte vectors, no atomics) into the slab of
// its voxel split in an fp32 workspace laid out [split][tap][ci][co]; a final kernel adds the slabs in split order
// (deterministic: the result is a function of the shapes only) and transposes into the checkpoint layout
// [co][ci][kd][kh][kw].
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (one elected thread), warps 2-5 = epilogue.
#include <cuda.h>
#include <cuda_bf16.h>

#include <stdlib.h>

#include "common.cuh"
#include "sm100_ptx.cuh"
#include "wgrad_finish.cuh"

namespace rg {

constexpr int HW_THREADS = 192;
constexpr int HW_TH = 16, HW_TW = 8;
constexpr int HW_TILE_BYTES = 128 * 128;     // 128 voxels x 64 channels bf16
constexpr int HW_MAX_STAGES = 6;

struct alignas(64) HaloWgradParams {
  CUtensorMap x_map;      // 5-D {C, W, H, D, N}; box {64, WB, HB, 1, 1}
  CUtensorMap dy_map;     // 5-D {C, W, H, D, N}; box {64, 8, 16, 1, 1}
  int kd, kh, kw, pd, ph, pw;
  int cin, cout;
  int q_chunks, groups;          // ci chunks; tap-pair groups per (kd, q)
  int pairs_total;               // ceil(kh*kw / 2)
  int n_pass, n_co_tiles, n_splits, chunks_per_split, total_chunks;
  int tiles_h, tiles_w, out_d, in_d, batch;
  int wb, slice_bytes, slice_tx, stages;
  int kd_first, kd_step, n_kd;   // depth taps this launch covers: kd = kd_first + i*kd_step
  int dpair;                     // 1: N = {dY[d], dY[d+1]} x 64 co -> taps kd and kd-1 in one MMA (BN = 128)
  int planes;                    // out_d + dpair (the chunk walk starts one plane early when dpair)
  float* ws;                     // [n_splits][taps][cin][cout] fp32 partial sums
};

template <int BN>
__global__ void __launch_bounds__(HW_THREADS, 1) halo_wgrad_kernel(const __grid_constant__ HaloWgradParams p) {
  constexpr int NACC = 512 / BN;            // accumulators (tap pairs) per CTA
  constexpr int DY_BYTES = (BN / 64) * HW_TILE_BYTES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);
  const int stage_bytes = p.slice_bytes + DY_BYTES;
  uint8_t* tail = base_ptr + (size_t)p.stages * stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);          // full[6] empty[6] done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 8 * (2 * HW_MAX_STAGES + 1));
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (HW_MAX_STAGES + s); };
  const uint32_t done_bar = bar0 + 8u * (2 * HW_MAX_STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int u = blockIdx.x;
  const int pass = u % p.n_pass; u /= p.n_pass;
  const int co_tile = u % p.n_co_tiles;
  const int split = u / p.n_co_tiles;
  const int gi = pass % p.groups;
  const int q = (pass / p.groups) % p.q_chunks;
  const int kd = p.kd_first + p.kd_step * (pass / (p.groups * p.q_chunks));
  const int j_begin = (int)(((long long)gi * p.pairs_total) / p.groups);
  const int j_end = (int)(((long long)(gi + 1) * p.pairs_total) / p.groups);
  const int c_begin = split * p.chunks_per_split;
  const int c_end = min(c_begin + p.chunks_per_split, p.total_chunks);
  const int ntaps_hw = p.kh * p.kw;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < p.stages; ++s) { ptx::mbar_init(full(s), 1); ptx::mbar_init(empty(s), 1); }
    ptx::mbar_init(done_bar, 1);
    ptx::fence_mbar_init();
    ptx::prefetch_tmap(&p.x_map);
    ptx::prefetch_tmap(&p.dy_map);
  }
  if (warp == 1) ptx::tmem_alloc<512>(ptx::smem_u32(tmem_slot));
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto decode = [&](int c, int& n, int& d, int& h0, int& w0) {
    const int iw = c % p.tiles_w; c /= p.tiles_w;
    const int ih = c % p.tiles_h; c /= p.tiles_h;
    d = c % p.planes - p.dpair; n = c / p.planes;
    h0 = ih * HW_TH; w0 = iw * HW_TW;
  };

  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer =================
      int st = 0; uint32_t ph = 0;
      for (int c = c_begin; c < c_end; ++c) {
        int n, d, h0, w0;
        decode(c, n, d, h0, w0);
        const int z = d + kd - p.pd;
        if (z < 0 || z >= p.in_d) continue;           // the whole X plane is padding: contributes nothing
        ptx::mbar_wait_backoff(empty(st), ph ^ 1u);
        ptx::mbar_arrive_expect_tx(full(st), (uint32_t)(p.slice_tx + DY_BYTES));
        const uint32_t sa = base + (uint32_t)(st * stage_bytes);
        ptx::tma_load_5d(sa, &p.x_map, full(st), q * 64, w0 - p.pw, h0 - p.ph, z, n);
#pragma unroll
        for (int s = 0; s < BN / 64; ++s)
          ptx::tma_load_5d(sa + (uint32_t)(p.slice_bytes + s * HW_TILE_BYTES), &p.dy_map, full(st),
                           p.dpair ? co_tile * 64 : co_tile * BN + s * 64, w0, h0, p.dpair ? d + s : d, n);
        if (++st == p.stages) { st = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer: one elected thread runs the loop (inside an elect.sync region ptxas computes the
    // descriptors in the uniform datapath; see the note in ring_conv_sm100.cu: the loop is issue-bound otherwise) ======
    if (ptx::elect_one_sync()) {
    constexpr uint32_t idesc = ptx::idesc_bf16(128, BN, 1, 1);          // both operands MN-major
    const uint32_t sbo_a = (uint32_t)(p.wb * 128);
    const uint32_t b_hi = ((1024u >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);
    const uint32_t a_hi = ((sbo_a >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);
    const uint32_t b_lbo = ((uint32_t)(HW_TILE_BYTES >> 4) & 0x3FFFu) << 16;
    // per tap pair (loop-invariant): row shift of the first tap and LBO = distance to the second, both >> 4
    uint32_t a_off[NACC];
    const int npairs = j_end - j_begin;
#pragma unroll
    for (int jj = 0; jj < NACC; ++jj) {
      const int j = j_begin + jj;
      const int ta = 2 * j, tb = (2 * j + 1 < ntaps_hw) ? 2 * j + 1 : 2 * j;       // unpaired last tap: duplicate
      const int sha = (ta / p.kw) * p.wb + (ta % p.kw);
      const int shb = (tb / p.kw) * p.wb + (tb % p.kw);
      a_off[jj] = (uint32_t)(sha * 8) | ((((uint32_t)((shb - sha) * 128) >> 4) & 0x3FFFu) << 16);
    }
    const uint32_t k_step_a = (uint32_t)((2 * p.wb * 128) >> 4);       // 16 voxels (K) = two h rows of the tile
    int st = 0; uint32_t ph = 0;
    uint32_t acc = 0;
    for (int c = c_begin; c < c_end; ++c) {
      int n, d, h0, w0;
      decode(c, n, d, h0, w0);
      const int z = d + kd - p.pd;
      if (z < 0 || z >= p.in_d) continue;
      ptx::mbar_wait(full(st), ph);
      ptx::tc_fence_after();
      const uint32_t sa = base + (uint32_t)(st * stage_bytes);
      const uint32_t a_base = (sa >> 4) & 0x3FFFu;
      const uint32_t b_lo0 = (((sa + (uint32_t)p.slice_bytes) >> 4) & 0x3FFFu) | b_lbo;
#pragma unroll
      for (int jj = 0; jj < NACC; ++jj) {
        if (jj < npairs) {
          const uint32_t a_lo0 = a_base + a_off[jj];
          const uint32_t dcol = tmem_base + (uint32_t)(jj * BN);
          ptx::umma_bf16_x8(dcol, a_lo0, k_step_a, a_hi, b_lo0, 128u, b_hi, idesc, acc);
        }
      }
      ptx::umma_commit(empty(st));
      if (++st == p.stages) { st = 0; ph ^= 1u; }
      acc = 1u;
    }
    ptx::umma_commit(done_bar);
    }
    __syncwarp();
  } else {
    // ================= epilogue: TMEM -> vector stores into ws[split][tap][ci][co] =================
    const int quad = warp & 3;
    ptx::mbar_wait_backoff(done_bar, 0);
    ptx::tc_fence_after();
    // did any chunk of this CTA see a real X plane?  (otherwise the accumulators were never written)
    bool any = false;
    {
      const int per_plane = p.tiles_h * p.tiles_w;
      const int p0 = c_begin / per_plane, p1 = (c_end - 1) / per_plane;
      for (int pl = p0; pl <= p1 && pl < p0 + p.planes; ++pl) {
        const int z = pl % p.planes - p.dpair + kd - p.pd;
        any = any || (z >= 0 && z < p.in_d);
      }
    }
    const bool live = any && c_end > c_begin;       // otherwise this CTA's slab entries are zero
    {
      float* ws_split = p.ws + (size_t)split * ((size_t)p.kd * ntaps_hw * p.cin * p.cout);
      for (int j = j_begin; j < j_end; ++j) {
        const int t = 2 * j + (quad >> 1);
        if (t >= ntaps_hw) continue;                          // unpaired second half (warp-uniform)
        const int ci = q * 64 + (quad & 1) * 32 + lane;
        float* dst0 = ws_split + ((size_t)(kd * ntaps_hw + t) * p.cin + ci) * p.cout;
#pragma unroll 1
        for (int cb = 0; cb < BN; cb += 32) {
          // dpair: columns 64..127 are the same 64 output channels against dY[d+1], i.e. depth tap kd-1
          float* dst = p.dpair ? dst0 - (size_t)(cb >> 6) * ntaps_hw * p.cin * p.cout + co_tile * 64 + (cb & 63) - cb
                               : dst0 + co_tile * BN;
          uint32_t v[32];
          if (live) {
            ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)((j - j_begin) * BN + cb), v);
            ptx::tmem_ld_wait();
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] = 0u;
          }
#pragma unroll
          for (int e = 0; e < 32; e += 4)
            *reinterpret_cast<uint4*>(dst + cb + e) = make_uint4(v[e], v[e + 1], v[e + 2], v[e + 3]);
        }
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace rg

// =================================================================================================
// Host side
// =================================================================================================
namespace rg {

typedef CUresult (*HwEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static HwEncodeFn hw_encode_fn() {
  static HwEncodeFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (HwEncodeFn)p;
  }
  return fn;
}

static int hw_num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

constexpr int HW_SMEM_LIMIT = 227 * 1024;
constexpr int HW_TAIL_BYTES = 8 * (2 * HW_MAX_STAGES + 1) + 64;

struct HwGeom { HaloWgradParams p; int bn; int smem_bytes; int grid; };

// variant 0: every depth tap, N = output channels.  variant 1 (cout == 64): depth taps 1,3,5,.. paired with their
// predecessor through N = {dY[d], dY[d+1]} (full-rate N = 128 MMAs).  variant 2: the last depth tap alone (KD odd).
static int hw_geometry(const rg_igemm_desc* d, HwGeom* g, int variant = 0) {
  RG_UNSUPPORTED(d->mode != 2, "halo wgrad: descriptor mode must be 2 (wgrad)");
  for (int i = 0; i < 3; ++i) {
    RG_UNSUPPORTED(d->stride[i] != 1, "halo wgrad: stride must be 1");
    RG_UNSUPPORTED(d->kernel[i] < 1 || d->kernel[i] > 8, "halo wgrad: kernel extent must be 1..8");
    RG_UNSUPPORTED(d->pad[i] < 0 || d->pad[i] > d->kernel[i] - 1, "halo wgrad: padding must be within the kernel");
    RG_UNSUPPORTED(d->out_dhw[i] != d->in_dhw[i] + 2 * d->pad[i] - d->kernel[i] + 1 || d->out_dhw[i] < 1,
                   "halo wgrad: extents inconsistent with geometry");
  }
  RG_UNSUPPORTED(d->cin % 64 != 0 || d->cout % 64 != 0, "halo wgrad: channel counts must be multiples of 64");
  RG_UNSUPPORTED(d->x_strides[3] % 8 != 0 || d->y_strides[3] % 8 != 0, "halo wgrad: voxel stride must be a multiple of 8");
  HaloWgradParams& p = g->p;
  memset(&p, 0, sizeof(p));
  p.kd = d->kernel[0]; p.kh = d->kernel[1]; p.kw = d->kernel[2];
  p.pd = d->pad[0]; p.ph = d->pad[1]; p.pw = d->pad[2];
  p.cin = d->cin; p.cout = d->cout;
  g->bn = d->cout % 256 == 0 ? 256 : (d->cout % 192 == 0 ? 192 : (d->cout % 128 == 0 ? 128 : 64));
  p.kd_first = 0; p.kd_step = 1; p.n_kd = p.kd; p.dpair = 0;
  if (variant == 1) { g->bn = 128; p.dpair = 1; p.kd_first = 1; p.kd_step = 2; p.n_kd = p.kd / 2; }
  if (variant == 2) { g->bn = 64; p.kd_first = p.kd - 1; p.n_kd = 1; }
  const int nacc = 512 / g->bn;
  p.q_chunks = d->cin / 64;
  p.pairs_total = (p.kh * p.kw + 1) / 2;
  p.groups = (p.pairs_total + nacc - 1) / nacc;
  p.n_pass = p.n_kd * p.q_chunks * p.groups;
  p.n_co_tiles = p.dpair ? d->cout / 64 : d->cout / g->bn;
  p.wb = HW_TW + p.kw - 1;
  const int hb = HW_TH + p.kh - 1;
  RG_UNSUPPORTED(p.wb > 256 || hb > 256, "halo wgrad: halo box too large");
  p.slice_tx = hb * p.wb * 128;
  p.slice_bytes = (p.slice_tx + 1023) & ~1023;
  const int stage_bytes = p.slice_bytes + (g->bn / 64) * HW_TILE_BYTES;
  p.stages = (HW_SMEM_LIMIT - 1024 - HW_TAIL_BYTES) / stage_bytes;
  if (p.stages > HW_MAX_STAGES) p.stages = HW_MAX_STAGES;
  RG_UNSUPPORTED(p.stages < 2, "halo wgrad: stage does not fit shared memory twice");
  g->smem_bytes = 1024 + p.stages * stage_bytes + HW_TAIL_BYTES;
  p.tiles_h = (d->out_dhw[1] + HW_TH - 1) / HW_TH;
  p.tiles_w = (d->out_dhw[2] + HW_TW - 1) / HW_TW;
  p.out_d = d->out_dhw[0]; p.in_d = d->in_dhw[0]; p.batch = d->n;
  p.planes = p.out_d + p.dpair;
  p.total_chunks = d->n * p.planes * p.tiles_h * p.tiles_w;
  // voxel splits: about four waves of CTAs, but at least 32 chunks per CTA so the reduction epilogue stays small
  const long long base_units = (long long)p.n_pass * p.n_co_tiles;
  long long splits = (4LL * hw_num_sms()) / base_units;
  const long long max_splits = (p.total_chunks + 31) / 32;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  p.chunks_per_split = (int)((p.total_chunks + splits - 1) / splits);
  p.n_splits = (p.total_chunks + p.chunks_per_split - 1) / p.chunks_per_split;
  g->grid = (int)(base_units * p.n_splits);
  return RG_OK;
}

template <int BN> static int hw_launch(const HwGeom& g, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(halo_wgrad_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, HW_SMEM_LIMIT);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(halo_wgrad)");
    attr_set = true;
  }
  halo_wgrad_kernel<BN><<<g.grid, HW_THREADS, g.smem_bytes, st>>>(g.p);
  return RG_OK;
}

}  // namespace rg

using namespace rg;

extern "C" {

int rg_halo_wgrad_supported(const rg_igemm_desc* d) {
  if (!d) return 0;
  HwGeom g;
  return hw_geometry(d, &g) == RG_OK ? 1 : 0;
}

static bool hw_use_dpair(const rg_igemm_desc* d) {
  static const char* nopair = getenv("RAYGUN_B200_WGRAD_NO_DPAIR");
  return d->cout == 64 && d->kernel[0] >= 4 &&       // (3-tap kernels: two short launches cost more than they gain)
         !(nopair && nopair[0] == '1');
}

size_t rg_halo_wgrad_workspace_bytes(const rg_igemm_desc* d) {
  if (!d) return 0;
  const bool dpair = hw_use_dpair(d);
  int max_splits = 1;
  for (int variant = dpair ? 1 : 0; variant <= (dpair ? 2 : 0); ++variant) {
    if (variant == 2 && d->kernel[0] % 2 == 0) break;
    HwGeom gv;
    if (hw_geometry(d, &gv, variant)) return 0;
    if (gv.p.n_splits > max_splits) max_splits = gv.p.n_splits;
  }
  return (size_t)max_splits * d->cout * d->cin * d->kernel[0] * d->kernel[1] * d->kernel[2] * sizeof(float);
}

int rg_halo_wgrad(const rg_igemm_desc* d, const void* x, const void* dy, float* dw, void* workspace, int accumulate,
                  void* stream) {
  RG_CHECK_ARG(d && x && dy && dw && workspace, "rg_halo_wgrad: null argument");
  HwGeom g;
  int rc = hw_geometry(d, &g);
  if (rc) return rc;
  HwEncodeFn enc = hw_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled unavailable"); return RG_E_CUDA; }
  RG_CHECK_ARG(((uintptr_t)x & 15) == 0 && ((uintptr_t)dy & 15) == 0 && ((uintptr_t)workspace & 15) == 0,
               "rg_halo_wgrad: tensors must be 16-byte aligned");
  HaloWgradParams& p = g.p;
  cudaStream_t st = (cudaStream_t)stream;
  for (int which = 0; which < 2; ++which) {
    const int64_t* str = which == 0 ? d->x_strides : d->y_strides;
    const int* ext = which == 0 ? d->in_dhw : d->out_dhw;
    const int ch = which == 0 ? d->cin : d->cout;
    for (int i = 0; i < 4; ++i) RG_UNSUPPORTED((str[i] * 2) % 16 != 0, "rg_halo_wgrad: strides must be multiples of 16 bytes");
    cuuint64_t dims[5] = {(cuuint64_t)ch, (cuuint64_t)ext[2], (cuuint64_t)ext[1], (cuuint64_t)ext[0], (cuuint64_t)d->n};
    cuuint64_t gs[4] = {(cuuint64_t)(str[3] * 2), (cuuint64_t)(str[2] * 2), (cuuint64_t)(str[1] * 2), (cuuint64_t)(str[0] * 2)};
    cuuint32_t bx[5] = {64, (cuuint32_t)(which == 0 ? p.wb : HW_TW), (cuuint32_t)(which == 0 ? HW_TH + p.kh - 1 : HW_TH), 1, 1};
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(which == 0 ? &p.x_map : &p.dy_map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, (void*)(which == 0 ? x : dy), dims, gs,
                     bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("halo wgrad: cuTensorMapEncodeTiled failed: %d", (int)r); return RG_E_CUDA; }
  }
  const int taps = p.kd * p.kh * p.kw;
  const bool dpair = hw_use_dpair(d);
  HwSplits splits;
  for (int i = 0; i < 8; ++i) splits.of_kd[i] = 0;
  for (int variant = dpair ? 1 : 0; variant <= (dpair ? 2 : 0); ++variant) {
    if (variant == 2 && p.kd % 2 == 0) break;
    HwGeom gv;
    rc = hw_geometry(d, &gv, variant);
    if (rc) return rc;
    gv.p.x_map = p.x_map; gv.p.dy_map = p.dy_map;
    gv.p.ws = (float*)workspace;
    for (int i = 0; i < gv.p.n_kd; ++i) {       // every depth tap this launch covers is written in all of its split slabs
      const int kd = gv.p.kd_first + i * gv.p.kd_step;
      splits.of_kd[kd] = gv.p.n_splits;
      if (gv.p.dpair) splits.of_kd[kd - 1] = gv.p.n_splits;
    }
    if (gv.bn == 256) rc = hw_launch<256>(gv, st);
    else if (gv.bn == 192) rc = hw_launch<192>(gv, st);
    else if (gv.bn == 128) rc = hw_launch<128>(gv, st);
    else rc = hw_launch<64>(gv, st);
    if (rc) return rc;
    RG_LAUNCH_CHECK("rg_halo_wgrad");
  }
  wgrad_finish_launch((const float*)workspace, splits, p.kh * p.kw, taps, d->cin, d->cout, dw, accumulate, st);
  RG_LAUNCH_CHECK("rg_halo_wgrad(finish)");
  return RG_OK;
}

}  // extern "C"
