// Small-K stem convolution on tcgen05 tensor cores: Conv3d layers whose whole reduction Cin*kD*kH*kW fits ONE 64-wide
// K block (the 1 -> ngf 3^3 convolution that opens every U-Net level 0, UNet.py:52-87 / 420-431: K = 27).
// The folded path (rg_im2col_hw + a (kD,1,1) ring convolution) writes and re-reads a 64-channel tensor of the full
// volume just to feed the tensor cores: 3.4 ms per 256^3 block.  Here the im2col row of every output voxel (its K
// input taps) is gathered straight into the 128B-swizzled K-major A tile in shared memory, so HBM sees the
// 1-channel input once and the output once: the kernel is HBM-write bound (128 B per voxel for 64 channels).
//
// Warp roles (544 threads): warp 0 = MMA issuer (one elected thread), warps 1-8 = A-tile builders (3^3 single-channel
// stems: one warp per 4 x 32 tile through a shared-memory input patch; other shapes: two groups of 128 threads that
// alternate 16 x 8 tiles, thread = output voxel), warps 9-16 = epilogue (thread = TMEM lane = voxel, two
// warps per lane quadrant that split the produced channels in groups of 32).
// Tile = 4 x 32 or 16 (h) x 8 (w) output voxels of one depth plane; persistent CTAs walk the tiles round-robin.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace rg {

constexpr int ST_THREADS = 544;
constexpr int ST_TH = 16, ST_TW = 8;
constexpr int ST_STAGES = 8;                 // A tiles in flight (16 KB each); >= the 8 builder warps, so a warp can never
                                             // run two uses ahead of the issuer on one stage (parity waits would alias)
constexpr int ST_A_BYTES = 128 * 128;
constexpr int ST_MAX_COUT = 256;
constexpr int ST_STAGE_OUT = 32 * 64;          // epilogue staging: 32 voxels x 32 channels bf16, two buffers per warp
constexpr int ST_OUT_BYTES = 8 * 2 * ST_STAGE_OUT;
constexpr int ST_PATCH_PITCH = 50;             // elements per patch row (34 used); 25 words: 4 consecutive rows -> disjoint banks
constexpr int ST_PATCH_BYTES = 2048;           // 18 rows x 50 x 2 B = 1800, one buffer per builder warp

struct alignas(64) StemParams {
  CUtensorMap y_map;                         // 5-D {Cout, W, H, D, N} view of y; box {32, 8, 4, 1, 1}, 64B swizzle
  const __nv_bfloat16* x;
  long long xs_n, xs_c, xs_d, xs_h, xs_w;    // element strides of the input (any layout)
  int cin, kd, kh, kw, pd, ph, pw, K;
  int in_d, in_h, in_w;
  int out_d, out_h, out_w, batch;
  int tiles_h, tiles_w;
  int pad_mode;                              // RG_PAD_ZEROS / REFLECT / REPLICATE: how taps outside the input are read
  int tile_h, tile_w;                        // 16 x 8 (generic builders) or 4 x 32 (3^3 single-channel row builders)
  int quad_h;                                // h rows per TMEM lane quadrant = rows of one epilogue store box (4 or 1)
  int afull_count;                           // arrivals that complete an A stage: 128 builder threads, or 1 warp leader
  long long total_tiles;
  const __nv_bfloat16* w;                    // [cout][64] bf16, K-major, zero padded
  int cout;
  const float* bias;
  int act; float slope;
};

__device__ __forceinline__ void stem_tile_coords(const StemParams& p, long long t, int& n, int& d, int& h0, int& w0) {
  unsigned u = (unsigned)t;                         // total_tiles < 2^31 (checked on the host): 32-bit divisions
  const unsigned iw = u % (unsigned)p.tiles_w; u /= (unsigned)p.tiles_w;
  const unsigned ih = u % (unsigned)p.tiles_h; u /= (unsigned)p.tiles_h;
  d = (int)(u % (unsigned)p.out_d); n = (int)(u / (unsigned)p.out_d);
  h0 = (int)ih * p.tile_h; w0 = (int)iw * p.tile_w;
}

// Source index of tap coordinate i along an axis of extent n.  Zero padding: out-of-range taps read 0 (ok = false);
// reflect / replicate (the 'same' padding of UNet.py:61-69 folded into the gather: no padded copy of the input): the
// mirrored / clamped index.  Tile overhang far outside the tensor is clamped too (those voxels are never stored).
__device__ __forceinline__ int stem_src(int i, int n, int mode, bool& ok) {
  if (mode == RG_PAD_ZEROS) { ok = i >= 0 && i < n; return i; }
  ok = true;
  if (mode == RG_PAD_REFLECT) { i = i < 0 ? -i : i; i = i >= n ? 2 * n - 2 - i : i; }
  return min(max(i, 0), n - 1);
}

// KSTEPS = ceil(K / 16): the builder writes 2*KSTEPS 16-byte chunks per row, the issuer KSTEPS MMAs per tile
// SPEC333: Cin = 1, 3 x 3 x 3 taps unrolled at compile time (the U-Net stem); otherwise generic loops.
template <int KSTEPS, bool SPEC333>
__global__ void __launch_bounds__(ST_THREADS, 1) stem_conv_kernel(const __grid_constant__ StemParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);
  const uint32_t smem_a = base;
  const uint32_t smem_w = base + ST_STAGES * ST_A_BYTES;
  uint8_t* w_ptr = base_ptr + ST_STAGES * ST_A_BYTES;
  uint8_t* out_stage = w_ptr + ST_MAX_COUT * 128;            // 1024-aligned
  uint8_t* patch_base = out_stage + ST_OUT_BYTES;
  uint8_t* tail = patch_base + 8 * ST_PATCH_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);      // afull[S] aempty[S] tfull[2] tempty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 8 * (2 * ST_STAGES + 4));
  float* bias_s = reinterpret_cast<float*>(tail + 8 * (2 * ST_STAGES + 4) + 64);      // [ST_MAX_COUT]
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto afull = [&](int s) { return bar0 + 8u * s; };
  auto aempty = [&](int s) { return bar0 + 8u * (ST_STAGES + s); };
  auto tfull = [&](int s) { return bar0 + 8u * (2 * ST_STAGES + s); };
  auto tempty = [&](int s) { return bar0 + 8u * (2 * ST_STAGES + 2 + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < ST_STAGES; ++s) { ptx::mbar_init(afull(s), (uint32_t)p.afull_count); ptx::mbar_init(aempty(s), 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(tfull(s), 1); ptx::mbar_init(tempty(s), 256); }
    ptx::fence_mbar_init();
    ptx::prefetch_tmap(&p.y_map);
  }
  if (warp == 0) ptx::tmem_alloc<512>(ptx::smem_u32(tmem_slot));
  // weights -> smem, K-major rows of 128 B with the 128B swizzle (16-byte chunk j of row r lands at chunk j ^ (r & 7))
  for (int i = threadIdx.x; i < p.cout * 8; i += ST_THREADS) {
    const int r = i >> 3, j = i & 7;
    const uint4 v = reinterpret_cast<const uint4*>(p.w)[i];
    *reinterpret_cast<uint4*>(w_ptr + r * 128 + ((j ^ (r & 7)) << 4)) = v;
  }
  for (int i = threadIdx.x; i < p.cout; i += ST_THREADS) bias_s[i] = p.bias ? p.bias[i] : 0.f;
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t acc_cols = (uint32_t)p.cout;               // accumulator stage pitch in TMEM columns

  if (warp == 0) {
    if (ptx::elect_one_sync()) {
      // ================= MMA issuer =================
      const uint32_t idesc = ptx::idesc_bf16(128, 0, 0, 0) | ((uint32_t)(p.cout >> 3) << 17);
      const uint32_t hi = ((1024u >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);        // SBO 1024 B | version | SW128
      const uint32_t b_lo = ((smem_w >> 4) & 0x3FFFu) | (1u << 16);
      int st = 0; uint32_t ph = 0;
      long long i = 0;
      for (long long t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
        const int acc = (int)(i & 1);
        if (i >= 2) ptx::mbar_wait(tempty(acc), (uint32_t)(((i >> 1) - 1) & 1));
        ptx::mbar_wait(afull(st), ph);
        ptx::tc_fence_after();
        const uint32_t a_lo = (((smem_a + (uint32_t)(st * ST_A_BYTES)) >> 4) & 0x3FFFu) | (1u << 16);
        const uint32_t d_tmem = tmem_base + (uint32_t)acc * acc_cols;
#pragma unroll
        for (int k = 0; k < KSTEPS; ++k)
          ptx::umma_bf16(d_tmem, ((uint64_t)hi << 32) | (uint64_t)(a_lo + 2 * k), ((uint64_t)hi << 32) | (uint64_t)(b_lo + 2 * k),
                         idesc, k > 0 ? 1u : 0u);
        ptx::umma_commit(aempty(st));
        ptx::umma_commit(tfull(acc));
        if (++st == ST_STAGES) { st = 0; ph ^= 1u; }
      }
    }
    __syncwarp();
  } else if (warp <= 8) {
    // ================= builders =================
    uint8_t* a_ptr = base_ptr;
    if constexpr (SPEC333) {
      // Tile = 4 (h) x 32 (w) output voxels; ONE WARP builds a whole tile.  The warp stages the tile's input patch
      // (3 d x 6 h x 34 w single-channel values, zero outside the input = the conv's zero padding) in shared memory
      // with 20 coalesced 2-byte loads per lane, then every lane assembles the 27-tap rows of its four voxels
      // (w = q + 8k) from the patch and writes them as swizzled 16-byte chunks.  Row r of the tile is voxel
      // (h = r / 32, w = r % 32), so a quarter-warp's chunks land in eight different swizzle positions
      // (conflict-free) and a TMEM lane quadrant is one 32-voxel row for the epilogue's store box.
      // The patch of the warp's NEXT tile is loaded into registers before the current one is assembled.
      static_assert(KSTEPS == 2, "row builder: K = 27 taps padded to 32");
      const int bw = warp - 1;                                   // 0..7
      const int q = lane & 7, vh = lane >> 3;
      uint16_t* patch = reinterpret_cast<uint16_t*>(patch_base + bw * ST_PATCH_BYTES);
      const int xsw = (int)p.xs_w;
      auto load_patch = [&](long long t, uint16_t* v) {
        int n, d, h0, w0;
        stem_tile_coords(p, t, n, d, h0, w0);
        const __nv_bfloat16* xn = p.x + (long long)n * p.xs_n;
        bool win, ein;
        const int xw = stem_src(w0 - p.pw + lane, p.in_w, p.pad_mode, win);
        // columns 32, 33 of patch rows 2i, 2i+1 .. : lane l covers row l >> 1 (rows 0..15), lanes 0..3 rows 16, 17
        const int xe = stem_src(w0 - p.pw + 32 + (lane & 1), p.in_w, p.pad_mode, ein);
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          bool zin;
          const int z = stem_src(d - p.pd + a, p.in_d, p.pad_mode, zin);
#pragma unroll
          for (int hh = 0; hh < 6; ++hh) {
            bool yin;
            const int y = stem_src(h0 - p.ph + hh, p.in_h, p.pad_mode, yin);
            const bool ok = zin && yin && win;
            uint16_t val = 0;
            if (ok) val = __ldg(reinterpret_cast<const uint16_t*>(xn + (long long)z * p.xs_d + (long long)y * p.xs_h) + xw * xsw);
            v[a * 6 + hh] = val;
          }
        }
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int row = e * 16 + (lane >> 1);                  // patch row a * 6 + hh
          const int a = row / 6, hh = row - a * 6;
          bool zin, yin;
          const int z = stem_src(d - p.pd + a, p.in_d, p.pad_mode, zin), y = stem_src(h0 - p.ph + hh, p.in_h, p.pad_mode, yin);
          const bool ok = row < 18 && zin && yin && ein;
          uint16_t val = 0;
          if (ok) val = __ldg(reinterpret_cast<const uint16_t*>(xn + (long long)z * p.xs_d + (long long)y * p.xs_h) + xe * xsw);
          v[18 + e] = val;
        }
      };
      uint16_t cur[20], nxt[20];
      const long long t_first = blockIdx.x + (long long)bw * gridDim.x, t_step = 8LL * gridDim.x;
      if (t_first < p.total_tiles) load_patch(t_first, cur);
      long long i = bw;
      for (long long t = t_first; t < p.total_tiles; t += t_step, i += 8) {
        // registers -> patch (pitch ST_PATCH_PITCH elements per row: the four h rows a warp reads sit in disjoint banks)
#pragma unroll
        for (int row = 0; row < 18; ++row) patch[row * ST_PATCH_PITCH + lane] = cur[row];
        patch[(lane >> 1) * ST_PATCH_PITCH + 32 + (lane & 1)] = cur[18];
        if (lane < 4) patch[(16 + (lane >> 1)) * ST_PATCH_PITCH + 32 + (lane & 1)] = cur[19];
        const bool more = t + t_step < p.total_tiles;
        if (more) load_patch(t + t_step, nxt);
        __syncwarp();
        const int st = (int)((unsigned)i % ST_STAGES);
        const uint32_t ph = (uint32_t)(((unsigned)i / ST_STAGES) & 1);
        ptx::mbar_wait_backoff(aempty(st), ph ^ 1u);
        uint8_t* tile_ptr = a_ptr + st * ST_A_BYTES;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int vw = q + 8 * k;
          const uint16_t* src = patch + vh * ST_PATCH_PITCH + vw;
          uint32_t wd[16];
#pragma unroll
          for (int m = 0; m < 16; ++m) {
            uint32_t lo = 0, hi = 0;
            if (2 * m < 27) {
              const int kk = 2 * m, a = kk / 9, b = (kk / 3) % 3, c = kk % 3;
              lo = src[(a * 6 + b) * ST_PATCH_PITCH + c];
            }
            if (2 * m + 1 < 27) {
              const int kk = 2 * m + 1, a = kk / 9, b = (kk / 3) % 3, c = kk % 3;
              hi = src[(a * 6 + b) * ST_PATCH_PITCH + c];
            }
            wd[m] = lo | (hi << 16);
          }
          const int r = vh * 32 + vw;
          uint8_t* rowp = tile_ptr + r * 128;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            *reinterpret_cast<uint4*>(rowp + ((j ^ (r & 7)) << 4)) = make_uint4(wd[4 * j], wd[4 * j + 1], wd[4 * j + 2], wd[4 * j + 3]);
        }
        ptx::fence_proxy_async_smem();             // generic-proxy writes -> visible to the tensor core (async proxy)
        __syncwarp();                              // also: every lane is done reading the patch before it is rewritten
        if (lane == 0) ptx::mbar_arrive(afull(st));
        if (more) {
#pragma unroll
          for (int kk = 0; kk < 20; ++kk) cur[kk] = nxt[kk];
        }
      }
    } else {
      // generic shapes: clear the row, then one 2-byte store per tap into its swizzled position
      const int grp = (warp - 1) >> 2;
      const int v = (threadIdx.x - 32) & 127;
      const int vh = v >> 3, vw = v & 7;
      const long long t_first = blockIdx.x + (long long)grp * gridDim.x, t_step = 2LL * gridDim.x;
      long long i = grp;
      for (long long t = t_first; t < p.total_tiles; t += t_step, i += 2) {
        int n, d, h0, w0;
        stem_tile_coords(p, t, n, d, h0, w0);
        const int st = (int)((unsigned)i % ST_STAGES);
        const uint32_t ph = (uint32_t)(((unsigned)i / ST_STAGES) & 1);
        const int hh = h0 + vh - p.ph, ww = w0 + vw - p.pw, dd = d - p.pd;
        const __nv_bfloat16* xn = p.x + (long long)n * p.xs_n;
        uint8_t* rowp = a_ptr + st * ST_A_BYTES + v * 128;
        ptx::mbar_wait_backoff(aempty(st), ph ^ 1u);
#pragma unroll
        for (int j = 0; j < 2 * KSTEPS; ++j) *reinterpret_cast<uint4*>(rowp + ((j ^ (v & 7)) << 4)) = make_uint4(0, 0, 0, 0);
        int k = 0;
        for (int ci = 0; ci < p.cin; ++ci)
          for (int a = 0; a < p.kd; ++a) {
            bool zin;
            const int z = stem_src(dd + a, p.in_d, p.pad_mode, zin);
            for (int b = 0; b < p.kh; ++b) {
              bool yin;
              const int y = stem_src(hh + b, p.in_h, p.pad_mode, yin);
              const bool rin = zin && yin;
              const __nv_bfloat16* row = xn + ci * p.xs_c + (long long)z * p.xs_d + (long long)y * p.xs_h;
              for (int c = 0; c < p.kw; ++c, ++k) {
                bool xin;
                const int xw = stem_src(ww + c, p.in_w, p.pad_mode, xin);
                if (rin && xin)
                  *reinterpret_cast<__nv_bfloat16*>(rowp + ((((k >> 3) ^ (v & 7)) << 4) | ((k & 7) << 1))) =
                      __ldg(row + (long long)xw * p.xs_w);
              }
            }
          }
        ptx::fence_proxy_async_smem();
        ptx::mbar_arrive(afull(st));
      }
    }
  } else {
    // ================= epilogue: thread = TMEM lane = output voxel; warp pair (quad, half) splits the channels ======
    const int quad = warp & 3;
    const int half = (warp - 9) >> 2;
    const int act = p.act;
    const float slope = p.slope;
    // a warp owns 4 (h) x 8 (w) voxels: 64 bytes per voxel and column group go through a 64B-swizzled staging
    // buffer and one TMA store (full segments, clipped at the tensor edge) instead of per-lane scattered stores
    uint8_t* my_stage = out_stage + (warp - 9) * 2 * ST_STAGE_OUT;
    const uint32_t my_stage_u32 = ptx::smem_u32(my_stage);
    const uint32_t row_off = (uint32_t)(lane * 64), sw = (uint32_t)((lane >> 1) & 3);
    uint32_t gcount = 0;
    long long i = 0;
    for (long long t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
      int n, d, h0, w0;
      stem_tile_coords(p, t, n, d, h0, w0);
      const int acc = (int)(i & 1);
      ptx::mbar_wait_backoff(tfull(acc), (uint32_t)((i >> 1) & 1));
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)acc * acc_cols;
      for (int cb = half * 32; cb < p.cout; cb += 64, ++gcount) {
        uint32_t r[32];
        ptx::tmem_ld_32x32(taddr + (uint32_t)cb, r);
        if (lane == 0) ptx::tma_store_wait_read<1>();       // the store issued two groups ago has read this buffer
        __syncwarp();
        ptx::tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int c = 0; c < 32; c += 4) {
          const float4 b4 = *reinterpret_cast<const float4*>(bias_s + cb + c);
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
        uint8_t* dst = my_stage + (gcount & 1u) * ST_STAGE_OUT + row_off;
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
          ptx::tma_store_5d(&p.y_map, my_stage_u32 + (gcount & 1u) * ST_STAGE_OUT, cb, w0, h0 + quad * p.quad_h, d, n);
          ptx::tma_store_commit();
        }
      }
      ptx::tc_fence_before();
      ptx::mbar_arrive(tempty(acc));
    }
    if (lane == 0) ptx::tma_store_wait_all<0>();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc<512>(tmem_base);
  }
}

static int stem_num_sms() {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

static int stem_check(const rg_igemm_desc* d) {
  RG_UNSUPPORTED(d->mode != 0, "stem conv: forward only");
  long long K = d->cin;
  for (int i = 0; i < 3; ++i) {
    RG_UNSUPPORTED(d->stride[i] != 1, "stem conv: stride must be 1");
    RG_UNSUPPORTED(d->kernel[i] < 1 || d->pad[i] < 0, "stem conv: bad geometry");
    RG_UNSUPPORTED(d->out_dhw[i] != d->in_dhw[i] + 2 * d->pad[i] - d->kernel[i] + 1 || d->out_dhw[i] < 1,
                   "stem conv: extents inconsistent with geometry");
    K *= d->kernel[i];
  }
  RG_UNSUPPORTED(K < 1 || K > 64, "stem conv: Cin * taps must be <= 64");
  RG_UNSUPPORTED((long long)d->n * d->out_dhw[0] * ((d->out_dhw[1] + 3) / 4) * ((d->out_dhw[2] + ST_TW - 1) / ST_TW) >=
                     (1LL << 31) || llabs(d->x_strides[3]) >= (1LL << 20), "stem conv: problem too large for 32-bit tile indices");
  RG_UNSUPPORTED(d->cout % 32 != 0 || d->cout < 32 || d->cout > ST_MAX_COUT, "stem conv: Cout must be a multiple of 32, <= 256");
  for (int i = 0; i < 4; ++i) RG_UNSUPPORTED(d->y_strides[i] % 8 != 0, "stem conv: output strides must be multiples of 16 bytes");
  return RG_OK;
}

template <int KSTEPS, bool SPEC>
static int stem_launch(const StemParams& p, int grid, size_t smem, cudaStream_t st) {
  static bool set = false;
  if (!set) {
    if (cudaFuncSetAttribute(stem_conv_kernel<KSTEPS, SPEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
      set_error("stem conv: cannot raise the dynamic shared memory limit");
      return RG_E_CUDA;
    }
    set = true;
  }
  stem_conv_kernel<KSTEPS, SPEC><<<grid, ST_THREADS, smem, st>>>(p);
  return RG_OK;
}

}  // namespace rg

using namespace rg;

extern "C" {

int rg_stem_conv_supported(const rg_igemm_desc* d) {
  if (!d) return 0;
  return stem_check(d) == RG_OK ? 1 : 0;
}

int rg_stem_conv_run(const rg_igemm_desc* d, const void* x, long long x_channel_stride, const void* w64, const float* bias,
                     void* y, int pad_mode, void* stream) {
  RG_CHECK_ARG(d && x && w64 && y, "rg_stem_conv_run: null argument");
  RG_CHECK_ARG(pad_mode >= RG_PAD_ZEROS && pad_mode <= RG_PAD_REPLICATE, "rg_stem_conv_run: bad pad mode %d", pad_mode);
  int rc = stem_check(d);
  if (rc) return rc;
  if (pad_mode == RG_PAD_REFLECT)
    for (int i = 0; i < 3; ++i) RG_CHECK_ARG(d->pad[i] < d->in_dhw[i], "rg_stem_conv_run: reflect padding must be smaller than the input");
  RG_CHECK_ARG(((uintptr_t)y & 15) == 0 && ((uintptr_t)w64 & 15) == 0, "rg_stem_conv_run: output / weights must be 16-byte aligned");
  StemParams p;
  memset(&p, 0, sizeof(p));
  p.x = (const __nv_bfloat16*)x;
  p.xs_n = d->x_strides[0]; p.xs_d = d->x_strides[1]; p.xs_h = d->x_strides[2]; p.xs_w = d->x_strides[3];
  p.xs_c = x_channel_stride;
  p.cin = d->cin; p.kd = d->kernel[0]; p.kh = d->kernel[1]; p.kw = d->kernel[2];
  p.pd = d->pad[0]; p.ph = d->pad[1]; p.pw = d->pad[2];
  p.K = d->cin * d->kernel[0] * d->kernel[1] * d->kernel[2];
  p.pad_mode = pad_mode;
  p.in_d = d->in_dhw[0]; p.in_h = d->in_dhw[1]; p.in_w = d->in_dhw[2];
  p.out_d = d->out_dhw[0]; p.out_h = d->out_dhw[1]; p.out_w = d->out_dhw[2];
  p.batch = d->n;
  const bool spec = d->cin == 1 && d->kernel[0] == 3 && d->kernel[1] == 3 && d->kernel[2] == 3;
  p.tile_h = spec ? 4 : ST_TH; p.tile_w = spec ? 32 : ST_TW;
  p.quad_h = spec ? 1 : 4;
  p.afull_count = spec ? 1 : 128;
  p.tiles_h = (p.out_h + p.tile_h - 1) / p.tile_h;
  p.tiles_w = (p.out_w + p.tile_w - 1) / p.tile_w;
  p.total_tiles = (long long)p.batch * p.out_d * p.tiles_h * p.tiles_w;
  p.w = (const __nv_bfloat16*)w64;
  p.cout = d->cout;
  p.bias = d->has_bias ? bias : nullptr;
  p.act = d->act; p.slope = d->slope;
  {
    typedef CUresult (*EncFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncFn enc = nullptr;
    if (!enc) {
      void* q = nullptr;
      cudaDriverEntryPointQueryResult r;
      if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &q, cudaEnableDefault, &r) == cudaSuccess && r == cudaDriverEntryPointSuccess)
        enc = (EncFn)q;
    }
    if (!enc) { set_error("cuTensorMapEncodeTiled unavailable"); return RG_E_CUDA; }
    cuuint64_t dims[5] = {(cuuint64_t)d->cout, (cuuint64_t)d->out_dhw[2], (cuuint64_t)d->out_dhw[1], (cuuint64_t)d->out_dhw[0], (cuuint64_t)d->n};
    cuuint64_t gs[4] = {(cuuint64_t)(d->y_strides[3] * 2), (cuuint64_t)(d->y_strides[2] * 2), (cuuint64_t)(d->y_strides[1] * 2),
                        (cuuint64_t)(d->y_strides[0] * 2)};
    cuuint32_t bx[5] = {32, (cuuint32_t)p.tile_w, (cuuint32_t)p.quad_h, 1, 1};     // 32 voxels of one TMEM lane quadrant
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(&p.y_map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, y, dims, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("stem conv: cuTensorMapEncodeTiled (output) failed: %d", (int)r); return RG_E_CUDA; }
  }
  const size_t smem = 1024 + (size_t)ST_STAGES * ST_A_BYTES + ST_MAX_COUT * 128 + ST_OUT_BYTES + 8 * ST_PATCH_BYTES +
                      8 * (2 * ST_STAGES + 4) + 64 + ST_MAX_COUT * 4;
  const int grid = (int)(p.total_tiles < stem_num_sms() ? p.total_tiles : stem_num_sms());
  cudaStream_t st = (cudaStream_t)stream;
  const int ksteps = (p.K + 15) / 16;
  if (spec) rc = stem_launch<2, true>(p, grid, smem, st);
  else if (ksteps == 1) rc = stem_launch<1, false>(p, grid, smem, st);
  else if (ksteps == 2) rc = stem_launch<2, false>(p, grid, smem, st);
  else if (ksteps == 3) rc = stem_launch<3, false>(p, grid, smem, st);
  else rc = stem_launch<4, false>(p, grid, smem, st);
  if (rc) return rc;
  RG_LAUNCH_CHECK("rg_stem_conv_run");
  return RG_OK;
}

}  // extern "C"
