// Slice-ring implicit-GEMM Conv3d (stride 1) on tcgen05 tensor cores: the kernel for layers that produce 64-channel
// blocks with many taps (the 7^3 64->64 ResNet head, ResNet.py:324-327, is 70 % of the generator's FLOPs; the 3^3
// 64-channel U-Net levels, UNet.py:52-87).  The plain implicit GEMM (igemm_sm100.cu) is L2-feed bound there: it pulls
// a fresh 16 KB activation tile and an 8 KB weight tile per 128x64x64 MMA block (192 B/clk/SM at tensor rate), and a
// 128 x 64 x 16 SS-mode MMA is itself shared-memory-read bound (48 clk instead of 32; profiles/r01_exp_umma_*.txt).
//
// Formulation (input-stationary along depth):
//   * a CTA owns an output column: a 16(h) x 8(w) tile, a depth segment [d0, d0+L);
//   * input slice z (one depth plane of the tile WITH its (kh,kw) halo, 64 channels) is loaded ONCE by TMA into
//     128B-swizzled smem; the A operand of tap (kh,kw) is the same buffer addressed through a UMMA descriptor whose
//     start is shifted by (kh*WB+kw) rows and whose 8-row-group stride is WB*128 B (swizzle is anchored on absolute
//     smem address bits, so row-shifted descriptors are exact -- verified by scripts/exp_umma.cu);
//   * slice z contributes to the KD output planes d = z+pd-kd, so one MMA multiplies the A tile with the weights of
//     ALL kd at once: N = 64*KD columns (split in <= 256-column pieces) -> full-rate MMAs, A read once per KD taps;
//   * TMEM holds a ring of eight 128x64 fp32 accumulators, one per output plane in flight; plane d lives in block
//     (d-d0) mod 8.  S consecutive input slices share every weight stage (S + KD - 1 <= 8), which divides the weight
//     traffic by S.  When a plane has received its last slice, the epilogue warps drain it (bias, activation, bf16,
//     Instance/BatchNorm sum / sum-of-squares), re-zero the block and hand it back.
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner, warps 2-9 = epilogue (two per TMEM lane
// quadrant, 32 columns each).
#include <cuda.h>
#include <cuda_bf16.h>

#include <stdlib.h>

#include "common.cuh"
#include "sm100_ptx.cuh"
#include "reduce_rows.cuh"

namespace rg {

constexpr int RING_THREADS = 320;       // warp 0 producer, warp 1 MMA issuer, warps 2-9 epilogue
constexpr int RING_BLOCKS = 8;            // accumulator blocks of 64 TMEM columns
constexpr int WBLK_BYTES = 64 * 128;      // weights of one (tap, kd): 64 produced channels x 64 reduction channels bf16
constexpr int STAGE_KD = 4;               // kd blocks per weight stage (<= 256 MMA columns)
constexpr int RING_TH = 16, RING_TW = 8;  // output tile (h, w); 8 = one swizzle-atom row group
constexpr int MAX_NA = 12, MAX_NB = 4;
constexpr int MAX_GS = 4;                                           // slices per group (S) the MMA warp unrolls
constexpr int RING_HEAD_MAX = 4;                                     // output channels of a fused 1x1x1 head
constexpr int RING_BAR_BYTES = 8 * (2 * MAX_NA + 2 * MAX_NB + 4) + 64;
constexpr int RING_HEAD_W_BYTES = RING_HEAD_MAX * 64 * 4;           // head weights [o][64] fp32
constexpr int RING_HEAD_X_BYTES = 4 * 2 * 32 * RING_HEAD_MAX * 4;   // partial dots: [quad][plane parity][lane][o]
constexpr int RING_TAIL_BYTES = RING_BAR_BYTES + RING_HEAD_W_BYTES + RING_HEAD_X_BYTES;

struct alignas(64) RingParams {
  CUtensorMap a_map;      // 5-D {C, W, H, D, N}; box {64, WB, HB, 1, 1}
  CUtensorMap b_map;      // 2-D {64, rows};      box {64, 64}
  int kd, kh, kw;
  int pd, ph, pw;         // coordinate offsets (zero padding through TMA out-of-bounds fill)
  int q_chunks;           // reduction channels / 64
  int n_tiles_n;          // produced channels / 64
  int S, lag, na, nb;
  int slice_bytes;        // smem pitch of a slice buffer (multiple of 1024)
  int slice_tx;           // bytes one slice TMA delivers: HB * WB * 128
  int stage_bytes;        // smem pitch of a weight stage
  int wb;                 // slice row pitch in voxels: RING_TW + kw - 1
  int tiles_h, tiles_w, segs, seg_len, batch;
  int in_d;               // depth extent of the input (fully out-of-range slices are skipped)
  int out_d, out_h, out_w;
  long long ys_n, ys_d, ys_h, ys_w;
  __nv_bfloat16* y;
  int cout_total;
  int act; float slope;
  const float* bias;
  float* stats_part;      // partial rows [4 * gridDim.x][2 * stats_G]: row = 4 * CTA + TMEM lane quadrant (reduce_rows.cuh)
  long long stats_G;      // batch * cout_total
  // fused 1x1x1 head behind a 64-channel layer (rg_ring_run_head): the activated, bf16-rounded channels of a voxel are
  // reduced against head_w in the epilogue and only head_n values per voxel leave the SM (y is not written)
  const float* head_w;    // [head_n][64] fp32, nullptr = no head
  const float* head_b;    // [head_n] or nullptr
  int head_n, head_act; float head_slope;
  __nv_bfloat16* head_y;
  long long hs_n, hs_d, hs_h, hs_w;
};

struct RingUnit { int nt, n, h0, w0, d0, len, z_begin, z_end, zg0; };

// Work partition: the (column, output plane) space is flattened column-major and cut into gridDim.x equal ranges, so
// every CTA does the same number of plane-taps (no wave quantisation); a range is walked as at most a few units
// (column, d0, len), each of which re-reads only the KD-1 halo slices of its column.
struct RingWalk {
  long long pos, end;
  __device__ __forceinline__ RingWalk(const RingParams& p) {
    if (p.segs > 0) {          // equal depth segments, round-robin over CTAs (neighbouring columns run concurrently)
      pos = blockIdx.x;
      end = (long long)p.batch * p.segs * p.tiles_h * p.tiles_w * p.n_tiles_n;
    } else {
      const long long total = (long long)p.batch * p.tiles_h * p.tiles_w * p.n_tiles_n * p.out_d;
      pos = total * blockIdx.x / gridDim.x;
      end = total * (blockIdx.x + 1) / gridDim.x;
    }
  }
  __device__ __forceinline__ bool next(const RingParams& p, RingUnit& r) {
    if (pos >= end) return false;
    int u;
    if (p.segs > 0) {
      u = (int)pos;
      pos += gridDim.x;
      r.nt = u % p.n_tiles_n; u /= p.n_tiles_n;
      const int iw = u % p.tiles_w; u /= p.tiles_w;
      const int ih = u % p.tiles_h; u /= p.tiles_h;
      const int sg = u % p.segs; r.n = u / p.segs;
      r.h0 = ih * RING_TH; r.w0 = iw * RING_TW; r.d0 = sg * p.seg_len;
      r.len = min(p.seg_len, p.out_d - r.d0);
    } else {
      u = (int)(pos / p.out_d);
      r.d0 = (int)(pos - (long long)u * p.out_d);
      r.len = (int)min((long long)(p.out_d - r.d0), end - pos);
      pos += r.len;
      r.nt = u % p.n_tiles_n; u /= p.n_tiles_n;
      const int iw = u % p.tiles_w; u /= p.tiles_w;
      const int ih = u % p.tiles_h; r.n = u / p.tiles_h;
      r.h0 = ih * RING_TH; r.w0 = iw * RING_TW;
    }
    r.z_begin = r.d0 - p.pd;
    r.z_end = r.d0 + r.len - 1 - p.pd + p.kd - 1;
    // Slice groups are anchored on the ABSOLUTE plane index (group boundaries where z + pd is a multiple of S), not on the
    // unit's first plane: the order in which an output voxel receives its (slice, chunk, tap) contributions -- hence its
    // fp32 rounding -- is then independent of how the depth axis was cut into units, and of the read-block geometry of a
    // blockwise prediction whenever block origins are multiples of S at this resolution level (S <= 3: ring_geometry).
    r.zg0 = r.z_begin - (r.d0 % p.S);
    return true;
  }
};

#ifdef RING_DEBUG_NARROW      // experiment: every MMA only 64 columns wide (wrong results; is the loop pipe- or issue-bound?)
#define RING_DBG_N(c) 1
#else
#define RING_DBG_N(c) (c)
#endif
#ifdef RING_DEBUG_TIMING
// developer instrumentation (scripts/ring_timing.py): cycles the MMA warp spends per wait site, per CTA
__device__ unsigned long long ring_dbg[148 * 8];
#define DBG_T0() const long long dbg_t0 = clock64()
#define DBG_ADD(k) dbg_acc[k] += clock64() - dbg_t0
#else
#define DBG_T0()
#define DBG_ADD(k)
#endif

__global__ void __launch_bounds__(RING_THREADS, 1) ring_conv_kernel(const __grid_constant__ RingParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);
  const uint32_t smem_a = base;
  const uint32_t smem_b = base + (uint32_t)(p.na * p.slice_bytes);
  uint8_t* tail = base_ptr + (size_t)p.na * p.slice_bytes + (size_t)p.nb * p.stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);     // afull[na] aempty[na] bfull[nb] bempty[nb] tfull[2] tempty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 8 * (2 * MAX_NA + 2 * MAX_NB + 4));
  float* head_ws = reinterpret_cast<float*>(tail + RING_BAR_BYTES);
  float* head_xs = reinterpret_cast<float*>(tail + RING_BAR_BYTES + RING_HEAD_W_BYTES);
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto afull = [&](int s) { return bar0 + 8u * s; };
  auto aempty = [&](int s) { return bar0 + 8u * (MAX_NA + s); };
  auto bfull = [&](int s) { return bar0 + 8u * (2 * MAX_NA + s); };
  auto bempty = [&](int s) { return bar0 + 8u * (2 * MAX_NA + MAX_NB + s); };
  auto tfull = [&](int s) { return bar0 + 8u * (2 * MAX_NA + 2 * MAX_NB + s); };
  auto tempty = [&](int s) { return bar0 + 8u * (2 * MAX_NA + 2 * MAX_NB + 2 + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntaps_hw = p.kh * p.kw;
  const int parts = (p.kd + STAGE_KD - 1) / STAGE_KD;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < p.na; ++s) { ptx::mbar_init(afull(s), 1); ptx::mbar_init(aempty(s), 1); }
    for (int s = 0; s < p.nb; ++s) { ptx::mbar_init(bfull(s), 1); ptx::mbar_init(bempty(s), 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(tfull(s), 1); ptx::mbar_init(tempty(s), 256); }
    ptx::fence_mbar_init();
    ptx::prefetch_tmap(&p.a_map);
    ptx::prefetch_tmap(&p.b_map);
  }
  if (warp == 1) ptx::tmem_alloc<512>(ptx::smem_u32(tmem_slot));
  if (p.head_w)
    for (int i = threadIdx.x; i < p.head_n * 64; i += RING_THREADS) head_ws[i] = p.head_w[i];
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (warp >= 2 && warp < 6) {   // all accumulator blocks start at zero: every MMA accumulates
    const uint32_t t = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
#pragma unroll 1
    for (int c = 0; c < 512; c += 32) ptx::tmem_st_32x32_fill(t + c, 0u);
    ptx::tmem_st_wait();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();

  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer =================
      int a_i = 0; uint32_t a_ph = 0;
      int b_i = 0; uint32_t b_ph = 0;
      RingUnit r;
      for (RingWalk walk(p); walk.next(p, r);) {
        for (int zg = r.zg0; zg <= r.z_end; zg += p.S) {
          for (int q = 0; q < p.q_chunks; ++q) {
            auto load_slice = [&](int s) {
              const int z = zg + s;
              if (z < r.z_begin || z > r.z_end || z < 0 || z >= p.in_d) return;
              ptx::mbar_wait_backoff(aempty(a_i), a_ph ^ 1u);
              ptx::mbar_arrive_expect_tx(afull(a_i), (uint32_t)p.slice_tx);
              ptx::tma_load_5d(smem_a + (uint32_t)(a_i * p.slice_bytes), &p.a_map, afull(a_i), q * 64, r.w0 - p.pw,
                               r.h0 - p.ph, z, r.n);
              if (++a_i == p.na) { a_i = 0; a_ph ^= 1u; }
            };
            // consumer order: slice 0, first weight stage, remaining slices, remaining weight stages
            load_slice(0);
            bool rest_done = false;
            for (int t = 0; t < ntaps_hw; ++t)
              for (int part = 0; part < parts; ++part) {
                const int nblk = min(STAGE_KD, p.kd - part * STAGE_KD);
                ptx::mbar_wait_backoff(bempty(b_i), b_ph ^ 1u);
                ptx::mbar_arrive_expect_tx(bfull(b_i), (uint32_t)(nblk * WBLK_BYTES));
                const int row0 = ((((r.nt * p.q_chunks + q) * ntaps_hw + t) * p.kd) + part * STAGE_KD) * 64;
                for (int j = 0; j < nblk; ++j)
                  ptx::tma_load_2d(smem_b + (uint32_t)(b_i * p.stage_bytes + j * WBLK_BYTES), &p.b_map, bfull(b_i), 0,
                                   row0 + j * 64);
                if (++b_i == p.nb) { b_i = 0; b_ph ^= 1u; }
                if (!rest_done) {
                  for (int s = 1; s < p.S; ++s) load_slice(s);
                  rest_done = true;
                }
              }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    // The loop is ISSUE-bound unless it is lean: a 128 x 192 x 16 MMA occupies the tensor pipe for 96 clk and a single
    // thread issues a dependent instruction every ~7-10 clk, so an MMA may cost ~10 instructions at most (measured with
    // scripts/ring_timing.py: 16 instructions per MMA, 8 of them R2UR, ran at 137 clk per MMA whatever its width).
    // tcgen05.mma takes every operand from UNIFORM registers.  Inside an elect.sync region ptxas knows that one
    // thread is active and computes descriptors directly in the uniform datapath; a predicated issue from a
    // convergent warp instead moves every operand R -> UR per MMA.  So: one elected thread runs the whole loop, the
    // (slice, weight part) -> accumulator block mapping is integer arithmetic done once per group, and the four
    // K steps of a chunk are one asm block.
    if (ptx::elect_one_sync()) {
    const uint32_t idesc_base = ptx::idesc_bf16(128, 0, 0, 0);
    const uint32_t a_hi = (((uint32_t)(p.wb * 128) >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);   // SBO | version | SW128
    const uint32_t b_hi = ((1024u >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);
    const uint32_t lbo_lo = 1u << 16;
    const uint32_t tmem_u = tmem_base;
    int a_i = 0; uint32_t a_ph = 0;
    int b_i = 0; uint32_t b_ph = 0;
    int g = 0, waited = -1;
    RingUnit r;
#ifdef RING_DEBUG_TIMING
    long long dbg_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long long dbg_start = clock64();
#endif
    for (RingWalk walk(p); walk.next(p, r);) {
      bool first = true;
      const int d0 = r.d0, d1 = d0 + r.len;
      const int z_begin = d0 - p.pd, z_end = d1 - 1 - p.pd + p.kd - 1;
      const int z_last = min(z_end, p.in_d - 1);
      for (int zg = r.zg0; zg <= z_end; zg += p.S) {
        // slice zg+s: weight block i (kd = KD-1-i) feeds output plane d = dbase + i, dbase = z + pd - KD + 1, which
        // lives in accumulator block (d - d0) & 7; only planes inside [d0, d1) exist: i in [ilo, ihi)
        // Per (weight part, slice) that is at most two MMA pieces: blocks up to the end of the ring, then the rest
        // from block 0.  Everything the tap loop needs is computed here, once per group.
        uint32_t tm1[2][MAX_GS], id1[2][MAX_GS], id2[2][MAX_GS], wo1[2][MAX_GS], wo2[2][MAX_GS];
        int ord[MAX_GS];
        int nvalid = 0;
        { DBG_T0();
#pragma unroll
        for (int s = 0; s < MAX_GS; ++s) {
          const int z = zg + s;
          const bool valid = s < p.S && z >= z_begin && z >= 0 && z <= z_last;
          const int dbase = z + p.pd - p.kd + 1;
          const int ilo = max(0, d0 - dbase), ihi = valid ? min(p.kd, d1 - dbase) : 0;
#pragma unroll
          for (int part = 0; part < 2; ++part) {
            const int p_lo = part * STAGE_KD, p_hi = min(p.kd, p_lo + STAGE_KD);
            const int lo = max(p_lo, ilo), cnt = min(p_hi, ihi) - lo;
            const int blk = (dbase - d0 + lo) & (RING_BLOCKS - 1);
            const int c1 = min(cnt, RING_BLOCKS - blk);       // blocks before the ring wraps
            tm1[part][s] = tmem_u + (uint32_t)(blk * 64);
            id1[part][s] = cnt > 0 ? idesc_base | ((uint32_t)(RING_DBG_N(c1) * 8) << 17) : 0u;
            id2[part][s] = cnt > c1 && cnt > 0 ? idesc_base | ((uint32_t)(RING_DBG_N(cnt - c1) * 8) << 17) : 0u;
            wo1[part][s] = (uint32_t)((lo - p_lo) * (WBLK_BYTES >> 4));
            wo2[part][s] = (uint32_t)((lo - p_lo + c1) * (WBLK_BYTES >> 4));
          }
          ord[s] = nvalid;
          nvalid += valid ? 1 : 0;
        }
        DBG_ADD(0); }
        const int need = first ? g - 1 : g - p.lag;
        { DBG_T0();
        while (waited < need) { ++waited; ptx::mbar_wait(tempty(waited & 1), (uint32_t)((waited >> 1) & 1)); }
        DBG_ADD(1); }
        ptx::tc_fence_after();
        for (int q = 0; q < p.q_chunks; ++q) {
          // slices of this chunk: wait for their TMA, then the A descriptor base of every slice
          { DBG_T0();
          for (int j = 0; j < nvalid; ++j) {
            int slot = a_i + j; uint32_t sph = a_ph;
            if (slot >= p.na) { slot -= p.na; sph ^= 1u; }
            ptx::mbar_wait(afull(slot), sph);
          }
          DBG_ADD(2); }
          ptx::tc_fence_after();
          uint32_t abase[MAX_GS];
#pragma unroll
          for (int s = 0; s < MAX_GS; ++s) {
            int slot = a_i + ord[s]; if (slot >= p.na) slot -= p.na;
            abase[s] = ((smem_a + (uint32_t)(slot * p.slice_bytes)) >> 4) | lbo_lo;
          }
          for (int tkh = 0; tkh < p.kh; ++tkh)
          for (int tkw = 0; tkw < p.kw; ++tkw) {
            const uint32_t a_shift = (uint32_t)((tkh * p.wb + tkw) * 8);        // rows * 128 B >> 4
#pragma unroll
            for (int part = 0; part < 2; ++part) {
              if (part < parts) {
                { DBG_T0();
                ptx::mbar_wait(bfull(b_i), b_ph);
                DBG_ADD(3); }
                ptx::tc_fence_after();
                const uint32_t b_lo = (((smem_b + (uint32_t)(b_i * p.stage_bytes)) >> 4) & 0x3FFFu) | lbo_lo;
#pragma unroll
                for (int s = 0; s < MAX_GS; ++s) {
                  if (id1[part][s]) {
                    const uint32_t a_lo = abase[s] + a_shift;
                    ptx::umma_bf16_x4(tm1[part][s], a_lo, a_hi, b_lo + wo1[part][s], b_hi, id1[part][s]);
                    if (id2[part][s]) ptx::umma_bf16_x4(tmem_u, a_lo, a_hi, b_lo + wo2[part][s], b_hi, id2[part][s]);
                  }
                }
                ptx::umma_commit(bempty(b_i));
                if (++b_i == p.nb) { b_i = 0; b_ph ^= 1u; }
              }
            }
          }
          for (int j = 0; j < nvalid; ++j) {
            ptx::umma_commit(aempty(a_i));
            if (++a_i == p.na) { a_i = 0; a_ph ^= 1u; }
          }
        }
        ptx::umma_commit(tfull(g & 1));
        ++g;
        first = false;
      }
    }
#ifdef RING_DEBUG_TIMING
    if (blockIdx.x < 148) {
      dbg_acc[4] = clock64() - dbg_start;
      for (int k = 0; k < 8; ++k) ring_dbg[blockIdx.x * 8 + k] = (unsigned long long)dbg_acc[k];
    }
#endif
    }
    __syncwarp();
  } else {
    // ================= epilogue (warps 2..9): warp pair (quad, half) owns TMEM lanes 32*quad.. and 32 of the 64 columns
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = quad * 32 + lane;
    const int rh = row >> 3, rw = row & 7;
    const uint32_t t_lane = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(half * 32);
    const bool want_stats = p.stats_part != nullptr;
    // no atomics: this warp is the only writer of its 32 columns in row 4 * CTA + quad
    float* st_row = want_stats ? p.stats_part + ((size_t)blockIdx.x * 4 + quad) * (size_t)(2 * p.stats_G) : nullptr;
    float st_sum = 0.f, st_sq = 0.f;          // lane l: channel half*32 + l of the current (n, channel tile)
    int st_n = -1, st_nt = -1;
    auto flush_stats = [&]() {
      if (!want_stats || st_n < 0) return;
      const size_t c = (size_t)st_n * p.cout_total + st_nt * 64 + half * 32 + lane;
      st_row[c] += st_sum;
      st_row[p.stats_G + c] += st_sq;
      st_sum = st_sq = 0.f;
    };
    const bool has_head = p.head_w != nullptr;
    uint32_t head_plane = 0;                    // planes this warp pair has drained (parity selects the exchange buffer)
    int g = 0;
    RingUnit r;
    for (RingWalk walk(p); walk.next(p, r);) {
      if (r.n != st_n || r.nt != st_nt) { flush_stats(); st_n = r.n; st_nt = r.nt; }
      const int oh = r.h0 + rh, ow = r.w0 + rw;
      const bool valid = oh < p.out_h && ow < p.out_w;
      float bias_r[32];
#pragma unroll
      for (int c = 0; c < 32; ++c) bias_r[c] = 0.f;
      if (p.bias) {
#pragma unroll
        for (int c = 0; c < 32; c += 4) {
          const float4 b4 = *reinterpret_cast<const float4*>(p.bias + r.nt * 64 + half * 32 + c);
          bias_r[c] = b4.x; bias_r[c + 1] = b4.y; bias_r[c + 2] = b4.z; bias_r[c + 3] = b4.w;
        }
      }
      __nv_bfloat16* ybase = p.y + (size_t)r.n * p.ys_n + (size_t)oh * p.ys_h + (size_t)ow * p.ys_w + r.nt * 64 + half * 32;
      for (int zg = r.zg0; zg <= r.z_end; zg += p.S) {
        ptx::mbar_wait_backoff(tfull(g & 1), (uint32_t)((g >> 1) & 1));
        ptx::tc_fence_after();
        // planes whose last contributing slice is in this group: d = z + pd - KD + 1
        const int d_first = max(r.d0, zg + p.pd - p.kd + 1);
        const int d_last = min(r.d0 + r.len - 1, min(zg + p.S - 1, r.z_end) + p.pd - p.kd + 1);
        if (d_last < d_first) {
          ptx::tc_fence_before();
          ptx::mbar_arrive(tempty(g & 1));
        }
        for (int d = d_first; d <= d_last; ++d) {
          const uint32_t taddr = t_lane + (uint32_t)(((d - r.d0) & (RING_BLOCKS - 1)) * 64);
          uint32_t v[32];
          ptx::tmem_ld_32x32(taddr, v);
          ptx::tmem_ld_wait();
          ptx::tmem_st_32x32_fill(taddr, 0u);
          if (d == d_last) {      // accumulators drained and re-zeroed: hand the blocks back before the arithmetic
            ptx::tmem_st_wait();
            ptx::tc_fence_before();
            ptx::mbar_arrive(tempty(g & 1));
          }
          float f[32];
#pragma unroll
          for (int c = 0; c < 32; ++c) f[c] = __uint_as_float(v[c]) + bias_r[c];
          if (p.act == RG_ACT_RELU) {
#pragma unroll
            for (int c = 0; c < 32; ++c) f[c] = fmaxf(f[c], 0.f);
          } else if (p.act == RG_ACT_LEAKY) {
#pragma unroll
            for (int c = 0; c < 32; ++c) f[c] = f[c] > 0.f ? f[c] : f[c] * p.slope;
          } else if (p.act != RG_ACT_NONE) {
#pragma unroll
            for (int c = 0; c < 32; ++c) f[c] = act_apply(f[c], p.act, p.slope);
          }
          uint32_t pk[16];
#pragma unroll
          for (int c = 0; c < 32; c += 2) {
            __nv_bfloat162 h2 = __floats2bfloat162_rn(valid ? f[c] : 0.f, valid ? f[c + 1] : 0.f);
            pk[c >> 1] = *reinterpret_cast<uint32_t*>(&h2);
          }
          if (has_head) {
            // thread = voxel: dot of this warp's 32 (bf16-rounded, as the unfused layer would have stored them) channels
            // with the head weights; the upper-half warp hands its partial sums to the lower-half warp of the same
            // lane quadrant through shared memory (one named barrier per plane, buffers alternate by plane parity)
            float hacc[RING_HEAD_MAX];
#pragma unroll
            for (int o = 0; o < RING_HEAD_MAX; ++o) {
              hacc[o] = 0.f;
              if (o < p.head_n) {
                const float4* w4 = reinterpret_cast<const float4*>(head_ws + o * 64 + half * 32);
#pragma unroll
                for (int c = 0; c < 32; c += 4) {
                  const float4 wv = w4[c >> 2];
                  const float2 x0 = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&pk[c >> 1]));
                  const float2 x1 = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&pk[(c >> 1) + 1]));
                  hacc[o] = fmaf(x0.x, wv.x, hacc[o]); hacc[o] = fmaf(x0.y, wv.y, hacc[o]);
                  hacc[o] = fmaf(x1.x, wv.z, hacc[o]); hacc[o] = fmaf(x1.y, wv.w, hacc[o]);
                }
              }
            }
            float* xch = head_xs + ((quad * 2 + (int)(head_plane & 1u)) * 32 + lane) * RING_HEAD_MAX;
            if (half == 1) *reinterpret_cast<float4*>(xch) = make_float4(hacc[0], hacc[1], hacc[2], hacc[3]);
            asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory");
            if (half == 0 && valid) {
              const float4 up = *reinterpret_cast<const float4*>(xch);
              __nv_bfloat16* hdst = p.head_y + (size_t)r.n * p.hs_n + (size_t)d * p.hs_d + (size_t)oh * p.hs_h + (size_t)ow * p.hs_w;
#pragma unroll
              for (int o = 0; o < RING_HEAD_MAX; ++o)
                if (o < p.head_n) {
                  const float other = o == 0 ? up.x : (o == 1 ? up.y : (o == 2 ? up.z : up.w));
                  hdst[o] = __float2bfloat16_rn(act_apply(hacc[o] + other + (p.head_b ? p.head_b[o] : 0.f), p.head_act, p.head_slope));
                }
            }
            ++head_plane;
          } else if (valid) {
            uint4* dst = reinterpret_cast<uint4*>(ybase + (size_t)d * p.ys_d);
#pragma unroll
            for (int j = 0; j < 4; ++j) dst[j] = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
          }
          if (want_stats) {
            // transpose-reduce over the warp's 32 rows: after 5 halving steps lane l holds channel l of this half
#pragma unroll
            for (int pass = 0; pass < 2; ++pass) {
              float w[32];
#pragma unroll
              for (int c = 0; c < 32; c += 2) {
                const float2 x = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&pk[c >> 1]));
                w[c] = pass ? x.x * x.x : x.x; w[c + 1] = pass ? x.y * x.y : x.y;
              }
#pragma unroll
              for (int m = 16, hf = 16; m >= 1; m >>= 1, hf >>= 1) {
                const bool up = (lane & m) != 0;
#pragma unroll
                for (int i = 0; i < hf; ++i) {
                  const float mine = up ? w[hf + i] : w[i];
                  const float other = up ? w[i] : w[hf + i];
                  w[i] = mine + __shfl_xor_sync(0xffffffffu, other, m);
                }
              }
              if (pass) st_sq += w[0]; else st_sum += w[0];
            }
          }
        }
        ++g;
      }
    }
    flush_stats();
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

typedef CUresult (*RingEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static RingEncodeFn ring_encode_fn() {
  static RingEncodeFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (RingEncodeFn)p;
  }
  return fn;
}

static int ring_num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

constexpr int RING_SMEM_LIMIT = 227 * 1024;

struct RingGeom {
  int kc, rows;                 // reduction / produced channels
  int in_ext[3], out_ext[3];    // d,h,w of the tensor read / written by this mode
  int pad[3];
  const int64_t* in_str; const int64_t* out_str;
  RingParams p;
  int smem_bytes;
};

// Fills everything except tensor maps and pointers; returns RG_OK / RG_E_UNSUPPORTED.
static int ring_geometry(const rg_igemm_desc* d, RingGeom* g) {
  RG_UNSUPPORTED(d->mode != 0 && d->mode != 1, "ring conv: fprop / dgrad only");
  for (int i = 0; i < 3; ++i) {
    RG_UNSUPPORTED(d->stride[i] != 1, "ring conv: stride must be 1");
    RG_UNSUPPORTED(d->kernel[i] < 1 || d->kernel[i] > 8, "ring conv: kernel extent must be 1..8");
    RG_UNSUPPORTED(d->pad[i] < 0 || d->pad[i] > d->kernel[i] - 1, "ring conv: padding must be within the kernel");
    RG_UNSUPPORTED(d->out_dhw[i] != d->in_dhw[i] + 2 * d->pad[i] - d->kernel[i] + 1 || d->out_dhw[i] < 1,
                   "ring conv: extents inconsistent with geometry");
  }
  g->kc = d->mode == 0 ? d->cin : d->cout;
  g->rows = d->mode == 0 ? d->cout : d->cin;
  RG_UNSUPPORTED(g->kc % 64 != 0 || g->rows % 64 != 0, "ring conv: channel counts must be multiples of 64");
  RG_UNSUPPORTED(d->x_strides[3] % 8 != 0 || d->y_strides[3] % 8 != 0, "ring conv: voxel stride must be a multiple of 8");
  for (int i = 0; i < 3; ++i) {
    g->in_ext[i] = d->mode == 0 ? d->in_dhw[i] : d->out_dhw[i];
    g->out_ext[i] = d->mode == 0 ? d->out_dhw[i] : d->in_dhw[i];
    g->pad[i] = d->mode == 0 ? d->pad[i] : d->kernel[i] - 1 - d->pad[i];
  }
  g->in_str = d->mode == 0 ? d->x_strides : d->y_strides;
  g->out_str = d->mode == 0 ? d->y_strides : d->x_strides;
  RingParams& p = g->p;
  memset(&p, 0, sizeof(p));
  p.kd = d->kernel[0]; p.kh = d->kernel[1]; p.kw = d->kernel[2];
  p.pd = g->pad[0]; p.ph = g->pad[1]; p.pw = g->pad[2];
  p.q_chunks = g->kc / 64; p.n_tiles_n = g->rows / 64;
  p.wb = RING_TW + p.kw - 1;
  const int hb = RING_TH + p.kh - 1;
  RG_UNSUPPORTED(p.wb > 256 || hb > 256, "ring conv: halo box too large");
  p.slice_tx = hb * p.wb * 128;
  p.slice_bytes = (p.slice_tx + 1023) & ~1023;
  p.stage_bytes = (p.kd < STAGE_KD ? p.kd : STAGE_KD) * WBLK_BYTES;
  // slices per weight stage (S + KD - 1 accumulator blocks in flight; one group of drain slack when 2S + KD - 1 <= 8).
  // Groups are anchored on absolute plane indices (RingWalk), so results are independent of the block geometry of a
  // blockwise prediction when block origins are multiples of S at every resolution level.  S <= 3 by default: with the
  // 24-voxel halo, any write size that is a multiple of 12 keeps the read origins multiples of 12 = 3 * 4 (S = 3 at the
  // coarsest of three U-Net levels; the 7^3 layers run S = 2).  RAYGUN_B200_RING_S = 2 relaxes that to multiples of 8
  // at ~5 % more weight traffic per slice (measured: 455 vs 480 Mvoxel/s on cfg4); 4 is for experiments.
  static const int s_cap = [] { const char* e = getenv("RAYGUN_B200_RING_S"); const int v = e ? atoi(e) : 3; return v < 1 ? 1 : (v > 4 ? 4 : v); }();
  p.S = RING_BLOCKS - (p.kd - 1);
  if (p.S > s_cap) p.S = s_cap;
  p.lag = (2 * p.S + p.kd - 1 <= RING_BLOCKS) ? 2 : 1;
  for (;;) {
    p.nb = 3;
    const int avail = RING_SMEM_LIMIT - 1024 - RING_TAIL_BYTES - p.nb * p.stage_bytes;
    int na = avail / p.slice_bytes;
    if (na > 2 * p.S) na = 2 * p.S;
    if (na > MAX_NA) na = MAX_NA;
    if (na >= p.S + 1 || (na >= p.S && p.S == 1)) { p.na = na; break; }
    RG_UNSUPPORTED(p.S == 1, "ring conv: slice does not fit shared memory");
    --p.S;
    p.lag = (2 * p.S + p.kd - 1 <= RING_BLOCKS) ? 2 : 1;
  }
  g->smem_bytes = 1024 + p.na * p.slice_bytes + p.nb * p.stage_bytes + RING_TAIL_BYTES;
  p.tiles_h = (g->out_ext[1] + RING_TH - 1) / RING_TH;
  p.tiles_w = (g->out_ext[2] + RING_TW - 1) / RING_TW;
  p.batch = d->n;
  p.in_d = g->in_ext[0];
  p.out_d = g->out_ext[0]; p.out_h = g->out_ext[1]; p.out_w = g->out_ext[2];
  // depth segment length: balance wave quantisation against the (KD-1) extra slices every segment re-reads
  {
    const long long cols = (long long)d->n * p.tiles_h * p.tiles_w * p.n_tiles_n;
    const int sms = ring_num_sms();
    double best = 1e300; int best_len = p.out_d;
    for (int segs = 1; segs <= p.out_d; ++segs) {
      const int len = (p.out_d + segs - 1) / segs;
      if ((p.out_d + len - 1) / len != segs) continue;
      if (len < 4 && segs > 1) break;
      const long long units = cols * segs;
      const double waves = (double)((units + sms - 1) / sms);
      const double cost = waves * (len + 0.35 * (p.kd - 1) + 1.0);
      if (cost < best) { best = cost; best_len = len; }
    }
    p.seg_len = best_len;
    p.segs = (p.out_d + best_len - 1) / best_len;
    static const char* mode = getenv("RAYGUN_B200_RING_PARTITION");
    if (!(mode && mode[0] == 's')) p.segs = 0;    // default "balanced": equal flattened plane ranges per CTA;
                                                  // RAYGUN_B200_RING_PARTITION=segments keeps round-robin segments
  }
  p.cout_total = g->rows;
  p.act = d->act; p.slope = d->slope;
  p.ys_n = g->out_str[0]; p.ys_d = g->out_str[1]; p.ys_h = g->out_str[2]; p.ys_w = g->out_str[3];
  return RG_OK;
}

// fp32 [Cout][Cin][kd][kh][kw] -> bf16 [nt][q][kh][kw][i][64 produced][64 reduction], kd = KD-1-i (see kernel header).
__global__ void ring_pack_kernel(const float* __restrict__ w, int cout, int cin, int kd, int kh, int kw, int mode,
                                 __nv_bfloat16* __restrict__ out) {
  const int rows = mode == 0 ? cout : cin, kc = mode == 0 ? cin : cout;
  const int Q = kc / 64;
  const long long total = (long long)cout * cin * kd * kh * kw;
  for (long long o = blockIdx.x * (long long)blockDim.x + threadIdx.x; o < total; o += (long long)gridDim.x * blockDim.x) {
    long long r = o;
    const int ck = (int)(r % 64); r /= 64;
    const int cn = (int)(r % 64); r /= 64;
    const int i = (int)(r % kd); r /= kd;
    const int b = (int)(r % kw); r /= kw;
    const int a = (int)(r % kh); r /= kh;
    const int q = (int)(r % Q); const int nt = (int)(r / Q);
    int co, ci, td, th, tw;
    if (mode == 0) { co = nt * 64 + cn; ci = q * 64 + ck; td = kd - 1 - i; th = a; tw = b; }
    else { ci = nt * 64 + cn; co = q * 64 + ck; td = i; th = kh - 1 - a; tw = kw - 1 - b; }
    (void)rows;
    out[o] = __float2bfloat16_rn(w[((((long long)co * cin + ci) * kd + td) * kh + th) * kw + tw]);
  }
}

}  // namespace rg

using namespace rg;

extern "C" {

int rg_ring_supported(const rg_igemm_desc* d) {
  if (!d) return 0;
  RingGeom g;
  return ring_geometry(d, &g) == RG_OK ? 1 : 0;
}

int rg_ring_pack_weights(const float* w, int cout, int cin, const int* kernel, int mode, void* out_bf16, void* stream) {
  RG_CHECK_ARG(w && out_bf16 && kernel && cout > 0 && cin > 0 && (mode == 0 || mode == 1), "rg_ring_pack_weights: bad argument");
  RG_UNSUPPORTED(cout % 64 != 0 || cin % 64 != 0, "rg_ring_pack_weights: channel counts must be multiples of 64");
  const long long total = (long long)cout * cin * kernel[0] * kernel[1] * kernel[2];
  ring_pack_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(w, cout, cin, kernel[0], kernel[1], kernel[2], mode,
                                                                          (__nv_bfloat16*)out_bf16);
  RG_LAUNCH_CHECK("rg_ring_pack_weights");
  return RG_OK;
}

size_t rg_ring_workspace_bytes(const rg_igemm_desc* d) {
  if (!d) return 0;
  const int rows = d->mode == 0 ? d->cout : d->cin;
  return (size_t)4 * ring_num_sms() * 2 * d->n * rows * sizeof(float);
}

}  // extern "C"

namespace rg {
struct RingHead { const float* w; const float* b; int n, act; float slope; void* out; const int64_t* strides; };
}

static int ring_run(const rg_igemm_desc* d, const void* in, const void* w_ring, const float* bias, void* out, float* stats_sum,
                    float* stats_sqsum, void* workspace, const rg::RingHead* head, void* stream) {
  RingGeom g;
  int rc = ring_geometry(d, &g);
  if (rc) return rc;
  RingEncodeFn enc = ring_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled unavailable"); return RG_E_CUDA; }
  RingParams& p = g.p;
  RG_CHECK_ARG(((uintptr_t)in & 15) == 0 && ((uintptr_t)out & 15) == 0, "rg_ring_run: tensors must be 16-byte aligned");
  for (int i = 0; i < 4; ++i)
    RG_UNSUPPORTED((g.in_str[i] * 2) % 16 != 0 || (g.out_str[i] * 2) % 16 != 0, "rg_ring_run: strides must be multiples of 16 bytes");
  {
    cuuint64_t dims[5] = {(cuuint64_t)g.kc, (cuuint64_t)g.in_ext[2], (cuuint64_t)g.in_ext[1], (cuuint64_t)g.in_ext[0], (cuuint64_t)d->n};
    cuuint64_t gs[4] = {(cuuint64_t)(g.in_str[3] * 2), (cuuint64_t)(g.in_str[2] * 2), (cuuint64_t)(g.in_str[1] * 2), (cuuint64_t)(g.in_str[0] * 2)};
    cuuint32_t bx[5] = {64, (cuuint32_t)p.wb, (cuuint32_t)(RING_TH + p.kh - 1), 1, 1};
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(&p.a_map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, (void*)in, dims, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("ring conv: cuTensorMapEncodeTiled (activation) failed: %d", (int)r); return RG_E_CUDA; }
  }
  {
    const long long rows = (long long)p.n_tiles_n * p.q_chunks * p.kh * p.kw * p.kd * 64;
    cuuint64_t dims[2] = {64, (cuuint64_t)rows};
    cuuint64_t gs[1] = {128};
    cuuint32_t bx[2] = {64, 64};
    cuuint32_t es[2] = {1, 1};
    CUresult r = enc(&p.b_map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)w_ring, dims, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("ring conv: cuTensorMapEncodeTiled (weights) failed: %d", (int)r); return RG_E_CUDA; }
  }
  p.y = (__nv_bfloat16*)out;
  p.bias = bias;
  if (head) {
    p.head_w = head->w; p.head_b = head->b; p.head_n = head->n; p.head_act = head->act; p.head_slope = head->slope;
    p.head_y = (__nv_bfloat16*)head->out;
    p.hs_n = head->strides[0]; p.hs_d = head->strides[1]; p.hs_h = head->strides[2]; p.hs_w = head->strides[3];
  }
  p.stats_part = stats_sum ? (float*)workspace : nullptr;
  p.stats_G = (long long)d->n * g.rows;
  if (stats_sum) {
    cudaError_t e = cudaMemsetAsync(workspace, 0, rg_ring_workspace_bytes(d), (cudaStream_t)stream);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(ring statistics rows)");
  }
  static int smem_set = 0;
  if (smem_set < g.smem_bytes) {
    cudaError_t e = cudaFuncSetAttribute(ring_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, RING_SMEM_LIMIT);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(ring_conv)");
    smem_set = RING_SMEM_LIMIT;
  }
  long long want;
  if (p.segs > 0) want = (long long)p.batch * p.segs * p.tiles_h * p.tiles_w * p.n_tiles_n;
  else want = ((long long)p.batch * p.tiles_h * p.tiles_w * p.n_tiles_n * p.out_d + 3) / 4;
  const int grid = (int)(want < ring_num_sms() ? (want < 1 ? 1 : want) : ring_num_sms());
  ring_conv_kernel<<<grid, RING_THREADS, g.smem_bytes, (cudaStream_t)stream>>>(p);
  RG_LAUNCH_CHECK("rg_ring_run");
  if (stats_sum)
    return launch_rows_sum((const float*)workspace, 4 * ring_num_sms(), 2 * p.stats_G, 2 * p.stats_G, stats_sum, stats_sqsum,
                           p.stats_G, 0, 1.0, (cudaStream_t)stream);
  return RG_OK;
}

extern "C" {

int rg_ring_run(const rg_igemm_desc* d, const void* in, const void* w_ring, const float* bias, void* out, float* stats_sum,
                float* stats_sqsum, void* workspace, void* stream) {
  RG_CHECK_ARG(d && in && w_ring && out, "rg_ring_run: null argument");
  RG_CHECK_ARG((stats_sum == nullptr) == (stats_sqsum == nullptr), "rg_ring_run: stats pointers must come together");
  RG_CHECK_ARG(!stats_sum || workspace, "rg_ring_run: statistics need rg_ring_workspace_bytes() of workspace");
  return ring_run(d, in, w_ring, bias, out, stats_sum, stats_sqsum, workspace, nullptr, stream);
}

int rg_ring_run_head(const rg_igemm_desc* d, const void* in, const void* w_ring, const float* bias, const float* head_w,
                     const float* head_b, int head_n, int head_act, float head_slope, void* head_out,
                     const int64_t* head_strides, void* stream) {
  RG_CHECK_ARG(d && in && w_ring && head_w && head_out && head_strides, "rg_ring_run_head: null argument");
  RG_UNSUPPORTED(d->mode != 0 || d->cout != 64, "rg_ring_run_head: the producing layer must be a forward convolution with 64 output channels");
  RG_UNSUPPORTED(head_n < 1 || head_n > RING_HEAD_MAX, "rg_ring_run_head: 1..4 head channels");
  RG_CHECK_ARG(((uintptr_t)head_w & 15) == 0, "rg_ring_run_head: head weights must be 16-byte aligned");
  const rg::RingHead head{head_w, head_b, head_n, head_act, head_slope, head_out, head_strides};
  return ring_run(d, in, w_ring, bias, nullptr, nullptr, nullptr, nullptr, &head, stream);
}

/* Introspection for tests / DESIGN.md: the configuration the host would pick for this layer. */
int rg_ring_config(const rg_igemm_desc* d, int* out8) {
  RG_CHECK_ARG(d && out8, "rg_ring_config: null argument");
  RingGeom g;
  int rc = ring_geometry(d, &g);
  if (rc) return rc;
  out8[0] = g.p.S; out8[1] = g.p.lag; out8[2] = g.p.na; out8[3] = g.p.nb; out8[4] = g.p.seg_len; out8[5] = g.p.segs;
  out8[6] = g.smem_bytes; out8[7] = g.p.slice_bytes;
  return RG_OK;
}

#ifdef RING_DEBUG_TIMING
int rg_ring_debug_read(unsigned long long* out) {
  return cudaMemcpyFromSymbol(out, rg::ring_dbg, sizeof(rg::ring_dbg)) == cudaSuccess ? 0 : -1;
}
#endif

}  // extern "C"
