// Implicit-GEMM Conv3d / ConvTranspose3d on tcgen05 tensor cores (sm_100a), bf16 NDHWC, fp32 TMEM accumulate.
//
// One persistent, warp-specialised kernel serves fprop, dgrad and the transposed-conv forward:
//   D[128 voxels x BN channels] = sum over (tap, 64-channel chunk) of A[128 x 64] * B[BN x 64]^T
// A tile  = one 5-D TMA box {64 ch, tw, th, td, 1} of the activation tensor shifted by the tap offset;
//           TMA out-of-bounds zero fill IS the zero padding and the ragged-edge handling.
//           Strided (stride-2) gathers read one of 8 parity-subsampled views (separate tensor maps).
// B tile  = 2-D TMA box {64, BN} of the K-major packed weights.
// Both land in 128B-swizzled smem and feed tcgen05.mma (UMMA 128 x BN x 16) through smem descriptors.
// Warp roles: warp0 = TMA producer, warp1 = MMA issuer (+TMEM owner), warps2-5 = epilogue
// (tcgen05.ld -> bias / activation -> bf16 -> swizzled smem -> TMA store, + per-(n,c) sum / sumsq for
// Instance/BatchNorm).  Two TMEM accumulator stages overlap the epilogue of tile i with the MMAs of i+1.
#include <cuda.h>
#include <cuda_bf16.h>

#include <vector>

#include "common.cuh"
#include "sm100_ptx.cuh"
#include "reduce_rows.cuh"
#include "wgrad_finish.cuh"

namespace rg {

constexpr int TILE_M = 128;     // voxels per tile (UMMA M)
constexpr int BLOCK_K = 64;     // channels per K block (128 B of bf16 = one swizzle row)
constexpr int UMMA_K = 16;
constexpr int A_BYTES = TILE_M * BLOCK_K * 2;       // 16 KB
constexpr int STAGE_OUT_BYTES = TILE_M * 64 * 2;    // epilogue staging slab: 128 rows x 64 ch bf16
constexpr int IGEMM_THREADS = 192;
constexpr int MAX_TAPS_DIM = 8;

struct alignas(64) IgemmParams {
  CUtensorMap a_maps[8];
  CUtensorMap b_map;
  CUtensorMap o_map;
  int ntap[3];
  signed char tap_off[3][MAX_TAPS_DIM];
  signed char tap_par[3][MAX_TAPS_DIM];
  signed char tap_kidx[3][MAX_TAPS_DIM];
  int kdims[3];
  int kc;             // reduction channels per tap (multiple of 64)
  int tiles_w, tiles_h, tiles_d, batch;
  int tw, th, td;
  int a_box_bytes;    // bytes one activation TMA box delivers: tw*th*td rows of 128 B (<= A_BYTES: free-form boxes)
  int n_tiles_n;      // output-channel tiles of BN
  int out_w, out_h, out_d;
  int cout_total;
  int act; float slope;
  const float* bias;
  float* stats_part;      // partial rows [4 * gridDim.x][2 * stats_G]: row = 4 * CTA + epilogue warp, columns [0,G) sum, [G,2G) sumsq
  long long stats_G;      // batch * cout_total
};

template <int BN> struct IgemmCfg {
  static constexpr int B_BYTES = BN * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN >= 192) ? 4 : (BN >= 128 ? 6 : 8);
  static constexpr int TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));
  static constexpr int SMEM_BYTES = 1024 /*align slack*/ + STAGES * STAGE_BYTES + 2 * STAGE_OUT_BYTES + 1024 /*bars+bias*/;
};

template <int BN>
__global__ void __launch_bounds__(IGEMM_THREADS, 1) igemm_conv_kernel(const __grid_constant__ IgemmParams p) {
  using Cfg = IgemmCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);
  const uint32_t smem_a = base;                                     // STAGES x 16 KB
  const uint32_t smem_b = base + STAGES * A_BYTES;                  // STAGES x B_BYTES
  const uint32_t smem_o = smem_b + STAGES * Cfg::B_BYTES;           // 2 x 16 KB
  uint8_t* tail = base_ptr + STAGES * Cfg::STAGE_BYTES + 2 * STAGE_OUT_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);               // full[S], empty[S], tfull[2], tempty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 8 * (2 * STAGES + 4));
  float* bias_s = reinterpret_cast<float*>(tail + 8 * (2 * STAGES + 4) + 16);   // BN floats
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (STAGES + s); };
  auto tfull_bar = [&](int s) { return bar0 + 8u * (2 * STAGES + s); };
  auto tempty_bar = [&](int s) { return bar0 + 8u * (2 * STAGES + 2 + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_m = p.tiles_w * p.tiles_h * p.tiles_d * p.batch;
  const int total_tiles = tiles_m * p.n_tiles_n;
  const int chunks = p.kc / BLOCK_K;
  const int ntaps = p.ntap[0] * p.ntap[1] * p.ntap[2];
  const int kblocks = ntaps * chunks;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(full_bar(s), 1); ptx::mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(tfull_bar(s), 1); ptx::mbar_init(tempty_bar(s), 128); }
    ptx::fence_mbar_init();
    ptx::prefetch_tmap(&p.b_map);
    ptx::prefetch_tmap(&p.o_map);
    ptx::prefetch_tmap(&p.a_maps[0]);
  }
  if (warp == 1) ptx::tmem_alloc<Cfg::TMEM_COLS>(ptx::smem_u32(tmem_slot));
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto decode_tile = [&](int tile, int& nt, int& n, int& d0, int& h0, int& w0) {
    nt = tile % p.n_tiles_n;
    int m = tile / p.n_tiles_n;
    int iw = m % p.tiles_w; m /= p.tiles_w;
    int ih = m % p.tiles_h; m /= p.tiles_h;
    int id = m % p.tiles_d; n = m / p.tiles_d;
    w0 = iw * p.tw; h0 = ih * p.th; d0 = id * p.td;
  };

  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer =================
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        int nt, n, d0, h0, w0;
        decode_tile(tile, nt, n, d0, h0, w0);
        for (int a = 0; a < p.ntap[0]; ++a)
          for (int b = 0; b < p.ntap[1]; ++b)
            for (int c = 0; c < p.ntap[2]; ++c) {
              const int mi = p.tap_par[0][a] * 4 + p.tap_par[1][b] * 2 + p.tap_par[2][c];
              const CUtensorMap* am = &p.a_maps[mi];
              const int cd = d0 + p.tap_off[0][a], ch = h0 + p.tap_off[1][b], cw = w0 + p.tap_off[2][c];
              const int koff = ((p.tap_kidx[0][a] * p.kdims[1] + p.tap_kidx[1][b]) * p.kdims[2] + p.tap_kidx[2][c]) * p.kc;
              for (int q = 0; q < chunks; ++q) {
                ptx::mbar_wait(empty_bar(stage), phase ^ 1u);
                ptx::mbar_arrive_expect_tx(full_bar(stage), (uint32_t)(p.a_box_bytes + Cfg::B_BYTES));
                ptx::tma_load_5d(smem_a + stage * A_BYTES, am, full_bar(stage), q * BLOCK_K, cw, ch, cd, n);
                ptx::tma_load_2d(smem_b + stage * Cfg::B_BYTES, &p.b_map, full_bar(stage), koff + q * BLOCK_K, nt * BN);
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
              }
            }
      }
    }
  } else if (warp == 1) {
    if (ptx::elect_one_sync()) {      // elect.sync region: ptxas keeps the descriptors in the uniform datapath
      // ================= MMA issuer =================
      constexpr uint32_t idesc = ptx::idesc_bf16(TILE_M, BN, 0, 0);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        ptx::mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < kblocks; ++kb) {
          ptx::mbar_wait(full_bar(stage), phase);
          ptx::tc_fence_after();
          const uint64_t adesc = ptx::smem_desc_sw128(smem_a + stage * A_BYTES, 16, 1024);
          const uint64_t bdesc = ptx::smem_desc_sw128(smem_b + stage * Cfg::B_BYTES, 16, 1024);
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            // advance 32 B along K inside the 128 B swizzle row: +2 in the (addr >> 4) field
            ptx::umma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
          }
          ptx::umma_commit(empty_bar(stage));
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        ptx::umma_commit(tfull_bar(acc));
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
    __syncwarp();
  } else {
    // ================= epilogue (warps 2..5, 128 threads) =================
    const int et = threadIdx.x - 64;            // 0..127
    const int quad = warp & 3;                  // TMEM lane quadrant this warp may access
    const int row = quad * 32 + lane;           // tile row == TMEM lane
    const int rw = row % p.tw, rh = (row / p.tw) % p.th, rd = row / (p.tw * p.th);
    constexpr int NSLAB = BN / 64;
    float st_acc[NSLAB][4];
#pragma unroll
    for (int s = 0; s < NSLAB; ++s) st_acc[s][0] = st_acc[s][1] = st_acc[s][2] = st_acc[s][3] = 0.f;
    int st_n = -1, st_nt = -1;
    const bool want_stats = p.stats_part != nullptr;
    // no atomics: this thread is the only writer of its columns in row 4 * CTA + (et >> 5) (reduce_rows.cuh)
    float* st_row = want_stats ? p.stats_part + ((size_t)blockIdx.x * 4 + (et >> 5)) * (size_t)(2 * p.stats_G) : nullptr;
    auto flush_stats = [&]() {
      if (!want_stats || st_n < 0) return;
#pragma unroll
      for (int s = 0; s < NSLAB; ++s) {
        const int c = st_nt * BN + s * 64 + 2 * (et & 31);
        float* ps = st_row + (size_t)st_n * p.cout_total + c;
        float* pq = ps + p.stats_G;
        ps[0] += st_acc[s][0]; ps[1] += st_acc[s][1];
        pq[0] += st_acc[s][2]; pq[1] += st_acc[s][3];
        st_acc[s][0] = st_acc[s][1] = st_acc[s][2] = st_acc[s][3] = 0.f;
      }
    };
    int acc = 0; uint32_t acc_phase = 0;
    int obuf = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int nt, n, d0, h0, w0;
      decode_tile(tile, nt, n, d0, h0, w0);
      if (n != st_n || nt != st_nt) { flush_stats(); st_n = n; st_nt = nt; }
      // rd >= td: a row past the end of a free-form box (tw*th*td < 128): the MMA read stale shared memory for it
      const bool valid = (rd < p.td) && (w0 + rw < p.out_w) && (h0 + rh < p.out_h) && (d0 + rd < p.out_d);
      // bias for this channel tile -> smem (visible after the first named barrier below)
      for (int i = et; i < BN; i += 128) bias_s[i] = p.bias ? p.bias[nt * BN + i] : 0.f;
      ptx::mbar_wait(tfull_bar(acc), acc_phase);
      ptx::tc_fence_after();
#pragma unroll
      for (int s = 0; s < NSLAB; ++s) {
        uint32_t v[64];
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + s * 64);
        ptx::tmem_ld_32x32(taddr, v);
        ptx::tmem_ld_32x32(taddr + 32, v + 32);
        ptx::tmem_ld_wait();
        if (s == NSLAB - 1) {   // accumulator fully drained: hand the TMEM stage back to the MMA warp
          ptx::tc_fence_before();
          ptx::mbar_arrive(tempty_bar(acc));
        }
        // staging buffer `obuf` was last read by the TMA store issued two slabs ago
        if (et == 0) ptx::tma_store_wait_read<1>();
        ptx::named_bar_sync(1, 128);
        const uint32_t so = smem_o + obuf * STAGE_OUT_BYTES + row * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          uint32_t pk[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int c = j * 8 + e * 2;
            float f0 = __uint_as_float(v[c]) + bias_s[s * 64 + c];
            float f1 = __uint_as_float(v[c + 1]) + bias_s[s * 64 + c + 1];
            f0 = act_apply(f0, p.act, p.slope); f1 = act_apply(f1, p.act, p.slope);
            if (!valid) { f0 = 0.f; f1 = 0.f; }
            __nv_bfloat162 h = __floats2bfloat162_rn(f0, f1);
            pk[e] = *reinterpret_cast<uint32_t*>(&h);
          }
          const uint32_t dst = so + (uint32_t)((j ^ (row & 7)) * 16);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]) : "memory");
        }
        ptx::fence_proxy_async_smem();
        ptx::named_bar_sync(2, 128);
        if (et == 0) {
          ptx::tma_store_5d(&p.o_map, smem_o + obuf * STAGE_OUT_BYTES, nt * BN + s * 64, w0, h0, d0, n);
          ptx::tma_store_commit();
        }
        if (want_stats) {
          // column sums over this thread's 32-row quarter, channel pair (et & 31); reads are conflict-free
          const int pr = et & 31, r0 = (et >> 5) * 32;
          const int chunk = pr >> 2, within = (pr & 3) * 4;
          float a0 = 0.f, a1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll 8
          for (int r = 0; r < 32; ++r) {
            const int rr = r0 + r;
            uint32_t wv;
            const uint32_t src = smem_o + obuf * STAGE_OUT_BYTES + rr * 128 + ((chunk ^ (rr & 7)) * 16) + within;
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(wv) : "r"(src));
            __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&wv);
            float2 f = __bfloat1622float2(h);
            a0 += f.x; a1 += f.y; q0 += f.x * f.x; q1 += f.y * f.y;
          }
          st_acc[s][0] += a0; st_acc[s][1] += a1; st_acc[s][2] += q0; st_acc[s][3] += q1;
        }
        obuf ^= 1;
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    flush_stats();
    if (et == 0) ptx::tma_store_wait_all<0>();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
  }
}


// -------------------------------------------------------------------------------------------------
// Weight gradient:  dW[co][ci][tap] += sum_vox X[vox (+) tap][ci] * dY[vox][co]
// GEMM view (both operands MN-major, i.e. the voxel/K index selects the 128 B smem row):
//   D[M = 2 slabs x 64 ci][N = Cout tile] += A[M][K = 128 voxels] * B[N][K]
//   A slab = TMA box {64 ci, tw, th, td, 1} of X shifted by one tap   (one slab per (tap, ci-chunk))
//   B      = TMA boxes {64 co, tw, th, td, 1} of dY (loaded once per voxel chunk, shared by all slabs)
// A CTA owns one (slab group, voxel split): all 512 TMEM columns hold 512/N accumulators of 128 x N,
// it streams its voxel range once and finally stores its partial sums into the slab of its voxel split
// (workspace [split][tap][ci][co]); halo_wgrad_finish_kernel adds the slabs in split order (wgrad_finish.cuh).
// -------------------------------------------------------------------------------------------------
struct alignas(64) WgradParams {
  CUtensorMap x_maps[8];
  CUtensorMap dy_map;
  int ntap[3];
  signed char tap_off[3][MAX_TAPS_DIM];
  signed char tap_par[3][MAX_TAPS_DIM];
  signed char tap_kidx[3][MAX_TAPS_DIM];
  int kdims[3];
  int cin, cout;
  int tiles_w, tiles_h, tiles_d, batch;
  int tw, th, td;
  int n_co_tiles, n_groups, n_splits, chunks_per_split, total_chunks;
  int n_slabs;        // ntaps * cin/64
  float* ws;          // [n_splits][taps][cin][cout] fp32 partial sums
};

template <int BN> struct WgradCfg {
  static constexpr int NACC = 512 / BN;                 // accumulators (slab pairs) per CTA
  static constexpr int B_BYTES = (BN / 64) * A_BYTES;   // dY slabs of one voxel chunk
  static constexpr int A_STAGE = 2 * A_BYTES;           // one slab pair
  static constexpr int A_STAGES = (BN >= 256) ? 3 : (BN >= 128 ? 4 : 5);     // BN 192: 96 + 128 KB
  static constexpr int SMEM_BYTES = 1024 + 2 * B_BYTES + A_STAGES * A_STAGE + 256;
};

template <int BN>
__global__ void __launch_bounds__(IGEMM_THREADS, 1) igemm_wgrad_kernel(const __grid_constant__ WgradParams p) {
  using Cfg = WgradCfg<BN>;
  constexpr int NACC = Cfg::NACC, AST = Cfg::A_STAGES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);
  const uint32_t smem_b = base;                              // 2 x B_BYTES
  const uint32_t smem_a = base + 2 * Cfg::B_BYTES;           // AST x 32 KB
  uint8_t* tail = base_ptr + 2 * Cfg::B_BYTES + AST * Cfg::A_STAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);        // afull[AST], aempty[AST], bfull[2], bempty[2], done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 8 * (2 * AST + 5));
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto afull = [&](int s) { return bar0 + 8u * s; };
  auto aempty = [&](int s) { return bar0 + 8u * (AST + s); };
  auto bfull = [&](int s) { return bar0 + 8u * (2 * AST + s); };
  auto bempty = [&](int s) { return bar0 + 8u * (2 * AST + 2 + s); };
  const uint32_t done_bar = bar0 + 8u * (2 * AST + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // unit decode: blockIdx.x = ((split * n_groups) + group) * n_co_tiles + co_tile
  int u = blockIdx.x;
  const int co_tile = u % p.n_co_tiles; u /= p.n_co_tiles;
  const int group = u % p.n_groups;
  const int split = u / p.n_groups;
  const int c_begin = split * p.chunks_per_split;
  const int c_end = min(c_begin + p.chunks_per_split, p.total_chunks);
  const int ci_chunks = p.cin / 64;
  const int first_slab = group * 2 * NACC;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < AST; ++s) { ptx::mbar_init(afull(s), 1); ptx::mbar_init(aempty(s), 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(bfull(s), 1); ptx::mbar_init(bempty(s), 1); }
    ptx::mbar_init(done_bar, 1);
    ptx::fence_mbar_init();
    ptx::prefetch_tmap(&p.dy_map);
    ptx::prefetch_tmap(&p.x_maps[0]);
  }
  if (warp == 1) ptx::tmem_alloc<512>(ptx::smem_u32(tmem_slot));
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ================= TMA producer =================
      int stage = 0; uint32_t phase = 0;
      int bb = 0; uint32_t bphase = 0;
      for (int chunk = c_begin; chunk < c_end; ++chunk) {
        int m = chunk;
        const int iw = m % p.tiles_w; m /= p.tiles_w;
        const int ih = m % p.tiles_h; m /= p.tiles_h;
        const int id = m % p.tiles_d; const int n = m / p.tiles_d;
        const int w0 = iw * p.tw, h0 = ih * p.th, d0 = id * p.td;
        ptx::mbar_wait(bempty(bb), bphase ^ 1u);
        ptx::mbar_arrive_expect_tx(bfull(bb), Cfg::B_BYTES);
#pragma unroll
        for (int s = 0; s < BN / 64; ++s)
          ptx::tma_load_5d(smem_b + bb * Cfg::B_BYTES + s * A_BYTES, &p.dy_map, bfull(bb), co_tile * BN + s * 64, w0, h0, d0, n);
        if (++bb == 2) { bb = 0; bphase ^= 1u; }
        for (int j = 0; j < NACC; ++j) {
          ptx::mbar_wait(aempty(stage), phase ^ 1u);
          ptx::mbar_arrive_expect_tx(afull(stage), Cfg::A_STAGE);
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            int slab = first_slab + 2 * j + half;
            if (slab >= p.n_slabs) slab = p.n_slabs - 1;      // duplicate; masked in the epilogue
            const int q = slab % ci_chunks;
            int t = slab / ci_chunks;
            const int c = t % p.ntap[2]; t /= p.ntap[2];
            const int b = t % p.ntap[1]; const int a = t / p.ntap[1];
            const int mi = p.tap_par[0][a] * 4 + p.tap_par[1][b] * 2 + p.tap_par[2][c];
            ptx::tma_load_5d(smem_a + stage * Cfg::A_STAGE + half * A_BYTES, &p.x_maps[mi], afull(stage), q * 64,
                             w0 + p.tap_off[2][c], h0 + p.tap_off[1][b], d0 + p.tap_off[0][a], n);
          }
          if (++stage == AST) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (ptx::elect_one_sync()) {      // elect.sync region: ptxas keeps the descriptors in the uniform datapath
      // ================= MMA issuer =================
      constexpr uint32_t idesc = ptx::idesc_bf16(TILE_M, BN, 1, 1);    // both operands MN-major
      int stage = 0; uint32_t phase = 0;
      int bb = 0; uint32_t bphase = 0;
      for (int chunk = c_begin; chunk < c_end; ++chunk) {
        ptx::mbar_wait(bfull(bb), bphase);
        const uint64_t bdesc = ptx::smem_desc_sw128(smem_b + bb * Cfg::B_BYTES, A_BYTES, 1024);
        for (int j = 0; j < NACC; ++j) {
          ptx::mbar_wait(afull(stage), phase);
          ptx::tc_fence_after();
          const uint64_t adesc = ptx::smem_desc_sw128(smem_a + stage * Cfg::A_STAGE, A_BYTES, 1024);
#pragma unroll
          for (int k = 0; k < TILE_M / UMMA_K; ++k) {
            // 16 voxels (K) = two 8-row groups = 2048 B further down the slab
            ptx::umma_bf16(tmem_base + (uint32_t)(j * BN), adesc + (uint64_t)(128 * k), bdesc + (uint64_t)(128 * k), idesc,
                           (chunk > c_begin || k > 0) ? 1u : 0u);
          }
          ptx::umma_commit(aempty(stage));
          if (++stage == AST) { stage = 0; phase ^= 1u; }
        }
        ptx::umma_commit(bempty(bb));
        if (++bb == 2) { bb = 0; bphase ^= 1u; }
      }
      ptx::umma_commit(done_bar);
    }
    __syncwarp();
  } else if (c_end > c_begin) {
    // ================= epilogue: TMEM -> vector stores into ws[split][tap][ci][co] =================
    const int quad = warp & 3;
    const int taps_total = p.kdims[0] * p.kdims[1] * p.kdims[2];
    ptx::mbar_wait(done_bar, 0);
    ptx::tc_fence_after();
    for (int j = 0; j < NACC; ++j) {
      const int slab = first_slab + 2 * j + (quad >> 1);
      const bool valid = slab < p.n_slabs;                    // warp-uniform
      int widx = 0, ci = 0;
      if (valid) {
        const int q = slab % ci_chunks;
        int t = slab / ci_chunks;
        const int c = t % p.ntap[2]; t /= p.ntap[2];
        const int b = t % p.ntap[1]; const int a = t / p.ntap[1];
        widx = (p.tap_kidx[0][a] * p.kdims[1] + p.tap_kidx[1][b]) * p.kdims[2] + p.tap_kidx[2][c];
        ci = q * 64 + (quad & 1) * 32 + lane;
      }
#pragma unroll 1
      for (int cb = 0; cb < BN; cb += 32) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(j * BN + cb), v);
        ptx::tmem_ld_wait();
        if (valid) {
          float* dst = p.ws + (((size_t)split * taps_total + widx) * p.cin + ci) * p.cout + co_tile * BN + cb;
#pragma unroll
          for (int e = 0; e < 32; e += 4)
            *reinterpret_cast<uint4*>(dst + e) = make_uint4(v[e], v[e + 1], v[e + 2], v[e + 3]);
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
// Host side: plans, tensor maps, launches
// =================================================================================================
namespace rg {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

static int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// 5-D bf16 NDHWC view {C, W, H, D, N}; strides in elements; `step` subsamples each spatial dim.
static int encode_act_map(CUtensorMap* m, const void* ptr, int channels, const int ext[3] /*d,h,w*/, int batch,
                          const int64_t strides[4] /*n,d,h,w*/, const int off[3], const int step[3], const int box[3] /*d,h,w*/) {
  EncodeTiledFn enc = get_encode();
  if (!enc) { set_error("cuTensorMapEncodeTiled unavailable"); return RG_E_CUDA; }
  int e[3];
  for (int i = 0; i < 3; ++i) {
    e[i] = (ext[i] - off[i] + step[i] - 1) / step[i];
    if (e[i] < 1) { set_error("tensor map: empty parity view"); return RG_E_UNSUPPORTED; }
  }
  const char* base = (const char*)ptr + 2 * (off[0] * strides[1] + off[1] * strides[2] + off[2] * strides[3]);
  cuuint64_t dims[5] = {(cuuint64_t)channels, (cuuint64_t)e[2], (cuuint64_t)e[1], (cuuint64_t)e[0], (cuuint64_t)batch};
  cuuint64_t gs[4] = {(cuuint64_t)(strides[3] * step[2] * 2), (cuuint64_t)(strides[2] * step[1] * 2),
                      (cuuint64_t)(strides[1] * step[0] * 2), (cuuint64_t)(strides[0] * 2)};
  cuuint32_t bx[5] = {64, (cuuint32_t)box[2], (cuuint32_t)box[1], (cuuint32_t)box[0], 1};
  cuuint32_t es[5] = {1, 1, 1, 1, 1};
  if (((uintptr_t)base & 15) != 0) { set_error("tensor map: base not 16-byte aligned"); return RG_E_BADARG; }
  for (int i = 0; i < 4; ++i)
    if (gs[i] % 16 != 0) { set_error("tensor map: stride %d not a multiple of 16 bytes", i); return RG_E_UNSUPPORTED; }
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, (void*)base, dims, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (activation) failed: %d", (int)r); return RG_E_CUDA; }
  return RG_OK;
}

static int encode_weight_map(CUtensorMap* m, const void* ptr, long long ktot, int rows, int bn) {
  EncodeTiledFn enc = get_encode();
  if (!enc) { set_error("cuTensorMapEncodeTiled unavailable"); return RG_E_CUDA; }
  cuuint64_t dims[2] = {(cuuint64_t)ktot, (cuuint64_t)rows};
  cuuint64_t gs[1] = {(cuuint64_t)(ktot * 2)};
  cuuint32_t bx[2] = {64, (cuuint32_t)bn};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)ptr, dims, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (weights) failed: %d", (int)r); return RG_E_CUDA; }
  return RG_OK;
}

// Pick the {d,h,w} output box of one 128-row MMA tile.
// exact (wgrad: the box is the GEMM-K dimension, every row must be data): power-of-two dims with product 128 that waste
// the fewest padded voxels; ties -> most cubic.
// free-form (fprop / dgrad: rows are independent outputs): ANY dims with product <= 128 -- the TMA box lands as
// consecutive 128-byte rows whatever its shape, rows past its end are masked in the epilogue -- chosen to minimise the
// waves of the persistent grid, then the tile count (odd extents: the 34^3 padded data gradient of a ResnetBlock
// takes 5 waves of (1,7,17) boxes instead of 6 waves of (4,4,8); the discriminator's 33^3 / 17^3 / 9^3 maps likewise).
static bool free_form_boxes() {
  static const bool on = [] { const char* e = getenv("RAYGUN_B200_FREE_BOX"); return !(e && e[0] == '0'); }();
  return on;
}

static void choose_box(const int ext[3], int box[3], bool exact, long long mult, int sms) {
  if (!exact) {
    long long best_w = -1, best_t = -1; int bscore = 0;
    for (int bd = 1; bd <= 128 && bd <= ext[0]; ++bd)
      for (int bh = 1; bd * bh <= 128 && bh <= ext[1]; ++bh) {
        int bw = 128 / (bd * bh);
        if (bw > ext[2]) bw = ext[2];
        if (bw > 256) bw = 256;
        const long long tiles = (long long)((ext[0] + bd - 1) / bd) * ((ext[1] + bh - 1) / bh) * ((ext[2] + bw - 1) / bw) * mult;
        const long long waves = (tiles + sms - 1) / sms;
        int mx = bd > bh ? bd : bh; mx = mx > bw ? mx : bw;
        const int score = -mx + (bw >= 8 ? 1000 : 0);
        if (best_w < 0 || waves < best_w || (waves == best_w && (tiles < best_t || (tiles == best_t && score > bscore)))) {
          best_w = waves; best_t = tiles; bscore = score; box[0] = bd; box[1] = bh; box[2] = bw;
        }
      }
    return;
  }
  long long best = -1; int bscore = 0;
  for (int bd = 1; bd <= 128; bd *= 2)
    for (int bh = 1; bd * bh <= 128; bh *= 2) {
      int bw = 128 / (bd * bh);
      if (bw > 256) continue;
      long long cover = 1;
      const int b[3] = {bd, bh, bw};
      for (int i = 0; i < 3; ++i) cover *= (long long)((ext[i] + b[i] - 1) / b[i]) * b[i];
      int mx = bd > bh ? bd : bh; mx = mx > bw ? mx : bw;
      int score = -mx + (bw >= 8 ? 1000 : 0);   // prefer >= 8 along w (full 1 KB swizzle atoms per h-row)
      if (best < 0 || cover < best || (cover == best && score > bscore)) {
        best = cover; bscore = score; box[0] = bd; box[1] = bh; box[2] = bw;
      }
    }
}

struct Launch {
  IgemmParams p;          // geometry filled at plan time; maps/pointers at run time
  int a_used[8];          // which parity maps are referenced
  int out_off[3], out_step[3], out_ext[3], box[3];
};

}  // namespace rg

struct rg_plan {
  rg_igemm_desc d;
  int bn;
  int kc, rows;           // reduction channels, output channels (of this mode)
  int taps;
  std::vector<rg::Launch> launches;
};

using namespace rg;

template <int BN> static int launch_wgrad(const WgradParams& p, int grid, cudaStream_t st) {
  using Cfg = WgradCfg<BN>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(igemm_wgrad_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(wgrad)");
    attr_set = true;
  }
  igemm_wgrad_kernel<BN><<<grid, IGEMM_THREADS, Cfg::SMEM_BYTES, st>>>(p);
  return RG_OK;
}

template <int BN> static int launch_bn(const IgemmParams& p, int grid, cudaStream_t st) {
  using Cfg = IgemmCfg<BN>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(igemm_conv_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(igemm)");
    attr_set = true;
  }
  igemm_conv_kernel<BN><<<grid, IGEMM_THREADS, Cfg::SMEM_BYTES, st>>>(p);
  return RG_OK;
}

static void wgrad_split_plan(const rg_plan* pl, int* n_groups, int* n_splits, int* cps, int* total_chunks) {
  const rg_igemm_desc& d = pl->d;
  const Launch& L = pl->launches[0];
  const int nacc = 512 / pl->bn;
  const int n_slabs = pl->taps * (d.cin / 64);
  *n_groups = (n_slabs + 2 * nacc - 1) / (2 * nacc);
  *total_chunks = L.p.tiles_w * L.p.tiles_h * L.p.tiles_d * d.n;
  const long long base_units = (long long)*n_groups * (d.cout / pl->bn);
  long long splits = (2LL * num_sms() + base_units - 1) / base_units;     // ~2 waves of CTAs
  const long long max_splits = (*total_chunks + 3) / 4;                    // >= 4 chunks per CTA
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  *cps = (int)((*total_chunks + splits - 1) / splits);
  *n_splits = (*total_chunks + *cps - 1) / *cps;
}

extern "C" {

int rg_plan_create(const rg_igemm_desc* d, rg_plan** out) {
  RG_CHECK_ARG(d && out, "rg_plan_create: null argument");
  RG_CHECK_ARG(d->mode >= 0 && d->mode <= 2, "rg_plan_create: bad mode");
  const int kc = d->mode == 1 ? d->cout : d->cin;      // reduction channels (wgrad: the slab channels)
  const int rows = d->mode == 1 ? d->cin : d->cout;    // produced channels  (wgrad: the N dimension)
  RG_UNSUPPORTED(kc % 64 != 0, "igemm: reduction channels (%d) must be a multiple of 64", kc);
  RG_UNSUPPORTED(rows % 64 != 0, "igemm: output channels (%d) must be a multiple of 64", rows);
  for (int i = 0; i < 3; ++i) {
    RG_UNSUPPORTED(d->stride[i] != 1 && d->stride[i] != 2, "igemm: stride must be 1 or 2");
    RG_UNSUPPORTED(d->kernel[i] < 1 || d->kernel[i] > MAX_TAPS_DIM, "igemm: kernel extent must be 1..8");
    RG_CHECK_ARG(d->pad[i] >= 0 && d->pad[i] < 64, "igemm: bad padding");
    long long lo = (long long)(d->out_dhw[i] - 1) * d->stride[i] + d->kernel[i] - 2 * d->pad[i];
    RG_CHECK_ARG(d->out_dhw[i] >= 1 && d->in_dhw[i] >= lo && d->in_dhw[i] < lo + d->stride[i],
                 "igemm: extents inconsistent with geometry (dim %d: in %d out %d)", i, d->in_dhw[i], d->out_dhw[i]);
  }
  RG_UNSUPPORTED(d->x_strides[3] % 8 != 0 || d->y_strides[3] % 8 != 0, "igemm: voxel stride must be a multiple of 8 elements");
  rg_plan* pl = new rg_plan();
  pl->d = *d; pl->kc = kc; pl->rows = rows;
  pl->taps = d->kernel[0] * d->kernel[1] * d->kernel[2];
  pl->bn = rows % 256 == 0 ? 256 : (rows % 192 == 0 ? 192 : (rows % 128 == 0 ? 128 : 64));   // 192: fmap_inc_factor 3 widths

  const int nclass = d->mode != 1 ? 1 : d->stride[0] * d->stride[1] * d->stride[2];
  for (int cls = 0; cls < nclass; ++cls) {
    Launch L;
    memset(&L, 0, sizeof(L));
    int q[3] = {0, 0, 0};
    if (d->mode == 1) { int c = cls; q[2] = c % d->stride[2]; c /= d->stride[2]; q[1] = c % d->stride[1]; q[0] = c / d->stride[1]; }
    bool empty = false;
    for (int i = 0; i < 3; ++i) {
      int nt = 0;
      const int s = d->stride[i], pd = d->pad[i];
      if (d->mode != 1) {
        for (int k = 0; k < d->kernel[i]; ++k) {
          int e = k - pd;
          int par = ((e % s) + s) % s;
          L.p.tap_off[i][nt] = (signed char)((e - par) / s);
          L.p.tap_par[i][nt] = (signed char)par;
          L.p.tap_kidx[i][nt] = (signed char)k;
          ++nt;
        }
        L.out_off[i] = 0; L.out_step[i] = 1; L.out_ext[i] = d->out_dhw[i];
      } else {
        for (int k = 0; k < d->kernel[i]; ++k) {
          int e = q[i] + pd - k;
          if (((e % s) + s) % s != 0) continue;
          L.p.tap_off[i][nt] = (signed char)(e / s);   // exact (e is a multiple of s, may be negative)
          L.p.tap_par[i][nt] = 0;
          L.p.tap_kidx[i][nt] = (signed char)k;
          ++nt;
        }
        L.out_off[i] = q[i]; L.out_step[i] = s;
        L.out_ext[i] = (d->in_dhw[i] - q[i] + s - 1) / s;
        if (L.out_ext[i] < 1) empty = true;
      }
      if (nt == 0) { delete pl; set_error("igemm: output class without contributing taps (kernel < stride)"); return RG_E_UNSUPPORTED; }
      L.p.ntap[i] = nt;
      L.p.kdims[i] = d->kernel[i];
    }
    if (empty) continue;
    for (int a = 0; a < L.p.ntap[0]; ++a)
      for (int b = 0; b < L.p.ntap[1]; ++b)
        for (int c = 0; c < L.p.ntap[2]; ++c) L.a_used[L.p.tap_par[0][a] * 4 + L.p.tap_par[1][b] * 2 + L.p.tap_par[2][c]] = 1;
    choose_box(L.out_ext, L.box, d->mode == 2 || !free_form_boxes(), (long long)d->n * (rows / pl->bn), num_sms());
    L.p.kc = kc;
    L.p.td = L.box[0]; L.p.th = L.box[1]; L.p.tw = L.box[2];
    L.p.a_box_bytes = L.box[0] * L.box[1] * L.box[2] * BLOCK_K * 2;
    L.p.tiles_d = (L.out_ext[0] + L.box[0] - 1) / L.box[0];
    L.p.tiles_h = (L.out_ext[1] + L.box[1] - 1) / L.box[1];
    L.p.tiles_w = (L.out_ext[2] + L.box[2] - 1) / L.box[2];
    L.p.batch = d->n;
    L.p.n_tiles_n = rows / pl->bn;
    L.p.out_d = L.out_ext[0]; L.p.out_h = L.out_ext[1]; L.p.out_w = L.out_ext[2];
    L.p.cout_total = rows;
    L.p.act = d->act; L.p.slope = d->slope;
    pl->launches.push_back(L);
  }
  *out = pl;
  return RG_OK;
}

int rg_plan_destroy(rg_plan* plan) {
  delete plan;
  return RG_OK;
}

size_t rg_plan_workspace_bytes(const rg_plan* plan) {
  if (!plan) return 0;
  if (plan->d.mode == 2) {
    int ng, ns, cps, tc;
    wgrad_split_plan(plan, &ng, &ns, &cps, &tc);
    return (size_t)ns * plan->d.cout * plan->d.cin * plan->taps * sizeof(float);
  }
  // statistics partial rows: 4 epilogue warps x one CTA per SM, 2 statistics x (n * produced channels)
  return (size_t)4 * num_sms() * 2 * plan->d.n * plan->rows * sizeof(float);
}

int rg_igemm_run(const rg_plan* pl, const void* in, const void* w_packed, const float* bias, void* outp,
                 float* stats_sum, float* stats_sqsum, void* workspace, void* stream) {
  RG_CHECK_ARG(pl && in && w_packed && outp, "rg_igemm_run: null argument");
  RG_CHECK_ARG(pl->d.mode != 2, "rg_igemm_run: plan was created for wgrad");
  RG_CHECK_ARG((stats_sum == nullptr) == (stats_sqsum == nullptr), "rg_igemm_run: stats pointers must come together");
  RG_CHECK_ARG(!stats_sum || workspace, "rg_igemm_run: statistics need the plan's workspace");
  const rg_igemm_desc& d = pl->d;
  cudaStream_t st = (cudaStream_t)stream;
  const long long stats_G = (long long)d.n * pl->rows;
  if (stats_sum) {
    cudaError_t e = cudaMemsetAsync(workspace, 0, rg_plan_workspace_bytes(pl), st);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(igemm statistics rows)");
  }
  // in-side / out-side tensors of this mode
  const int64_t* in_str = d.mode == 0 ? d.x_strides : d.y_strides;
  const int64_t* out_str = d.mode == 0 ? d.y_strides : d.x_strides;
  const int* in_ext = d.mode == 0 ? d.in_dhw : d.out_dhw;
  const int* out_ext_full = d.mode == 0 ? d.out_dhw : d.in_dhw;
  for (size_t li = 0; li < pl->launches.size(); ++li) {
    IgemmParams p = pl->launches[li].p;
    const Launch& L = pl->launches[li];
    int rc;
    for (int mi = 0; mi < 8; ++mi) {
      if (!L.a_used[mi]) continue;
      const int off[3] = {(mi >> 2) & 1, (mi >> 1) & 1, mi & 1};
      const int step[3] = {d.mode == 0 ? d.stride[0] : 1, d.mode == 0 ? d.stride[1] : 1, d.mode == 0 ? d.stride[2] : 1};
      rc = encode_act_map(&p.a_maps[mi], in, pl->kc, in_ext, d.n, in_str, off, step, L.box);
      if (rc) return rc;
    }
    rc = encode_act_map(&p.o_map, outp, pl->rows, out_ext_full, d.n, out_str, L.out_off, L.out_step, L.box);
    if (rc) return rc;
    rc = encode_weight_map(&p.b_map, w_packed, (long long)pl->taps * pl->kc, pl->rows, pl->bn);
    if (rc) return rc;
    p.bias = bias; p.stats_part = stats_sum ? (float*)workspace : nullptr; p.stats_G = stats_G;
    const long long total = (long long)p.tiles_w * p.tiles_h * p.tiles_d * p.batch * p.n_tiles_n;
    const int grid = (int)(total < num_sms() ? total : num_sms());
    if (pl->bn == 256) rc = launch_bn<256>(p, grid, st);
    else if (pl->bn == 192) rc = launch_bn<192>(p, grid, st);
    else if (pl->bn == 128) rc = launch_bn<128>(p, grid, st);
    else rc = launch_bn<64>(p, grid, st);
    if (rc) return rc;
    RG_LAUNCH_CHECK("rg_igemm_run");
  }
  if (stats_sum)
    return launch_rows_sum((const float*)workspace, 4 * num_sms(), 2 * stats_G, 2 * stats_G, stats_sum, stats_sqsum, stats_G, 0,
                           1.0, st);
  return RG_OK;
}

int rg_igemm_wgrad(const rg_plan* pl, const void* x, const void* dy, float* dw, float* dbias, void* workspace,
                   int accumulate, void* stream) {
  RG_CHECK_ARG(pl && x && dy && dw && workspace, "rg_igemm_wgrad: null argument");
  RG_CHECK_ARG(pl->d.mode == 2, "rg_igemm_wgrad: plan was not created for wgrad");
  RG_CHECK_ARG(dbias == nullptr, "rg_igemm_wgrad: dbias is produced by rg_norm_stats(kind=batch) on dy");
  const rg_igemm_desc& d = pl->d;
  cudaStream_t st = (cudaStream_t)stream;
  const Launch& L = pl->launches[0];
  WgradParams p;
  memset(&p, 0, sizeof(p));
  memcpy(p.ntap, L.p.ntap, sizeof(p.ntap));
  memcpy(p.tap_off, L.p.tap_off, sizeof(p.tap_off));
  memcpy(p.tap_par, L.p.tap_par, sizeof(p.tap_par));
  memcpy(p.tap_kidx, L.p.tap_kidx, sizeof(p.tap_kidx));
  memcpy(p.kdims, L.p.kdims, sizeof(p.kdims));
  p.cin = d.cin; p.cout = d.cout;
  p.tiles_w = L.p.tiles_w; p.tiles_h = L.p.tiles_h; p.tiles_d = L.p.tiles_d; p.batch = d.n;
  p.tw = L.p.tw; p.th = L.p.th; p.td = L.p.td;
  int rc;
  for (int mi = 0; mi < 8; ++mi) {
    if (!L.a_used[mi]) continue;
    const int off[3] = {(mi >> 2) & 1, (mi >> 1) & 1, mi & 1};
    rc = encode_act_map(&p.x_maps[mi], x, d.cin, d.in_dhw, d.n, d.x_strides, off, d.stride, L.box);
    if (rc) return rc;
  }
  const int zero[3] = {0, 0, 0}, one[3] = {1, 1, 1};
  rc = encode_act_map(&p.dy_map, dy, d.cout, d.out_dhw, d.n, d.y_strides, zero, one, L.box);
  if (rc) return rc;
  const int bn = pl->bn;
  p.n_co_tiles = d.cout / bn;
  p.n_slabs = pl->taps * (d.cin / 64);
  wgrad_split_plan(pl, &p.n_groups, &p.n_splits, &p.chunks_per_split, &p.total_chunks);
  const long long base_units = (long long)p.n_groups * p.n_co_tiles;
  p.ws = (float*)workspace;
  const int grid = (int)(base_units * p.n_splits);
  if (bn == 256) rc = launch_wgrad<256>(p, grid, st);
  else if (bn == 192) rc = launch_wgrad<192>(p, grid, st);
  else if (bn == 128) rc = launch_wgrad<128>(p, grid, st);
  else rc = launch_wgrad<64>(p, grid, st);
  if (rc) return rc;
  RG_LAUNCH_CHECK("rg_igemm_wgrad");
  HwSplits splits;
  for (int i = 0; i < 8; ++i) splits.of_kd[i] = p.n_splits;
  wgrad_finish_launch((const float*)workspace, splits, d.kernel[1] * d.kernel[2], pl->taps, d.cin, d.cout, dw, accumulate, st);
  RG_LAUNCH_CHECK("rg_igemm_wgrad(finish)");
  return RG_OK;
}

}  // extern "C"

// ---- weight packing ---------------------------------------------------------------------------------
namespace rg {
__global__ void pack_weights_kernel(const float* __restrict__ w, int cout, int cin, int taps, __nv_bfloat16* __restrict__ wf,
                                    __nv_bfloat16* __restrict__ wd) {
  long long total = (long long)cout * cin * taps;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int t = (int)(i % taps); long long r = i / taps;
    int ci = (int)(r % cin); int co = (int)(r / cin);
    __nv_bfloat16 v = __float2bfloat16_rn(w[i]);
    if (wf) wf[((long long)co * taps + t) * cin + ci] = v;
    if (wd) wd[((long long)ci * taps + t) * cout + co] = v;
  }
}
}  // namespace rg

extern "C" int rg_pack_weights(const float* w, int cout, int cin, int taps, void* w_fprop_bf16, void* w_dgrad_bf16,
                               void* stream) {
  RG_CHECK_ARG(w && cout > 0 && cin > 0 && taps > 0 && (w_fprop_bf16 || w_dgrad_bf16), "rg_pack_weights: bad argument");
  long long total = (long long)cout * cin * taps;
  rg::pack_weights_kernel<<<rg::grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
      w, cout, cin, taps, (__nv_bfloat16*)w_fprop_bf16, (__nv_bfloat16*)w_dgrad_bf16);
  RG_LAUNCH_CHECK("rg_pack_weights");
  return RG_OK;
}
