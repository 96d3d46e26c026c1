// Second stage of the tensor-core weight-gradient kernels (halo_wgrad_sm100.cu, igemm_sm100.cu): the CTAs of voxel split s
// STORE their partial sums into slab s of an fp32 workspace [split][tap][ci][co] (coalesced along co, no atomics); this
// kernel adds the slabs in ascending split order -- so dW is bit-reproducible -- and transposes into the checkpoint
// layout dw[co][ci][kd][kh][kw] (overwrite, or accumulate into an existing gradient bucket).
#pragma once
#include "common.cuh"

namespace rg {

// sum over splits of ws[split][tap][ci][co] -> dw[co][ci][tap]  (= or +=).
// grid (ceil(cout/32), cin, taps), block (32 co, 8 split lanes): split lane l adds slabs l, l+8, ... in ascending order (two
// chains so loads overlap), the 8 lanes are combined in ascending order -- a fixed order whatever the launch.  Reads are
// 128-byte rows along co; the 32 results of a block are scattered into the [co][ci][tap] layout (4 bytes each: negligible
// next to the n_splits slabs read).  (One 32 x 32 (tap, co) tile per block left a 64 -> 64 3^3 layer with 128 blocks each
// walking 197 slabs serially: 1.85 TB/s.)
struct HwSplits { int of_kd[8]; };
static __global__ void __launch_bounds__(256) halo_wgrad_finish_kernel(const float* __restrict__ ws, HwSplits splits, int taps_hw,
                                                                       int taps, int cin, int cout, float* __restrict__ dw,
                                                                       int accumulate) {
  __shared__ float part[8][33];
  const int co = blockIdx.x * 32 + threadIdx.x, ci = blockIdx.y, t = blockIdx.z, l = threadIdx.y;
  float a0 = 0.f, a1 = 0.f;
  if (co < cout) {
    const float* src = ws + ((size_t)t * cin + ci) * cout + co;
    const size_t slab = (size_t)taps * cin * cout;
    const int n_splits = splits.of_kd[t / taps_hw];
    int sidx = l;
    for (; sidx + 8 < n_splits; sidx += 16) { a0 += src[sidx * slab]; a1 += src[(sidx + 8) * slab]; }
    if (sidx < n_splits) a0 += src[sidx * slab];
  }
  part[l][threadIdx.x] = a0 + a1;
  __syncthreads();
  if (l == 0 && co < cout) {
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) v += part[k][threadIdx.x];
    float* o = dw + ((size_t)co * cin + ci) * taps + t;
    *o = accumulate ? *o + v : v;
  }
}

// Few splits: one 32 x 32 (tap, co) tile per block, every thread adds its element's slabs in ascending order (two chains) and
// the tile is transposed through shared memory so the writes along `tap` are contiguous.  grid (ceil(cout/32), cin, ceil(taps/32)),
// block (32, 8).  Each of the two kernels has its own fixed summation order; which one runs is a function of the layer shape only
// (`wgrad_finish_launch`), so results stay reproducible.
static __global__ void __launch_bounds__(256) halo_wgrad_finish_tile_kernel(const float* __restrict__ ws, HwSplits splits, int taps_hw,
                                                                            int taps, int cin, int cout, float* __restrict__ dw,
                                                                            int accumulate) {
  __shared__ float tile[32][33];
  const int co0 = blockIdx.x * 32, ci = blockIdx.y, t0 = blockIdx.z * 32;
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int t = t0 + r, co = co0 + threadIdx.x;
    float a0 = 0.f, a1 = 0.f;
    if (t < taps && co < cout) {
      const float* src = ws + ((size_t)t * cin + ci) * cout + co;
      const size_t slab = (size_t)taps * cin * cout;
      const int n_splits = splits.of_kd[t / taps_hw];
      int sidx = 0;
      for (; sidx + 1 < n_splits; sidx += 2) { a0 += src[sidx * slab]; a1 += src[(sidx + 1) * slab]; }
      if (sidx < n_splits) a0 += src[sidx * slab];
    }
    tile[r][threadIdx.x] = a0 + a1;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int co = co0 + r, t = t0 + threadIdx.x;
    if (co < cout && t < taps) {
      float* o = dw + ((size_t)co * cin + ci) * taps + t;
      *o = accumulate ? *o + tile[threadIdx.x][r] : tile[threadIdx.x][r];
    }
  }
}

// Many slabs over a small weight tensor (64 -> 64 3^3 at 8 x 64^3: 197 slabs of 442 KB) want split-parallel blocks; few slabs
// over a large one (256 -> 256 3^3: 16 slabs of 7 MB) want one thread per element.
static inline void wgrad_finish_launch(const float* ws, const HwSplits& splits, int taps_hw, int taps, int cin, int cout, float* dw,
                                       int accumulate, cudaStream_t st) {
  int max_splits = 0;
  for (int i = 0; i < 8; ++i) max_splits = splits.of_kd[i] > max_splits ? splits.of_kd[i] : max_splits;
  if (max_splits >= 96) {
    dim3 fg((cout + 31) / 32, cin, taps);
    halo_wgrad_finish_kernel<<<fg, dim3(32, 8), 0, st>>>(ws, splits, taps_hw, taps, cin, cout, dw, accumulate);
  } else {
    dim3 fg((cout + 31) / 32, cin, (taps + 31) / 32);
    halo_wgrad_finish_tile_kernel<<<fg, dim3(32, 8), 0, st>>>(ws, splits, taps_hw, taps, cin, cout, dw, accumulate);
  }
}

}  // namespace rg
