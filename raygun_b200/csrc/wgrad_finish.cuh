// Second stage of the tensor-core weight-gradient kernels (halo_wgrad_sm100.cu, igemm_sm100.cu): the CTAs of voxel split s
// STORE their partial sums into slab s of an fp32 workspace [split][tap][ci][co] (coalesced along co, no atomics); this
// kernel adds the slabs in ascending split order -- so dW is bit-reproducible -- and transposes into the checkpoint
// layout dw[co][ci][kd][kh][kw] (overwrite, or accumulate into an existing gradient bucket).
#pragma once
#include "common.cuh"

namespace rg {

// sum over splits (ascending) of ws[split][tap][ci][co] -> dw[co][ci][tap]  (= or +=)
struct HwSplits { int of_kd[8]; };
static __global__ void halo_wgrad_finish_kernel(const float* __restrict__ ws, HwSplits splits, int taps_hw, int taps, int cin, int cout,
                                         float* __restrict__ dw, int accumulate) {
  __shared__ float tile[32][33];
  // grid: (ceil(cout/32), cin, ceil(taps/32)); block (32, 8)
  const int co0 = blockIdx.x * 32, ci = blockIdx.y, t0 = blockIdx.z * 32;
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int t = t0 + r, co = co0 + threadIdx.x;
    float a = 0.f;
    if (t < taps && co < cout) {
      const float* src = ws + ((size_t)t * cin + ci) * cout + co;
      const size_t slab = (size_t)taps * cin * cout;
      const int n_splits = splits.of_kd[t / taps_hw];
      float a1 = 0.f, a2 = 0.f, a3 = 0.f;       // four loads in flight per thread; fixed combination order
      int sidx = 0;
      for (; sidx + 3 < n_splits; sidx += 4) {
        a += src[sidx * slab]; a1 += src[(sidx + 1) * slab]; a2 += src[(sidx + 2) * slab]; a3 += src[(sidx + 3) * slab];
      }
      for (; sidx < n_splits; ++sidx) a += src[sidx * slab];
      a = (a + a1) + (a2 + a3);
    }
    tile[r][threadIdx.x] = a;
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

}  // namespace rg
