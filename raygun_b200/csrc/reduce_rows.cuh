// Deterministic two-stage reductions.
// No kernel of this library finishes a floating-point reduction with atomics: the order in which atomics land varies
// from run to run, and a 1-ulp difference in one InstanceNorm statistic or weight gradient grows into percent-level
// loss differences within a few hundred optimiser steps (the reference's CPU path and cuDNN's deterministic
// algorithms do not have that property).  Instead every reducing block writes its partial result to its OWN row of
// a caller-provided workspace, and `rows_sum_kernel` adds the rows in ascending order (double accumulation).
// The row a block writes, and hence the result, is a pure function of the tensor shape.
#pragma once
#include "common.cuh"

namespace rg {

constexpr long long RED_MAX_BLOCKS = 148LL * 8;     // upper bound of the grids of the elementwise reducers

// fp32 elements of workspace the normalisation / bias reducers need for an (N, C, ...) tensor (both statistics)
inline long long red_ws_floats(long long n, long long c) { return 2 * c * (RED_MAX_BLOCKS + 2 * n + 2) + 4 * n * c + 64; }

// out[g] (+)= scale * sum_{r < rows} part[r * ld + g].  Columns g < split go to out0[g], the others to out1[g - split]
// (two statistics kept side by side in one row).  Block = 32 columns x 32 row lanes: row lane l adds rows l, l+32, ... in
// ascending order (two independent chains so loads overlap), the lanes are combined in ascending order by one thread
// per column -- a fixed order, whatever the grid.  (8 row lanes left this kernel latency-bound: ~20 us per launch for the
// 592 statistics rows of a convolution epilogue, 400+ launches per training step.)
constexpr int ROWS_SUM_LANES = 32;
static __global__ void __launch_bounds__(32 * ROWS_SUM_LANES) rows_sum_kernel(const float* __restrict__ part, int rows,
                                                                               long long groups, long long ld,
                                                                               float* __restrict__ out0, float* __restrict__ out1,
                                                                               long long split, int accumulate, double scale) {
  __shared__ double sh[ROWS_SUM_LANES][33];
  const int gx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const long long g = blockIdx.x * 32LL + gx;
  double a0 = 0.0, a1 = 0.0;
  if (g < groups) {
    const float* p = part + g;
    int r = ry;
    for (; r + ROWS_SUM_LANES < rows; r += 2 * ROWS_SUM_LANES) {
      a0 += (double)p[(long long)r * ld];
      a1 += (double)p[(long long)(r + ROWS_SUM_LANES) * ld];
    }
    if (r < rows) a0 += (double)p[(long long)r * ld];
  }
  sh[ry][gx] = a0 + a1;
  __syncthreads();
  if (ry == 0 && g < groups) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < ROWS_SUM_LANES; ++k) t += sh[k][gx];
    float* dst = g < split ? out0 + g : out1 + (g - split);
    const float v = (float)(t * scale);
    *dst = accumulate ? *dst + v : v;
  }
}

static inline int launch_rows_sum(const float* part, int rows, long long groups, long long ld, float* out0, float* out1,
                                  long long split, int accumulate, double scale, cudaStream_t st) {
  if (groups <= 0) return RG_OK;
  rows_sum_kernel<<<(unsigned)((groups + 31) / 32), 32 * ROWS_SUM_LANES, 0, st>>>(part, rows, groups, ld, out0, out1 ? out1 : out0, split,
                                                                 accumulate, scale);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "rows_sum");
  return RG_OK;
}

}  // namespace rg
