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

// out[g] (+)= scale * sum_{r < rows} part[r * ld + g], r ascending.  Columns g < split go to out0[g], the others to
// out1[g - split] (two statistics kept side by side in one row).  Block = 32 columns x 8 row lanes.
static __global__ void __launch_bounds__(256) rows_sum_kernel(const float* __restrict__ part, int rows, long long groups,
                                                              long long ld, float* __restrict__ out0,
                                                              float* __restrict__ out1, long long split, int accumulate,
                                                              double scale) {
  __shared__ double sh[8][33];
  const int gx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const long long g = blockIdx.x * 32LL + gx;
  // four independent partial sums per row lane (loads in flight instead of one dependent chain), combined in a fixed order
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  if (g < groups) {
    const float* p = part + g;
    int r = ry;
    for (; r + 24 < rows; r += 32) {
      a0 += (double)p[(long long)r * ld];
      a1 += (double)p[(long long)(r + 8) * ld];
      a2 += (double)p[(long long)(r + 16) * ld];
      a3 += (double)p[(long long)(r + 24) * ld];
    }
    for (; r < rows; r += 8) a0 += (double)p[(long long)r * ld];
  }
  const double a = (a0 + a1) + (a2 + a3);
  sh[ry][gx] = a;
  __syncthreads();
  if (ry == 0 && g < groups) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sh[k][gx];
    float* dst = g < split ? out0 + g : out1 + (g - split);
    const float v = (float)(t * scale);
    *dst = accumulate ? *dst + v : v;
  }
}

static inline int launch_rows_sum(const float* part, int rows, long long groups, long long ld, float* out0, float* out1,
                                  long long split, int accumulate, double scale, cudaStream_t st) {
  if (groups <= 0) return RG_OK;
  rows_sum_kernel<<<(unsigned)((groups + 31) / 32), 256, 0, st>>>(part, rows, groups, ld, out0, out1 ? out1 : out0, split,
                                                                 accumulate, scale);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "rows_sum");
  return RG_OK;
}

}  // namespace rg
