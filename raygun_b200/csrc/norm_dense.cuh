// Dense NDHWC bf16 fast paths of the normalisation kernels (the layout every tensor-core convolution writes).
// The generic vec_* kernels unravel a 64-bit linear index per 16-byte access (four 64-bit div/mod ~ 300 instructions)
// and re-read the per-channel statistics for every element: they ran at 20-30 % of HBM bandwidth (profiles/
// r01_step_kernel_table_v4.txt).  When every operand is dense (sw = C, sh = W*C, sd = H*W*C) the offset of voxel v is
// v*C, so a thread keeps ONE channel group of 8 for its whole life, hoists the folded per-channel coefficients into
// registers and walks voxels with a constant stride: ~30 instructions per 16 bytes, no division.
// Thread layout: blockIdx.y = sample n, blockIdx.x = voxel chunk; tid = lv * cg + cgi  (cg = C/8 channel groups,
// 256 % cg == 0), so a warp covers 32/cg... consecutive voxels x all channel groups = contiguous 16-byte accesses.
#pragma once
#include "common.cuh"
#include "vec_ndhwc.cuh"

namespace rg {

inline bool dense_ok(const rg_tensor5* t, const void* p) {
  return vec_ok(t, p) && t->sw == t->c && t->sh == t->w * t->sw && t->sd == t->h * t->sh;
}
inline bool dense_cg_ok(const rg_tensor5* t) {
  const long long cg = t->c >> 3;
  return cg >= 1 && cg <= 256 && 256 % cg == 0 && t->d * t->h * t->w < (1LL << 31) / t->c;
}

__device__ __forceinline__ void act8(float* v, int act, float slope) {
  if (act == RG_ACT_RELU) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
  } else if (act == RG_ACT_LEAKY) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = v[k] > 0.f ? v[k] : v[k] * slope;
  } else if (act != RG_ACT_NONE) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = act_apply(v[k], act, slope);
  }
}
__device__ __forceinline__ void act_grad8(float* g, const float* y, int act, float slope) {
  if (act == RG_ACT_RELU) {
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = y[k] > 0.f ? g[k] : 0.f;
  } else if (act == RG_ACT_LEAKY) {
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = y[k] > 0.f ? g[k] : g[k] * slope;
  } else if (act != RG_ACT_NONE) {
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] *= act_grad_from_y(y[k], act, slope);
  }
}

struct DenseGeom { int C, cg, lanes_v; unsigned vox, vox_per_block; long long xn, gn, yn, on; };   // per-sample strides

// y = act((x - mean) * rstd * gamma + beta + residual)
__global__ void __launch_bounds__(256) dense_norm_apply_kernel(NormCtx cx, const float* __restrict__ beta, DenseGeom gm,
                                                               const __nv_bfloat16* __restrict__ x,
                                                               const __nv_bfloat16* __restrict__ res,
                                                               __nv_bfloat16* __restrict__ y) {
  const int cgi = threadIdx.x % gm.cg, lv = threadIdx.x / gm.cg;
  const int c = cgi << 3, n = blockIdx.y;
  float sc[8], sh[8];
  const int g0 = group_of(cx.kind, n, c, gm.C);
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float r = cx.mean ? cx.rstd[g0 + k] : 1.f, m = cx.mean ? cx.mean[g0 + k] : 0.f;
    const float ga = cx.gamma ? cx.gamma[c + k] : 1.f, be = beta ? beta[c + k] : 0.f;
    sc[k] = r * ga; sh[k] = be - m * r * ga;
  }
  const unsigned v0 = blockIdx.x * gm.vox_per_block, v1 = min(v0 + gm.vox_per_block, gm.vox);
  const __nv_bfloat16* xp = x + (size_t)n * gm.xn + c;
  const __nv_bfloat16* rp = res ? res + (size_t)n * gm.gn + c : nullptr;
  __nv_bfloat16* yp = y + (size_t)n * gm.on + c;
#pragma unroll 2
  for (unsigned v = v0 + lv; v < v1; v += gm.lanes_v) {
    const size_t o = (size_t)v * gm.C;
    float f[8];
    ld8(xp + o, f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], sh[k]);
    if (rp) {
      float r[8];
      ld8(rp + o, r);
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] += r[k];
    }
    act8(f, cx.act, cx.slope);
    st8(yp + o, f);
  }
}

// dx = gamma*rstd * (g - mean(g) - xhat * mean(g*xhat)),  g = dy * act'(y);  eval mode: dx = gamma*rstd*g.
// Folded per channel: dx = c1*g + c2*x + c3.
__global__ void __launch_bounds__(256) dense_norm_bwd_dx_kernel(NormCtx cx, int training, const float* __restrict__ red,
                                                                float inv_count, int groups, DenseGeom gm,
                                                                const __nv_bfloat16* __restrict__ x,
                                                                const __nv_bfloat16* __restrict__ dy,
                                                                const __nv_bfloat16* __restrict__ y,
                                                                __nv_bfloat16* __restrict__ dx) {
  const int cgi = threadIdx.x % gm.cg, lv = threadIdx.x / gm.cg;
  const int c = cgi << 3, n = blockIdx.y;
  const int g0 = group_of(cx.kind, n, c, gm.C);
  float c1[8], c2[8], c3[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float r = cx.rstd[g0 + k], m = cx.mean[g0 + k];
    const float ga = cx.gamma ? cx.gamma[c + k] : 1.f;
    c1[k] = ga * r;
    if (training) {
      const float A = red[g0 + k] * inv_count, B = red[groups + g0 + k] * inv_count;
      c2[k] = -c1[k] * B * r;
      c3[k] = -c1[k] * A - c2[k] * m;
    } else {
      c2[k] = 0.f; c3[k] = 0.f;
    }
  }
  const bool has_act = cx.act != RG_ACT_NONE;
  const unsigned v0 = blockIdx.x * gm.vox_per_block, v1 = min(v0 + gm.vox_per_block, gm.vox);
  const __nv_bfloat16* xp = x + (size_t)n * gm.xn + c;
  const __nv_bfloat16* gp = dy + (size_t)n * gm.gn + c;
  const __nv_bfloat16* yp = has_act ? y + (size_t)n * gm.yn + c : nullptr;
  __nv_bfloat16* op = dx + (size_t)n * gm.on + c;
#pragma unroll 2
  for (unsigned v = v0 + lv; v < v1; v += gm.lanes_v) {
    const size_t o = (size_t)v * gm.C;
    float g[8], out[8];
    ld8(gp + o, g);
    if (has_act) {
      float yy[8];
      ld8(yp + o, yy);
      act_grad8(g, yy, cx.act, cx.slope);
    }
    if (training) {
      float xv[8];
      ld8(xp + o, xv);
#pragma unroll
      for (int k = 0; k < 8; ++k) out[k] = fmaf(c1[k], g[k], fmaf(c2[k], xv[k], c3[k]));
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) out[k] = c1[k] * g[k];
    }
    st8(op + o, out);
  }
}

// MODE 0: (sum x, sum x^2).  MODE 1: (sum g, sum g*xhat), g = dy * act'(y).  One partial-row entry per (block, channel).
template <int MODE>
__global__ void __launch_bounds__(256) dense_reduce2_kernel(NormCtx cx, DenseGeom gm, const __nv_bfloat16* __restrict__ x,
                                                            const __nv_bfloat16* __restrict__ dy,
                                                            const __nv_bfloat16* __restrict__ y, RedRows ro) {
  const int cgi = threadIdx.x % gm.cg, lv = threadIdx.x / gm.cg;
  const int c = cgi << 3, n = blockIdx.y;
  const int g0 = group_of(cx.kind, n, c, gm.C);
  float mr[8], rs[8];            // xhat = x*rs - mr
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    rs[k] = MODE ? cx.rstd[g0 + k] : 1.f;
    mr[k] = MODE ? cx.mean[g0 + k] * rs[k] : 0.f;
  }
  float s0[8], s1[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) { s0[k] = 0.f; s1[k] = 0.f; }
  const bool has_act = MODE && cx.act != RG_ACT_NONE;
  const unsigned v0 = blockIdx.x * gm.vox_per_block, v1 = min(v0 + gm.vox_per_block, gm.vox);
  const __nv_bfloat16* xp = x + (size_t)n * gm.xn + c;
  const __nv_bfloat16* gp = MODE ? dy + (size_t)n * gm.gn + c : nullptr;
  const __nv_bfloat16* yp = has_act ? y + (size_t)n * gm.yn + c : nullptr;
#pragma unroll 2
  for (unsigned v = v0 + lv; v < v1; v += gm.lanes_v) {
    const size_t o = (size_t)v * gm.C;
    float xv[8];
    ld8(xp + o, xv);
    if (MODE == 0) {
#pragma unroll
      for (int k = 0; k < 8; ++k) { s0[k] += xv[k]; s1[k] = fmaf(xv[k], xv[k], s1[k]); }
    } else {
      float g[8];
      ld8(gp + o, g);
      if (has_act) {
        float yy[8];
        ld8(yp + o, yy);
        act_grad8(g, yy, cx.act, cx.slope);
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) { s0[k] += g[k]; s1[k] = fmaf(g[k], fmaf(xv[k], rs[k], -mr[k]), s1[k]); }
    }
  }
  // reduce over the voxel lanes of the block (threads with the same cgi): smem transpose, 8 channels at a time
  __shared__ float red0[256], red1[256];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    red0[threadIdx.x] = s0[k];
    red1[threadIdx.x] = s1[k];
    __syncthreads();
    if (threadIdx.x < gm.cg) {
      float a = 0.f, b = 0.f;
      for (int l = 0; l < gm.lanes_v; ++l) { a += red0[threadIdx.x + gm.cg * l]; b += red1[threadIdx.x + gm.cg * l]; }
      const int grp = group_of(cx.kind, n, (threadIdx.x << 3) + k, gm.C);
      float* row = red_row(ro, n);
      row[grp] = a;
      row[ro.G + grp] = b;
    }
    __syncthreads();
  }
}

// grid: about 8 blocks per SM in total, at least `min_iters` voxels per voxel lane (reductions pay a fixed smem +
// atomics tail per block, so they want fat blocks; streaming kernels only need enough loads in flight)
inline DenseGeom dense_geom(const rg_tensor5* t, dim3* grid, int min_iters = 4) {
  DenseGeom g;
  g.C = (int)t->c; g.cg = (int)(t->c >> 3); g.lanes_v = 256 / g.cg;
  g.vox = (unsigned)(t->d * t->h * t->w);
  long long blocks_x = (148LL * 8 + t->n - 1) / t->n;
  long long vpb = (g.vox + blocks_x - 1) / blocks_x;
  const long long min_vpb = (long long)min_iters * g.lanes_v;
  if (vpb < min_vpb) vpb = min_vpb;
  vpb = (vpb + g.lanes_v - 1) / g.lanes_v * g.lanes_v;
  g.vox_per_block = (unsigned)vpb;
  blocks_x = (g.vox + vpb - 1) / vpb;
  *grid = dim3((unsigned)blocks_x, (unsigned)t->n);
  g.xn = g.gn = g.yn = g.on = t->sn;
  return g;
}

}  // namespace rg
