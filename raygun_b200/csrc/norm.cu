// InstanceNorm3d / BatchNorm3d: statistics, apply (fused activation + residual), backward.
// HBM-bound: stats 1 read; apply 1 read + 1 write; backward = 1 reduction pass (x, dy, y) + 1 pass.
#include "common.cuh"
#include "reduce_rows.cuh"
#include "vec_ndhwc.cuh"

namespace rg {

struct NormCtx {
  int kind;            // RG_NORM_INSTANCE | RG_NORM_BATCH
  int act; float slope;
  const float* mean; const float* rstd; const float* gamma;
};

__device__ __forceinline__ int group_of(int kind, long long n, long long c, long long C) {
  return kind == RG_NORM_INSTANCE ? (int)(n * C + c) : (int)c;
}

// Partial-row output of the reducing kernels (reduce_rows.cuh): block (bx, n) writes row bx (instance statistics: the
// samples' groups are disjoint columns of one row) or row n * gridDim.x + bx (batch statistics: every sample adds to the
// same C columns).  A row holds both statistics: columns [0, G) and [G, 2G).
struct RedRows { float* part; long long G; int per_n; };
__device__ __forceinline__ float* red_row(const RedRows& ro, long long n) {
  return ro.part + ((ro.per_n ? n * (long long)gridDim.x : 0) + blockIdx.x) * (2 * ro.G);
}

}  // namespace rg
#include "norm_dense.cuh"
namespace rg {

// MODE 0: (x, x^2).  MODE 1: (g, g*xhat) with g = dy * act'(y).
template <int MODE, typename TX, typename TG>
__device__ __forceinline__ void red_terms(const NormCtx& cx, int grp, const T5& xd, const TX* x, const T5& gd,
                                          const TG* dy, const T5& yd, const TG* y, long long n, long long c,
                                          long long d, long long h, long long w, float& a, float& b) {
  float xv = ldf(x + xd.off(n, c, d, h, w));
  if (MODE == 0) {
    a = xv; b = xv * xv;
  } else {
    float g = ldf(dy + gd.off(n, c, d, h, w));
    if (cx.act != RG_ACT_NONE) g *= act_grad_from_y(ldf(y + yd.off(n, c, d, h, w)), cx.act, cx.slope);
    float xhat = (xv - cx.mean[grp]) * cx.rstd[grp];
    a = g; b = g * xhat;
  }
}

// W-innermost storage: blockIdx.y = (n*C+c), blockIdx.x = chunk of the D*H*W extent.
template <int MODE, typename TX, typename TG>
__global__ void reduce2_w_inner(NormCtx cx, T5 xd, const TX* __restrict__ x, T5 gd, const TG* __restrict__ dy, T5 yd,
                                const TG* __restrict__ y, RedRows ro) {
  long long n = blockIdx.y / xd.c, c = blockIdx.y % xd.c;
  int grp = group_of(cx.kind, n, c, xd.c);
  long long vox = xd.d * xd.h * xd.w;
  float s0 = 0.f, s1 = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < vox;
       i += (long long)gridDim.x * blockDim.x) {
    long long w = i % xd.w, t = i / xd.w;
    long long h = t % xd.h, d = t / xd.h;
    float a, b;
    red_terms<MODE>(cx, grp, xd, x, gd, dy, yd, y, n, c, d, h, w, a, b);
    s0 += a; s1 += b;
  }
  __shared__ float sh0[32], sh1[32];
  s0 = warp_sum(s0); s1 = warp_sum(s1);
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { sh0[wid] = s0; sh1[wid] = s1; }
  __syncthreads();
  if (wid == 0) {
    int nw = blockDim.x >> 5;
    s0 = lane < nw ? sh0[lane] : 0.f; s1 = lane < nw ? sh1[lane] : 0.f;
    s0 = warp_sum(s0); s1 = warp_sum(s1);
    if (lane == 0) { float* row = red_row(ro, n); row[grp] = s0; row[ro.G + grp] = s1; }
  }
}

// C-innermost storage (NDHWC): blockIdx.y = n, blockIdx.x = chunk of voxels.  Lane = channel
// within a 32-channel tile, warps stride over voxels; cross-warp reduce in smem; one atomic per
// (block, channel).
template <int MODE, typename TX, typename TG>
__global__ void reduce2_c_inner(NormCtx cx, T5 xd, const TX* __restrict__ x, T5 gd, const TG* __restrict__ dy, T5 yd,
                                const TG* __restrict__ y, RedRows ro, long long vox_per_block) {
  long long n = blockIdx.y;
  long long vox = xd.d * xd.h * xd.w;
  long long v0 = blockIdx.x * vox_per_block;
  long long v1 = v0 + vox_per_block < vox ? v0 + vox_per_block : vox;
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __shared__ float sh0[8][33], sh1[8][33];
  for (long long c0 = 0; c0 < xd.c; c0 += 32) {
    long long c = c0 + lane;
    float s0 = 0.f, s1 = 0.f;
    if (c < xd.c) {
      int grp = group_of(cx.kind, n, c, xd.c);
      for (long long i = v0 + wid; i < v1; i += nw) {
        long long w = i % xd.w, t = i / xd.w;
        long long h = t % xd.h, d = t / xd.h;
        float a, b;
        red_terms<MODE>(cx, grp, xd, x, gd, dy, yd, y, n, c, d, h, w, a, b);
        s0 += a; s1 += b;
      }
    }
    sh0[wid][lane] = s0; sh1[wid][lane] = s1;
    __syncthreads();
    if (wid == 0 && c < xd.c) {
      for (int k = 1; k < nw; ++k) { s0 += sh0[k][lane]; s1 += sh1[k][lane]; }
      int grp = group_of(cx.kind, n, c, xd.c);
      float* row = red_row(ro, n);
      row[grp] = s0; row[ro.G + grp] = s1;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// vectorised NDHWC bf16 fast paths (8 channels / thread)
// ------------------------------------------------------------------------------------------
__global__ void vec_norm_apply_kernel(NormCtx cx, const float* __restrict__ beta, T5 xd, const __nv_bfloat16* __restrict__ x,
                                      T5 rd, const __nv_bfloat16* __restrict__ res, T5 yd, __nv_bfloat16* __restrict__ y) {
  const long long total = yd.numel() >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    VIdx p = vunravel(i, yd);
    float v[8];
    ld8(x + voff(xd, p.n, p.d, p.h, p.w, p.c), v);
    if (cx.mean) {
      const int g0 = group_of(cx.kind, p.n, p.c, xd.c);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = (v[k] - __ldg(cx.mean + g0 + k)) * __ldg(cx.rstd + g0 + k);
    }
    if (cx.gamma) {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] *= __ldg(cx.gamma + p.c + k);
    }
    if (beta) {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] += __ldg(beta + p.c + k);
    }
    if (res) {
      float r[8];
      ld8(res + voff(rd, p.n, p.d, p.h, p.w, p.c), r);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] += r[k];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = act_apply(v[k], cx.act, cx.slope);
    st8(y + voff(yd, p.n, p.d, p.h, p.w, p.c), v);
  }
}

__global__ void vec_norm_bwd_dx_kernel(NormCtx cx, int training, const float* __restrict__ red, float inv_count, T5 xd,
                                       const __nv_bfloat16* __restrict__ x, T5 gd, const __nv_bfloat16* __restrict__ dy,
                                       T5 yd, const __nv_bfloat16* __restrict__ y, T5 dxd, __nv_bfloat16* __restrict__ dx,
                                       int groups) {
  const long long total = dxd.numel() >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    VIdx p = vunravel(i, dxd);
    const int g0 = group_of(cx.kind, p.n, p.c, xd.c);
    float g[8], out[8];
    ld8(dy + voff(gd, p.n, p.d, p.h, p.w, p.c), g);
    if (cx.act != RG_ACT_NONE) {
      float yy[8];
      ld8(y + voff(yd, p.n, p.d, p.h, p.w, p.c), yy);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] *= act_grad_from_y(yy[k], cx.act, cx.slope);
    }
    if (training) {
      float xv[8];
      ld8(x + voff(xd, p.n, p.d, p.h, p.w, p.c), xv);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float r = __ldg(cx.rstd + g0 + k);
        const float gam = cx.gamma ? __ldg(cx.gamma + p.c + k) : 1.f;
        const float xhat = (xv[k] - __ldg(cx.mean + g0 + k)) * r;
        out[k] = gam * r * (g[k] - __ldg(red + g0 + k) * inv_count - xhat * __ldg(red + groups + g0 + k) * inv_count);
      }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        out[k] = (cx.gamma ? __ldg(cx.gamma + p.c + k) : 1.f) * __ldg(cx.rstd + g0 + k) * g[k];
    }
    st8(dx + voff(dxd, p.n, p.d, p.h, p.w, p.c), out);
  }
}

// blockIdx.y = n, blockIdx.x = voxel chunk.  Thread = (channel group of 8, voxel lane); the block walks
// its voxel range with 16-byte loads, reduces across voxel lanes in smem, one atomic per (block, channel).
template <int MODE>
__global__ void __launch_bounds__(256) vec_reduce2_kernel(NormCtx cx, T5 xd, const __nv_bfloat16* __restrict__ x, T5 gd,
                                                          const __nv_bfloat16* __restrict__ dy, T5 yd,
                                                          const __nv_bfloat16* __restrict__ y, RedRows ro,
                                                          long long vox_per_block) {
  const long long n = blockIdx.y;
  const long long vox = xd.d * xd.h * xd.w;
  const long long v0 = blockIdx.x * vox_per_block;
  const long long v1 = v0 + vox_per_block < vox ? v0 + vox_per_block : vox;
  const int cgroups = (int)(xd.c >> 3);
  __shared__ float red0[256], red1[256];
  // process channel groups in tiles of up to 32 groups (256 channels) so that 256 threads = 32 cg x 8 voxel lanes
  for (int cg0 = 0; cg0 < cgroups; cg0 += 32) {
    const int ncg = min(32, cgroups - cg0);
    const int lanes_v = 256 / ncg;                 // voxel lanes (>= 8)
    const int cgi = threadIdx.x % ncg, lv = threadIdx.x / ncg;
    float s0[8], s1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { s0[k] = 0.f; s1[k] = 0.f; }
    if (lv < lanes_v) {
      const int c = (cg0 + cgi) << 3;
      const int g0 = group_of(cx.kind, n, c, xd.c);
      for (long long i = v0 + lv; i < v1; i += lanes_v) {
        const long long w = i % xd.w, t = i / xd.w;
        const long long h = t % xd.h, d = t / xd.h;
        float xv[8];
        ld8(x + voff(xd, n, d, h, w, c), xv);
        if (MODE == 0) {
#pragma unroll
          for (int k = 0; k < 8; ++k) { s0[k] += xv[k]; s1[k] += xv[k] * xv[k]; }
        } else {
          float g[8];
          ld8(dy + voff(gd, n, d, h, w, c), g);
          if (cx.act != RG_ACT_NONE) {
            float yy[8];
            ld8(y + voff(yd, n, d, h, w, c), yy);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] *= act_grad_from_y(yy[k], cx.act, cx.slope);
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float xhat = (xv[k] - __ldg(cx.mean + g0 + k)) * __ldg(cx.rstd + g0 + k);
            s0[k] += g[k]; s1[k] += g[k] * xhat;
          }
        }
      }
    }
    // reduce over voxel lanes: lanes with the same cgi are threadIdx.x = cgi + ncg*lv
    // 8 rounds (one per channel of the group) through a 256-entry staging array
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      red0[threadIdx.x] = s0[k];
      red1[threadIdx.x] = s1[k];
      __syncthreads();
      if (threadIdx.x < ncg) {
        float a = 0.f, b = 0.f;
        for (int l = 0; l < lanes_v; ++l) { a += red0[threadIdx.x + ncg * l]; b += red1[threadIdx.x + ncg * l]; }
        const int c = ((cg0 + threadIdx.x) << 3) + k;
        const int grp = group_of(cx.kind, n, c, xd.c);
        float* row = red_row(ro, n);
        row[grp] = a;
        row[ro.G + grp] = b;
      }
      __syncthreads();
    }
  }
}

__global__ void finalize_kernel(const float* __restrict__ sum, const float* __restrict__ sq, int groups, double count,
                                float eps, float* __restrict__ mean, float* __restrict__ rstd, float* __restrict__ var_out) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= groups) return;
  double m = (double)sum[g] / count;
  double v = (double)sq[g] / count - m * m;
  if (v < 0.0) v = 0.0;
  mean[g] = (float)m;
  rstd[g] = (float)(1.0 / sqrt(v + (double)eps));
  if (var_out) var_out[g] = (float)v;
}

template <typename TX, typename TR, typename TY>
__global__ void norm_apply_kernel(NormCtx cx, const float* __restrict__ beta, T5 xd, const TX* __restrict__ x, T5 rd,
                                  const TR* __restrict__ res, T5 yd, TY* __restrict__ y, bool c_inner) {
  long long total = yd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, yd, c_inner);
    float v = ldf(x + xd.off(p.n, p.c, p.d, p.h, p.w));
    if (cx.mean) {
      int grp = group_of(cx.kind, p.n, p.c, xd.c);
      v = (v - cx.mean[grp]) * cx.rstd[grp];
    }
    if (cx.gamma) v = v * cx.gamma[p.c];
    if (beta) v += beta[p.c];
    if (res) v += ldf(res + rd.off(p.n, p.c, p.d, p.h, p.w));
    stf(y + yd.off(p.n, p.c, p.d, p.h, p.w), act_apply(v, cx.act, cx.slope));
  }
}

template <typename TX, typename TG, typename TD>
__global__ void norm_bwd_dx_kernel(NormCtx cx, int training, const float* __restrict__ red, float inv_count, T5 xd,
                                   const TX* __restrict__ x, T5 gd, const TG* __restrict__ dy, T5 yd,
                                   const TG* __restrict__ y, T5 dxd, TD* __restrict__ dx, int groups, bool c_inner) {
  long long total = dxd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, dxd, c_inner);
    int grp = group_of(cx.kind, p.n, p.c, xd.c);
    float g = ldf(dy + gd.off(p.n, p.c, p.d, p.h, p.w));
    if (cx.act != RG_ACT_NONE) g *= act_grad_from_y(ldf(y + yd.off(p.n, p.c, p.d, p.h, p.w)), cx.act, cx.slope);
    float gam = cx.gamma ? cx.gamma[p.c] : 1.f;
    float r = cx.rstd[grp];
    float out;
    if (training) {
      float xhat = (ldf(x + xd.off(p.n, p.c, p.d, p.h, p.w)) - cx.mean[grp]) * r;
      out = gam * r * (g - red[grp] * inv_count - xhat * red[groups + grp] * inv_count);
    } else {
      out = gam * r * g;
    }
    stf(dx + dxd.off(p.n, p.c, p.d, p.h, p.w), out);
  }
}

// dgamma[c] = sum_n S2[n,c]; dbeta[c] = sum_n S1[n,c]
__global__ void affine_grad_kernel(const float* __restrict__ red, int kind, int N, int C, int groups,
                                   float* __restrict__ dgamma, float* __restrict__ dbeta) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float a = 0.f, b = 0.f;
  if (kind == RG_NORM_INSTANCE) {
    for (int n = 0; n < N; ++n) { a += red[n * C + c]; b += red[groups + n * C + c]; }
  } else {
    a = red[c]; b = red[groups + c];
  }
  if (dbeta) dbeta[c] = a;
  if (dgamma) dgamma[c] = b;
}

template <int MODE, typename TX, typename TG>
static int launch_reduce2(const NormCtx& cx, const rg_tensor5* xd, const TX* x, const rg_tensor5* gd, const TG* dy,
                          const rg_tensor5* yd, const TG* y, float* out0, float* out1, float* ws, cudaStream_t st) {
  T5 xv = view(xd), gv = gd ? view(gd) : xv, yv = yd ? view(yd) : xv;
  long long vox = xv.d * xv.h * xv.w;
  const long long G = cx.kind == RG_NORM_INSTANCE ? xv.n * xv.c : xv.c;
  RedRows ro{ws, G, cx.kind == RG_NORM_INSTANCE ? 0 : 1};
  long long rows_x;
  if (dense_ok(xd, x) && dense_cg_ok(xd) && xv.n <= 65535 &&
      (MODE == 0 || (dense_ok(gd, dy) && (cx.act == RG_ACT_NONE || dense_ok(yd, y))))) {
    dim3 grid;
    DenseGeom gm = dense_geom(xd, &grid, 16);     // >= 16 voxels per voxel lane (48 left 1 block per SM on 32^3 maps)
    if (MODE) { gm.gn = gd->sn; if (yd) gm.yn = yd->sn; }
    dense_reduce2_kernel<MODE><<<grid, 256, 0, st>>>(cx, gm, (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy,
                                                     (const __nv_bfloat16*)y, ro);
    rows_x = grid.x;
  } else if (vec_ok(xd, x) && (MODE == 0 || (vec_ok(gd, dy) && (cx.act == RG_ACT_NONE || vec_ok(yd, y))))) {
    long long blocks_x = (148LL * 4 + xv.n - 1) / xv.n;
    long long vpb = (vox + blocks_x - 1) / blocks_x;
    if (vpb < 256) vpb = 256;
    blocks_x = (vox + vpb - 1) / vpb;
    dim3 grid((unsigned)blocks_x, (unsigned)xv.n);
    vec_reduce2_kernel<MODE><<<grid, 256, 0, st>>>(cx, xv, (const __nv_bfloat16*)x, gv, (const __nv_bfloat16*)dy, yv,
                                                   (const __nv_bfloat16*)y, ro, vpb);
    rows_x = blocks_x;
  } else if (channels_inner(xd)) {
    long long blocks_x = (148LL * 8 + xv.n - 1) / xv.n;
    long long vpb = (vox + blocks_x - 1) / blocks_x;
    if (vpb < 64) vpb = 64;
    blocks_x = (vox + vpb - 1) / vpb;
    dim3 grid((unsigned)blocks_x, (unsigned)xv.n);
    reduce2_c_inner<MODE, TX, TG><<<grid, 256, 0, st>>>(cx, xv, x, gv, dy, yv, y, ro, vpb);
    rows_x = blocks_x;
  } else {
    long long groups = xv.n * xv.c;
    long long bx = (148LL * 8 + groups - 1) / groups;
    long long maxbx = (vox + 1023) / 1024;
    if (bx > maxbx) bx = maxbx;
    if (bx < 1) bx = 1;
    RG_CHECK_ARG(groups <= 65535, "norm reduce: N*C=%lld exceeds grid.y limit", groups);
    dim3 grid((unsigned)bx, (unsigned)groups);
    reduce2_w_inner<MODE, TX, TG><<<grid, 256, 0, st>>>(cx, xv, x, gv, dy, yv, y, ro);
    rows_x = bx;
  }
  RG_LAUNCH_CHECK("norm reduce");
  const long long rows = ro.per_n ? rows_x * xv.n : rows_x;
  RG_CHECK_ARG(rows * 2 * G <= red_ws_floats(xv.n, xv.c), "norm reduce: workspace bound violated (rows %lld)", rows);
  return launch_rows_sum(ws, (int)rows, 2 * G, 2 * G, out0, out1, G, 0, 1.0, st);
}

// dx = dy * act'(y) AND sum over (n, voxels) of dx per channel in the same pass: the activation derivative and the
// bias gradient of a Conv3d + activation layer (ConvPass, UNet.py:61-69) need the same read of dy.  Dense NDHWC bf16.
__global__ void __launch_bounds__(256) dense_act_bwd_sum_kernel(int act, float slope, int C, int cg, int lanes_v, long long vox,
                                                                long long vox_per_block, const __nv_bfloat16* __restrict__ y,
                                                                const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx,
                                                                float* __restrict__ sums) {
  const int cgi = threadIdx.x % cg, lv = threadIdx.x / cg;
  const int c = cgi << 3;
  float s[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s[k] = 0.f;
  const long long v0 = blockIdx.x * vox_per_block, v1 = min(v0 + vox_per_block, vox);
#pragma unroll 2
  for (long long v = v0 + lv; v < v1; v += lanes_v) {
    const size_t o = (size_t)v * C + c;
    float g[8], yy[8];
    ld8(dy + o, g);
    ld8(y + o, yy);
    act_grad8(g, yy, act, slope);
    st8(dx + o, g);
#pragma unroll
    for (int k = 0; k < 8; ++k) s[k] += g[k];
  }
  __shared__ float red[256];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    red[threadIdx.x] = s[k];
    __syncthreads();
    if (threadIdx.x < cg) {
      float a = 0.f;
      for (int l = 0; l < lanes_v; ++l) a += red[threadIdx.x + cg * l];
      sums[(size_t)blockIdx.x * C + (threadIdx.x << 3) + k] = a;
    }
    __syncthreads();
  }
}

// Row-strided twin of dense_act_bwd_sum_kernel: y and dy may be interiors of padded buffers (a pad-arena output, a
// gradient folded in place by rg_pad3d_fold) -- channels contiguous, w stride = C, arbitrary h / d / n strides.
// A block owns a contiguous range of (n, d, h) rows; one index decode per row, constant strides inside it.
__global__ void __launch_bounds__(256) row_act_bwd_sum_kernel(int act, float slope, int cg, int lanes_v, T5 yd, T5 gd, T5 xd,
                                                              long long rows, long long rows_per_block,
                                                              const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ dy,
                                                              __nv_bfloat16* __restrict__ dx, float* __restrict__ sums) {
  const int cgi = threadIdx.x % cg, lv = threadIdx.x / cg;
  const int c = cgi << 3, W = (int)yd.w, H = (int)yd.h, D = (int)yd.d;
  float s[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s[k] = 0.f;
  const long long r0 = blockIdx.x * rows_per_block, r1 = min(r0 + rows_per_block, rows);
  for (long long r = r0; r < r1; ++r) {
    const int h = (int)(r % H);
    const long long t = r / H;
    const int d = (int)(t % D);
    const long long n = t / D;
    const __nv_bfloat16* yr = y + n * yd.sn + d * yd.sd + h * yd.sh + c;
    const __nv_bfloat16* gr = dy + n * gd.sn + d * gd.sd + h * gd.sh + c;
    __nv_bfloat16* xr = dx + n * xd.sn + d * xd.sd + h * xd.sh + c;
#pragma unroll 2
    for (int w = lv; w < W; w += lanes_v) {
      float g[8], yy[8];
      ld8(gr + (long long)w * gd.sw, g);
      ld8(yr + (long long)w * yd.sw, yy);
      act_grad8(g, yy, act, slope);
      st8(xr + (long long)w * xd.sw, g);
#pragma unroll
      for (int k = 0; k < 8; ++k) s[k] += g[k];
    }
  }
  __shared__ float red[256];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    red[threadIdx.x] = s[k];
    __syncthreads();
    if (threadIdx.x < cg) {
      float a = 0.f;
      for (int l = 0; l < lanes_v; ++l) a += red[threadIdx.x + cg * l];
      sums[(size_t)blockIdx.x * (cg << 3) + (threadIdx.x << 3) + k] = a;
    }
    __syncthreads();
  }
}

}  // namespace rg

using namespace rg;

extern "C" {

long long rg_reduce_ws_floats(long long n, long long c) { return red_ws_floats(n, c); }

int rg_rows_sum(const float* part, int rows, long long groups, float* out, int accumulate, void* stream) {
  RG_CHECK_ARG(part && out && rows >= 1 && groups >= 0, "rg_rows_sum: bad argument");
  return launch_rows_sum(part, rows, groups, groups, out, nullptr, groups, accumulate, 1.0, (cudaStream_t)stream);
}

int rg_norm_stats(const rg_tensor5* xd, const void* x, int kind, float* sum, float* sqsum, float* ws, void* stream) {
  RG_CHECK_ARG(xd && x && sum && sqsum && ws, "rg_norm_stats: null argument");
  RG_CHECK_ARG(kind == RG_NORM_INSTANCE || kind == RG_NORM_BATCH, "rg_norm_stats: bad kind");
  if (view(xd).numel() == 0) return RG_OK;
  NormCtx cx{kind, RG_ACT_NONE, 0.f, nullptr, nullptr, nullptr};
  cudaStream_t st = (cudaStream_t)stream;
  int rc = RG_OK;
  RG_DISPATCH_FLOAT(xd->dtype, TX, {
    rc = launch_reduce2<0, TX, TX>(cx, xd, (const TX*)x, nullptr, nullptr, nullptr, nullptr, sum, sqsum, ws, st);
  })
  if (rc) return rc;
  RG_LAUNCH_CHECK("rg_norm_stats");
  return RG_OK;
}

// ---- rg_norm_stats_static: the same statistics with NO caller workspace and no allocation ----------------------------------
// For callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): the partial
// rows of the two-stage reduction live in a file-scope device array instead of a caller-provided `ws`.  Every GPU that loads
// the library holds RG_NORM_STATS_STATIC_MAX_FLOATS * 4 bytes (4 MiB) until exit.  In HOST code `norm_static_scratch` is only
// a stand-in for the symbol, not a device address: its address is resolved once per device, with that device current, by
// cudaGetSymbolAddress, and that pointer is what the kernels receive.  All calls on a device share the array: calls that
// overlap on the GPU (two streams) race.  Nothing in it has to start at zero -- every partial row the second stage reads
// is written by the first stage of the same call -- so stale contents of an earlier call are harmless.
static __device__ __align__(256) float norm_static_scratch[RG_NORM_STATS_STATIC_MAX_FLOATS];

static int norm_static_scratch_ptr(float** out) {
  constexpr int MAX_DEV = 64;
  static float* table[MAX_DEV] = {nullptr};      // one resolved address per device; written once (idempotent if raced)
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "rg_norm_stats_static: cudaGetDevice");
  if (dev < 0 || dev >= MAX_DEV) { set_error("rg_norm_stats_static: device index %d out of range", dev); return RG_E_BADARG; }
  if (!table[dev]) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, norm_static_scratch);      // of the CURRENT device's copy of the symbol
    if (e != cudaSuccess) return cuda_fail(e, "rg_norm_stats_static: cudaGetSymbolAddress");
    if (!p) { set_error("rg_norm_stats_static: cudaGetSymbolAddress returned a null address"); return RG_E_CUDA; }
    table[dev] = p;
  }
  *out = table[dev];
  return RG_OK;
}

int rg_norm_stats_static(const rg_tensor5* xd, const void* x, int kind, float* sum, float* sqsum, void* stream) {
  RG_CHECK_ARG(xd && x && sum && sqsum, "rg_norm_stats_static: null argument");
  RG_CHECK_ARG(kind == RG_NORM_INSTANCE || kind == RG_NORM_BATCH, "rg_norm_stats_static: bad kind");
  const long long need = red_ws_floats(xd->n, xd->c);
  RG_CHECK_ARG(xd->n >= 0 && xd->c >= 0 && need <= (long long)RG_NORM_STATS_STATIC_MAX_FLOATS,
               "rg_norm_stats_static: N = %lld, C = %lld needs %lld floats of scratch, the built-in array holds %lld "
               "(use rg_norm_stats with a caller workspace)", (long long)xd->n, (long long)xd->c, need,
               (long long)RG_NORM_STATS_STATIC_MAX_FLOATS);
  if (view(xd).numel() == 0) return RG_OK;
  float* ws = nullptr;
  int rc = norm_static_scratch_ptr(&ws);
  if (rc) return rc;
  NormCtx cx{kind, RG_ACT_NONE, 0.f, nullptr, nullptr, nullptr};
  cudaStream_t st = (cudaStream_t)stream;
  RG_DISPATCH_FLOAT(xd->dtype, TX, {
    rc = launch_reduce2<0, TX, TX>(cx, xd, (const TX*)x, nullptr, nullptr, nullptr, nullptr, sum, sqsum, ws, st);
  })
  if (rc) return rc;
  RG_LAUNCH_CHECK("rg_norm_stats_static");
  return RG_OK;
}

int rg_norm_finalize(const float* sum, const float* sqsum, int groups, double count, float eps, float* mean,
                     float* rstd, float* var_out, void* stream) {
  RG_CHECK_ARG(sum && sqsum && mean && rstd && groups > 0 && count > 0, "rg_norm_finalize: bad argument");
  finalize_kernel<<<(groups + 127) / 128, 128, 0, (cudaStream_t)stream>>>(sum, sqsum, groups, count, eps, mean, rstd, var_out);
  RG_LAUNCH_CHECK("rg_norm_finalize");
  return RG_OK;
}

int rg_norm_apply(const rg_tensor5* xd, const void* x, int kind, const float* mean, const float* rstd,
                  const float* gamma, const float* beta, const rg_tensor5* rd, const void* residual, int act,
                  float slope, const rg_tensor5* yd, void* y, void* stream) {
  RG_CHECK_ARG(xd && x && yd && y, "rg_norm_apply: null argument");
  RG_CHECK_ARG(same_shape(xd, yd), "rg_norm_apply: shape mismatch");
  RG_CHECK_ARG((mean == nullptr) == (rstd == nullptr), "rg_norm_apply: mean/rstd must come together");
  if (residual) RG_CHECK_ARG(rd && same_shape(xd, rd), "rg_norm_apply: residual shape mismatch");
  long long total = view(yd).numel();
  if (total == 0) return RG_OK;
  NormCtx cx{kind, act, slope, mean, rstd, gamma};
  cudaStream_t st = (cudaStream_t)stream;
  int grid = grid_for(total, 256);
  bool ci = channels_inner(yd);
  int rdt = residual ? rd->dtype : xd->dtype;
  T5 rv = residual ? view(rd) : view(xd);
  if (dense_ok(xd, x) && dense_ok(yd, y) && (!residual || dense_ok(rd, residual)) && dense_cg_ok(xd) && xd->n <= 65535) {
    dim3 dgrid;
    DenseGeom gm = dense_geom(xd, &dgrid);
    gm.on = yd->sn;
    if (residual) gm.gn = rd->sn;
    dense_norm_apply_kernel<<<dgrid, 256, 0, st>>>(cx, beta, gm, (const __nv_bfloat16*)x, (const __nv_bfloat16*)residual,
                                                   (__nv_bfloat16*)y);
    RG_LAUNCH_CHECK("rg_norm_apply(dense)");
    return RG_OK;
  }
  if (vec_ok(xd, x) && vec_ok(yd, y) && (!residual || vec_ok(rd, residual))) {
    vec_norm_apply_kernel<<<grid_for(total >> 3, 256), 256, 0, st>>>(cx, beta, view(xd), (const __nv_bfloat16*)x, rv,
                                                                     (const __nv_bfloat16*)residual, view(yd), (__nv_bfloat16*)y);
    RG_LAUNCH_CHECK("rg_norm_apply(vec)");
    return RG_OK;
  }
  RG_DISPATCH_FLOAT(xd->dtype, TX, RG_DISPATCH_FLOAT(rdt, TR, RG_DISPATCH_FLOAT(yd->dtype, TY, {
    norm_apply_kernel<TX, TR, TY><<<grid, 256, 0, st>>>(cx, beta, view(xd), (const TX*)x, rv, (const TR*)residual, view(yd), (TY*)y, ci);
  })))
  RG_LAUNCH_CHECK("rg_norm_apply");
  return RG_OK;
}

int rg_norm_bwd(const rg_tensor5* xd, const void* x, const rg_tensor5* dyd, const void* dy, const rg_tensor5* yd,
                const void* y, int kind, int training, const float* mean, const float* rstd, const float* gamma,
                int act, float slope, float* red, float* dgamma, float* dbeta, const rg_tensor5* dxd, void* dx,
                float* ws, void* stream) {
  RG_CHECK_ARG(xd && x && dyd && dy && mean && rstd && red && ws, "rg_norm_bwd: null argument");
  RG_CHECK_ARG(same_shape(xd, dyd), "rg_norm_bwd: shape mismatch");
  if (act != RG_ACT_NONE) RG_CHECK_ARG(yd && y && same_shape(xd, yd) && yd->dtype == dyd->dtype, "rg_norm_bwd: y required for fused activation");
  if (dx) RG_CHECK_ARG(dxd && same_shape(xd, dxd), "rg_norm_bwd: dx shape mismatch");
  T5 xv = view(xd);
  if (xv.numel() == 0) return RG_OK;
  int groups = kind == RG_NORM_INSTANCE ? (int)(xv.n * xv.c) : (int)xv.c;
  double count = kind == RG_NORM_INSTANCE ? (double)(xv.d * xv.h * xv.w) : (double)(xv.n * xv.d * xv.h * xv.w);
  NormCtx cx{kind, act, slope, mean, rstd, gamma};
  cudaStream_t st = (cudaStream_t)stream;
  int rc = RG_OK;
  bool need_red = training || dgamma || dbeta;
  if (need_red) {
    RG_DISPATCH_FLOAT(xd->dtype, TX, RG_DISPATCH_FLOAT(dyd->dtype, TG, {
      rc = launch_reduce2<1, TX, TG>(cx, xd, (const TX*)x, dyd, (const TG*)dy, yd, (const TG*)y, red, red + groups, ws, st);
    }))
    if (rc) return rc;
    RG_LAUNCH_CHECK("rg_norm_bwd reduce");
    if (dgamma || dbeta) {
      affine_grad_kernel<<<(int)((xv.c + 127) / 128), 128, 0, st>>>(red, kind, (int)xv.n, (int)xv.c, groups, dgamma, dbeta);
      RG_LAUNCH_CHECK("rg_norm_bwd affine");
    }
  }
  if (dx) {
    long long total = xv.numel();
    int grid = grid_for(total, 256);
    bool ci = channels_inner(dxd);
    T5 yv = (yd && y) ? view(yd) : xv;
    if (dense_ok(xd, x) && dense_ok(dyd, dy) && dense_ok(dxd, dx) && (act == RG_ACT_NONE || dense_ok(yd, y)) &&
        dense_cg_ok(xd) && xd->n <= 65535) {
      dim3 dgrid;
      DenseGeom gm = dense_geom(xd, &dgrid);
      gm.gn = dyd->sn; gm.on = dxd->sn;
      if (act != RG_ACT_NONE) gm.yn = yd->sn;
      dense_norm_bwd_dx_kernel<<<dgrid, 256, 0, st>>>(cx, training, red, (float)(1.0 / count), groups, gm,
                                                      (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy,
                                                      (const __nv_bfloat16*)y, (__nv_bfloat16*)dx);
      RG_LAUNCH_CHECK("rg_norm_bwd dx(dense)");
      return RG_OK;
    }
    if (vec_ok(xd, x) && vec_ok(dyd, dy) && vec_ok(dxd, dx) && (act == RG_ACT_NONE || vec_ok(yd, y))) {
      vec_norm_bwd_dx_kernel<<<grid_for(total >> 3, 256), 256, 0, st>>>(
          cx, training, red, (float)(1.0 / count), xv, (const __nv_bfloat16*)x, view(dyd), (const __nv_bfloat16*)dy, yv,
          (const __nv_bfloat16*)y, view(dxd), (__nv_bfloat16*)dx, groups);
      RG_LAUNCH_CHECK("rg_norm_bwd dx(vec)");
      return RG_OK;
    }
    RG_DISPATCH_FLOAT(xd->dtype, TX, RG_DISPATCH_FLOAT(dyd->dtype, TG, RG_DISPATCH_FLOAT(dxd->dtype, TD, {
      norm_bwd_dx_kernel<TX, TG, TD><<<grid, 256, 0, st>>>(cx, training, red, (float)(1.0 / count), xv, (const TX*)x, view(dyd), (const TG*)dy, yv, (const TG*)y, view(dxd), (TD*)dx, groups, ci);
    })))
    RG_LAUNCH_CHECK("rg_norm_bwd dx");
  }
  return RG_OK;
}

int rg_act_bwd_bias(const rg_tensor5* yd, const void* y, const rg_tensor5* dyd, const void* dy, int act, float slope,
                    const rg_tensor5* dxd, void* dx, float* bias_sum, float* ws, void* stream) {
  RG_CHECK_ARG(yd && dyd && dxd && y && dy && dx && bias_sum && ws, "rg_act_bwd_bias: null argument");
  RG_CHECK_ARG(same_shape(yd, dyd) && same_shape(yd, dxd), "rg_act_bwd_bias: shape mismatch");
  const long long per = yd->d * yd->h * yd->w * yd->c;
  const long long vox = yd->n * yd->d * yd->h * yd->w;
  if (vox == 0) return RG_OK;
  const int cg = (int)(yd->c >> 3), lanes_v = cg ? 256 / cg : 0;
  if (!dense_ok(yd, y) || !dense_ok(dyd, dy) || !dense_ok(dxd, dx) || yd->sn != per || dyd->sn != per || dxd->sn != per) {
    // row-strided operands (pad-arena interiors, gradients folded in place): channels contiguous, w stride = C
    RG_UNSUPPORTED(!vec_ok(yd, y) || !vec_ok(dyd, dy) || !vec_ok(dxd, dx) || !dense_cg_ok(yd),
                   "rg_act_bwd_bias: needs bf16 NDHWC tensors with C %% 8 == 0 and 256 %% (C/8) == 0");
    const long long rows = yd->n * yd->d * yd->h;
    long long rpb = (rows + 148LL * 8 - 1) / (148LL * 8);
    const long long min_rpb = (12LL * lanes_v + yd->w - 1) / yd->w;     // >= 12 voxels per voxel lane
    if (rpb < min_rpb) rpb = min_rpb;
    const long long blocks = (rows + rpb - 1) / rpb;
    row_act_bwd_sum_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(act, slope, cg, lanes_v, view(yd), view(dyd), view(dxd),
                                                                              rows, rpb, (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy,
                                                                              (__nv_bfloat16*)dx, ws);
    RG_LAUNCH_CHECK("rg_act_bwd_bias(row)");
    return launch_rows_sum(ws, (int)blocks, yd->c, yd->c, bias_sum, nullptr, yd->c, 0, 1.0, (cudaStream_t)stream);
  }
  RG_UNSUPPORTED(!dense_cg_ok(yd), "rg_act_bwd_bias: needs C %% 8 == 0 and 256 %% (C/8) == 0");
  long long blocks = 148LL * 8;
  long long vpb = (vox + blocks - 1) / blocks;
  const long long min_vpb = 16LL * lanes_v;                 // a fixed smem reduction tail per block: >= 16 voxels per lane
  if (vpb < min_vpb) vpb = min_vpb;
  vpb = (vpb + lanes_v - 1) / lanes_v * lanes_v;
  blocks = (vox + vpb - 1) / vpb;
  dense_act_bwd_sum_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(act, slope, (int)yd->c, cg, lanes_v, vox, vpb,
                                                                              (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy,
                                                                              (__nv_bfloat16*)dx, ws);
  RG_LAUNCH_CHECK("rg_act_bwd_bias");
  return launch_rows_sum(ws, (int)blocks, yd->c, yd->c, bias_sum, nullptr, yd->c, 0, 1.0, (cudaStream_t)stream);
}

}  // extern "C"
