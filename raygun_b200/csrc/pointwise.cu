// 1x1x1 convolutions onto a handful of output channels (the generator's final Conv3d(ngf, output_nc, 1)
// of systems/BaseSystem.py:209-224, with its Tanh fused): pure HBM-bound streaming over a bf16 NDHWC
// tensor, 8 channels (16 bytes) per thread.  fprop = per-voxel dot products reduced with warp shuffles;
// dgrad = rank-Cout outer product; wgrad = reduction over all voxels (block partials + fp32 atomics).
#include "common.cuh"
#include "reduce_rows.cuh"
#include "vec_ndhwc.cuh"

namespace rg {

constexpr int PW_MAX_COUT = 4;

struct PwW { const float* w; const float* bias; int cin, cout; };

// One block per row (n, d, h); threads j*cg .. j*cg+cg-1 (cg = Cin/8, a power of two <= 32) share one voxel of the
// row.  A thread keeps its channel group for the whole kernel, so its weights live in registers and the only index
// arithmetic per 16-byte load is a shift (rows may be strided views: pad arenas, crops).
template <typename TY, int COUT>
__global__ void __launch_bounds__(256) pw_fprop_kernel(PwW pw, T5 xd, const __nv_bfloat16* __restrict__ x, T5 yd,
                                                       TY* __restrict__ y, int act, float slope) {
  const int cg = pw.cin >> 3;
  const int lg = 31 - __clz(cg);
  const int cgi = threadIdx.x & (cg - 1), c = cgi << 3;
  float wr[COUT][8];
#pragma unroll
  for (int co = 0; co < COUT; ++co)
#pragma unroll
    for (int k = 0; k < 8; ++k) wr[co][k] = pw.w[co * pw.cin + c + k];
  float br[COUT];
#pragma unroll
  for (int co = 0; co < COUT; ++co) br[co] = pw.bias ? pw.bias[co] : 0.f;
  const int W = (int)xd.w, H = (int)xd.h, D = (int)xd.d;
  const long long rows = xd.n * D * H;
  const int items = W << lg, items_up = (items + 31) & ~31;        // warp-uniform trip count for the shuffles
  const int step = (int)blockDim.x;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int h = (int)(row % H);
    const long long t = row / H;
    const int d = (int)(t % D), n = (int)(t / D);
    const __nv_bfloat16* xrow = x + n * xd.sn + d * xd.sd + h * xd.sh + c;
    TY* yrow = y + n * yd.sn + d * yd.sd + h * yd.sh;
    // two independent 16-byte loads in flight per thread and trip
    for (int j0 = threadIdx.x; j0 < items_up; j0 += 2 * step) {
      const int j1 = j0 + step;
      const bool live0 = j0 < items, live1 = j1 < items;
      const bool run1 = j1 < items_up;                              // warp-uniform
      const int w0 = j0 >> lg, w1 = j1 >> lg;
      float v0[8], v1[8];
      if (live0) ld8(xrow + (long long)w0 * xd.sw, v0);
      if (live1) ld8(xrow + (long long)w1 * xd.sw, v1);
      float a0[COUT], a1[COUT];
#pragma unroll
      for (int co = 0; co < COUT; ++co) { a0[co] = 0.f; a1[co] = 0.f; }
      if (live0) {
#pragma unroll
        for (int co = 0; co < COUT; ++co)
#pragma unroll
          for (int k = 0; k < 8; ++k) a0[co] = fmaf(v0[k], wr[co][k], a0[co]);
      }
      if (live1) {
#pragma unroll
        for (int co = 0; co < COUT; ++co)
#pragma unroll
          for (int k = 0; k < 8; ++k) a1[co] = fmaf(v1[k], wr[co][k], a1[co]);
      }
      for (int o = cg >> 1; o > 0; o >>= 1) {
#pragma unroll
        for (int co = 0; co < COUT; ++co) a0[co] += __shfl_xor_sync(0xffffffffu, a0[co], o);
        if (run1) {
#pragma unroll
          for (int co = 0; co < COUT; ++co) a1[co] += __shfl_xor_sync(0xffffffffu, a1[co], o);
        }
      }
      if (cgi == 0) {
        if (live0) {
#pragma unroll
          for (int co = 0; co < COUT; ++co) stf(yrow + co * yd.sc + (long long)w0 * yd.sw, act_apply(a0[co] + br[co], act, slope));
        }
        if (live1) {
#pragma unroll
          for (int co = 0; co < COUT; ++co) stf(yrow + co * yd.sc + (long long)w1 * yd.sw, act_apply(a1[co] + br[co], act, slope));
        }
      }
    }
  }
}

template <typename TG>
__global__ void pw_dgrad_kernel(PwW pw, T5 gd, const TG* __restrict__ dy, T5 dxd, __nv_bfloat16* __restrict__ dx) {
  const long long total = dxd.numel() >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    VIdx p = vunravel(i, dxd);
    float out[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int co = 0; co < pw.cout; ++co) {
      const float g = ldf(dy + gd.off(p.n, co, p.d, p.h, p.w));
#pragma unroll
      for (int k = 0; k < 8; ++k) out[k] = fmaf(g, __ldg(pw.w + co * pw.cin + p.c + k), out[k]);
    }
    st8(dx + voff(dxd, p.n, p.d, p.h, p.w, p.c), out);
  }
}

// part[block][co][c] = sum over the block's voxels of dy[vox][co] * x[vox][c];  part[block][cout*cin + co] = sum dy[vox][co]
// (one partial row per block, added in block order by rows_sum_kernel: deterministic)
template <typename TG>
__global__ void __launch_bounds__(256) pw_wgrad_kernel(int cin, int cout, T5 xd, const __nv_bfloat16* __restrict__ x, T5 gd,
                                                       const TG* __restrict__ dy, float* __restrict__ part,
                                                       long long vox_per_block) {
  const int cg = cin >> 3;                       // <= 32
  const int lanes_v = 256 / cg;
  const int cgi = threadIdx.x % cg, lv = threadIdx.x / cg;
  const long long vox_total = xd.n * xd.d * xd.h * xd.w;
  const long long v0 = blockIdx.x * vox_per_block;
  const long long v1 = v0 + vox_per_block < vox_total ? v0 + vox_per_block : vox_total;
  float acc[PW_MAX_COUT][8];
  float bsum[PW_MAX_COUT];
#pragma unroll
  for (int co = 0; co < PW_MAX_COUT; ++co) {
    bsum[co] = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[co][k] = 0.f;
  }
  if (lv < lanes_v) {
    for (long long i = v0 + lv; i < v1; i += lanes_v) {
      long long t = i;
      const long long w = t % xd.w; t /= xd.w;
      const long long h = t % xd.h; t /= xd.h;
      const long long d = t % xd.d; const long long n = t / xd.d;
      float v[8];
      ld8(x + voff(xd, n, d, h, w, cgi << 3), v);
#pragma unroll
      for (int co = 0; co < PW_MAX_COUT; ++co) {
        if (co < cout) {
          const float g = ldf(dy + gd.off(n, co, d, h, w));
          bsum[co] += g;
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[co][k] = fmaf(g, v[k], acc[co][k]);
        }
      }
    }
  }
  __shared__ float red[256];
  for (int co = 0; co < cout; ++co) {
#pragma unroll
    for (int k = 0; k <= 8; ++k) {          // k == 8: the bias partial
      red[threadIdx.x] = (lv < lanes_v) ? (k < 8 ? acc[co][k] : bsum[co]) : 0.f;
      __syncthreads();
      if (threadIdx.x < cg) {
        float a = 0.f;
        for (int l = 0; l < lanes_v; ++l) a += red[threadIdx.x + cg * l];
        float* row = part + (size_t)blockIdx.x * (cout * (cin + 1));
        if (k < 8) row[co * cin + (threadIdx.x << 3) + k] = a;
        else if (threadIdx.x == 0) row[cout * cin + co] = a;
      }
      __syncthreads();
    }
  }
}

static bool pw_shape_ok(int cin, int cout) {
  const int cg = cin >> 3;
  return cin % 8 == 0 && cg >= 1 && cg <= 32 && (cg & (cg - 1)) == 0 && cout >= 1 && cout <= PW_MAX_COUT;
}


// Data gradient of a convolution with <= 4 input channels (the discriminator's first layer, 1 -> ngf, k2 s2 p1,
// NLayerDiscriminator.py:137-145, reached by the generator loss): dx[n,ci,i] = sum_{tap, co} dy[n,co,(i+p-k)/s] w[co,ci,k].
// One thread per input voxel; dy rows (channels innermost, bf16) are read as 16-byte vectors; the weights sit in
// shared memory as [tap][ci][co].  HBM-bound: Cout*2 bytes read per contributing tap.
struct SmallDgradP { int cin, cout, kd, kh, kw, sd, sh, sw, pd, ph, pw; };

template <typename TD>
__global__ void __launch_bounds__(256) smallcin_dgrad_kernel(SmallDgradP g, T5 yd, const __nv_bfloat16* __restrict__ dy,
                                                             const float* __restrict__ w, T5 xd, TD* __restrict__ dx) {
  extern __shared__ float ws[];                       // [taps][cin][cout]
  const int taps = g.kd * g.kh * g.kw;
  for (int i = threadIdx.x; i < taps * g.cin * g.cout; i += blockDim.x) {
    const int co = i % g.cout, ci = (i / g.cout) % g.cin, t = i / (g.cout * g.cin);
    ws[i] = w[((long long)co * g.cin + ci) * taps + t];
  }
  __syncthreads();
  const long long vox = xd.d * xd.h * xd.w, total = xd.n * vox;
  for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < total; v += (long long)gridDim.x * blockDim.x) {
    const long long n = v / vox;
    long long r = v - n * vox;
    const int iw = (int)(r % xd.w); r /= xd.w;
    const int ih = (int)(r % xd.h), id = (int)(r / xd.h);
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int a = 0; a < g.kd; ++a) {
      const int td = id + g.pd - a;
      if (td < 0 || td % g.sd) continue;
      const int od = td / g.sd;
      if (od >= yd.d) continue;
      for (int b = 0; b < g.kh; ++b) {
        const int th = ih + g.ph - b;
        if (th < 0 || th % g.sh) continue;
        const int oh = th / g.sh;
        if (oh >= yd.h) continue;
        for (int c = 0; c < g.kw; ++c) {
          const int tw = iw + g.pw - c;
          if (tw < 0 || tw % g.sw) continue;
          const int ow = tw / g.sw;
          if (ow >= yd.w) continue;
          const __nv_bfloat16* row = dy + n * yd.sn + od * yd.sd + oh * yd.sh + ow * yd.sw;
          const float* wt = ws + ((a * g.kh + b) * g.kw + c) * g.cin * g.cout;
          for (int c8 = 0; c8 < g.cout; c8 += 8) {
            float q[8];
            ld8(row + c8, q);
#pragma unroll
            for (int ci = 0; ci < 4; ++ci) {          // static indices keep acc[] in registers
              if (ci < g.cin) {
                const float4* wc = reinterpret_cast<const float4*>(wt + ci * g.cout + c8);
                const float4 w0 = wc[0], w1 = wc[1];
                acc[ci] = fmaf(q[0], w0.x, acc[ci]); acc[ci] = fmaf(q[1], w0.y, acc[ci]);
                acc[ci] = fmaf(q[2], w0.z, acc[ci]); acc[ci] = fmaf(q[3], w0.w, acc[ci]);
                acc[ci] = fmaf(q[4], w1.x, acc[ci]); acc[ci] = fmaf(q[5], w1.y, acc[ci]);
                acc[ci] = fmaf(q[6], w1.z, acc[ci]); acc[ci] = fmaf(q[7], w1.w, acc[ci]);
              }
            }
          }
        }
      }
    }
#pragma unroll
    for (int ci = 0; ci < 4; ++ci)
      if (ci < g.cin) stf(dx + xd.off(n, ci, id, ih, iw), acc[ci]);
  }
}

}  // namespace rg

using namespace rg;

extern "C" {

int rg_pw_conv_fprop(const rg_tensor5* xd, const void* x, const float* w, const float* bias, int act, float slope,
                     const rg_tensor5* yd, void* y, void* stream) {
  RG_CHECK_ARG(xd && x && w && yd && y, "rg_pw_conv_fprop: null argument");
  RG_UNSUPPORTED(!vec_ok(xd, x) || !pw_shape_ok((int)xd->c, (int)yd->c), "rg_pw_conv_fprop: unsupported layout / channel counts");
  RG_CHECK_ARG(xd->n == yd->n && xd->d == yd->d && xd->h == yd->h && xd->w == yd->w, "rg_pw_conv_fprop: shape mismatch");
  PwW pw{w, bias, (int)xd->c, (int)yd->c};
  long long total = (view(xd).numel() >> 3);
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  RG_DISPATCH_FLOAT(yd->dtype, TY, {
    const int grid = grid_for(xd->n * xd->d * xd->h, 1, 148 * 32);
    switch (pw.cout) {
      case 1: pw_fprop_kernel<TY, 1><<<grid, 256, 0, st>>>(pw, view(xd), (const __nv_bfloat16*)x, view(yd), (TY*)y, act, slope); break;
      case 2: pw_fprop_kernel<TY, 2><<<grid, 256, 0, st>>>(pw, view(xd), (const __nv_bfloat16*)x, view(yd), (TY*)y, act, slope); break;
      case 3: pw_fprop_kernel<TY, 3><<<grid, 256, 0, st>>>(pw, view(xd), (const __nv_bfloat16*)x, view(yd), (TY*)y, act, slope); break;
      default: pw_fprop_kernel<TY, 4><<<grid, 256, 0, st>>>(pw, view(xd), (const __nv_bfloat16*)x, view(yd), (TY*)y, act, slope); break;
    }
  })
  RG_LAUNCH_CHECK("rg_pw_conv_fprop");
  return RG_OK;
}

int rg_pw_conv_dgrad(const rg_tensor5* dyd, const void* dy, const float* w, const rg_tensor5* dxd, void* dx, void* stream) {
  RG_CHECK_ARG(dyd && dy && w && dxd && dx, "rg_pw_conv_dgrad: null argument");
  RG_UNSUPPORTED(!vec_ok(dxd, dx) || !pw_shape_ok((int)dxd->c, (int)dyd->c), "rg_pw_conv_dgrad: unsupported layout / channel counts");
  RG_CHECK_ARG(dxd->n == dyd->n && dxd->d == dyd->d && dxd->h == dyd->h && dxd->w == dyd->w, "rg_pw_conv_dgrad: shape mismatch");
  PwW pw{w, nullptr, (int)dxd->c, (int)dyd->c};
  long long total = (view(dxd).numel() >> 3);
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  RG_DISPATCH_FLOAT(dyd->dtype, TG, {
    pw_dgrad_kernel<TG><<<grid_for(total, 256), 256, 0, st>>>(pw, view(dyd), (const TG*)dy, view(dxd), (__nv_bfloat16*)dx);
  })
  RG_LAUNCH_CHECK("rg_pw_conv_dgrad");
  return RG_OK;
}

size_t rg_pw_conv_wgrad_workspace_bytes(int cin, int cout) { return (size_t)148 * 4 * cout * (cin + 1) * sizeof(float); }

int rg_pw_conv_wgrad(const rg_tensor5* xd, const void* x, const rg_tensor5* dyd, const void* dy, float* dw, float* dbias,
                     float* ws, int accumulate, void* stream) {
  RG_CHECK_ARG(xd && x && dyd && dy && dw && ws, "rg_pw_conv_wgrad: null argument");
  RG_UNSUPPORTED(!vec_ok(xd, x) || !pw_shape_ok((int)xd->c, (int)dyd->c), "rg_pw_conv_wgrad: unsupported layout / channel counts");
  RG_CHECK_ARG(xd->n == dyd->n && xd->d == dyd->d && xd->h == dyd->h && xd->w == dyd->w, "rg_pw_conv_wgrad: shape mismatch");
  long long vox = xd->n * xd->d * xd->h * xd->w;
  if (vox == 0) return RG_OK;
  long long blocks = 148 * 4;
  long long vpb = (vox + blocks - 1) / blocks;
  if (vpb < 256) vpb = 256;
  blocks = (vox + vpb - 1) / vpb;
  cudaStream_t st = (cudaStream_t)stream;
  RG_DISPATCH_FLOAT(dyd->dtype, TG, {
    pw_wgrad_kernel<TG><<<(unsigned)blocks, 256, 0, st>>>((int)xd->c, (int)dyd->c, view(xd), (const __nv_bfloat16*)x, view(dyd),
                                                          (const TG*)dy, ws, vpb);
  })
  RG_LAUNCH_CHECK("rg_pw_conv_wgrad");
  const long long nw = (long long)xd->c * dyd->c;
  int rc = launch_rows_sum(ws, (int)blocks, nw, nw + dyd->c, dw, nullptr, nw, accumulate, 1.0, st);
  if (rc || !dbias) return rc;
  return launch_rows_sum(ws + nw, (int)blocks, dyd->c, nw + dyd->c, dbias, nullptr, dyd->c, 0, 1.0, st);   // bias: always overwritten
}

int rg_smallcin_dgrad(const rg_tensor5* dyd, const void* dy, const float* w, const rg_tensor5* dxd, void* dx,
                      const rg_conv_geom* g, void* stream) {
  RG_CHECK_ARG(dyd && dy && w && dxd && dx && g, "rg_smallcin_dgrad: null argument");
  RG_UNSUPPORTED(dxd->c < 1 || dxd->c > 4 || dyd->c % 8 != 0 || g->groups != 1, "rg_smallcin_dgrad: needs Cin <= 4, Cout % 8 == 0");
  RG_UNSUPPORTED(dyd->dtype != RG_BF16 || !vec_ok(dyd, dy), "rg_smallcin_dgrad: dy must be bf16 with channels innermost");
  const int taps = g->kernel[0] * g->kernel[1] * g->kernel[2];
  const size_t smem = (size_t)taps * dxd->c * dyd->c * sizeof(float);
  RG_UNSUPPORTED(smem > 48 * 1024, "rg_smallcin_dgrad: weights do not fit shared memory");
  SmallDgradP p{(int)dxd->c, (int)dyd->c, g->kernel[0], g->kernel[1], g->kernel[2], g->stride[0], g->stride[1], g->stride[2],
                g->pad[0], g->pad[1], g->pad[2]};
  const long long total = dxd->n * dxd->d * dxd->h * dxd->w;
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  RG_DISPATCH_FLOAT(dxd->dtype, TD, {
    smallcin_dgrad_kernel<TD><<<grid_for(total, 256), 256, smem, st>>>(p, view(dyd), (const __nv_bfloat16*)dy, w, view(dxd), (TD*)dx);
  })
  RG_LAUNCH_CHECK("rg_smallcin_dgrad");
  return RG_OK;
}


}  // extern "C"
