// Max-pool, nearest up-sampling, loss reductions (warp-shuffle), u8 normalise / denormalise, Adam.
#include "common.cuh"
#include "reduce_rows.cuh"
#include "vec_ndhwc.cuh"

namespace rg {

// ---- MaxPool3d(f, stride=f), floor mode (UNet.py:209-218) ---------------------------------
template <typename TS, typename TD>
__global__ void maxpool_fwd_kernel(T5 xd, const TS* __restrict__ x, T5 yd, TD* __restrict__ y, int fd, int fh,
                                   int fw, bool c_inner) {
  long long total = yd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, yd, c_inner);
    float m = -INFINITY;
    for (int a = 0; a < fd; ++a)
      for (int b = 0; b < fh; ++b)
        for (int c = 0; c < fw; ++c) {
          float v = ldf(x + xd.off(p.n, p.c, p.d * fd + a, p.h * fh + b, p.w * fw + c));
          m = (v > m || v != v) ? v : m;  // NaN propagates like ATen
        }
    stf(y + yd.off(p.n, p.c, p.d, p.h, p.w), m);
  }
}

// NDHWC bf16 fast path: one block per output row (n, d, h); a thread reduces the window of 8 channels of one output
// voxel with 16-byte loads (HBM-bound: reads the input once, 32-bit index arithmetic only).
__global__ void __launch_bounds__(256) vec_maxpool_fwd_kernel(T5 xd, const __nv_bfloat16* __restrict__ x, T5 yd,
                                                              __nv_bfloat16* __restrict__ y, int fd, int fh, int fw) {
  const int cg = (int)(yd.c >> 3), W = (int)yd.w, H = (int)yd.h, D = (int)yd.d;
  const long long rows = yd.n * D * H;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int h = (int)(row % H);
    const long long t = row / H;
    const int d = (int)(t % D), n = (int)(t / D);
    const __nv_bfloat16* xrow = x + n * xd.sn + (long long)d * fd * xd.sd + (long long)h * fh * xd.sh;
    __nv_bfloat16* yrow = y + n * yd.sn + d * yd.sd + h * yd.sh;
    for (int j = threadIdx.x; j < W * cg; j += blockDim.x) {
      const int w = j / cg, c = (j - w * cg) << 3;
      float m[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) m[k] = -INFINITY;
      for (int a = 0; a < fd; ++a)
        for (int b = 0; b < fh; ++b)
          for (int e = 0; e < fw; ++e) {
            float v[8];
            ld8(xrow + a * xd.sd + b * xd.sh + (long long)(w * fw + e) * xd.sw + c, v);
#pragma unroll
            for (int k = 0; k < 8; ++k) m[k] = (v[k] > m[k] || v[k] != v[k]) ? v[k] : m[k];
          }
      st8(yrow + (long long)w * yd.sw + c, m);
    }
  }
}

// Row-based NDHWC bf16 backward: one block per dy row (n, d, h), a thread owns (w, 8 channels) windows; two passes
// over the window (find the first maximal element in ATen order d,h,w -- NaN wins -- then write dy or 0).
__global__ void __launch_bounds__(256) vec_maxpool_bwd_kernel(T5 xd, const __nv_bfloat16* __restrict__ x, T5 gd,
                                                              const __nv_bfloat16* __restrict__ dy, T5 dxd,
                                                              __nv_bfloat16* __restrict__ dx, int fd, int fh, int fw) {
  const int cg = (int)(gd.c >> 3), W = (int)gd.w, H = (int)gd.h, D = (int)gd.d;
  const long long rows = gd.n * D * H;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int h = (int)(row % H);
    const long long t = row / H;
    const int d = (int)(t % D), n = (int)(t / D);
    const __nv_bfloat16* xrow = x + n * xd.sn + (long long)d * fd * xd.sd + (long long)h * fh * xd.sh;
    __nv_bfloat16* dxrow = dx + n * dxd.sn + (long long)d * fd * dxd.sd + (long long)h * fh * dxd.sh;
    const __nv_bfloat16* grow = dy + n * gd.sn + d * gd.sd + h * gd.sh;
    for (int j = threadIdx.x; j < W * cg; j += blockDim.x) {
      const int w = j / cg, c = (j - w * cg) << 3;
      float m[8];
      int best[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) { m[k] = -INFINITY; best[k] = -1; }
      int idx = 0;
      for (int a = 0; a < fd; ++a)
        for (int b = 0; b < fh; ++b)
          for (int e = 0; e < fw; ++e, ++idx) {
            float v[8];
            ld8(xrow + a * xd.sd + b * xd.sh + (long long)(w * fw + e) * xd.sw + c, v);
#pragma unroll
            for (int k = 0; k < 8; ++k)
              if (best[k] < 0 || v[k] > m[k] || (v[k] != v[k] && m[k] == m[k])) { m[k] = v[k]; best[k] = idx; }
          }
      float g[8];
      ld8(grow + (long long)w * gd.sw + c, g);
      idx = 0;
      for (int a = 0; a < fd; ++a)
        for (int b = 0; b < fh; ++b)
          for (int e = 0; e < fw; ++e, ++idx) {
            float o[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = best[k] == idx ? g[k] : 0.f;
            st8(dxrow + a * dxd.sd + b * dxd.sh + (long long)(w * fw + e) * dxd.sw + c, o);
          }
    }
  }
}

// backward: one thread per OUTPUT window; the first maximal element (ATen order d,h,w) gets dy.
// Windows are disjoint (stride == kernel) so plain stores are race-free.
template <typename TS, typename TD>
__global__ void maxpool_bwd_kernel(T5 xd, const TS* __restrict__ x, T5 gd, const TS* __restrict__ dy, T5 dxd,
                                   TD* __restrict__ dx, int fd, int fh, int fw, bool c_inner) {
  long long total = gd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, gd, c_inner);
    float m = -INFINITY;
    int ba = 0, bb = 0, bc = 0;
    bool first = true;
    for (int a = 0; a < fd; ++a)
      for (int b = 0; b < fh; ++b)
        for (int c = 0; c < fw; ++c) {
          float v = ldf(x + xd.off(p.n, p.c, p.d * fd + a, p.h * fh + b, p.w * fw + c));
          if (first || v > m || v != v) { m = v; ba = a; bb = b; bc = c; first = false; }
        }
    float g = ldf(dy + gd.off(p.n, p.c, p.d, p.h, p.w));
    for (int a = 0; a < fd; ++a)
      for (int b = 0; b < fh; ++b)
        for (int c = 0; c < fw; ++c)
          stf(dx + dxd.off(p.n, p.c, p.d * fd + a, p.h * fh + b, p.w * fw + c),
              (a == ba && b == bb && c == bc) ? g : 0.f);
  }
}

// ---- nn.Upsample(scale_factor=f, mode="nearest") (UNet.py:273) ------------------------------
template <typename TS, typename TD>
__global__ void upsample_fwd_kernel(T5 xd, const TS* __restrict__ x, T5 yd, TD* __restrict__ y, int fd, int fh,
                                    int fw, bool c_inner) {
  long long total = yd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, yd, c_inner);
    stf(y + yd.off(p.n, p.c, p.d, p.h, p.w), ldf(x + xd.off(p.n, p.c, p.d / fd, p.h / fh, p.w / fw)));
  }
}

template <typename TS, typename TD>
__global__ void upsample_bwd_kernel(T5 gd, const TS* __restrict__ dy, T5 dxd, TD* __restrict__ dx, int fd, int fh,
                                    int fw, bool c_inner) {
  long long total = dxd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, dxd, c_inner);
    float acc = 0.f;
    for (int a = 0; a < fd; ++a)
      for (int b = 0; b < fh; ++b)
        for (int c = 0; c < fw; ++c) acc += ldf(dy + gd.off(p.n, p.c, p.d * fd + a, p.h * fh + b, p.w * fw + c));
    stf(dx + dxd.off(p.n, p.c, p.d, p.h, p.w), acc);
  }
}

// ---- losses: mean reduction + gradient in one pass ------------------------------------------
// Each thread accumulates a private fp32 partial, warp-shuffle reduce, one smem slot per warp,
// one partial per block (grid is <= 148*8 blocks) summed in block order by rows_sum_kernel (reduce_rows.cuh).
template <typename TA, typename TB, typename TG>
__global__ void loss_kernel(int kind, T5 ad, const TA* __restrict__ a, T5 bd, const TB* __restrict__ b, float target,
                            float beta, float scale_loss, float* __restrict__ loss, float scale_grad, T5 gd,
                            TG* __restrict__ grad, bool c_inner) {
  long long total = ad.numel();
  float inv = 1.f / (float)total;
  float part = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, ad, c_inner);
    float av = ldf(a + ad.off(p.n, p.c, p.d, p.h, p.w));
    float l, g;
    if (kind == RG_LOSS_MSE_CONST) {
      float d = av - target; l = d * d; g = 2.f * d;
    } else if (kind == RG_LOSS_MEAN) {
      l = av; g = 1.f;
    } else {
      float d = av - ldf(b + bd.off(p.n, p.c, p.d, p.h, p.w));
      float ad_ = fabsf(d);
      if (kind == RG_LOSS_SMOOTHL1 && ad_ < beta) {
        l = 0.5f * d * d / beta; g = d / beta;
      } else {
        l = (kind == RG_LOSS_SMOOTHL1) ? ad_ - 0.5f * beta : ad_;
        g = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
      }
    }
    part += l;
    if (grad) stf(grad + gd.off(p.n, p.c, p.d, p.h, p.w), g * inv * scale_grad);
  }
  __shared__ float ws[32];
  part = warp_sum(part);
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) ws[wid] = part;
  __syncthreads();
  if (wid == 0) {
    float v = lane < (blockDim.x >> 5) ? ws[lane] : 0.f;
    v = warp_sum(v);
    if (lane == 0) loss[blockIdx.x] = v;        // `loss` = this launch's partial rows
  }
}

// ---- predict/worker.py:100-114 / :133-147 -----------------------------------------------------
template <typename TD>
__global__ void u8_normalise_kernel(const uint8_t* __restrict__ x, long long count, float scale, float shift,
                                    TD* __restrict__ y) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count;
       i += (long long)gridDim.x * blockDim.x) {
    float v = (float)x[i];
    v = v - 0.f;      // data -= iinfo.min
    v = v / 255.f;    // data /= iinfo.max
    v = v * scale;    // scaleShift_input
    v = v + shift;
    stf(y + i, v);
  }
}

template <typename TS>
__global__ void denorm_u8_kernel(T5 xd, const TS* __restrict__ x, uint8_t* __restrict__ out) {
  long long total = xd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, xd, false);
    float v = ldf(x + xd.off(p.n, p.c, p.d, p.h, p.w)) * 255.f;
    v = fminf(fmaxf(v, 0.f), 255.f);
    out[i] = (uint8_t)v;  // truncation toward zero, like numpy astype (worker.py:147)
  }
}

// ---- Adam (torch.optim.Adam, no amsgrad / weight decay) ----------------------------------------
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, long long count, float lr, float b1, float b2, float eps,
                            float bc1, float bc2_sqrt, float gscale) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count;
       i += (long long)gridDim.x * blockDim.x) {
    float gg = g[i] * gscale;
    float mm = m[i] * b1 + gg * (1.f - b1);        // lerp form differs from torch by <=1ulp
    float vv = v[i] * b2 + gg * gg * (1.f - b2);
    m[i] = mm; v[i] = vv;
    float denom = sqrtf(vv) / bc2_sqrt + eps;
    p[i] = p[i] - (lr / bc1) * (mm / denom);
  }
}

// Same update over a flat bucket with the step counter ON THE DEVICE (a captured CUDA graph replays with frozen kernel
// arguments: the bias corrections must come from memory).  step_dev[0] is incremented by adam_step_inc_kernel first.
__global__ void adam_step_inc_kernel(long long* step_dev) { step_dev[0] += 1; }

__global__ void adam_flat_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                 float* __restrict__ v, long long count, float lr, float b1, float b2, float eps,
                                 const long long* __restrict__ step_dev, float gscale) {
  const double step = (double)step_dev[0];
  const float bc1 = (float)(1.0 - pow((double)b1, step));
  const float bc2_sqrt = (float)sqrt(1.0 - pow((double)b2, step));
  const float step_size = lr / bc1;
  const long long n4 = count >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg4 = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    float* pe = &pp.x; const float* ge = &gg4.x; float* me = &mm.x; float* ve = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gg = ge[k] * gscale;
      me[k] = me[k] * b1 + gg * (1.f - b1);
      ve[k] = ve[k] * b2 + gg * gg * (1.f - b2);
      pe[k] = pe[k] - step_size * (me[k] / (sqrtf(ve[k]) / bc2_sqrt + eps));
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  for (long long i = (n4 << 2) + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
    const float gg = g[i] * gscale;
    const float mm = m[i] * b1 + gg * (1.f - b1), vv = v[i] * b2 + gg * gg * (1.f - b2);
    m[i] = mm; v[i] = vv;
    p[i] = p[i] - step_size * (mm / (sqrtf(vv) / bc2_sqrt + eps));
  }
}

}  // namespace rg

using namespace rg;

extern "C" {

static int check_pool(const rg_tensor5* big, const rg_tensor5* small, const int32_t f[3], const char* who) {
  RG_CHECK_ARG(f[0] > 0 && f[1] > 0 && f[2] > 0, "%s: bad factor", who);
  RG_CHECK_ARG(big->n == small->n && big->c == small->c, "%s: N/C mismatch", who);
  RG_CHECK_ARG(small->d == big->d / f[0] && small->h == big->h / f[1] && small->w == big->w / f[2],
               "%s: pooled shape mismatch", who);
  return RG_OK;
}

int rg_maxpool3d_fwd(const rg_tensor5* xd, const void* x, const int32_t f[3], const rg_tensor5* yd, void* y,
                     void* stream) {
  RG_CHECK_ARG(xd && yd && x && y && f, "rg_maxpool3d_fwd: null argument");
  int rc = check_pool(xd, yd, f, "rg_maxpool3d_fwd");
  if (rc) return rc;
  long long total = view(yd).numel();
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (vec_ok(xd, x) && vec_ok(yd, y)) {
    vec_maxpool_fwd_kernel<<<grid_for(yd->n * yd->d * yd->h, 1, 148 * 32), 256, 0, st>>>(
        view(xd), (const __nv_bfloat16*)x, view(yd), (__nv_bfloat16*)y, f[0], f[1], f[2]);
    RG_LAUNCH_CHECK("rg_maxpool3d_fwd(vec)");
    return RG_OK;
  }
  int grid = grid_for(total, 256);
  bool ci = channels_inner(yd);
  RG_DISPATCH_FLOAT(xd->dtype, TS, RG_DISPATCH_FLOAT(yd->dtype, TD, {
    maxpool_fwd_kernel<TS, TD><<<grid, 256, 0, st>>>(view(xd), (const TS*)x, view(yd), (TD*)y, f[0], f[1], f[2], ci);
  }))
  RG_LAUNCH_CHECK("rg_maxpool3d_fwd");
  return RG_OK;
}

int rg_maxpool3d_bwd(const rg_tensor5* xd, const void* x, const rg_tensor5* yd, const void* y,
                     const rg_tensor5* dyd, const void* dy, const int32_t f[3], const rg_tensor5* dxd, void* dx,
                     void* stream) {
  (void)yd; (void)y;
  RG_CHECK_ARG(xd && dyd && dxd && x && dy && dx && f, "rg_maxpool3d_bwd: null argument");
  int rc = check_pool(xd, dyd, f, "rg_maxpool3d_bwd");
  if (rc) return rc;
  RG_CHECK_ARG(same_shape(xd, dxd), "rg_maxpool3d_bwd: dx shape mismatch");
  RG_CHECK_ARG(xd->dtype == dyd->dtype, "rg_maxpool3d_bwd: x / dy dtype mismatch");
  long long total = view(dyd).numel();
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (vec_ok(xd, x) && vec_ok(dyd, dy) && vec_ok(dxd, dx)) {
    vec_maxpool_bwd_kernel<<<grid_for(dyd->n * dyd->d * dyd->h, 1, 148 * 32), 256, 0, st>>>(
        view(xd), (const __nv_bfloat16*)x, view(dyd), (const __nv_bfloat16*)dy, view(dxd), (__nv_bfloat16*)dx, f[0], f[1], f[2]);
    RG_LAUNCH_CHECK("rg_maxpool3d_bwd(vec)");
    return RG_OK;
  }
  int grid = grid_for(total, 256);
  bool ci = channels_inner(dyd);
  RG_DISPATCH_FLOAT(xd->dtype, TS, RG_DISPATCH_FLOAT(dxd->dtype, TD, {
    maxpool_bwd_kernel<TS, TD><<<grid, 256, 0, st>>>(view(xd), (const TS*)x, view(dyd), (const TS*)dy, view(dxd), (TD*)dx, f[0], f[1], f[2], ci);
  }))
  RG_LAUNCH_CHECK("rg_maxpool3d_bwd");
  return RG_OK;
}

int rg_upsample_nearest_fwd(const rg_tensor5* xd, const void* x, const int32_t f[3], const rg_tensor5* yd, void* y,
                            void* stream) {
  RG_CHECK_ARG(xd && yd && x && y && f, "rg_upsample_nearest_fwd: null argument");
  RG_CHECK_ARG(yd->d == xd->d * f[0] && yd->h == xd->h * f[1] && yd->w == xd->w * f[2] && yd->n == xd->n && yd->c == xd->c,
               "rg_upsample_nearest_fwd: shape mismatch");
  long long total = view(yd).numel();
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int grid = grid_for(total, 256);
  bool ci = channels_inner(yd);
  RG_DISPATCH_FLOAT(xd->dtype, TS, RG_DISPATCH_FLOAT(yd->dtype, TD, {
    upsample_fwd_kernel<TS, TD><<<grid, 256, 0, st>>>(view(xd), (const TS*)x, view(yd), (TD*)y, f[0], f[1], f[2], ci);
  }))
  RG_LAUNCH_CHECK("rg_upsample_nearest_fwd");
  return RG_OK;
}

int rg_upsample_nearest_bwd(const rg_tensor5* dyd, const void* dy, const int32_t f[3], const rg_tensor5* dxd,
                            void* dx, void* stream) {
  RG_CHECK_ARG(dyd && dxd && dy && dx && f, "rg_upsample_nearest_bwd: null argument");
  RG_CHECK_ARG(dyd->d == dxd->d * f[0] && dyd->h == dxd->h * f[1] && dyd->w == dxd->w * f[2] && dyd->n == dxd->n && dyd->c == dxd->c,
               "rg_upsample_nearest_bwd: shape mismatch");
  long long total = view(dxd).numel();
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int grid = grid_for(total, 256);
  bool ci = channels_inner(dxd);
  RG_DISPATCH_FLOAT(dyd->dtype, TS, RG_DISPATCH_FLOAT(dxd->dtype, TD, {
    upsample_bwd_kernel<TS, TD><<<grid, 256, 0, st>>>(view(dyd), (const TS*)dy, view(dxd), (TD*)dx, f[0], f[1], f[2], ci);
  }))
  RG_LAUNCH_CHECK("rg_upsample_nearest_bwd");
  return RG_OK;
}

int rg_loss_fwd_grad(int kind, const rg_tensor5* ad, const void* a, const rg_tensor5* bd, const void* b, float target,
                     float beta, float scale_loss, float* loss, float scale_grad, const rg_tensor5* gd, void* grad_a,
                     float* ws, void* stream) {
  RG_CHECK_ARG(ad && a && loss && ws, "rg_loss_fwd_grad: null argument");
  RG_CHECK_ARG(kind >= RG_LOSS_MSE_CONST && kind <= RG_LOSS_MEAN, "rg_loss_fwd_grad: bad kind %d", kind);
  bool two = (kind == RG_LOSS_SMOOTHL1 || kind == RG_LOSS_L1);
  if (two) {
    RG_CHECK_ARG(bd && b && same_shape(ad, bd), "rg_loss_fwd_grad: second operand missing or shape mismatch");
    RG_CHECK_ARG(kind != RG_LOSS_SMOOTHL1 || beta > 0.f, "rg_loss_fwd_grad: beta must be > 0");
  }
  if (grad_a) RG_CHECK_ARG(gd && same_shape(ad, gd), "rg_loss_fwd_grad: grad shape mismatch");
  long long total = view(ad).numel();
  RG_CHECK_ARG(total > 0, "rg_loss_fwd_grad: empty input");
  cudaStream_t st = (cudaStream_t)stream;
  int grid = grid_for(total, 256, 148 * 8);
  bool ci = channels_inner(ad);
  T5 bv = two ? view(bd) : view(ad);
  T5 gv = grad_a ? view(gd) : view(ad);
  int bdt = two ? bd->dtype : ad->dtype;
  int gdt = grad_a ? gd->dtype : ad->dtype;
  RG_DISPATCH_FLOAT(ad->dtype, TA, RG_DISPATCH_FLOAT(bdt, TB, RG_DISPATCH_FLOAT(gdt, TG, {
    loss_kernel<TA, TB, TG><<<grid, 256, 0, st>>>(kind, view(ad), (const TA*)a, bv, (const TB*)b, target, beta, scale_loss, ws, scale_grad, gv, (TG*)grad_a, ci);
  })))
  RG_LAUNCH_CHECK("rg_loss_fwd_grad");
  return launch_rows_sum(ws, grid, 1, 1, loss, nullptr, 1, 0, (double)scale_loss / (double)total, st);
}

int rg_u8_normalise(const uint8_t* x, int64_t count, float scale, float shift, int out_dtype, void* y, void* stream) {
  RG_CHECK_ARG(x && y && count >= 0, "rg_u8_normalise: bad argument");
  if (count == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int grid = grid_for(count, 256);
  RG_DISPATCH_FLOAT(out_dtype, TD, { u8_normalise_kernel<TD><<<grid, 256, 0, st>>>(x, count, scale, shift, (TD*)y); })
  RG_LAUNCH_CHECK("rg_u8_normalise");
  return RG_OK;
}

int rg_denorm_clamp_u8(const rg_tensor5* xd, const void* x, uint8_t* out, void* stream) {
  RG_CHECK_ARG(xd && x && out, "rg_denorm_clamp_u8: null argument");
  long long total = view(xd).numel();
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int grid = grid_for(total, 256);
  RG_DISPATCH_FLOAT(xd->dtype, TS, { denorm_u8_kernel<TS><<<grid, 256, 0, st>>>(view(xd), (const TS*)x, out); })
  RG_LAUNCH_CHECK("rg_denorm_clamp_u8");
  return RG_OK;
}

int rg_adam_step(float* p, const float* g, float* m, float* v, int64_t count, float lr, float beta1, float beta2,
                 float eps, int step, float grad_scale, void* stream) {
  RG_CHECK_ARG(p && g && m && v && count >= 0 && step >= 1, "rg_adam_step: bad argument");
  if (count == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  double bc1 = 1.0 - pow((double)beta1, (double)step);
  double bc2 = 1.0 - pow((double)beta2, (double)step);
  adam_kernel<<<grid_for(count, 256), 256, 0, st>>>(p, g, m, v, count, lr, beta1, beta2, eps, (float)bc1,
                                                    (float)sqrt(bc2), grad_scale);
  RG_LAUNCH_CHECK("rg_adam_step");
  return RG_OK;
}

int rg_adam_flat(float* p, const float* g, float* m, float* v, int64_t count, float lr, float beta1, float beta2, float eps,
                 long long* step_dev, float grad_scale, void* stream) {
  RG_CHECK_ARG(p && g && m && v && step_dev && count >= 0, "rg_adam_flat: bad argument");
  RG_CHECK_ARG((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0, "rg_adam_flat: buckets must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  adam_step_inc_kernel<<<1, 1, 0, st>>>(step_dev);
  RG_LAUNCH_CHECK("rg_adam_flat(step)");
  if (count == 0) return RG_OK;
  adam_flat_kernel<<<grid_for((count + 3) / 4, 256, 148 * 8), 256, 0, st>>>(p, g, m, v, count, lr, beta1, beta2, eps, step_dev,
                                                                           grad_scale);
  RG_LAUNCH_CHECK("rg_adam_flat");
  return RG_OK;
}

}  // extern "C"
