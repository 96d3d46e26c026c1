// Direct (CUDA-core) convolutions: the shape-generic, fp32-accumulate path.
//
// Serves (a) the fp32 parity mode (max-abs <= 1e-4 per layer vs the reference; tcgen05 has no IEEE
// fp32 MMA), and (b) in bf16 mode the layers the tensor-core path does not take: Cin=1 stems,
// Cout=1 tails / 1x1x1 heads, grouped 1x1x1 maps, odd channel counts.  Those are bandwidth-bound.
//
// All three kernels are one smem-tiled implicit GEMM (64x64 tile, BK=16, 4x4 register micro-tile per
// thread, 256 threads) that differs only in how the A / B tiles are gathered:
//   fprop: M = output voxels, N = Cout, K = Cin*taps          (A gathered from x)
//   dgrad: M = input  voxels, N = Cin,  K = Cout*taps         (A gathered from dy; also convT fwd)
//   wgrad: M = Cout,          N = Cin*taps, K = voxels (split-K over gridDim.z, fp32 atomics)
// Tensors are stride-generic (NCDHW fp32 or NDHWC bf16 ...). Zero padding only (see rg_pad3d_*).
#include "common.cuh"
#include "reduce_rows.cuh"

namespace rg {

constexpr int BM = 64, BN = 64, BK = 16, PADT = 4;

struct ConvP {
  T5 x, y;            // fprop-input-side tensor and fprop-output-side tensor (x/dx and y/dy)
  int cin, cout;      // per group
  int kd, kh, kw, sd, sh, sw, pd, ph, pw;
  int act; float slope;
};

#define RG_MICRO_KERNEL()                                                  \
  _Pragma("unroll") for (int kk = 0; kk < BK; ++kk) {                      \
    float a[4], b[4];                                                      \
    _Pragma("unroll") for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i]; \
    _Pragma("unroll") for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j]; \
    _Pragma("unroll") for (int i = 0; i < 4; ++i)                          \
      _Pragma("unroll") for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]); \
  }

template <typename TX, typename TY>
__global__ void __launch_bounds__(256) direct_fprop_kernel(ConvP p, const TX* __restrict__ x, const float* __restrict__ w,
                                                           const float* __restrict__ bias, TY* __restrict__ y) {
  __shared__ float As[BK][BM + PADT], Bs[BK][BN + PADT];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int taps = p.kd * p.kh * p.kw;
  const long long K = (long long)p.cin * taps;
  const long long ovox = p.y.d * p.y.h * p.y.w, Mtot = p.y.n * ovox;
  const long long m0 = (long long)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  // this thread's A-gather row (fixed voxel for all its loads)
  const int am = tid & 63, ak0 = tid >> 6;
  long long mm = m0 + am;
  bool mvalid = mm < Mtot;
  long long an = 0, od = 0, oh = 0, ow = 0;
  if (mvalid) { an = mm / ovox; long long r = mm % ovox; ow = r % p.y.w; r /= p.y.w; oh = r % p.y.h; od = r / p.y.h; }
  const long long id0 = od * p.sd - p.pd, ih0 = oh * p.sh - p.ph, iw0 = ow * p.sw - p.pw;
  float acc[4][4] = {};
  for (long long k0 = 0; k0 < K; k0 += BK) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int kk = ak0 + j * 4;
      long long k = k0 + kk;
      float v = 0.f;
      if (mvalid && k < K) {
        int ci = (int)(k / taps), t = (int)(k % taps);
        int kw_ = t % p.kw; t /= p.kw; int kh_ = t % p.kh; int kd_ = t / p.kh;
        long long id = id0 + kd_, ih = ih0 + kh_, iw = iw0 + kw_;
        if (id >= 0 && id < p.x.d && ih >= 0 && ih < p.x.h && iw >= 0 && iw < p.x.w)
          v = ldf(x + p.x.off(an, ci, id, ih, iw));
      }
      As[kk][am] = v;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int idx = tid + j * 256;
      int kk = idx & 15, nn = idx >> 4;
      long long k = k0 + kk;
      int co = n0 + nn;
      Bs[kk][nn] = (k < K && co < p.cout) ? w[(long long)co * K + k] : 0.f;
    }
    __syncthreads();
    RG_MICRO_KERNEL()
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    long long m = m0 + ty * 4 + i;
    if (m >= Mtot) continue;
    long long n = m / ovox, r = m % ovox;
    long long qw = r % p.y.w; r /= p.y.w; long long qh = r % p.y.h, qd = r / p.y.h;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int co = n0 + tx * 4 + j;
      if (co >= p.cout) continue;
      float v = acc[i][j] + (bias ? bias[co] : 0.f);
      stf(y + p.y.off(n, co, qd, qh, qw), act_apply(v, p.act, p.slope));
    }
  }
}

// dx[n,ci,i] = sum_{co,k : (i+p-k)%s==0} dy[n,co,(i+p-k)/s] * w[co,ci,k]   (+ bias[ci] for convT fwd)
template <typename TG, typename TD>
__global__ void __launch_bounds__(256) direct_dgrad_kernel(ConvP p, const TG* __restrict__ dy, const float* __restrict__ w,
                                                           const float* __restrict__ bias, TD* __restrict__ dx) {
  __shared__ float As[BK][BM + PADT], Bs[BK][BN + PADT];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int taps = p.kd * p.kh * p.kw;
  const long long K = (long long)p.cout * taps;
  const long long ivox = p.x.d * p.x.h * p.x.w, Mtot = p.x.n * ivox;
  const long long m0 = (long long)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int am = tid & 63, ak0 = tid >> 6;
  long long mm = m0 + am;
  bool mvalid = mm < Mtot;
  long long an = 0, id = 0, ih = 0, iw = 0;
  if (mvalid) { an = mm / ivox; long long r = mm % ivox; iw = r % p.x.w; r /= p.x.w; ih = r % p.x.h; id = r / p.x.h; }
  float acc[4][4] = {};
  for (long long k0 = 0; k0 < K; k0 += BK) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int kk = ak0 + j * 4;
      long long k = k0 + kk;
      float v = 0.f;
      if (mvalid && k < K) {
        int co = (int)(k / taps), t = (int)(k % taps);
        int kw_ = t % p.kw; t /= p.kw; int kh_ = t % p.kh; int kd_ = t / p.kh;
        long long td = id + p.pd - kd_, th = ih + p.ph - kh_, tw = iw + p.pw - kw_;
        if (td >= 0 && th >= 0 && tw >= 0 && td % p.sd == 0 && th % p.sh == 0 && tw % p.sw == 0) {
          td /= p.sd; th /= p.sh; tw /= p.sw;
          if (td < p.y.d && th < p.y.h && tw < p.y.w) v = ldf(dy + p.y.off(an, co, td, th, tw));
        }
      }
      As[kk][am] = v;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int idx = tid + j * 256;
      int nn = idx & 63, kk = idx >> 6;
      long long k = k0 + kk;
      int ci = n0 + nn;
      float v = 0.f;
      if (k < K && ci < p.cin) {
        int co = (int)(k / taps), t = (int)(k % taps);
        v = w[((long long)co * p.cin + ci) * taps + t];
      }
      Bs[kk][nn] = v;
    }
    __syncthreads();
    RG_MICRO_KERNEL()
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    long long m = m0 + ty * 4 + i;
    if (m >= Mtot) continue;
    long long n = m / ivox, r = m % ivox;
    long long qw = r % p.x.w; r /= p.x.w; long long qh = r % p.x.h, qd = r / p.x.h;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int ci = n0 + tx * 4 + j;
      if (ci >= p.cin) continue;
      float v = acc[i][j] + (bias ? bias[ci] : 0.f);
      stf(dx + p.x.off(n, ci, qd, qh, qw), act_apply(v, p.act, p.slope));
    }
  }
}

// part[split][co,ci,k] = sum over the split's (n,o) of dy[n,co,o] * x[n,ci,o*s-p+k]; split-K over blockIdx.z, one
// partial row per split (+ the bias partial behind it), added in split order by rows_sum_kernel
template <typename TX, typename TG>
__global__ void __launch_bounds__(256) direct_wgrad_kernel(ConvP p, const TX* __restrict__ x, const TG* __restrict__ dy,
                                                           float* __restrict__ part, long long ld, int want_bias,
                                                           long long k_per_split) {
  __shared__ float As[BK][BM + PADT], Bs[BK][BN + PADT];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int taps = p.kd * p.kh * p.kw;
  const long long NN = (long long)p.cin * taps;      // GEMM N extent
  const long long ovox = p.y.d * p.y.h * p.y.w, Ktot = p.y.n * ovox;
  const int m0 = blockIdx.x * BM;                    // cout tile
  const long long n0 = (long long)blockIdx.y * BN;   // (ci,tap) tile
  const long long kbeg = (long long)blockIdx.z * k_per_split;
  long long kend = kbeg + k_per_split; if (kend > Ktot) kend = Ktot;
  // B-gather column (fixed (ci,tap) per thread)
  const int bn = tid & 63, bk0 = tid >> 6;
  long long nn_g = n0 + bn;
  bool nvalid = nn_g < NN;
  int bci = 0, bkd = 0, bkh = 0, bkw = 0;
  if (nvalid) { bci = (int)(nn_g / taps); int t = (int)(nn_g % taps); bkw = t % p.kw; t /= p.kw; bkh = t % p.kh; bkd = t / p.kh; }
  float acc[4][4] = {};
  float rowsum[4] = {0.f, 0.f, 0.f, 0.f};
  const bool do_bias = want_bias && blockIdx.y == 0;
  const bool co_fast = p.y.sc == 1 && BM == 64 && BK == 16;
  // output voxel index -> (n, d, h, w): 32-bit divisions whenever the index space allows (a 64-bit division is ~60
  // instructions; eight of them per thread and K step dominated this kernel on the discriminator's 1 -> 64 stem)
  const bool small = Ktot < (1LL << 31);
  const unsigned ovox32 = (unsigned)ovox, yw32 = (unsigned)p.y.w, yh32 = (unsigned)p.y.h;
  auto decode = [&](long long k, long long& n, long long& qd, long long& qh, long long& qw) {
    if (small) {
      const unsigned k32 = (unsigned)k, n32 = k32 / ovox32;
      unsigned r = k32 - n32 * ovox32;
      const unsigned t = r / yw32;
      qw = r - t * yw32; qd = t / yh32; qh = t - (unsigned)qd * yh32; n = n32;
    } else {
      n = k / ovox; long long r = k % ovox;
      qw = r % p.y.w; r /= p.y.w; qh = r % p.y.h; qd = r / p.y.h;
    }
  };
  float* dw = part + (long long)blockIdx.z * ld;
  float* dbias = dw + (long long)p.cout * NN;
  for (long long k0 = kbeg; k0 < kend; k0 += BK) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {   // A: dy[vox][co]; threads fastest over vox (kk) for W-inner coalescing (NCDHW), or over
      int idx = tid + j * 256;      // co when the channels are innermost (NDHWC bf16: a warp reads 64 contiguous bytes)
      int kk = co_fast ? idx >> 6 : idx & 15, mm = co_fast ? idx & 63 : idx >> 4;
      long long k = k0 + kk;
      int co = m0 + mm;
      float v = 0.f;
      if (k < kend && co < p.cout) {
        long long n, qd, qh, qw;
        decode(k, n, qd, qh, qw);
        v = ldf(dy + p.y.off(n, co, qd, qh, qw));
      }
      As[kk][mm] = v;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int kk = bk0 + j * 4;
      long long k = k0 + kk;
      float v = 0.f;
      if (nvalid && k < kend) {
        long long n, qd, qh, qw;
        decode(k, n, qd, qh, qw);
        long long id = qd * p.sd - p.pd + bkd, ih = qh * p.sh - p.ph + bkh, iw = qw * p.sw - p.pw + bkw;
        if (id >= 0 && id < p.x.d && ih >= 0 && ih < p.x.h && iw >= 0 && iw < p.x.w)
          v = ldf(x + p.x.off(n, bci, id, ih, iw));
      }
      Bs[kk][bn] = v;
    }
    __syncthreads();
    RG_MICRO_KERNEL()
    if (do_bias && tx == 0) {
#pragma unroll
      for (int kk = 0; kk < BK; ++kk)
#pragma unroll
        for (int i = 0; i < 4; ++i) rowsum[i] += As[kk][ty * 4 + i];
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int co = m0 + ty * 4 + i;
    if (co >= p.cout) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      long long nn = n0 + tx * 4 + j;
      if (nn >= NN) continue;
      dw[(long long)co * NN + nn] = acc[i][j];
    }
    if (do_bias && tx == 0) dbias[co] = rowsum[i];
  }
}

static int fill_params(ConvP& p, const rg_tensor5* xd, const rg_tensor5* yd, const rg_conv_geom* g, const char* who) {
  RG_CHECK_ARG(g->groups >= 1, "%s: groups must be >= 1", who);
  RG_CHECK_ARG(xd->c % g->groups == 0 && yd->c % g->groups == 0, "%s: channels not divisible by groups", who);
  RG_CHECK_ARG(xd->n == yd->n, "%s: batch mismatch", who);
  for (int i = 0; i < 3; ++i) RG_CHECK_ARG(g->kernel[i] >= 1 && g->stride[i] >= 1 && g->pad[i] >= 0, "%s: bad geometry", who);
  const long long in[3] = {xd->d, xd->h, xd->w}, out[3] = {yd->d, yd->h, yd->w};
  for (int i = 0; i < 3; ++i) {
    long long lo = (out[i] - 1) * g->stride[i] + g->kernel[i] - 2 * g->pad[i];  // smallest legal input
    RG_CHECK_ARG(out[i] >= 1 && in[i] >= lo && in[i] < lo + g->stride[i],
                 "%s: spatial extents inconsistent with geometry (dim %d: in %lld out %lld)", who, i, in[i], out[i]);
  }
  p.x = view(xd); p.y = view(yd);
  p.cin = (int)(xd->c / g->groups); p.cout = (int)(yd->c / g->groups);
  p.kd = g->kernel[0]; p.kh = g->kernel[1]; p.kw = g->kernel[2];
  p.sd = g->stride[0]; p.sh = g->stride[1]; p.sw = g->stride[2];
  p.pd = g->pad[0]; p.ph = g->pad[1]; p.pw = g->pad[2];
  p.act = g->act; p.slope = g->slope;
  return RG_OK;
}

}  // namespace rg

using namespace rg;

extern "C" {

int rg_conv3d_direct_fprop(const rg_tensor5* xd, const void* x, const float* w, const float* bias,
                           const rg_tensor5* yd, void* y, const rg_conv_geom* g, void* stream) {
  RG_CHECK_ARG(xd && x && w && yd && y && g, "rg_conv3d_direct_fprop: null argument");
  ConvP p;
  int rc = fill_params(p, xd, yd, g, "rg_conv3d_direct_fprop");
  if (rc) return rc;
  long long Mtot = p.y.n * p.y.d * p.y.h * p.y.w;
  if (Mtot == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const long long taps = (long long)p.kd * p.kh * p.kw;
  dim3 grid((unsigned)((Mtot + BM - 1) / BM), (unsigned)((p.cout + BN - 1) / BN));
  for (int gi = 0; gi < g->groups; ++gi) {
    RG_DISPATCH_FLOAT(xd->dtype, TX, RG_DISPATCH_FLOAT(yd->dtype, TY, {
      const TX* xg = (const TX*)x + (long long)gi * p.cin * p.x.sc;
      TY* yg = (TY*)y + (long long)gi * p.cout * p.y.sc;
      direct_fprop_kernel<TX, TY><<<grid, 256, 0, st>>>(p, xg, w + (long long)gi * p.cout * p.cin * taps,
                                                        bias ? bias + gi * p.cout : nullptr, yg);
    }))
  }
  RG_LAUNCH_CHECK("rg_conv3d_direct_fprop");
  return RG_OK;
}

int rg_conv3d_direct_dgrad(const rg_tensor5* dyd, const void* dy, const float* w, const float* bias,
                           const rg_tensor5* dxd, void* dx, const rg_conv_geom* g, void* stream) {
  RG_CHECK_ARG(dyd && dy && w && dxd && dx && g, "rg_conv3d_direct_dgrad: null argument");
  ConvP p;
  int rc = fill_params(p, dxd, dyd, g, "rg_conv3d_direct_dgrad");
  if (rc) return rc;
  long long Mtot = p.x.n * p.x.d * p.x.h * p.x.w;
  if (Mtot == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const long long taps = (long long)p.kd * p.kh * p.kw;
  dim3 grid((unsigned)((Mtot + BM - 1) / BM), (unsigned)((p.cin + BN - 1) / BN));
  for (int gi = 0; gi < g->groups; ++gi) {
    RG_DISPATCH_FLOAT(dyd->dtype, TG, RG_DISPATCH_FLOAT(dxd->dtype, TD, {
      const TG* dyg = (const TG*)dy + (long long)gi * p.cout * p.y.sc;
      TD* dxg = (TD*)dx + (long long)gi * p.cin * p.x.sc;
      direct_dgrad_kernel<TG, TD><<<grid, 256, 0, st>>>(p, dyg, w + (long long)gi * p.cout * p.cin * taps,
                                                        bias ? bias + gi * p.cin : nullptr, dxg);
    }))
  }
  RG_LAUNCH_CHECK("rg_conv3d_direct_dgrad");
  return RG_OK;
}

static long long direct_wgrad_splits(const ConvP& p, long long* kps_out) {
  const long long taps = (long long)p.kd * p.kh * p.kw;
  long long Ktot = p.y.n * p.y.d * p.y.h * p.y.w;
  long long NN = (long long)p.cin * taps;
  unsigned gx = (unsigned)((p.cout + BM - 1) / BM), gy = (unsigned)((NN + BN - 1) / BN);
  long long want = (148LL * 4 + (long long)gx * gy - 1) / ((long long)gx * gy);   // ~4 blocks per SM
  long long maxsplit = (Ktot + 255) / 256;
  long long splits = want < 1 ? 1 : (want > maxsplit ? maxsplit : want);
  if (splits > 65535) splits = 65535;
  if (splits < 1) splits = 1;
  long long kps = (Ktot + splits - 1) / splits;
  kps = (kps + BK - 1) / BK * BK;
  if (kps < BK) kps = BK;
  splits = Ktot > 0 ? (Ktot + kps - 1) / kps : 1;
  if (kps_out) *kps_out = kps;
  return splits;
}

size_t rg_conv3d_direct_wgrad_workspace_bytes(const rg_tensor5* xd, const rg_tensor5* dyd, const rg_conv_geom* g) {
  if (!xd || !dyd || !g) return 0;
  ConvP p;
  if (fill_params(p, xd, dyd, g, "rg_conv3d_direct_wgrad_workspace_bytes")) return 0;
  const long long taps = (long long)p.kd * p.kh * p.kw;
  const long long ld = (long long)p.cout * p.cin * taps + p.cout;
  return (size_t)(direct_wgrad_splits(p, nullptr) * ld * g->groups) * sizeof(float);
}

int rg_conv3d_direct_wgrad(const rg_tensor5* xd, const void* x, const rg_tensor5* dyd, const void* dy, float* dw,
                           float* dbias, const rg_conv_geom* g, int accumulate, float* ws, void* stream) {
  RG_CHECK_ARG(xd && x && dyd && dy && dw && g && ws, "rg_conv3d_direct_wgrad: null argument");
  ConvP p;
  int rc = fill_params(p, xd, dyd, g, "rg_conv3d_direct_wgrad");
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const long long taps = (long long)p.kd * p.kh * p.kw;
  const long long wcount_g = (long long)p.cout * p.cin * taps;
  long long Ktot = p.y.n * p.y.d * p.y.h * p.y.w;
  if (Ktot == 0) {
    if (!accumulate) {
      cudaError_t e = cudaMemsetAsync(dw, 0, wcount_g * g->groups * sizeof(float), st);
      if (e != cudaSuccess) return cuda_fail(e, "rg_conv3d_direct_wgrad memset");
      if (dbias) {
        e = cudaMemsetAsync(dbias, 0, dyd->c * sizeof(float), st);
        if (e != cudaSuccess) return cuda_fail(e, "rg_conv3d_direct_wgrad memset bias");
      }
    }
    return RG_OK;
  }
  long long NN = (long long)p.cin * taps;
  unsigned gx = (unsigned)((p.cout + BM - 1) / BM), gy = (unsigned)((NN + BN - 1) / BN);
  long long kps;
  const long long splits = direct_wgrad_splits(p, &kps);
  const long long ld = wcount_g + p.cout;
  dim3 grid(gx, gy, (unsigned)splits);
  for (int gi = 0; gi < g->groups; ++gi) {
    float* part = ws + (long long)gi * splits * ld;
    RG_DISPATCH_FLOAT(xd->dtype, TX, RG_DISPATCH_FLOAT(dyd->dtype, TG, {
      const TX* xg = (const TX*)x + (long long)gi * p.cin * p.x.sc;
      const TG* dyg = (const TG*)dy + (long long)gi * p.cout * p.y.sc;
      direct_wgrad_kernel<TX, TG><<<grid, 256, 0, st>>>(p, xg, dyg, part, ld, dbias ? 1 : 0, kps);
    }))
    RG_LAUNCH_CHECK("rg_conv3d_direct_wgrad");
    rc = launch_rows_sum(part, (int)splits, wcount_g, ld, dw + (long long)gi * wcount_g, nullptr, wcount_g, accumulate, 1.0, st);
    if (rc) return rc;
    if (dbias) {      // the bias gradient is always overwritten (accumulate concerns the weight-gradient bucket only)
      rc = launch_rows_sum(part + wcount_g, (int)splits, p.cout, ld, dbias + gi * p.cout, nullptr, p.cout, 0, 1.0, st);
      if (rc) return rc;
    }
  }
  return RG_OK;
}

}  // extern "C"
