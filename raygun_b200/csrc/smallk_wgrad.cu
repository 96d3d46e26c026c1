// Weight gradient of small-reduction convolutions: Cin * kD * kH * kW <= 64 (the 1 -> ngf 3^3 stem of a U-Net, UNet.py:61-69;
// the first, stride-2 convolution of NLayerDiscriminator, NLayerDiscriminator.py:137-145).
//   dw[co][k] = sum over (n, output voxel o) of dy[n, o, co] * x[n, ci(k), o * stride - pad + tap(k)],   k = (ci, kd, kh, kw)
// 64 x 27 outputs over millions of voxels: ~2 KFLOP per 128 bytes of dy, i.e. bound by the CUDA cores' FMA rate, not by the
// tensor cores' operand feed.  The tensor-core route (fold the taps into 64 channels with an im2col pass, then a (kD,1,1)
// 64 -> 64 halo weight gradient) moved a 64-channel intermediate through HBM and ran 2.4x the useful FLOPs:
// 0.5 ms per 8 x 64^3 stem; the generic direct kernel took 0.38 ms on the discriminator's 8 x 32^3 stem output.
//
// Block = 64 output channels x 2 voxel lanes (eight blocks per SM).  Per tile (a run of <= 64 output voxels along w) the
// block stages the dy rows (64 channels, 16-byte loads) and the gathered taps of every voxel (fp32 [voxel][k]) in shared
// memory; thread (co, lane) then walks the voxels lane, lane + 2, .. with one dy value and KP broadcast taps per voxel: KP independent FMAs.
// The kernel is bound by instruction issue (FMAs are ~75 % of the inner loop), so the tap decode (ci, kd, kh, kw of every k
// and its element offset) is done once per block, not per gathered value.
// Deterministic: tiles are assigned to blocks round-robin, every (block, lane) writes its own partial row, and
// rows_sum_kernel adds the rows in ascending order (reduce_rows.cuh).
#include <cuda_bf16.h>

#include "common.cuh"
#include "reduce_rows.cuh"

namespace rg {

constexpr int SK_V = 64;            // output voxels per tile
constexpr int SK_LANES = 2;         // voxel lanes per block
constexpr int SK_THREADS = 64 * SK_LANES;
constexpr int SK_BLOCKS = 148 * 8;  // 8 small blocks per SM; partial rows = SK_BLOCKS * SK_LANES = 2368

struct SkP {
  int cin, cout, kd, kh, kw, sd, sh, sw, pd, ph, pw, K;
  T5 x, y;
};

template <int KP, typename TX>
__global__ void __launch_bounds__(SK_THREADS) smallk_wgrad_kernel(SkP p, const TX* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                                                   float* __restrict__ part, long long ld, int want_bias) {
  __shared__ __align__(16) float xt[SK_V][KP];
  __shared__ __align__(16) __nv_bfloat16 dys[SK_V][64];
  const int tid = threadIdx.x, co = tid & 63, vl = tid >> 6;
  const int co0 = blockIdx.y * 64;
  const int OW = (int)p.y.w, OH = (int)p.y.h, OD = (int)p.y.d;
  const int wtiles = (OW + SK_V - 1) / SK_V;
  const long long total = p.y.n * OD * OH * (long long)wtiles;
  float acc[KP];
#pragma unroll
  for (int k = 0; k < KP; ++k) acc[k] = 0.f;
  float bsum = 0.f;
  for (int i = tid; i < SK_V * KP; i += SK_THREADS) (&xt[0][0])[i] = 0.f;      // columns k >= K stay zero
  __shared__ int4 ktap[64];            // (a, b, c, -) of reduction index k
  __shared__ long long koff[64];       // its element offset ci * sc + a * sd + b * sh + c * sw in x
  if (tid < p.K) {
    const int khw = p.kh * p.kw;
    const int ci = tid / (p.kd * khw);
    int t = tid - ci * p.kd * khw;
    const int a = t / khw; t -= a * khw;
    const int b = t / p.kw, c = t - b * p.kw;
    ktap[tid] = make_int4(a, b, c, 0);
    koff[tid] = ci * p.x.sc + a * p.x.sd + b * p.x.sh + c * p.x.sw;
  }
  const int XD = (int)p.x.d, XH = (int)p.x.h, XW = (int)p.x.w;
  for (long long tile = blockIdx.x; tile < total; tile += gridDim.x) {
    const int wt = (int)(tile % wtiles);
    long long row = tile / wtiles;
    const int oh = (int)(row % OH); row /= OH;
    const int od = (int)(row % OD);
    const long long n = row / OD;
    const int w0 = wt * SK_V, nv = min(SK_V, OW - w0);
    const int id0 = od * p.sd - p.pd, ih0 = oh * p.sh - p.ph;
    const TX* xrow = x + n * p.x.sn + (long long)id0 * p.x.sd + (long long)ih0 * p.x.sh;
    __syncthreads();                     // the previous tile's readers are done (and the fills above are visible)
    for (int i = tid; i < nv * 8; i += SK_THREADS) {
      const int v = i >> 3, c8 = (i & 7) << 3;
      *reinterpret_cast<uint4*>(&dys[v][c8]) = *reinterpret_cast<const uint4*>(dy + p.y.off(n, co0 + c8, od, oh, w0 + v));
    }
    for (int i = tid; i < p.K * SK_V; i += SK_THREADS) {
      const int v = i & (SK_V - 1), k = i >> 6;                  // SK_V == 64: k is uniform over two warps
      if (v >= nv) continue;
      const int4 tp = ktap[k];
      const int iw0 = (w0 + v) * p.sw - p.pw;
      const int id = id0 + tp.x, ih = ih0 + tp.y, iw = iw0 + tp.z;
      float val = 0.f;
      if ((unsigned)id < (unsigned)XD && (unsigned)ih < (unsigned)XH && (unsigned)iw < (unsigned)XW)
        val = ldf(xrow + koff[k] + (long long)iw0 * p.x.sw);
      xt[v][k] = val;
    }
    __syncthreads();
    for (int v = vl; v < nv; v += SK_LANES) {
      const float d = __bfloat162float(dys[v][co]);
      bsum += d;
      const float4* t4 = reinterpret_cast<const float4*>(&xt[v][0]);
#pragma unroll
      for (int k = 0; k < KP; k += 4) {
        const float4 tv = t4[k >> 2];
        acc[k] = fmaf(d, tv.x, acc[k]); acc[k + 1] = fmaf(d, tv.y, acc[k + 1]);
        acc[k + 2] = fmaf(d, tv.z, acc[k + 2]); acc[k + 3] = fmaf(d, tv.w, acc[k + 3]);
      }
    }
  }
  float* rowp = part + ((long long)blockIdx.x * SK_LANES + vl) * ld;
  float* dst = rowp + (long long)(co0 + co) * p.K;
#pragma unroll
  for (int k = 0; k < KP; ++k)
    if (k < p.K) dst[k] = acc[k];
  if (want_bias) rowp[(long long)p.cout * p.K + co0 + co] = bsum;
}

static int sk_fill(SkP& p, const rg_tensor5* xd, const rg_tensor5* dyd, const rg_conv_geom* g) {
  RG_UNSUPPORTED(g->groups != 1, "small-K wgrad: groups must be 1");
  RG_UNSUPPORTED(dyd->dtype != RG_BF16 || dyd->sc != 1 || dyd->c % 64 != 0, "small-K wgrad: dy must be bf16, channels innermost, Cout % 64 == 0");
  RG_UNSUPPORTED(dyd->sn % 8 || dyd->sd % 8 || dyd->sh % 8 || dyd->sw % 8, "small-K wgrad: dy strides must be multiples of 16 bytes");
  RG_UNSUPPORTED(xd->dtype != RG_BF16 && xd->dtype != RG_F32, "small-K wgrad: x must be bf16 or fp32");
  p.cin = (int)xd->c; p.cout = (int)dyd->c;
  p.kd = g->kernel[0]; p.kh = g->kernel[1]; p.kw = g->kernel[2];
  p.sd = g->stride[0]; p.sh = g->stride[1]; p.sw = g->stride[2];
  p.pd = g->pad[0]; p.ph = g->pad[1]; p.pw = g->pad[2];
  const long long K = (long long)p.cin * p.kd * p.kh * p.kw;
  RG_UNSUPPORTED(K < 1 || K > 64, "small-K wgrad: Cin * taps must be <= 64");
  p.K = (int)K;
  RG_CHECK_ARG(xd->n == dyd->n, "small-K wgrad: batch mismatch");
  const int64_t in_[3] = {xd->d, xd->h, xd->w}, out_[3] = {dyd->d, dyd->h, dyd->w};
  for (int i = 0; i < 3; ++i) {
    RG_CHECK_ARG(g->stride[i] >= 1 && g->kernel[i] >= 1 && g->pad[i] >= 0, "small-K wgrad: bad geometry");
    RG_CHECK_ARG(out_[i] == (in_[i] + 2 * g->pad[i] - g->kernel[i]) / g->stride[i] + 1, "small-K wgrad: extents inconsistent with geometry");
  }
  p.x = view(xd); p.y = view(dyd);
  return RG_OK;
}

}  // namespace rg

using namespace rg;

extern "C" {

int rg_smallk_wgrad_supported(const rg_tensor5* xd, const rg_tensor5* dyd, const rg_conv_geom* g) {
  if (!xd || !dyd || !g) return 0;
  SkP p;
  return sk_fill(p, xd, dyd, g) == RG_OK ? 1 : 0;
}

size_t rg_smallk_wgrad_workspace_bytes(int cout, int k) {
  return (size_t)SK_BLOCKS * SK_LANES * ((size_t)cout * k + cout) * sizeof(float);
}

int rg_smallk_wgrad(const rg_tensor5* xd, const void* x, const rg_tensor5* dyd, const void* dy, const rg_conv_geom* g, float* dw,
                    float* dbias, float* ws, int accumulate, void* stream) {
  RG_CHECK_ARG(xd && x && dyd && dy && g && dw && ws, "rg_smallk_wgrad: null argument");
  SkP p;
  int rc = sk_fill(p, xd, dyd, g);
  if (rc) return rc;
  RG_CHECK_ARG(((uintptr_t)dy & 15) == 0, "rg_smallk_wgrad: dy must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const long long nw = (long long)p.cout * p.K, ld = nw + p.cout;
  const long long tiles = p.y.n * p.y.d * p.y.h * ((p.y.w + SK_V - 1) / SK_V);
  if (tiles == 0) {
    if (!accumulate) {
      cudaError_t e = cudaMemsetAsync(dw, 0, nw * sizeof(float), st);
      if (e != cudaSuccess) return cuda_fail(e, "rg_smallk_wgrad memset");
    }
    if (dbias) {
      cudaError_t e = cudaMemsetAsync(dbias, 0, p.cout * sizeof(float), st);
      if (e != cudaSuccess) return cuda_fail(e, "rg_smallk_wgrad memset bias");
    }
    return RG_OK;
  }
  const int blocks = (int)(tiles < SK_BLOCKS ? tiles : SK_BLOCKS);
  dim3 grid((unsigned)blocks, (unsigned)(p.cout / 64));
  const int wb = dbias ? 1 : 0;
  if (p.K <= 28) {
    if (xd->dtype == RG_BF16) smallk_wgrad_kernel<28, __nv_bfloat16><<<grid, SK_THREADS, 0, st>>>(p, (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, ws, ld, wb);
    else smallk_wgrad_kernel<28, float><<<grid, SK_THREADS, 0, st>>>(p, (const float*)x, (const __nv_bfloat16*)dy, ws, ld, wb);
  } else if (p.K <= 32) {
    if (xd->dtype == RG_BF16) smallk_wgrad_kernel<32, __nv_bfloat16><<<grid, SK_THREADS, 0, st>>>(p, (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, ws, ld, wb);
    else smallk_wgrad_kernel<32, float><<<grid, SK_THREADS, 0, st>>>(p, (const float*)x, (const __nv_bfloat16*)dy, ws, ld, wb);
  } else {
    if (xd->dtype == RG_BF16) smallk_wgrad_kernel<64, __nv_bfloat16><<<grid, SK_THREADS, 0, st>>>(p, (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, ws, ld, wb);
    else smallk_wgrad_kernel<64, float><<<grid, SK_THREADS, 0, st>>>(p, (const float*)x, (const __nv_bfloat16*)dy, ws, ld, wb);
  }
  RG_LAUNCH_CHECK("rg_smallk_wgrad");
  const int rows = blocks * SK_LANES;
  rc = launch_rows_sum(ws, rows, nw, ld, dw, nullptr, nw, accumulate, 1.0, st);
  if (rc || !dbias) return rc;
  return launch_rows_sum(ws + nw, rows, p.cout, ld, dbias, nullptr, p.cout, 0, 1.0, st);      // bias: always overwritten
}

}  // extern "C"
