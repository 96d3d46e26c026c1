// Trilinear resize with align_corners = True and its adjoint: the `sampling_bottleneck` of CycleModel
// (models/CycleModel.py:24-30: F.interpolate(x, scale_factor=s, mode="trilinear", align_corners=True) followed by
// F.interpolate(down, size=x.shape[-3:], ...)).  The operand is a generator OUTPUT (one channel, <= a few million
// voxels): latency-bound, so one thread per destination voxel with the eight-corner gather; the adjoint is a GATHER too
// (every source voxel sums the destination voxels that sampled it, in a fixed order) -- no atomics, bit-reproducible.
// Index arithmetic follows ATen's area_pixel_compute_source_index / compute_source_index_and_lambda in fp32.
#include "common.cuh"

namespace rg {

struct Axis { int in, out; float scale; };     // scale = (in - 1) / (out - 1), 0 when out == 1

__device__ __forceinline__ void src_index(const Axis& a, int j, int& i0, int& i1, float& l0, float& l1) {
  const float real = a.scale * (float)j;
  i0 = (int)real;
  if (i0 > a.in - 1) i0 = a.in - 1;
  i1 = i0 + (i0 < a.in - 1 ? 1 : 0);
  l1 = real - (float)i0;
  if (l1 < 0.f) l1 = 0.f;
  if (l1 > 1.f) l1 = 1.f;
  l0 = 1.f - l1;
}

template <typename TS, typename TD>
__global__ void resize_trilinear_fwd_kernel(T5 xd, const TS* __restrict__ x, T5 yd, TD* __restrict__ y, Axis ad, Axis ah, Axis aw) {
  const long long total = yd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, yd, false);
    int d0, d1, h0, h1, w0, w1;
    float ld0, ld1, lh0, lh1, lw0, lw1;
    src_index(ad, (int)p.d, d0, d1, ld0, ld1);
    src_index(ah, (int)p.h, h0, h1, lh0, lh1);
    src_index(aw, (int)p.w, w0, w1, lw0, lw1);
    auto at = [&](int d, int h, int w) { return ldf(x + xd.off(p.n, p.c, d, h, w)); };
    const float v = ld0 * (lh0 * (lw0 * at(d0, h0, w0) + lw1 * at(d0, h0, w1)) + lh1 * (lw0 * at(d0, h1, w0) + lw1 * at(d0, h1, w1))) +
                    ld1 * (lh0 * (lw0 * at(d1, h0, w0) + lw1 * at(d1, h0, w1)) + lh1 * (lw0 * at(d1, h1, w0) + lw1 * at(d1, h1, w1)));
    stf(y + yd.off(p.n, p.c, p.d, p.h, p.w), v);
  }
}

// weight with which destination index j samples source index i along one axis (0 when it does not)
__device__ __forceinline__ float axis_weight(const Axis& a, int j, int i) {
  int i0, i1; float l0, l1;
  src_index(a, j, i0, i1, l0, l1);
  float w = 0.f;
  if (i0 == i) w += l0;
  if (i1 == i) w += l1;      // i1 == i0 at the far edge: l1 == 0 there unless scale overshoots, and ATen adds both too
  return w;
}

// conservative destination range whose samples can touch source index i
__device__ __forceinline__ void dst_range(const Axis& a, int i, int& lo, int& hi) {
  if (a.scale <= 0.f) { lo = 0; hi = a.out - 1; return; }
  lo = (int)floorf((float)(i - 1) / a.scale) - 1;
  hi = (int)ceilf((float)(i + 1) / a.scale) + 1;
  if (lo < 0) lo = 0;
  if (hi > a.out - 1) hi = a.out - 1;
}

template <typename TS, typename TD>
__global__ void resize_trilinear_bwd_kernel(T5 gd, const TS* __restrict__ dy, T5 xd, TD* __restrict__ dx, Axis ad, Axis ah, Axis aw) {
  const long long total = xd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, xd, false);
    int dlo, dhi, hlo, hhi, wlo, whi;
    dst_range(ad, (int)p.d, dlo, dhi);
    dst_range(ah, (int)p.h, hlo, hhi);
    dst_range(aw, (int)p.w, wlo, whi);
    float acc = 0.f;
    for (int jd = dlo; jd <= dhi; ++jd) {
      const float wd = axis_weight(ad, jd, (int)p.d);
      if (wd == 0.f) continue;
      for (int jh = hlo; jh <= hhi; ++jh) {
        const float wh = axis_weight(ah, jh, (int)p.h);
        if (wh == 0.f) continue;
        float row = 0.f;
        for (int jw = wlo; jw <= whi; ++jw) {
          const float ww = axis_weight(aw, jw, (int)p.w);
          if (ww != 0.f) row += ww * ldf(dy + gd.off(p.n, p.c, jd, jh, jw));
        }
        acc += wd * wh * row;
      }
    }
    stf(dx + xd.off(p.n, p.c, p.d, p.h, p.w), acc);
  }
}

static Axis make_axis(long long in, long long out) {
  Axis a;
  a.in = (int)in; a.out = (int)out;
  a.scale = out > 1 ? (float)(in - 1) / (float)(out - 1) : 0.f;
  return a;
}

}  // namespace rg

using namespace rg;

extern "C" {

int rg_resize_trilinear_fwd(const rg_tensor5* xd, const void* x, const rg_tensor5* yd, void* y, void* stream) {
  RG_CHECK_ARG(xd && x && yd && y, "rg_resize_trilinear_fwd: null argument");
  RG_CHECK_ARG(xd->n == yd->n && xd->c == yd->c, "rg_resize_trilinear_fwd: N/C mismatch");
  RG_CHECK_ARG(xd->d >= 1 && xd->h >= 1 && xd->w >= 1 && yd->d >= 1 && yd->h >= 1 && yd->w >= 1, "rg_resize_trilinear_fwd: empty extent");
  const long long total = view(yd).numel();
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  RG_DISPATCH_FLOAT(xd->dtype, TS, RG_DISPATCH_FLOAT(yd->dtype, TD, {
    resize_trilinear_fwd_kernel<TS, TD><<<grid_for(total, 256), 256, 0, st>>>(view(xd), (const TS*)x, view(yd), (TD*)y,
        make_axis(xd->d, yd->d), make_axis(xd->h, yd->h), make_axis(xd->w, yd->w));
  }))
  RG_LAUNCH_CHECK("rg_resize_trilinear_fwd");
  return RG_OK;
}

int rg_resize_trilinear_bwd(const rg_tensor5* dyd, const void* dy, const rg_tensor5* dxd, void* dx, void* stream) {
  RG_CHECK_ARG(dyd && dy && dxd && dx, "rg_resize_trilinear_bwd: null argument");
  RG_CHECK_ARG(dxd->n == dyd->n && dxd->c == dyd->c, "rg_resize_trilinear_bwd: N/C mismatch");
  const long long total = view(dxd).numel();
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  RG_DISPATCH_FLOAT(dyd->dtype, TS, RG_DISPATCH_FLOAT(dxd->dtype, TD, {
    resize_trilinear_bwd_kernel<TS, TD><<<grid_for(total, 256), 256, 0, st>>>(view(dyd), (const TS*)dy, view(dxd), (TD*)dx,
        make_axis(dxd->d, dyd->d), make_axis(dxd->h, dyd->h), make_axis(dxd->w, dyd->w));
  }))
  RG_LAUNCH_CHECK("rg_resize_trilinear_bwd");
  return RG_OK;
}

}  // extern "C"
