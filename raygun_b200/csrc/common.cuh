// Shared helpers for libraygun_b200.so (sm_100a).  Not part of the public ABI.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/raygun_b200.h"

namespace rg {

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* where);

#define RG_CHECK_ARG(cond, ...)           \
  do {                                    \
    if (!(cond)) {                        \
      rg::set_error(__VA_ARGS__);         \
      return RG_E_BADARG;                 \
    }                                     \
  } while (0)

#define RG_UNSUPPORTED(cond, ...)         \
  do {                                    \
    if (cond) {                           \
      rg::set_error(__VA_ARGS__);         \
      return RG_E_UNSUPPORTED;            \
    }                                     \
  } while (0)

#define RG_LAUNCH_CHECK(where)                       \
  do {                                               \
    cudaError_t e__ = cudaGetLastError();            \
    if (e__ != cudaSuccess) return rg::cuda_fail(e__, where); \
  } while (0)

// Device-side view of an rg_tensor5 (by value in kernel params).
struct T5 {
  long long n, c, d, h, w;
  long long sn, sc, sd, sh, sw;
  __host__ __device__ long long numel() const { return n * c * d * h * w; }
  __device__ __forceinline__ long long off(long long in, long long ic, long long id, long long ih,
                                           long long iw) const {
    return in * sn + ic * sc + id * sd + ih * sh + iw * sw;
  }
};

inline T5 view(const rg_tensor5* t) {
  T5 v;
  v.n = t->n; v.c = t->c; v.d = t->d; v.h = t->h; v.w = t->w;
  v.sn = t->sn; v.sc = t->sc; v.sd = t->sd; v.sh = t->sh; v.sw = t->sw;
  return v;
}

inline bool same_shape(const rg_tensor5* a, const rg_tensor5* b) {
  return a->n == b->n && a->c == b->c && a->d == b->d && a->h == b->h && a->w == b->w;
}

// channels-innermost (NDHWC-like) storage?  Used to pick the loop order of elementwise kernels.
inline bool channels_inner(const rg_tensor5* t) { return t->sc == 1 && t->c > 1; }

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) {
  return __bfloat162float(*p);
}
template <> __device__ __forceinline__ float ldf<uint8_t>(const uint8_t* p) { return (float)*p; }

template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, float v) {
  *p = __float2bfloat16_rn(v);
}

__device__ __forceinline__ float act_apply(float v, int act, float slope) {
  switch (act) {
    case RG_ACT_RELU: return v > 0.f ? v : 0.f;
    case RG_ACT_LEAKY: return v > 0.f ? v : v * slope;
    case RG_ACT_TANH: return tanhf(v);
    case RG_ACT_SIGMOID: return 1.f / (1.f + expf(-v));
    default: return v;
  }
}
// derivative expressed through the OUTPUT y of the activation
__device__ __forceinline__ float act_grad_from_y(float y, int act, float slope) {
  switch (act) {
    case RG_ACT_RELU: return y > 0.f ? 1.f : 0.f;
    case RG_ACT_LEAKY: return y > 0.f ? 1.f : slope;
    case RG_ACT_TANH: return 1.f - y * y;
    case RG_ACT_SIGMOID: return y * (1.f - y);
    default: return 1.f;
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Decompose a linear index over (n,c,d,h,w) either W-innermost or C-innermost.
struct Idx5 { long long n, c, d, h, w; };
__device__ __forceinline__ Idx5 unravel(long long i, const T5& t, bool c_inner) {
  Idx5 r;
  if (c_inner) {
    r.c = i % t.c; i /= t.c;
    r.w = i % t.w; i /= t.w;
    r.h = i % t.h; i /= t.h;
    r.d = i % t.d; i /= t.d;
    r.n = i;
  } else {
    r.w = i % t.w; i /= t.w;
    r.h = i % t.h; i /= t.h;
    r.d = i % t.d; i /= t.d;
    r.c = i % t.c; i /= t.c;
    r.n = i;
  }
  return r;
}

inline int grid_for(long long work, int block, int max_blocks = 148 * 16) {
  long long b = (work + block - 1) / block;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return (int)b;
}

// Dispatch on an (input dtype, output dtype) pair restricted to {f32, bf16}.
#define RG_DISPATCH_FLOAT(DT, T, ...)                                   \
  if ((DT) == RG_F32) { using T = float; __VA_ARGS__ }                  \
  else if ((DT) == RG_BF16) { using T = __nv_bfloat16; __VA_ARGS__ }    \
  else { rg::set_error("unsupported dtype %d", (int)(DT)); return RG_E_UNSUPPORTED; }

}  // namespace rg
