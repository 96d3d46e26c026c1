// Vectorised fast paths for bf16 channels-last (NDHWC) tensors: every thread moves 8 channels
// (one 16-byte access) of one voxel.  Used by the HBM-bound kernels (norm apply / backward / stats,
// padding and its adjoint, activations, copies) whenever all operands qualify; the generic
// stride-walking kernels remain the fallback for every other layout / dtype.
#pragma once
#include "common.cuh"

namespace rg {

struct alignas(16) BF8 { __nv_bfloat162 v[4]; };

__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float* f) {
  BF8 b = *reinterpret_cast<const BF8*>(p);
#pragma unroll
  for (int i = 0; i < 4; ++i) { float2 t = __bfloat1622float2(b.v[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float* f) {
  BF8 b;
#pragma unroll
  for (int i = 0; i < 4; ++i) b.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  *reinterpret_cast<BF8*>(p) = b;
}

// Eligibility: bf16, channel stride 1, C % 8 == 0, every other stride a multiple of 8 elements,
// 16-byte aligned base.
inline bool vec_ok(const rg_tensor5* t, const void* p) {
  return t && p && t->dtype == RG_BF16 && t->sc == 1 && t->c % 8 == 0 && t->sn % 8 == 0 && t->sd % 8 == 0 &&
         t->sh % 8 == 0 && t->sw % 8 == 0 && (((uintptr_t)p) & 15) == 0;
}

// linear index over (n, d, h, w, c/8), channel groups fastest
struct VIdx { long long n, d, h, w; int c; };
__device__ __forceinline__ VIdx vunravel(long long i, const T5& t) {
  VIdx r;
  const int cg = (int)(t.c >> 3);
  if (i < 0x7fffffffLL) {      // 32-bit index arithmetic (a 64-bit div/mod costs ~5x more instructions)
    unsigned j = (unsigned)i, q;
    q = j / (unsigned)cg; r.c = (int)(j - q * (unsigned)cg) << 3; j = q;
    q = j / (unsigned)t.w; r.w = j - q * (unsigned)t.w; j = q;
    q = j / (unsigned)t.h; r.h = j - q * (unsigned)t.h; j = q;
    q = j / (unsigned)t.d; r.d = j - q * (unsigned)t.d; r.n = q;
    return r;
  }
  r.c = (int)(i % cg) << 3; i /= cg;
  r.w = i % t.w; i /= t.w;
  r.h = i % t.h; i /= t.h;
  r.d = i % t.d; r.n = i / t.d;
  return r;
}
__device__ __forceinline__ long long voff(const T5& t, long long n, long long d, long long h, long long w, int c) {
  return n * t.sn + d * t.sd + h * t.sh + w * t.sw + c;
}

}  // namespace rg
