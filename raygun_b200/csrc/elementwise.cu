// HBM-bound kernels of the hot path: strided copy/cast/crop, padding (+adjoint), activations,
// max-pool, nearest up-sampling, u8 normalise / denormalise, fused Adam.
// All kernels are grid-stride loops sized to a multiple of the 148 SMs; the loop order follows the
// innermost storage dimension of the destination so that stores (and usually loads) coalesce.
#include "common.cuh"
#include "vec_ndhwc.cuh"

namespace rg {

// ------------------------------------------------------------------------------------------
// copy5: dst (+)= src over dst's logical shape
// ------------------------------------------------------------------------------------------
template <typename TS, typename TD>
__global__ void copy5_kernel(T5 dd, TD* __restrict__ dst, T5 sd, const TS* __restrict__ src, bool c_inner,
                             bool accumulate) {
  long long total = dd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, dd, c_inner);
    float v = ldf(src + sd.off(p.n, p.c, p.d, p.h, p.w));
    TD* q = dst + dd.off(p.n, p.c, p.d, p.h, p.w);
    if (accumulate) v += ldf(q);
    stf(q, v);
  }
}

// ------------------------------------------------------------------------------------------
// padding
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ long long pad_src(long long o, long long lo, long long size, int mode, bool& valid) {
  long long i = o - lo;
  valid = true;
  if (i >= 0 && i < size) return i;
  if (mode == RG_PAD_REFLECT) {
    if (i < 0) i = -i;
    if (i >= size) i = 2 * (size - 1) - i;
    return i;
  }
  if (mode == RG_PAD_REPLICATE) return i < 0 ? 0 : size - 1;
  valid = false;
  return 0;
}

template <typename TS, typename TD>
__global__ void pad_fwd_kernel(T5 xd, const TS* __restrict__ x, T5 yd, TD* __restrict__ y, int lo_d, int lo_h,
                               int lo_w, int mode, bool c_inner) {
  long long total = yd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, yd, c_inner);
    bool vd, vh, vw;
    long long id = pad_src(p.d, lo_d, xd.d, mode, vd);
    long long ih = pad_src(p.h, lo_h, xd.h, mode, vh);
    long long iw = pad_src(p.w, lo_w, xd.w, mode, vw);
    float v = (vd && vh && vw) ? ldf(x + xd.off(p.n, p.c, id, ih, iw)) : 0.f;
    stf(y + yd.off(p.n, p.c, p.d, p.h, p.w), v);
  }
}

// Adjoint of pad: gather formulation (deterministic).  For every input voxel, sum the up-to-27
// padded-output positions that read it.  Per dimension an input index i is read by o = i+lo and, for
// reflect, by the mirror images about both faces; for replicate, border voxels absorb the whole shell.
__device__ __forceinline__ int pad_preimages(long long i, long long lo, long long size, long long osize, int mode,
                                             long long* out, int* rep_lo, int* rep_hi) {
  int cnt = 0;
  *rep_lo = 0; *rep_hi = 0;
  long long o = i + lo;
  if (o >= 0 && o < osize) out[cnt++] = o;
  if (mode == RG_PAD_REFLECT) {
    long long o1 = lo - i;                    // mirror about the low face (i>0)
    if (i > 0 && o1 >= 0 && o1 < lo) out[cnt++] = o1;
    long long o2 = lo + 2 * (size - 1) - i;   // mirror about the high face (i<size-1)
    if (i < size - 1 && o2 >= lo + size && o2 < osize) out[cnt++] = o2;
  } else if (mode == RG_PAD_REPLICATE) {
    if (i == 0) *rep_lo = (int)lo;
    if (i == size - 1) *rep_hi = (int)(osize - lo - size);
  }
  return cnt;
}

template <typename TS, typename TD>
__global__ void pad_bwd_kernel(T5 dyd, const TS* __restrict__ dy, T5 dxd, TD* __restrict__ dx, int lo_d, int lo_h,
                               int lo_w, int mode, bool c_inner) {
  long long total = dxd.numel();
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(idx, dxd, c_inner);
    // enumerate pre-images per dimension (explicit lists; replicate shells expanded as ranges)
    long long od[3 + 0], oh[3], ow[3];
    int rl, rh;
    long long ld[64], lh[64], lw[64];
    int nd = 0, nh = 0, nw = 0;
    int c = pad_preimages(p.d, lo_d, dxd.d, dyd.d, mode, od, &rl, &rh);
    for (int k = 0; k < c; ++k) ld[nd++] = od[k];
    for (int k = 0; k < rl && nd < 64; ++k) ld[nd++] = k;
    for (int k = 0; k < rh && nd < 64; ++k) ld[nd++] = lo_d + dxd.d + k;
    c = pad_preimages(p.h, lo_h, dxd.h, dyd.h, mode, oh, &rl, &rh);
    for (int k = 0; k < c; ++k) lh[nh++] = oh[k];
    for (int k = 0; k < rl && nh < 64; ++k) lh[nh++] = k;
    for (int k = 0; k < rh && nh < 64; ++k) lh[nh++] = lo_h + dxd.h + k;
    c = pad_preimages(p.w, lo_w, dxd.w, dyd.w, mode, ow, &rl, &rh);
    for (int k = 0; k < c; ++k) lw[nw++] = ow[k];
    for (int k = 0; k < rl && nw < 64; ++k) lw[nw++] = k;
    for (int k = 0; k < rh && nw < 64; ++k) lw[nw++] = lo_w + dxd.w + k;
    float acc = 0.f;
    for (int a = 0; a < nd; ++a)
      for (int b = 0; b < nh; ++b)
        for (int e = 0; e < nw; ++e) acc += ldf(dy + dyd.off(p.n, p.c, ld[a], lh[b], lw[e]));
    TD* q = dx + dxd.off(p.n, p.c, p.d, p.h, p.w);
    stf(q, acc);
  }
}

// ------------------------------------------------------------------------------------------
// activations
// ------------------------------------------------------------------------------------------
template <typename TS, typename TD>
__global__ void act_fwd_kernel(T5 xd, const TS* __restrict__ x, T5 yd, TD* __restrict__ y, int act, float slope,
                               bool c_inner) {
  long long total = yd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, yd, c_inner);
    float v = ldf(x + xd.off(p.n, p.c, p.d, p.h, p.w));
    stf(y + yd.off(p.n, p.c, p.d, p.h, p.w), act_apply(v, act, slope));
  }
}

template <typename TY, typename TG, typename TD>
__global__ void act_bwd_kernel(T5 yd, const TY* __restrict__ y, T5 gd, const TG* __restrict__ dy, T5 dxd,
                               TD* __restrict__ dx, int act, float slope, bool c_inner) {
  long long total = dxd.numel();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    Idx5 p = unravel(i, dxd, c_inner);
    float yy = ldf(y + yd.off(p.n, p.c, p.d, p.h, p.w));
    float g = ldf(dy + gd.off(p.n, p.c, p.d, p.h, p.w));
    stf(dx + dxd.off(p.n, p.c, p.d, p.h, p.w), g * act_grad_from_y(yy, act, slope));
  }
}


// ------------------------------------------------------------------------------------------
// vectorised NDHWC bf16 fast paths (8 channels / thread)
// ------------------------------------------------------------------------------------------
__global__ void vec_copy_kernel(T5 dd, __nv_bfloat16* __restrict__ dst, T5 sd, const __nv_bfloat16* __restrict__ src,
                                bool accumulate) {
  const long long total = dd.numel() >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    VIdx p = vunravel(i, dd);
    float v[8];
    ld8(src + voff(sd, p.n, p.d, p.h, p.w, p.c), v);
    __nv_bfloat16* q = dst + voff(dd, p.n, p.d, p.h, p.w, p.c);
    if (accumulate) {
      float o[8];
      ld8(q, o);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] += o[k];
    }
    st8(q, v);
  }
}

__global__ void vec_pad_fwd_kernel(T5 xd, const __nv_bfloat16* __restrict__ x, T5 yd, __nv_bfloat16* __restrict__ y,
                                   int lo_d, int lo_h, int lo_w, int mode) {
  const long long total = yd.numel() >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    VIdx p = vunravel(i, yd);
    bool vd, vh, vw;
    const long long id = pad_src(p.d, lo_d, xd.d, mode, vd);
    const long long ih = pad_src(p.h, lo_h, xd.h, mode, vh);
    const long long iw = pad_src(p.w, lo_w, xd.w, mode, vw);
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (vd && vh && vw) ld8(x + voff(xd, p.n, id, ih, iw, p.c), v);
    st8(y + voff(yd, p.n, p.d, p.h, p.w, p.c), v);
  }
}

// k-th padded-output index that reads input index i (k = 0 is always the interior image i + lo).
__device__ __forceinline__ int pad_npre(long long i, long long lo, long long hi, long long size, int mode) {
  if (mode == RG_PAD_REFLECT) return 1 + (i >= 1 && i <= lo ? 1 : 0) + (i <= size - 2 && i >= size - 1 - hi ? 1 : 0);
  if (mode == RG_PAD_REPLICATE) return 1 + (i == 0 ? (int)lo : 0) + (i == size - 1 ? (int)hi : 0);
  return 1;
}
__device__ __forceinline__ long long pad_pre(long long i, int k, long long lo, long long hi, long long size, int mode) {
  if (k == 0) return i + lo;
  if (mode == RG_PAD_REFLECT) {
    const bool low = (i >= 1 && i <= lo);
    if (low && k == 1) return lo - i;
    return lo + 2 * (size - 1) - i;
  }
  // replicate: border voxels absorb the shell
  if (i == 0 && k <= lo) return k - 1;
  const int kk = k - 1 - (i == 0 ? (int)lo : 0);
  return lo + size + kk;
}

__global__ void vec_pad_bwd_kernel(T5 gd, const __nv_bfloat16* __restrict__ dy, T5 dxd, __nv_bfloat16* __restrict__ dx,
                                   int lo_d, int lo_h, int lo_w, int mode) {
  const long long total = dxd.numel() >> 3;
  const long long hi_d = gd.d - dxd.d - lo_d, hi_h = gd.h - dxd.h - lo_h, hi_w = gd.w - dxd.w - lo_w;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    VIdx p = vunravel(i, dxd);
    const int nd = pad_npre(p.d, lo_d, hi_d, dxd.d, mode), nh = pad_npre(p.h, lo_h, hi_h, dxd.h, mode),
              nw = pad_npre(p.w, lo_w, hi_w, dxd.w, mode);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    __nv_bfloat16* q = dx + voff(dxd, p.n, p.d, p.h, p.w, p.c);
    for (int a = 0; a < nd; ++a) {
      const long long od = pad_pre(p.d, a, lo_d, hi_d, dxd.d, mode);
      for (int b = 0; b < nh; ++b) {
        const long long oh = pad_pre(p.h, b, lo_h, hi_h, dxd.h, mode);
        for (int c = 0; c < nw; ++c) {
          const long long ow = pad_pre(p.w, c, lo_w, hi_w, dxd.w, mode);
          float v[8];
          ld8(dy + voff(gd, p.n, od, oh, ow, p.c), v);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += v[k];
        }
      }
    }
    st8(q, acc);
  }
}

// Row-based adjoint of the padding for NDHWC bf16 (the reflect pads of every 'same' ConvPass / ResnetBlock in a
// training step): one block per dx row (n, d, h).  The (d, h) pre-image rows are found once per block; a thread owns
// (w, 8 channels) items of the row and adds the <= 3 w pre-images of each source row.  No per-element index division.
__global__ void __launch_bounds__(256) row_pad_bwd_kernel(T5 gd, const __nv_bfloat16* __restrict__ dy, T5 dxd,
                                                          __nv_bfloat16* __restrict__ dx, int lo_d, int lo_h, int lo_w, int mode) {
  const long long hi_d = gd.d - dxd.d - lo_d, hi_h = gd.h - dxd.h - lo_h, hi_w = gd.w - dxd.w - lo_w;
  const int W = (int)dxd.w, H = (int)dxd.h, D = (int)dxd.d, cg = (int)(dxd.c >> 3);
  const long long rows = dxd.n * D * H;
  const int items = W * cg;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int h = (int)(row % H);
    const long long t = row / H;
    const int d = (int)(t % D);
    const long long n = t / D;
    const int nd = pad_npre(d, lo_d, hi_d, D, mode), nh = pad_npre(h, lo_h, hi_h, H, mode);
    __nv_bfloat16* xrow = dx + n * dxd.sn + d * dxd.sd + h * dxd.sh;
    for (int j = threadIdx.x; j < items; j += blockDim.x) {
      const int w = j / cg, c = (j - w * cg) << 3;
      const int nw = pad_npre(w, lo_w, hi_w, W, mode);
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int a = 0; a < nd; ++a) {
        const long long od = pad_pre(d, a, lo_d, hi_d, D, mode);
        for (int b = 0; b < nh; ++b) {
          const long long oh = pad_pre(h, b, lo_h, hi_h, H, mode);
          const __nv_bfloat16* srow = dy + n * gd.sn + od * gd.sd + oh * gd.sh + c;
          for (int e = 0; e < nw; ++e) {
            const long long ow = pad_pre(w, e, lo_w, hi_w, W, mode);
            float v[8];
            ld8(srow + ow * gd.sw, v);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] += v[k];
          }
        }
      }
      st8(xrow + (long long)w * dxd.sw + c, acc);
    }
  }
}

// In-place adjoint of the padding: the shell of a padded NDHWC bf16 gradient is folded into the interior voxels whose
// images it holds, and the INTERIOR VIEW of the same buffer then is the pad's input gradient -- O(surface) bytes instead
// of the read + write of the whole volume an explicit adjoint copy costs (the backward twin of border_fill_kernel).
// Only voxels with more than one pre-image are rewritten; every other pre-image of such a voxel lies in the shell,
// which is never written, so the gather is race-free and deterministic.  One block per interior row (n, d, h).
__global__ void __launch_bounds__(256, 4) pad_fold_kernel(T5 gd, __nv_bfloat16* __restrict__ g, int lo_d, int lo_h, int lo_w,
                                                       int hi_d, int hi_h, int hi_w, int mode) {
  const int D = (int)gd.d - lo_d - hi_d, H = (int)gd.h - lo_h - hi_h, W = (int)gd.w - lo_w - hi_w, cg = (int)(gd.c >> 3);
  const long long rows = gd.n * D * H;
  const int ncand = (lo_w + 1) + (hi_w + 1);            // w candidates of a row with a single (d, h) pre-image
  const bool edge_only_ok = ncand < W;
  // A block owns 8 consecutive interior rows.  Most rows only own their few edge voxels (ncand * cg 16-byte items): one WARP
  // takes such a row.  A row with more than one (d, h) pre-image (d or h next to a face: whole planes of them) owns all W * cg
  // items: the whole BLOCK splits it -- with one warp per row these rows took 16x longer than their neighbours and set the
  // kernel's duration (0.65-0.9 TB/s under ncu).
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
  auto fold_item = [&](__nv_bfloat16* base, int d, int h, int nd, int nh, bool full, int j) {
    const int wi = j / cg, c = (j - wi * cg) << 3;
    const int w = full ? wi : (wi <= lo_w ? wi : W - 1 - hi_w + (wi - lo_w - 1));
    const int nw = pad_npre(w, lo_w, hi_w, W, mode);
    if (nd * nh * nw == 1) return;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int a = 0; a < nd; ++a) {
      const long long od = pad_pre(d, a, lo_d, hi_d, D, mode);
      for (int b = 0; b < nh; ++b) {
        const long long oh = pad_pre(h, b, lo_h, hi_h, H, mode);
        const __nv_bfloat16* srow = base + od * gd.sd + oh * gd.sh + c;
        for (int e = 0; e < nw; ++e) {
          const long long ow = pad_pre(w, e, lo_w, hi_w, W, mode);
          float v[8];
          ld8(srow + ow * gd.sw, v);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += v[k];
        }
      }
    }
    st8(base + (long long)(d + lo_d) * gd.sd + (long long)(h + lo_h) * gd.sh + (long long)(w + lo_w) * gd.sw + c, acc);
  };
  for (long long row0 = (long long)blockIdx.x * warps; row0 < rows; row0 += (long long)gridDim.x * warps) {
    // one 64-bit decode per block; the 8 rows step (h, d, n) incrementally
    int h = (int)(row0 % H);
    const long long t0 = row0 / H;
    int d = (int)(t0 % D);
    long long n = t0 / D;
    const int nrows = (int)min((long long)warps, rows - row0);
    for (int r = 0; r < nrows; ++r) {
      const int nd = pad_npre(d, lo_d, hi_d, D, mode), nh = pad_npre(h, lo_h, hi_h, H, mode);
      const bool full = (nd * nh > 1) || !edge_only_ok;
      if (full || warp == r) {
        __nv_bfloat16* base = g + n * gd.sn;
        if (full) {
          for (int j = threadIdx.x; j < W * cg; j += blockDim.x) fold_item(base, d, h, nd, nh, true, j);
        } else {
          for (int j = lane; j < ncand * cg; j += 32) fold_item(base, d, h, nd, nh, false, j);
        }
      }
      if (++h == H) { h = 0; if (++d == D) { d = 0; ++n; } }
    }
  }
}

__global__ void vec_act_fwd_kernel(T5 xd, const __nv_bfloat16* __restrict__ x, T5 yd, __nv_bfloat16* __restrict__ y,
                                   int act, float slope) {
  const long long total = yd.numel() >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    VIdx p = vunravel(i, yd);
    float v[8];
    ld8(x + voff(xd, p.n, p.d, p.h, p.w, p.c), v);
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = act_apply(v[k], act, slope);
    st8(y + voff(yd, p.n, p.d, p.h, p.w, p.c), v);
  }
}

__global__ void vec_act_bwd_kernel(T5 yd, const __nv_bfloat16* __restrict__ y, T5 gd, const __nv_bfloat16* __restrict__ dy,
                                   T5 dxd, __nv_bfloat16* __restrict__ dx, int act, float slope) {
  const long long total = dxd.numel() >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    VIdx p = vunravel(i, dxd);
    float yy[8], g[8];
    ld8(y + voff(yd, p.n, p.d, p.h, p.w, p.c), yy);
    ld8(dy + voff(gd, p.n, p.d, p.h, p.w, p.c), g);
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] *= act_grad_from_y(yy[k], act, slope);
    st8(dx + voff(dxd, p.n, p.d, p.h, p.w, p.c), g);
  }
}


// Folded stem input: y[n, d, h, w, (ci*kh + a)*kw + b] = x[n, ci, d, h + a - ph, w + b - pw] (zero outside / beyond
// Cin*kh*kw).  One block per output row (n, d, h): no per-element index division; the 8 channels of a thread walk
// the (ci, a, b) odometer incrementally.
template <typename TS>
__global__ void __launch_bounds__(256) im2col_hw_kernel(T5 xd, const TS* __restrict__ x, int kh, int kw, int ph, int pw,
                                                        T5 yd, __nv_bfloat16* __restrict__ y) {
  const int cg = (int)(yd.c >> 3);
  const int cvalid = (int)xd.c * kh * kw;
  const int W = (int)yd.w, H = (int)yd.h, D = (int)yd.d;
  const long long rows = yd.n * D * H;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int h = (int)(row % H);
    const long long t = row / H;
    const int d = (int)(t % D), n = (int)(t / D);
    const TS* xrow = x + n * xd.sn + d * xd.sd;
    __nv_bfloat16* yrow = y + n * yd.sn + d * yd.sd + h * yd.sh;
    for (int j = threadIdx.x; j < W * cg; j += blockDim.x) {
      const int w = j / cg, c0 = (j - w * cg) << 3;
      if (c0 >= cvalid) {             // padding channels of the folded layout: plain zero store
        *reinterpret_cast<uint4*>(yrow + (long long)w * yd.sw + c0) = make_uint4(0u, 0u, 0u, 0u);
        continue;
      }
      int b = c0 % kw, tt = c0 / kw, a = tt % kh, ci = tt / kh;
      float v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float val = 0.f;
        const int ih = h + a - ph, iw = w + b - pw;
        if (c0 + k < cvalid && ih >= 0 && ih < xd.h && iw >= 0 && iw < xd.w) val = ldf(xrow + ci * xd.sc + ih * xd.sh + iw * xd.sw);
        v[k] = val;
        if (++b == kw) { b = 0; if (++a == kh) { a = 0; ++ci; } }
      }
      st8(yrow + (long long)w * yd.sw + c0, v);
    }
  }
}

// In-place border fill of a padded NDHWC bf16 buffer whose INTERIOR a convolution epilogue has already written:
// mirrors (reflect) or clamps (replicate) or zeroes the `m`-voxel shell, touching only O(surface) bytes instead of
// the read + write of the whole volume an explicit pad copy costs.  One block per padded row (n, d, h).
__global__ void __launch_bounds__(256) border_fill_kernel(T5 t, __nv_bfloat16* __restrict__ p, int md, int mh, int mw, int mode) {
  const int cg = (int)(t.c >> 3), W = (int)t.w, H = (int)t.h, D = (int)t.d;
  const int iD = D - 2 * md, iH = H - 2 * mh, iW = W - 2 * mw;          // interior extents
  const long long rows = t.n * D * H;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const int h = (int)(row % H);
    const long long q = row / H;
    const int d = (int)(q % D), n = (int)(q / D);
    bool vd, vh, vw;
    const int sd_ = (int)pad_src(d, md, iD, mode, vd) + md, sh_ = (int)pad_src(h, mh, iH, mode, vh) + mh;
    const bool row_interior = (sd_ == d && sh_ == h);
    __nv_bfloat16* dst = p + n * t.sn + d * t.sd + h * t.sh;
    const __nv_bfloat16* src = p + n * t.sn + sd_ * t.sd + sh_ * t.sh;
    // interior rows only need their 2*mw edge voxels; border rows need the full width
    const int nw = row_interior ? 2 * mw : W;
    for (int j = threadIdx.x; j < nw * cg; j += blockDim.x) {
      int w = j / cg;
      const int c = (j - w * cg) << 3;
      if (row_interior && w >= mw) w = W - 2 * mw + w;                 // right edge
      const int sw_ = (int)pad_src(w, mw, iW, mode, vw) + mw;
      float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (vd && vh && vw) ld8(src + (long long)sw_ * t.sw + c, v);
      st8(dst + (long long)w * t.sw + c, v);
    }
  }
}

}  // namespace rg

using namespace rg;

extern "C" {

int rg_copy5(const rg_tensor5* dd, void* dst, const rg_tensor5* sd, const void* src, int accumulate,
             void* stream) {
  RG_CHECK_ARG(dd && sd && dst && src, "rg_copy5: null argument");
  RG_CHECK_ARG(same_shape(dd, sd), "rg_copy5: shape mismatch");
  long long total = view(dd).numel();
  if (total == 0) return RG_OK;
  bool ci = channels_inner(dd);
  cudaStream_t st = (cudaStream_t)stream;
  if (vec_ok(dd, dst) && vec_ok(sd, src)) {
    vec_copy_kernel<<<grid_for(total >> 3, 256), 256, 0, st>>>(view(dd), (__nv_bfloat16*)dst, view(sd), (const __nv_bfloat16*)src, accumulate != 0);
    RG_LAUNCH_CHECK("rg_copy5(vec)");
    return RG_OK;
  }
  int grid = grid_for(total, 256);
  RG_DISPATCH_FLOAT(sd->dtype, TS, RG_DISPATCH_FLOAT(dd->dtype, TD, {
    copy5_kernel<TS, TD><<<grid, 256, 0, st>>>(view(dd), (TD*)dst, view(sd), (const TS*)src, ci, accumulate != 0);
  }))
  RG_LAUNCH_CHECK("rg_copy5");
  return RG_OK;
}

int rg_pad3d_fwd(const rg_tensor5* xd, const void* x, const rg_tensor5* yd, void* y, const int32_t lo[3],
                 int mode, void* stream) {
  RG_CHECK_ARG(xd && yd && x && y && lo, "rg_pad3d_fwd: null argument");
  RG_CHECK_ARG(xd->n == yd->n && xd->c == yd->c, "rg_pad3d_fwd: N/C mismatch");
  RG_CHECK_ARG(mode >= RG_PAD_ZEROS && mode <= RG_PAD_REPLICATE, "rg_pad3d_fwd: bad mode %d", mode);
  if (mode == RG_PAD_REFLECT) {
    // nn.ReflectionPad3d requires pad < size in every dimension
    RG_CHECK_ARG(lo[0] < xd->d && lo[1] < xd->h && lo[2] < xd->w &&
                 yd->d - xd->d - lo[0] < xd->d && yd->h - xd->h - lo[1] < xd->h && yd->w - xd->w - lo[2] < xd->w,
                 "rg_pad3d_fwd: reflect padding must be smaller than the input");
  }
  long long total = view(yd).numel();
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (vec_ok(xd, x) && vec_ok(yd, y)) {
    vec_pad_fwd_kernel<<<grid_for(total >> 3, 256), 256, 0, st>>>(view(xd), (const __nv_bfloat16*)x, view(yd), (__nv_bfloat16*)y, lo[0], lo[1], lo[2], mode);
    RG_LAUNCH_CHECK("rg_pad3d_fwd(vec)");
    return RG_OK;
  }
  int grid = grid_for(total, 256);
  bool ci = channels_inner(yd);
  RG_DISPATCH_FLOAT(xd->dtype, TS, RG_DISPATCH_FLOAT(yd->dtype, TD, {
    pad_fwd_kernel<TS, TD><<<grid, 256, 0, st>>>(view(xd), (const TS*)x, view(yd), (TD*)y, lo[0], lo[1], lo[2], mode, ci);
  }))
  RG_LAUNCH_CHECK("rg_pad3d_fwd");
  return RG_OK;
}

int rg_pad3d_bwd(const rg_tensor5* dyd, const void* dy, const rg_tensor5* dxd, void* dx, const int32_t lo[3],
                 int mode, void* stream) {
  RG_CHECK_ARG(dyd && dxd && dy && dx && lo, "rg_pad3d_bwd: null argument");
  RG_CHECK_ARG(dxd->n == dyd->n && dxd->c == dyd->c, "rg_pad3d_bwd: N/C mismatch");
  RG_CHECK_ARG(lo[0] < 30 && lo[1] < 30 && lo[2] < 30, "rg_pad3d_bwd: pad too large");
  long long total = view(dxd).numel();
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (vec_ok(dyd, dy) && vec_ok(dxd, dx) && dxd->w * (dxd->c >> 3) >= 64 && dxd->w * (dxd->c >> 3) < (1LL << 30)) {
    row_pad_bwd_kernel<<<grid_for(dxd->n * dxd->d * dxd->h, 1, 148 * 16), 256, 0, st>>>(view(dyd), (const __nv_bfloat16*)dy, view(dxd),
                                                                                     (__nv_bfloat16*)dx, lo[0], lo[1], lo[2], mode);
    RG_LAUNCH_CHECK("rg_pad3d_bwd(row)");
    return RG_OK;
  }
  if (vec_ok(dyd, dy) && vec_ok(dxd, dx)) {
    vec_pad_bwd_kernel<<<grid_for(total >> 3, 256), 256, 0, st>>>(view(dyd), (const __nv_bfloat16*)dy, view(dxd), (__nv_bfloat16*)dx, lo[0], lo[1], lo[2], mode);
    RG_LAUNCH_CHECK("rg_pad3d_bwd(vec)");
    return RG_OK;
  }
  int grid = grid_for(total, 128);
  bool ci = channels_inner(dxd);
  RG_DISPATCH_FLOAT(dyd->dtype, TS, RG_DISPATCH_FLOAT(dxd->dtype, TD, {
    pad_bwd_kernel<TS, TD><<<grid, 128, 0, st>>>(view(dyd), (const TS*)dy, view(dxd), (TD*)dx, lo[0], lo[1], lo[2], mode, ci);
  }))
  RG_LAUNCH_CHECK("rg_pad3d_bwd");
  return RG_OK;
}

int rg_act_fwd(const rg_tensor5* xd, const void* x, int act, float slope, const rg_tensor5* yd, void* y,
               void* stream) {
  RG_CHECK_ARG(xd && yd && x && y, "rg_act_fwd: null argument");
  RG_CHECK_ARG(same_shape(xd, yd), "rg_act_fwd: shape mismatch");
  long long total = view(yd).numel();
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (vec_ok(xd, x) && vec_ok(yd, y)) {
    vec_act_fwd_kernel<<<grid_for(total >> 3, 256), 256, 0, st>>>(view(xd), (const __nv_bfloat16*)x, view(yd), (__nv_bfloat16*)y, act, slope);
    RG_LAUNCH_CHECK("rg_act_fwd(vec)");
    return RG_OK;
  }
  int grid = grid_for(total, 256);
  bool ci = channels_inner(yd);
  RG_DISPATCH_FLOAT(xd->dtype, TS, RG_DISPATCH_FLOAT(yd->dtype, TD, {
    act_fwd_kernel<TS, TD><<<grid, 256, 0, st>>>(view(xd), (const TS*)x, view(yd), (TD*)y, act, slope, ci);
  }))
  RG_LAUNCH_CHECK("rg_act_fwd");
  return RG_OK;
}

int rg_act_bwd(const rg_tensor5* yd, const void* y, const rg_tensor5* dyd, const void* dy, int act, float slope,
               const rg_tensor5* dxd, void* dx, void* stream) {
  RG_CHECK_ARG(yd && dyd && dxd && y && dy && dx, "rg_act_bwd: null argument");
  RG_CHECK_ARG(same_shape(yd, dyd) && same_shape(yd, dxd), "rg_act_bwd: shape mismatch");
  RG_CHECK_ARG(yd->dtype == dyd->dtype, "rg_act_bwd: y / dy dtype mismatch");
  long long total = view(dxd).numel();
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (vec_ok(yd, y) && vec_ok(dyd, dy) && vec_ok(dxd, dx)) {
    vec_act_bwd_kernel<<<grid_for(total >> 3, 256), 256, 0, st>>>(view(yd), (const __nv_bfloat16*)y, view(dyd), (const __nv_bfloat16*)dy, view(dxd), (__nv_bfloat16*)dx, act, slope);
    RG_LAUNCH_CHECK("rg_act_bwd(vec)");
    return RG_OK;
  }
  int grid = grid_for(total, 256);
  bool ci = channels_inner(dxd);
  RG_DISPATCH_FLOAT(yd->dtype, TS, RG_DISPATCH_FLOAT(dxd->dtype, TD, {
    act_bwd_kernel<TS, TS, TD><<<grid, 256, 0, st>>>(view(yd), (const TS*)y, view(dyd), (const TS*)dy, view(dxd), (TD*)dx, act, slope, ci);
  }))
  RG_LAUNCH_CHECK("rg_act_bwd");
  return RG_OK;
}

int rg_im2col_hw(const rg_tensor5* xd, const void* x, int kh, int kw, int ph, int pw, const rg_tensor5* yd, void* y,
                 void* stream) {
  RG_CHECK_ARG(xd && x && yd && y && kh >= 1 && kw >= 1 && ph >= 0 && pw >= 0, "rg_im2col_hw: bad argument");
  RG_CHECK_ARG(vec_ok(yd, y), "rg_im2col_hw: output must be bf16 NDHWC with C %% 8 == 0");
  RG_CHECK_ARG(yd->c >= xd->c * kh * kw, "rg_im2col_hw: output has too few channels");
  RG_CHECK_ARG(yd->n == xd->n && yd->d == xd->d && yd->h == xd->h + 2 * ph - kh + 1 && yd->w == xd->w + 2 * pw - kw + 1,
               "rg_im2col_hw: shape mismatch");
  long long total = view(yd).numel() >> 3;
  if (total == 0) return RG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  RG_DISPATCH_FLOAT(xd->dtype, TS, {
    im2col_hw_kernel<TS><<<grid_for(yd->n * yd->d * yd->h, 1, 148 * 32), 256, 0, st>>>(view(xd), (const TS*)x, kh, kw, ph, pw, view(yd), (__nv_bfloat16*)y);
  })
  RG_LAUNCH_CHECK("rg_im2col_hw");
  return RG_OK;
}

int rg_pad3d_fold(const rg_tensor5* gd, void* g, const int32_t lo[3], const int32_t hi[3], int mode, void* stream) {
  RG_CHECK_ARG(gd && g && lo && hi, "rg_pad3d_fold: null argument");
  RG_CHECK_ARG(mode >= RG_PAD_ZEROS && mode <= RG_PAD_REPLICATE, "rg_pad3d_fold: bad mode %d", mode);
  RG_UNSUPPORTED(!vec_ok(gd, g), "rg_pad3d_fold: buffer must be bf16 NDHWC with C %% 8 == 0");
  for (int i = 0; i < 3; ++i) RG_CHECK_ARG(lo[i] >= 0 && lo[i] < 30 && hi[i] >= 0 && hi[i] < 30, "rg_pad3d_fold: bad pad");
  const long long id = gd->d - lo[0] - hi[0], ih = gd->h - lo[1] - hi[1], iw = gd->w - lo[2] - hi[2];
  RG_CHECK_ARG(id >= 1 && ih >= 1 && iw >= 1, "rg_pad3d_fold: pad larger than the buffer");
  if (mode == RG_PAD_REFLECT)
    RG_CHECK_ARG(lo[0] < id && hi[0] < id && lo[1] < ih && hi[1] < ih && lo[2] < iw && hi[2] < iw,
                 "rg_pad3d_fold: reflect padding must be smaller than the interior");
  if (mode == RG_PAD_ZEROS || gd->n * gd->c == 0) return RG_OK;       // zero padding: the interior view already is the adjoint
  pad_fold_kernel<<<grid_for((gd->n * id * ih + 7) / 8, 1, 148 * 32), 256, 0, (cudaStream_t)stream>>>(
      view(gd), (__nv_bfloat16*)g, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2], mode);
  RG_LAUNCH_CHECK("rg_pad3d_fold");
  return RG_OK;
}

int rg_pad_border_fill(const rg_tensor5* td, void* buf, const int32_t margin[3], int mode, void* stream) {
  RG_CHECK_ARG(td && buf && margin, "rg_pad_border_fill: null argument");
  RG_CHECK_ARG(mode >= RG_PAD_ZEROS && mode <= RG_PAD_REPLICATE, "rg_pad_border_fill: bad mode %d", mode);
  RG_UNSUPPORTED(!vec_ok(td, buf), "rg_pad_border_fill: buffer must be bf16 NDHWC with C %% 8 == 0");
  for (int i = 0; i < 3; ++i) RG_CHECK_ARG(margin[i] >= 0 && margin[i] < 30, "rg_pad_border_fill: bad margin");
  const long long id = td->d - 2 * margin[0], ih = td->h - 2 * margin[1], iw = td->w - 2 * margin[2];
  RG_CHECK_ARG(id >= 1 && ih >= 1 && iw >= 1, "rg_pad_border_fill: margin larger than the buffer");
  if (mode == RG_PAD_REFLECT)
    RG_CHECK_ARG(margin[0] < id && margin[1] < ih && margin[2] < iw, "rg_pad_border_fill: reflect margin must be smaller than the interior");
  if (td->n * td->c == 0) return RG_OK;
  border_fill_kernel<<<grid_for(td->n * td->d * td->h, 1, 148 * 32), 256, 0, (cudaStream_t)stream>>>(
      view(td), (__nv_bfloat16*)buf, margin[0], margin[1], margin[2], mode);
  RG_LAUNCH_CHECK("rg_pad_border_fill");
  return RG_OK;
}

}  // extern "C"
