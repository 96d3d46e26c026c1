"""Raw (non-autograd) operator layer: torch tensors in, C-ABI calls out.

Every function here allocates its outputs with torch (PyTorch owns all memory), describes the
tensors as ``rg_tensor5`` views and launches hand-written sm_100a kernels from
``libraygun_b200.so`` on the current CUDA stream.  Nothing here calls a torch compute kernel.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os
import threading

import torch

from . import _C
from ._C import (ACT_NONE, NORM_BATCH, NORM_INSTANCE, PAD_REFLECT, PAD_REPLICATE, PAD_ZEROS, check, count_launch, i3,
                 lib, ptr, require_device, stream_ptr, t5)

# --------------------------------------------------------------------------------------------
# precision mode
# --------------------------------------------------------------------------------------------
# The mode is PROCESS-global, not thread-local: autograd executes backward on its own per-device worker thread, and the
# backward-side dispatch (conv_wgrad / conv_dgrad choose tensor-core, pointwise or direct kernels by mode) must see the mode
# the forward ran in.  (It was thread-local until round 2: inside `with precision("fp32")` the backward thread saw the
# default "bf16" and took bf16 weight-gradient paths for eligible layers.)
_precision = None


def get_precision() -> str:
    """'bf16' (tensor-core path, bf16 NDHWC activations) or 'fp32' (CUDA-core parity path)."""
    return _precision or os.environ.get("RAYGUN_B200_PRECISION", "bf16")


def set_precision(p: str):
    global _precision
    if p not in ("bf16", "fp32"):
        raise ValueError(f"precision must be 'bf16' or 'fp32', got {p!r}")
    _precision = p


@contextlib.contextmanager
def precision(p: str):
    """`with precision("fp32"):` -- forward AND the backward calls made inside the block run in mode `p`."""
    global _precision
    old = _precision
    set_precision(p)
    try:
        yield
    finally:
        _precision = old


def use_tensor_cores() -> bool:
    return os.environ.get("RAYGUN_B200_IGEMM", "1") != "0"


def compute_dtype():
    return torch.bfloat16 if get_precision() == "bf16" else torch.float32


def empty5(shape, dtype, device):
    """Allocate an activation: bf16 -> NDHWC (channels_last_3d) storage, fp32 -> NCDHW."""
    if dtype == torch.bfloat16:
        return torch.empty(shape, dtype=dtype, device=device, memory_format=torch.channels_last_3d)
    return torch.empty(shape, dtype=dtype, device=device)


# ---- pad arenas (inference) -------------------------------------------------------------------------
# Under torch.no_grad() the bf16 convolutions write their output into the INTERIOR of a buffer that is one voxel
# larger on every side.  A following convolution with reflect / replicate padding <= 1 then only mirrors the thin
# shell in place (rg_pad_border_fill) instead of copying the whole tensor into a padded one -- the explicit
# `F.pad` + conv of the reference (UNet.py:61-69) costs a read + write of every activation.
ARENA_MARGIN = 1


_arena_flag = None      # set by functional.* around autograd.Function.apply: Function.forward always runs with grad mode
                        # off, so the caller's grad mode has to be handed down explicitly


class arena_scope:
    """with arena_scope(allow): kernels called inside see arena_enabled() == allow (and bf16, and the env switch)."""

    def __init__(self, allow: bool):
        self.allow = bool(allow)

    def __enter__(self):
        global _arena_flag
        self.prev, _arena_flag = _arena_flag, self.allow

    def __exit__(self, *exc):
        global _arena_flag
        _arena_flag = self.prev


def arena_enabled() -> bool:
    """Margin-allocated conv outputs are an inference-only layout (training keeps dense tensors for the norm kernels)."""
    import os
    if get_precision() != "bf16" or os.environ.get("RAYGUN_B200_ARENA", "1") == "0":
        return False
    return _arena_flag if _arena_flag is not None else not torch.is_grad_enabled()


def empty5_arena(shape, dtype, device, margin=ARENA_MARGIN):
    """Interior view (logical `shape`) of a freshly allocated NDHWC buffer with `margin` extra voxels per side."""
    n, c, d, h, w = shape
    m = margin
    full = empty5((n, c, d + 2 * m, h + 2 * m, w + 2 * m), dtype, device)
    full._rg_arena = m
    return full[:, :, m:m + d, m:m + h, m:m + w]


def arena_of(x, lo, hi):
    """(full_buffer, margin) when x is exactly the interior of a pad arena whose margin covers lo / hi, else None."""
    b = x._base
    m = getattr(b, "_rg_arena", None) if b is not None else None
    if m is None or max(max(lo), max(hi)) > m:
        return None
    n, c, d, h, w = x.shape
    if tuple(b.shape) != (n, c, d + 2 * m, h + 2 * m, w + 2 * m) or x.stride() != b.stride():
        return None
    bs = b.stride()
    if x.storage_offset() != b.storage_offset() + m * (bs[2] + bs[3] + bs[4]):
        return None
    return b, m


def border_fill(full, margin, mode):
    check(lib().rg_pad_border_fill(C.byref(t5(full)), ptr(full), i3((margin,) * 3), C.c_int(mode), stream_ptr(full.device)),
          "rg_pad_border_fill")
    count_launch()


def is_ndhwc(t: torch.Tensor) -> bool:
    n, c, d, h, w = t.shape
    return t.dtype == torch.bfloat16 and tuple(t.stride()) == (d * h * w * c, 1, h * w * c, w * c, c)


# --------------------------------------------------------------------------------------------
# data movement
# --------------------------------------------------------------------------------------------
def copy_into(dst: torch.Tensor, src: torch.Tensor, accumulate: bool = False):
    require_device(src)
    check(lib().rg_copy5(C.byref(t5(dst)), ptr(dst), C.byref(t5(src)), ptr(src), C.c_int(int(accumulate)),
                         stream_ptr(src.device)), "rg_copy5")
    count_launch()
    return dst


def convert(x: torch.Tensor, dtype) -> torch.Tensor:
    """Cast / re-layout to the canonical storage of `dtype` (no-op if already there)."""
    if x.dtype == dtype and (is_ndhwc(x) if dtype == torch.bfloat16 else x.is_contiguous()):
        return x
    return copy_into(empty5(x.shape, dtype, x.device), x)


def channels_inner_bf16(x: torch.Tensor) -> bool:
    """bf16, channels innermost, every other stride and the base 16-byte aligned: what the TMA-fed kernels and the
    16-byte vector kernels address directly (dense NDHWC, pad-arena interiors, channel slices, crops)."""
    return (x.dtype == torch.bfloat16 and x.dim() == 5 and x.stride(1) == 1 and x.data_ptr() % 16 == 0
            and all(s % 8 == 0 for i, s in enumerate(x.stride()) if i != 1))


def conv_operand(x: torch.Tensor) -> torch.Tensor:
    """`x` as a bf16 operand of the tensor-core / pointwise convolutions: strided channel-inner views are used in
    place (no densifying copy), anything else is converted to dense NDHWC bf16."""
    return x if channels_inner_bf16(x) else convert(x, torch.bfloat16)


def im2col_hw(x, kh, kw, ph, pw, channels=64):
    """Fold the H/W taps of a small-Cin convolution into `channels` bf16 NDHWC channels."""
    require_device(x)
    n, c, d, h, w = x.shape
    y = empty5((n, channels, d, h + 2 * ph - kh + 1, w + 2 * pw - kw + 1), torch.bfloat16, x.device)
    check(lib().rg_im2col_hw(C.byref(t5(x)), ptr(x), C.c_int(kh), C.c_int(kw), C.c_int(ph), C.c_int(pw),
                             C.byref(t5(y)), ptr(y), stream_ptr(x.device)), "rg_im2col_hw")
    count_launch()
    return y


def pad3d(x, lo, hi, mode):
    require_device(x)
    n, c, d, h, w = x.shape
    y = empty5((n, c, d + lo[0] + hi[0], h + lo[1] + hi[1], w + lo[2] + hi[2]), x.dtype, x.device)
    check(lib().rg_pad3d_fwd(C.byref(t5(x)), ptr(x), C.byref(t5(y)), ptr(y), i3(lo), C.c_int(mode),
                             stream_ptr(x.device)), "rg_pad3d_fwd")
    count_launch()
    return y


def is_row_ndhwc(t: torch.Tensor) -> bool:
    """bf16, channels contiguous, 16-byte aligned rows: dense NDHWC tensors and interiors / channel-complete crops of them."""
    st = t.stride()
    return (t.dtype == torch.bfloat16 and t.dim() == 5 and st[1] == 1 and t.shape[1] % 8 == 0
            and all(v % 8 == 0 for v in (st[0], st[2], st[3], st[4])) and t.data_ptr() % 16 == 0)


def pad3d_fold(dy, lo, hi, mode):
    """Adjoint of the padding IN PLACE: folds the shell of the padded gradient `dy` into its interior and returns the
    interior VIEW (shape of the pad's input).  `dy` must be a bf16 NDHWC buffer this call may overwrite."""
    require_device(dy)
    check(lib().rg_pad3d_fold(C.byref(t5(dy)), ptr(dy), i3(lo), i3(hi), C.c_int(mode), stream_ptr(dy.device)), "rg_pad3d_fold")
    count_launch()
    d, h, w = dy.shape[2:]
    return dy[:, :, lo[0]: d - hi[0], lo[1]: h - hi[1], lo[2]: w - hi[2]]


def pad3d_bwd(dy, x_shape, lo, mode):
    dx = empty5(x_shape, dy.dtype, dy.device)          # fully written by the kernel
    check(lib().rg_pad3d_bwd(C.byref(t5(dy)), ptr(dy), C.byref(t5(dx)), ptr(dx), i3(lo), C.c_int(mode),
                             stream_ptr(dy.device)), "rg_pad3d_bwd")
    count_launch()
    return dx


# --------------------------------------------------------------------------------------------
# activations / pooling / upsampling
# --------------------------------------------------------------------------------------------
def act_fwd(x, act, slope=0.0, out_dtype=None):
    require_device(x)
    y = empty5(x.shape, out_dtype or x.dtype, x.device)
    check(lib().rg_act_fwd(C.byref(t5(x)), ptr(x), C.c_int(act), C.c_float(slope), C.byref(t5(y)), ptr(y),
                           stream_ptr(x.device)), "rg_act_fwd")
    count_launch()
    return y


def act_bwd(y, dy, act, slope=0.0):
    if dy.dtype != y.dtype:
        dy = convert(dy, y.dtype)
    dx = empty5(y.shape, y.dtype, y.device)
    check(lib().rg_act_bwd(C.byref(t5(y)), ptr(y), C.byref(t5(dy)), ptr(dy), C.c_int(act), C.c_float(slope),
                           C.byref(t5(dx)), ptr(dx), stream_ptr(y.device)), "rg_act_bwd")
    count_launch()
    return dx


def act_bwd_bias(y, dy, act, slope=0.0):
    """(dy * act'(y), per-channel sum of it) in one pass for bf16 NDHWC operands (dense, or row-strided interiors of
    padded buffers), else (act_bwd, None)."""
    if (is_row_ndhwc(y) and is_row_ndhwc(dy) and y.shape[1] <= 2048 and 256 % (y.shape[1] // 8) == 0
            and y.shape[2] * y.shape[3] * y.shape[4] * y.shape[1] < 2 ** 31):
        dx = empty5(y.shape, y.dtype, y.device)
        db = torch.empty(y.shape[1], dtype=torch.float32, device=y.device)
        ws = reduce_ws(1, y.shape[1], y.device)
        check(lib().rg_act_bwd_bias(C.byref(t5(y)), ptr(y), C.byref(t5(dy)), ptr(dy), C.c_int(act), C.c_float(slope),
                                    C.byref(t5(dx)), ptr(dx), ptr(db), ptr(ws), stream_ptr(y.device)), "rg_act_bwd_bias")
        count_launch(2)
        return dx, db
    return act_bwd(y, dy, act, slope), None


def maxpool_fwd(x, f, arena=False):
    require_device(x)
    n, c, d, h, w = x.shape
    shape = (n, c, d // f[0], h // f[1], w // f[2])
    y = empty5_arena(shape, x.dtype, x.device) if arena else empty5(shape, x.dtype, x.device)
    check(lib().rg_maxpool3d_fwd(C.byref(t5(x)), ptr(x), i3(f), C.byref(t5(y)), ptr(y), stream_ptr(x.device)),
          "rg_maxpool3d_fwd")
    count_launch()
    return y


def maxpool_bwd(x, dy, f):
    if dy.dtype != x.dtype:
        dy = convert(dy, x.dtype)
    dx = empty5(x.shape, x.dtype, x.device)
    if any(s % k for s, k in zip(x.shape[2:], f)):
        dx.zero_()
    check(lib().rg_maxpool3d_bwd(C.byref(t5(x)), ptr(x), None, None, C.byref(t5(dy)), ptr(dy), i3(f),
                                 C.byref(t5(dx)), ptr(dx), stream_ptr(x.device)), "rg_maxpool3d_bwd")
    count_launch()
    return dx


def upsample_fwd(x, f):
    require_device(x)
    n, c, d, h, w = x.shape
    y = empty5((n, c, d * f[0], h * f[1], w * f[2]), x.dtype, x.device)
    check(lib().rg_upsample_nearest_fwd(C.byref(t5(x)), ptr(x), i3(f), C.byref(t5(y)), ptr(y),
                                        stream_ptr(x.device)), "rg_upsample_nearest_fwd")
    count_launch()
    return y


def upsample_bwd(dy, f):
    n, c, d, h, w = dy.shape
    dx = empty5((n, c, d // f[0], h // f[1], w // f[2]), dy.dtype, dy.device)
    check(lib().rg_upsample_nearest_bwd(C.byref(t5(dy)), ptr(dy), i3(f), C.byref(t5(dx)), ptr(dx),
                                        stream_ptr(dy.device)), "rg_upsample_nearest_bwd")
    count_launch()
    return dx


def resize_trilinear_fwd(x, size):
    """F.interpolate(x, size=size, mode="trilinear", align_corners=True) (CycleModel.py:24-30)."""
    require_device(x)
    y = torch.empty((x.shape[0], x.shape[1], *size), dtype=x.dtype, device=x.device)
    check(lib().rg_resize_trilinear_fwd(C.byref(t5(x)), ptr(x), C.byref(t5(y)), ptr(y), stream_ptr(x.device)),
          "rg_resize_trilinear_fwd")
    count_launch()
    return y


def resize_trilinear_bwd(dy, in_size):
    """Adjoint of resize_trilinear_fwd: gradient w.r.t. an input of spatial extents `in_size`."""
    require_device(dy)
    dx = torch.empty((dy.shape[0], dy.shape[1], *in_size), dtype=dy.dtype, device=dy.device)
    check(lib().rg_resize_trilinear_bwd(C.byref(t5(dy)), ptr(dy), C.byref(t5(dx)), ptr(dx), stream_ptr(dy.device)),
          "rg_resize_trilinear_bwd")
    count_launch()
    return dx


# --------------------------------------------------------------------------------------------
# normalisation
# --------------------------------------------------------------------------------------------
RED_MAX_BLOCKS = 148 * 8


def reduce_ws(n, c, device):
    """Uninitialised fp32 scratch for the two-stage (partial rows + ordered sum) reductions: rg_reduce_ws_floats(n, c)."""
    return torch.empty(2 * c * (RED_MAX_BLOCKS + 2 * n + 2) + 4 * n * c + 64, dtype=torch.float32, device=device)


def ws_bytes(nbytes, device):
    return torch.empty(max(int(nbytes) // 4, 1), dtype=torch.float32, device=device)


def rows_sum(part, rows, groups, out=None, accumulate=False):
    """out[g] (+)= sum_r part[r, g] in ascending r (deterministic second stage of a caller-owned partial-row buffer)."""
    if out is None:
        out = torch.empty(groups, dtype=torch.float32, device=part.device)
    check(lib().rg_rows_sum(ptr(part), C.c_int(rows), C.c_longlong(groups), ptr(out), C.c_int(int(accumulate)),
                            stream_ptr(part.device)), "rg_rows_sum")
    count_launch()
    return out


def norm_stats(x, kind):
    """sum / sum-of-squares per (n,c) [instance] or per c [batch]; fp32."""
    require_device(x)
    groups = x.shape[0] * x.shape[1] if kind == NORM_INSTANCE else x.shape[1]
    s = torch.empty(2, groups, dtype=torch.float32, device=x.device)
    ws = reduce_ws(x.shape[0], x.shape[1], x.device)
    check(lib().rg_norm_stats(C.byref(t5(x)), ptr(x), C.c_int(kind), ptr(s[0]), ptr(s[1]), ptr(ws), stream_ptr(x.device)),
          "rg_norm_stats")
    count_launch(2)
    return s[0], s[1]


def norm_finalize(ssum, ssq, count, eps):
    groups = ssum.numel()
    out = torch.empty(3, groups, dtype=torch.float32, device=ssum.device)
    check(lib().rg_norm_finalize(ptr(ssum), ptr(ssq), C.c_int(groups), C.c_double(float(count)), C.c_float(eps),
                                 ptr(out[0]), ptr(out[1]), ptr(out[2]), stream_ptr(ssum.device)), "rg_norm_finalize")
    count_launch()
    return out[0], out[1], out[2]  # mean, rstd, biased var


def norm_apply(x, kind, mean, rstd, gamma, beta, residual, act, slope, out_dtype=None):
    require_device(x)
    y = empty5(x.shape, out_dtype or x.dtype, x.device)
    rd = C.byref(t5(residual)) if residual is not None else None
    check(lib().rg_norm_apply(C.byref(t5(x)), ptr(x), C.c_int(kind), ptr(mean), ptr(rstd), ptr(gamma), ptr(beta),
                              rd, ptr(residual), C.c_int(act), C.c_float(slope), C.byref(t5(y)), ptr(y),
                              stream_ptr(x.device)), "rg_norm_apply")
    count_launch()
    return y


def norm_bwd(x, dy, y, kind, training, mean, rstd, gamma, act, slope, want_affine, need_dx=True):
    """Returns dx, dgamma, dbeta (the latter two None unless want_affine)."""
    if act != ACT_NONE and dy.dtype != y.dtype:
        dy = convert(dy, y.dtype)
    groups = mean.numel()
    c = x.shape[1]
    red = torch.empty(2 * groups, dtype=torch.float32, device=x.device)
    ws = reduce_ws(x.shape[0], c, x.device)
    dg = torch.empty(2, c, dtype=torch.float32, device=x.device) if want_affine else None
    dx = empty5(x.shape, x.dtype, x.device) if need_dx else None
    yd = C.byref(t5(y)) if y is not None else None
    check(lib().rg_norm_bwd(C.byref(t5(x)), ptr(x), C.byref(t5(dy)), ptr(dy), yd, ptr(y), C.c_int(kind),
                            C.c_int(int(training)), ptr(mean), ptr(rstd), ptr(gamma), C.c_int(act), C.c_float(slope),
                            ptr(red), ptr(dg[0]) if want_affine else None, ptr(dg[1]) if want_affine else None,
                            C.byref(t5(dx)) if need_dx else None, ptr(dx), ptr(ws), stream_ptr(x.device)), "rg_norm_bwd")
    count_launch(3 if training or want_affine else 1)
    return dx, (dg[0] if want_affine else None), (dg[1] if want_affine else None)


# --------------------------------------------------------------------------------------------
# losses / prediction arithmetic / optimiser
# --------------------------------------------------------------------------------------------
def loss_fwd_grad(kind, a, b=None, target=0.0, beta=1.0, want_grad=True):
    """mean-reduced loss (fp32 scalar tensor) and d loss / d a (same storage as a)."""
    require_device(a)
    loss = torch.empty(1, dtype=torch.float32, device=a.device)
    ws = torch.empty(RED_MAX_BLOCKS, dtype=torch.float32, device=a.device)
    grad = empty5(a.shape, a.dtype, a.device) if want_grad else None
    check(lib().rg_loss_fwd_grad(C.c_int(kind), C.byref(t5(a)), ptr(a), C.byref(t5(b)) if b is not None else None,
                                 ptr(b), C.c_float(target), C.c_float(beta), C.c_float(1.0), ptr(loss), C.c_float(1.0),
                                 C.byref(t5(grad)) if want_grad else None, ptr(grad), ptr(ws), stream_ptr(a.device)),
          "rg_loss_fwd_grad")
    count_launch(2)
    return loss[0], grad


def u8_normalise(u8: torch.Tensor, scale=2.0, shift=-1.0, dtype=None):
    require_device(u8)
    dtype = dtype or compute_dtype()
    u8 = u8.contiguous()
    y = torch.empty(u8.shape, dtype=dtype, device=u8.device)
    check(lib().rg_u8_normalise(ptr(u8), C.c_int64(u8.numel()), C.c_float(scale), C.c_float(shift),
                                C.c_int(_C.RG_BF16 if dtype == torch.bfloat16 else _C.RG_F32), ptr(y),
                                stream_ptr(u8.device)), "rg_u8_normalise")
    count_launch()
    return y


def denorm_clamp_u8(x: torch.Tensor, out: torch.Tensor | None = None):
    """x: 5-D view (any strides, e.g. a cropped slice) -> contiguous uint8 of the same shape (`out`: preallocated)."""
    require_device(x)
    if out is None:
        out = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    elif tuple(out.shape) != tuple(x.shape) or out.dtype != torch.uint8 or not out.is_contiguous():
        raise ValueError("denorm_clamp_u8: out must be a contiguous uint8 tensor of x's shape")
    check(lib().rg_denorm_clamp_u8(C.byref(t5(x)), ptr(x), ptr(out), stream_ptr(x.device)), "rg_denorm_clamp_u8")
    count_launch()
    return out


def adam_step(p, g, m, v, lr, beta1, beta2, eps, step, grad_scale=1.0):
    require_device(p)
    check(lib().rg_adam_step(ptr(p), ptr(g), ptr(m), ptr(v), C.c_int64(p.numel()), C.c_float(lr), C.c_float(beta1),
                             C.c_float(beta2), C.c_float(eps), C.c_int(step), C.c_float(grad_scale),
                             stream_ptr(p.device)), "rg_adam_step")
    count_launch()
