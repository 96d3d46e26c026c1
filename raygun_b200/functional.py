"""Autograd layer: ``torch.autograd.Function`` wrappers around the sm_100a kernels.

Each Function is the replacement of one torch.nn call site of the reference (cited per class).
Forward and backward both run hand-written kernels only; autograd is used for graph bookkeeping.
"""
from __future__ import annotations

import torch

from . import conv as CV
from . import kernels as K
from .optim import bucket_grad
from ._C import (ACT_LEAKY, ACT_NONE, ACT_RELU, ACT_SIGMOID, ACT_TANH, LOSS_L1, LOSS_MEAN, LOSS_MSE_CONST, LOSS_SMOOTHL1,
                 NORM_BATCH, NORM_INSTANCE, PAD_REFLECT, PAD_REPLICATE, PAD_ZEROS)

import os as _os
TRAIN_ARENA = _os.environ.get("RAYGUN_B200_TRAIN_ARENA", "1") != "0"     # pad arenas for conv+activation outputs in training
PAD_FOLD = _os.environ.get("RAYGUN_B200_PAD_FOLD", "1") != "0"           # in-place pad adjoint behind our convolutions
TRAIN_CAT = _os.environ.get("RAYGUN_B200_TRAIN_CAT", "1") != "0"         # training-time in-place U-Net skip concat

PAD_MODES = {"zeros": PAD_ZEROS, "constant": PAD_ZEROS, "reflect": PAD_REFLECT, "replicate": PAD_REPLICATE}


def _t3(v):
    return (int(v),) * 3 if isinstance(v, int) else tuple(int(a) for a in v)


# ---------------------------------------------------------------------------------------------
# casts
# ---------------------------------------------------------------------------------------------
class _Cast(torch.autograd.Function):
    """dtype / layout change (reference: implicit; predict/worker.py:101-107 .float())."""

    @staticmethod
    def forward(ctx, x, dtype):
        ctx.src_dtype = x.dtype
        return K.convert(x, dtype)

    @staticmethod
    def backward(ctx, dy):
        return K.convert(dy, ctx.src_dtype), None


def cast(x, dtype):
    if x.dtype == dtype and (K.is_ndhwc(x) if dtype == torch.bfloat16 else x.is_contiguous()):
        return x
    if x.dtype == dtype == torch.bfloat16 and x.dim() == 5 and x.shape[1] == 1 and not (x.requires_grad and torch.is_grad_enabled()):
        return x          # one channel: every layout is "channels innermost"; the kernels that take it read element strides
    return _Cast.apply(x, dtype)


def to_compute(x):
    """Bring an input into the storage of the active precision mode."""
    return cast(x, K.compute_dtype())


# ---------------------------------------------------------------------------------------------
# convolution
# ---------------------------------------------------------------------------------------------
class _Out:
    """A destination view handed to an autograd.Function WITHOUT being one of its tensor inputs: the kernel writes into it
    and the Function returns it as a fresh output (training-time in-place skip concat; a tensor argument that is returned
    would be treated by autograd as an aliased input)."""
    __slots__ = ("t",)

    def __init__(self, t):
        self.t = t


def _unbox(out):
    return out.t if isinstance(out, _Out) else out


def _dest(out):
    """`out` as the Functions take it: dropped in training unless the caller marked it with `mark_train_out`."""
    if out is None or not torch.is_grad_enabled():
        return out
    return _Out(out) if getattr(out, "_rg_train_out", False) else None


def mark_train_out(view):
    """Allow a convolution to write its TRAINING-mode output into `view` (a channel slice of a concat buffer whose
    consumer is wired up with `cat_join`)."""
    view._rg_train_out = True
    return view


class _Conv3d(torch.autograd.Function):
    """nn.Conv3d forward/backward with zero padding and a fused activation epilogue
    (UNet.py:61-69, ResNet.py:271-298,407-416, NLayerDiscriminator.py:137-174, BaseSystem.py:209-224).
    Returns (y, sum, sqsum): the statistics feed the following norm layer (non-differentiable)."""

    @staticmethod
    def _narrow_out(x, w, stride, groups, want_stats):
        """Few produced channels over a wide reduction (the discriminator's last conv, 512 -> 1, 3^3,
        NLayerDiscriminator.py:166-174): run it on the tensor cores with the produced channels zero-padded to 64
        instead of the CUDA-core kernel (64x the FLOPs of a 5832-voxel layer is still microseconds)."""
        cout, cin = w.shape[0], w.shape[1]
        return (cout < 64 and cin % 64 == 0 and groups == 1 and tuple(stride) == (1, 1, 1) and not want_stats
                and w.shape[2] * w.shape[3] * w.shape[4] > 1 and K.compute_dtype() == torch.bfloat16
                and K.use_tensor_cores() and x.is_cuda)

    @staticmethod
    def forward(ctx, x, w, bias, stride, pad, groups, act, slope, want_stats, out=None):
        out = _unbox(out)
        ctx.narrow = _Conv3d._narrow_out(x, w, stride, groups, want_stats)
        if ctx.narrow:
            cout = w.shape[0]
            wp = CV.WeightSrc(w, lambda p_: CV._pad_rows(p_, 64), (64, *w.shape[1:]), "rows64")
            bp = None
            if bias is not None:
                bp = torch.zeros(64, dtype=torch.float32, device=w.device)
                bp[:cout] = bias.detach()
            yp, _, _ = CV.conv_fprop(x, wp, bp, stride, pad, groups, act, slope, False)
            y = K.copy_into(K.empty5((yp.shape[0], cout, *yp.shape[2:]), yp.dtype, yp.device), yp[:, :cout])
            s = q = None
        else:
            y, s, q = CV.conv_fprop(x, w, bias, stride, pad, groups, act, slope, want_stats, out=out)
        ctx.geom = (stride, pad, groups, act, slope)
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x, w, y if act != ACT_NONE else None)
        if s is None:
            s = q = x.new_empty(0, dtype=torch.float32)
        ctx.mark_non_differentiable(s, q)
        return y, s, q

    @staticmethod
    def backward(ctx, dy, _ds, _dq):
        x, w, y = ctx.saved_tensors
        stride, pad, groups, act, slope = ctx.geom
        db_pre = None
        if act != ACT_NONE:
            if ctx.has_bias and ctx.needs_input_grad[2] and not ctx.narrow:
                dy, db_pre = K.act_bwd_bias(y, dy, act, slope)          # activation derivative + bias gradient, one pass
            else:
                dy = K.act_bwd(y, dy, act, slope)
        dx = dw = db = None
        if ctx.narrow:
            cout = w.shape[0]
            dyp = torch.zeros(dy.shape[0], 64, *dy.shape[2:], dtype=torch.bfloat16, device=dy.device).contiguous(
                memory_format=torch.channels_last_3d)
            K.copy_into(dyp[:, :cout], dy)
            wp = CV.WeightSrc(w, lambda p_: CV._pad_rows(p_, 64), (64, *w.shape[1:]), "rows64")
            if ctx.needs_input_grad[0]:
                dx, _, _ = CV.conv_dgrad(dyp, wp, x.shape, stride, pad, groups, out_dtype=x.dtype)
            if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
                dwp, dbp = CV.conv_wgrad(x, dyp, wp.shape, stride, pad, groups, want_bias=ctx.has_bias)
                dw = dwp[:cout].contiguous() if ctx.needs_input_grad[1] else None
                db = dbp[:cout].contiguous() if dbp is not None else None
            return dx, dw, db, None, None, None, None, None, None, None
        if ctx.needs_input_grad[0]:
            dx, _, _ = CV.conv_dgrad(dy, w, x.shape, stride, pad, groups, out_dtype=x.dtype)
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            # a live gradient bucket (optim.FlatAdam): the kernel accumulates into its view, autograd gets None
            bg = bucket_grad(w) if ctx.needs_input_grad[1] else None
            dw, db = CV.conv_wgrad(x, dy, w.shape, stride, pad, groups, want_bias=ctx.has_bias, db=db_pre, dw_out=bg,
                                   accumulate=bg is not None)
            if not ctx.needs_input_grad[1] or bg is not None:
                dw = None
        return dx, dw, db, None, None, None, None, None, None, None



class _StemConv3d(torch.autograd.Function):
    """Small-Cin convolution (the 7^3 1->64 ResNet stem, ResNet.py:283; the first U-Net ConvPass conv,
    UNet.py:61-69) on the tensor cores: the H/W taps are folded into 64 bf16 channels by an im2col kernel,
    leaving a (KD,1,1) convolution with Cin = 64 for the implicit-GEMM kernels (fprop and wgrad).
    The folded input is recomputed in backward instead of being saved."""

    @staticmethod
    def forward(ctx, x, w, bias, pad, act, slope, want_stats, out=None):
        out = _unbox(out)
        cout, cin, kd, kh, kw = w.shape
        cf = cin * kh * kw
        if CV.stem_direct_ok(w, want_stats):
            # whole reduction <= 64 (U-Net stem): im2col rows are built in shared memory, no folded intermediate
            y, s, q = CV.stem_conv_fprop(x, w, bias, pad, act, slope, out=out), None, None
        else:
            xc = K.im2col_hw(x, kh, kw, pad[1], pad[2], 64)

            def fold(p_):          # [co][ci][kd][kh][kw] -> [co][(ci,kh,kw) padded to 64][kd][1][1]
                w2 = p_.detach().permute(0, 2, 1, 3, 4).reshape(cout, kd, cf)
                return torch.nn.functional.pad(w2, (0, 64 - cf)).permute(0, 2, 1).reshape(cout, 64, kd, 1, 1).contiguous()

            w2 = CV.WeightSrc(w, fold, (cout, 64, kd, 1, 1), "stemfold")
            y, s, q = CV.conv_fprop(xc, w2, bias, (1, 1, 1), (pad[0], 0, 0), 1, act, slope, want_stats, out=out)
        ctx.cfg = (pad, act, slope, tuple(w.shape), (cout, 64, kd, 1, 1))
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x, w, y if act != ACT_NONE else None)
        if s is None:
            s = q = x.new_empty(0, dtype=torch.float32)
        ctx.mark_non_differentiable(s, q)
        return y, s, q

    @staticmethod
    def backward(ctx, dy, _ds, _dq):
        x, w, y = ctx.saved_tensors
        pad, act, slope, wshape, w2shape = ctx.cfg
        cout, cin, kd, kh, kw = wshape
        db_pre = None
        if act != ACT_NONE:
            if ctx.has_bias and ctx.needs_input_grad[2]:
                dy, db_pre = K.act_bwd_bias(y, dy, act, slope)
            else:
                dy = K.act_bwd(y, dy, act, slope)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx, _, _ = CV.conv_dgrad(dy, w, x.shape, (1, 1, 1), pad, 1, out_dtype=x.dtype)
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            if CV.smallk_wgrad_ok(x, dy, wshape):
                # whole reduction <= 64: FMA-bound on the CUDA cores, straight into the gradient bucket when there is one
                bg = bucket_grad(w) if ctx.needs_input_grad[1] else None
                dw, db = CV.conv_wgrad(x, dy, wshape, (1, 1, 1), pad, 1, want_bias=ctx.has_bias, db=db_pre, dw_out=bg,
                                       accumulate=bg is not None)
                if not ctx.needs_input_grad[1] or bg is not None:
                    dw = None
            else:
                xc = K.im2col_hw(x, kh, kw, pad[1], pad[2], 64)
                dw2, db = CV.conv_wgrad(xc, dy, w2shape, (1, 1, 1), (pad[0], 0, 0), 1, want_bias=ctx.has_bias, db=db_pre)
                dw = dw2.reshape(cout, 64, kd)[:, : cin * kh * kw].reshape(cout, cin, kh, kw, kd).permute(0, 1, 4, 2, 3)
                dw = dw.contiguous() if ctx.needs_input_grad[1] else None
        return dx, dw, db, None, None, None, None, None


def _stem_eligible(x, w, stride, groups):
    cout, cin, kd, kh, kw = w.shape
    return (K.get_precision() == "bf16" and K.use_tensor_cores() and groups == 1 and stride == (1, 1, 1)
            and cin % 64 != 0 and cin * kh * kw <= 64 and cout % 64 == 0 and kd <= 8)


def conv3d(x, w, bias=None, stride=1, padding=0, groups=1, act=ACT_NONE, slope=0.0, want_stats=False, out=None):
    """`out` (inference only): a preallocated NDHWC view the kernel should write into, e.g. a channel slice of the
    skip-connection concat buffer; the result IS `out` when the kernel could honour it (callers check identity)."""
    out = _dest(out)
    # training: a convolution + activation with no norm behind it (every G-UNet ConvPass conv) also writes into a pad
    # arena -- the next reflect-padded conv mirrors the shell in place, the activation backward reads the interior view
    train_arena = TRAIN_ARENA and act != ACT_NONE and not want_stats
    with K.arena_scope(not torch.is_grad_enabled() or train_arena):
        if _stem_eligible(x, w, _t3(stride), int(groups)):
            y, s, q = _StemConv3d.apply(x, w, bias, _t3(padding), act, float(slope), bool(want_stats), out)
        else:
            y, s, q = _Conv3d.apply(x, w, bias, _t3(stride), _t3(padding), int(groups), act, float(slope), bool(want_stats), out)
    if want_stats:
        return y, (s, q) if s.numel() else None
    return y


class _ConvTranspose3d(torch.autograd.Function):
    """nn.ConvTranspose3d (UNet.py:265-269 k=s=factor; ResNet.py:325-333 k3 s2 p1 op1).
    Forward is the data-gradient kernel of the conv that shares the weight [Cin_T, Cout_T, k..]."""

    @staticmethod
    def forward(ctx, x, w, bias, stride, pad, out_pad, act, slope, want_stats, out=None):
        out = _unbox(out)
        n, cin, d, h, wd_ = x.shape
        k = w.shape[2:]
        osz = [(i - 1) * s - 2 * p + kk + op for i, s, p, kk, op in zip((d, h, wd_), stride, pad, k, out_pad)]
        if CV.upconv2_ok(x, w, stride, pad, out_pad, want_stats):
            y, s, q = CV.upconv2_fprop(x, w, bias, act, slope, out=out), None, None      # k = s = 2: one fused launch
        else:
            y, s, q = CV.conv_dgrad(x, w, (n, w.shape[1], *osz), stride, pad, 1, bias=bias, act=act, slope=slope,
                                    want_stats=want_stats, out=out)
        ctx.geom = (stride, pad, act, slope)
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x, w, y if act != ACT_NONE else None)
        if s is None:
            s = q = x.new_empty(0, dtype=torch.float32)
        ctx.mark_non_differentiable(s, q)
        return y, s, q

    @staticmethod
    def backward(ctx, dy, _ds, _dq):
        x, w, y = ctx.saved_tensors
        stride, pad, act, slope = ctx.geom
        if act != ACT_NONE:
            dy = K.act_bwd(y, dy, act, slope)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx, _, _ = CV.conv_fprop(dy, w, None, stride, pad, 1, out_dtype=x.dtype)
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            # conv' : dy (Cout_T) -> x (Cin_T); its wgrad has "x" = dy and "dy" = x; bias grad = sum(dy)
            bg = bucket_grad(w) if ctx.needs_input_grad[1] else None
            dw, _ = CV.conv_wgrad(dy, x, w.shape, stride, pad, 1, want_bias=False, dw_out=bg, accumulate=bg is not None)
            if ctx.has_bias:
                s, _ = K.norm_stats(dy, NORM_BATCH)
                db = s
            if not ctx.needs_input_grad[1] or bg is not None:
                dw = None
        return dx, dw, db, None, None, None, None, None, None, None


def conv_transpose3d(x, w, bias=None, stride=1, padding=0, output_padding=0, act=ACT_NONE, slope=0.0,
                     want_stats=False, out=None):
    out = _dest(out)
    with K.arena_scope(not torch.is_grad_enabled()):
        y, s, q = _ConvTranspose3d.apply(x, w, bias, _t3(stride), _t3(padding), _t3(output_padding), act, float(slope),
                                         bool(want_stats), out)
    if want_stats:
        return y, (s, q) if s.numel() else None
    return y


# ---------------------------------------------------------------------------------------------
# padding
# ---------------------------------------------------------------------------------------------
class _Pad3d(torch.autograd.Function):
    """nn.ReflectionPad3d / ReplicationPad3d / F.pad (ResNet.py:270-279,392-401; Conv3d padding_mode).
    fold=True: the adjoint folds the shell of the incoming gradient IN PLACE and hands its interior view upstream
    (O(surface) bytes); only where the single consumer of the padded tensor is one of our convolutions, whose
    data gradient is a fresh buffer."""

    @staticmethod
    def forward(ctx, x, lo, hi, mode, fold=False):
        ctx.args = (tuple(x.shape), lo, hi, mode, fold)
        return K.pad3d(x, lo, hi, mode)

    @staticmethod
    def backward(ctx, dy):
        shape, lo, hi, mode, fold = ctx.args
        if fold and K.is_ndhwc(dy) and dy.shape[1] % 8 == 0:
            return K.pad3d_fold(dy, lo, hi, mode), None, None, None, None
        return K.pad3d_bwd(dy, shape, lo, mode), None, None, None, None


class _PadArena(torch.autograd.Function):
    """Padding of a tensor that IS the interior of a pad arena (kernels.empty5_arena): the shell is mirrored in place
    and the padded view of the same buffer is returned (no copy); the adjoint folds the gradient's shell in place.
    Same call sites as _Pad3d (UNet.py:61-69 padding_mode="reflect")."""

    @staticmethod
    def forward(ctx, x, lo, hi, mode):
        full, m = K.arena_of(x, lo, hi)
        K.border_fill(full, m, mode)
        d, h, w = x.shape[2:]
        ctx.args = (tuple(x.shape), lo, hi, mode)
        return full[:, :, m - lo[0]: m + d + hi[0], m - lo[1]: m + h + hi[1], m - lo[2]: m + w + hi[2]]

    @staticmethod
    def backward(ctx, dy):
        shape, lo, hi, mode = ctx.args
        if K.is_ndhwc(dy) and dy.shape[1] % 8 == 0:
            return K.pad3d_fold(dy, lo, hi, mode), None, None, None
        return K.pad3d_bwd(dy, shape, lo, mode), None, None, None


def pad3d(x, lo, hi=None, mode="reflect", fold=False):
    lo = _t3(lo)
    hi = lo if hi is None else _t3(hi)
    if not any(lo) and not any(hi):
        return x
    mode = PAD_MODES[mode] if isinstance(mode, str) else mode
    if x.dim() == 5 and x.is_cuda and K.arena_of(x, lo, hi) is not None:
        return _PadArena.apply(x, lo, hi, mode)
    return _Pad3d.apply(x, lo, hi, mode, bool(fold) and x.dtype == torch.bfloat16)


# ---------------------------------------------------------------------------------------------
# normalisation (+ fused activation and residual add)
# ---------------------------------------------------------------------------------------------
class _NormAct(torch.autograd.Function):
    """y = act(norm(x) * gamma + beta [+ residual]).  nn.InstanceNorm3d / nn.BatchNorm3d
    (ResNet.py:236-417, NLayerDiscriminator.py:146-168, UNet.py:84-88) fused with the activation
    that follows it and with the ResnetBlock skip add (ResNet.py:436-443).
    kind None = no normalisation (plain `act(x + residual)`)."""

    @staticmethod
    def forward(ctx, x, kind, mean, rstd, gamma, beta, residual, act, slope, training):
        y = K.norm_apply(x, kind if kind is not None else NORM_INSTANCE, mean, rstd, gamma, beta, residual, act, slope)
        ctx.cfg = (kind, act, slope, training, gamma is not None, residual is not None)
        ctx.save_for_backward(x if kind is not None else None, mean, rstd, gamma, y if act != ACT_NONE else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, mean, rstd, gamma, y = ctx.saved_tensors
        kind, act, slope, training, affine, has_res = ctx.cfg
        if kind is None:
            g = K.act_bwd(y, dy, act, slope) if act != ACT_NONE else dy
            return (g if ctx.needs_input_grad[0] else None, None, None, None, None, None,
                    g if has_res and ctx.needs_input_grad[6] else None, None, None, None)
        dres = None
        if has_res and ctx.needs_input_grad[6]:
            dres = K.act_bwd(y, dy, act, slope) if act != ACT_NONE else dy
        dx, dgamma, dbeta = K.norm_bwd(x, dy, y, kind, training, mean, rstd, gamma, act, slope, want_affine=affine,
                                       need_dx=ctx.needs_input_grad[0])
        return dx, None, None, None, dgamma, dbeta, dres, None, None, None


def norm_act(x, kind, mean, rstd, gamma=None, beta=None, residual=None, act=ACT_NONE, slope=0.0, training=True):
    y = _NormAct.apply(x, kind, mean, rstd, gamma, beta, residual, act, float(slope), bool(training))
    y._rg_dense_grad = True      # the norm backward kernels are fastest on dense gradients: no in-place pad fold upstream
    return y


def add_act(x, residual, act=ACT_NONE, slope=0.0):
    """act(x + residual): residual adds of ResNet.py:436-443, UNet.py:111-117, ResidualUNet.py:689-701."""
    return _NormAct.apply(x, None, None, None, None, None, residual, act, float(slope), False)


def batch_stats(x, kind, eps, precomputed=None):
    """mean / rstd / biased var per group from fresh or epilogue-provided sums."""
    n, c = x.shape[0], x.shape[1]
    vox = x.shape[2] * x.shape[3] * x.shape[4]
    if precomputed is not None:
        s, q = precomputed
        if kind == NORM_BATCH:  # epilogue sums are per (n,c)
            s, q = s.view(n, c).sum(0), q.view(n, c).sum(0)
    else:
        s, q = K.norm_stats(x.detach(), kind)
    count = vox if kind == NORM_INSTANCE else n * vox
    return K.norm_finalize(s, q, count, eps)


# ---------------------------------------------------------------------------------------------
# activation, pooling, up-sampling
# ---------------------------------------------------------------------------------------------
class _Act(torch.autograd.Function):
    """nn.ReLU / LeakyReLU / Tanh / Sigmoid (UNet.py:41-47, NLayerDiscriminator.py:138, BaseSystem.py:175-176)."""

    @staticmethod
    def forward(ctx, x, act, slope):
        y = K.act_fwd(x, act, slope)
        ctx.cfg = (act, slope)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return K.act_bwd(y, dy, *ctx.cfg), None, None


def activation(x, act, slope=0.0):
    return x if act == ACT_NONE else _Act.apply(x, act, float(slope))


class _MaxPool3d(torch.autograd.Function):
    """nn.MaxPool3d(f, stride=f) (UNet.py:209-218)."""

    @staticmethod
    def forward(ctx, x, f, arena=False):
        ctx.f = f
        ctx.save_for_backward(x)
        return K.maxpool_fwd(x, f, arena=arena)

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        return K.maxpool_bwd(x, dy, ctx.f), None, None


def max_pool3d(x, f):
    # training: the pooled tensor feeds the next level's reflect-padded ConvPass -> allocate it as a pad arena
    arena = (TRAIN_ARENA and torch.is_grad_enabled() and x.dtype == torch.bfloat16 and x.shape[1] % 8 == 0
             and K.get_precision() == "bf16")
    return _MaxPool3d.apply(x, _t3(f), arena)


class _UpsampleNearest(torch.autograd.Function):
    """nn.Upsample(scale_factor=f, mode='nearest') (UNet.py:273)."""

    @staticmethod
    def forward(ctx, x, f):
        ctx.f = f
        return K.upsample_fwd(x, f)

    @staticmethod
    def backward(ctx, dy):
        return K.upsample_bwd(dy, ctx.f), None


def upsample_nearest(x, f):
    return _UpsampleNearest.apply(x, _t3(f))


class _ResizeTrilinear(torch.autograd.Function):
    """F.interpolate(x, size, mode="trilinear", align_corners=True) and its adjoint (models/CycleModel.py:24-30)."""

    @staticmethod
    def forward(ctx, x, size):
        ctx.in_size = tuple(x.shape[2:])
        return K.resize_trilinear_fwd(x, size)

    @staticmethod
    def backward(ctx, dy):
        return K.resize_trilinear_bwd(dy, ctx.in_size), None


def resize_trilinear(x, size):
    size = tuple(int(s) for s in size)
    if tuple(x.shape[2:]) == size:
        return x
    return _ResizeTrilinear.apply(x, size)


def sampling_bottleneck(x, scale_factor):
    """CycleModel.sampling_bottleneck for a 3-D network: down by `scale_factor` (output extent floor(in * s), as
    F.interpolate computes it), then back up to the original extents."""
    import math
    sf = (scale_factor,) * 3 if not isinstance(scale_factor, (tuple, list)) else tuple(scale_factor)
    size = tuple(x.shape[2:])
    down = tuple(max(1, int(math.floor(float(s) * float(f)))) for s, f in zip(size, sf))
    return resize_trilinear(resize_trilinear(x, down), size)


# ---------------------------------------------------------------------------------------------
# crop / concat
# ---------------------------------------------------------------------------------------------
def center_crop_view(x, spatial):
    """Centre crop as a strided VIEW (no copy): UNet.py:97-106,327-336; ResNet.py:436-440."""
    sl = [slice(None), slice(None)]
    for have, want in zip(x.shape[2:], spatial):
        o = (have - want) // 2
        sl.append(slice(o, o + want))
    return x[tuple(sl)]


class _CropCat(torch.autograd.Function):
    """torch.cat([crop(f_left), crop(g_up)], dim=1) (UNet.py:338-351) as two strided copies into one
    buffer; backward hands out strided views / zero-padded slices of dy."""

    @staticmethod
    def forward(ctx, a, b, spatial, train_arena=False):
        ctx.train_arena = train_arena
        av, bv = center_crop_view(a, spatial), center_crop_view(b, spatial)
        n = a.shape[0]
        dt = a.dtype if a.dtype == b.dtype else K.compute_dtype()
        shape = (n, a.shape[1] + b.shape[1], *spatial)
        # training: the concat buffer is a pad arena too (the decoder ConvPass behind it pads by mirroring its shell)
        arena = ctx.train_arena and dt == torch.bfloat16 and shape[1] % 8 == 0 and K.get_precision() == "bf16"
        out = K.empty5_arena(shape, dt, a.device) if arena else K.empty5(shape, dt, a.device)
        K.copy_into(out[:, : a.shape[1]], av)
        K.copy_into(out[:, a.shape[1]:], bv)
        ctx.shapes = (tuple(a.shape), tuple(b.shape), tuple(spatial), a.dtype, b.dtype)
        return out

    @staticmethod
    def backward(ctx, dy):
        sa, sb, spatial, da, db = ctx.shapes

        def part(sl, shape, dt):
            g = dy[:, sl]
            if tuple(shape[2:]) == spatial:
                # densified: handing the channel-slice view upstream was measured slower (autograd's skip-gradient add
                # and the up-conv bias reduction leave their vectorised paths on strided operands: +3.8 ms vs 2.6 ms of
                # copies per cfg3 step)
                return K.convert(g, dt)
            full = K.empty5(shape, dt, dy.device)
            full.zero_()
            K.copy_into(center_crop_view(full, spatial), g)
            return full

        ga = part(slice(0, sa[1]), sa, da) if ctx.needs_input_grad[0] else None
        gb = part(slice(sa[1], sa[1] + sb[1]), sb, db) if ctx.needs_input_grad[1] else None
        return ga, gb, None, None


class _CatJoin(torch.autograd.Function):
    """torch.cat([f_left, g_up], dim=1) (UNet.py:349-351, 'same' padding: no crop) when both producers have ALREADY written
    their outputs into the two channel slices of one concat buffer (`mark_train_out`): no data moves in forward; this node
    only ties the buffer to both producers.  Backward = `_CropCat`'s: the two channel slices of dy, densified."""

    @staticmethod
    def forward(ctx, a, b, cat):
        cat = _unbox(cat)
        ca = a.shape[1]
        assert a.data_ptr() == cat[:, :ca].data_ptr() and b.data_ptr() == cat[:, ca:].data_ptr()
        ctx.meta = (ca, b.shape[1], a.dtype, b.dtype)
        return cat

    @staticmethod
    def backward(ctx, dy):
        ca, cb, da, db = ctx.meta
        ga = K.convert(dy[:, :ca], da) if ctx.needs_input_grad[0] else None
        gb = K.convert(dy[:, ca: ca + cb], db) if ctx.needs_input_grad[1] else None
        return ga, gb, None


def cat_join(a, b, cat):
    return _CatJoin.apply(a, b, _Out(cat))


def crop_cat(a, b, spatial):
    return _CropCat.apply(a, b, tuple(int(s) for s in spatial), TRAIN_ARENA and torch.is_grad_enabled())


class _Crop(torch.autograd.Function):
    """centre crop that materialises (used where a contiguous tensor is needed downstream)."""

    @staticmethod
    def forward(ctx, x, spatial):
        ctx.shape = tuple(x.shape)
        ctx.spatial = spatial
        v = center_crop_view(x, spatial)
        return K.copy_into(K.empty5(v.shape, x.dtype, x.device), v)

    @staticmethod
    def backward(ctx, dy):
        full = K.empty5(ctx.shape, dy.dtype, dy.device)
        full.zero_()
        K.copy_into(center_crop_view(full, ctx.spatial), dy)
        return full, None


def center_crop(x, spatial):
    spatial = tuple(int(s) for s in spatial)
    if tuple(x.shape[2:]) == spatial:
        return x
    return _Crop.apply(x, spatial)


# ---------------------------------------------------------------------------------------------
# losses
# ---------------------------------------------------------------------------------------------
class _Loss(torch.autograd.Function):
    """Mean-reduced losses with the gradient produced in the same pass.
    MSE vs a constant label: GANLoss.py:24-25,48-64.  SmoothL1 / L1: LinkCycleLoss.py:22,99-105."""

    @staticmethod
    def forward(ctx, kind, a, b, target, beta):
        want = a.requires_grad or (b is not None and b.requires_grad)
        loss, grad = K.loss_fwd_grad(kind, a.detach(), b.detach() if b is not None else None, target, beta, want)
        ctx.save_for_backward(grad)
        ctx.b_dtype = b.dtype if b is not None else None
        return loss

    @staticmethod
    def backward(ctx, dl):
        (grad,) = ctx.saved_tensors
        ga = gb = None
        if grad is not None:
            # scale by the incoming scalar on the device (no host sync): y = act-less norm_apply w/ gamma=dl
            scale = dl.to(torch.float32).reshape(1)
            if ctx.needs_input_grad[1]:
                ga = _scale(grad, scale)
            if ctx.needs_input_grad[2]:
                gb = _scale(grad, -scale)
        return None, ga, gb, None, None


def _scale(x, scale1):
    """x * scalar (device scalar), via the norm-apply kernel with a per-channel gamma."""
    gamma = scale1.expand(x.shape[1]).contiguous()
    return K.norm_apply(x, NORM_INSTANCE, None, None, gamma, None, None, ACT_NONE, 0.0)


def mse_const_loss(pred, target: float):
    return _Loss.apply(LOSS_MSE_CONST, pred, None, float(target), 1.0)


def mean_loss(pred):
    return _Loss.apply(LOSS_MEAN, pred, None, 0.0, 1.0)


def smooth_l1_loss(a, b, beta: float = 1.0):
    return _Loss.apply(LOSS_SMOOTHL1, a, b, 0.0, float(beta))


def l1_loss(a, b):
    return _Loss.apply(LOSS_L1, a, b, 0.0, 1.0)
