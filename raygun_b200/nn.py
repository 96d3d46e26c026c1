"""Layer modules: subclasses of the torch.nn parameter containers the reference instantiates
(same class lineage, same ``state_dict`` keys, same attributes) whose ``forward`` runs the
hand-written sm_100a kernels.  ``run_layers`` executes a list of such modules with peephole
fusion (pad -> conv -> norm -> activation [+ residual]).
"""
from __future__ import annotations

import functools

import torch
import torch.nn as tnn

from . import conv as CV
from . import functional as F
from . import kernels as K
from ._C import ACT_LEAKY, ACT_NONE, ACT_RELU, ACT_SIGMOID, ACT_TANH, NORM_BATCH, NORM_INSTANCE


def _t3(v):
    return (int(v),) * 3 if isinstance(v, int) else tuple(int(a) for a in v)


# ---------------------------------------------------------------------------------------------
# activations
# ---------------------------------------------------------------------------------------------
def act_code(m):
    """(code, slope) of a recognised activation module, else None (opaque torch module)."""
    if m is None or isinstance(m, tnn.Identity):
        return ACT_NONE, 0.0
    t = type(m)
    if t in (ReLU, tnn.ReLU):
        return ACT_RELU, 0.0
    if t in (LeakyReLU, tnn.LeakyReLU):
        return ACT_LEAKY, float(m.negative_slope)
    if t in (Tanh, tnn.Tanh):
        return ACT_TANH, 0.0
    if t in (Sigmoid, tnn.Sigmoid):
        return ACT_SIGMOID, 0.0
    return None


class ReLU(tnn.ReLU):
    def forward(self, x):
        return F.activation(x, ACT_RELU) if x.dim() == 5 else super().forward(x)


class LeakyReLU(tnn.LeakyReLU):
    def forward(self, x):
        if x.dim() == 5:
            return F.activation(x, ACT_LEAKY, self.negative_slope)
        return super().forward(x)


class Tanh(tnn.Tanh):
    def forward(self, x):
        return F.activation(x, ACT_TANH) if x.dim() == 5 else super().forward(x)


class Sigmoid(tnn.Sigmoid):
    def forward(self, x):
        return F.activation(x, ACT_SIGMOID) if x.dim() == 5 else super().forward(x)


_ACT_CLASSES = {tnn.ReLU: ReLU, tnn.LeakyReLU: LeakyReLU, tnn.Tanh: Tanh, tnn.Sigmoid: Sigmoid}


def make_activation(activation, *args, **kwargs):
    """Reference semantics (UNet.py:41-47): a name in torch.nn, a class/callable, or None.
    Known classes are swapped for the kernel-backed subclasses; anything else runs as given."""
    if activation is None:
        return tnn.Identity()
    cls = getattr(tnn, activation) if isinstance(activation, str) else activation
    if isinstance(cls, type) and cls in _ACT_CLASSES:
        return _ACT_CLASSES[cls](*args, **kwargs)
    if isinstance(cls, functools.partial) and cls.func in _ACT_CLASSES:
        return _ACT_CLASSES[cls.func](*cls.args, *args, **{**cls.keywords, **kwargs})
    return cls(*args, **kwargs)


# ---------------------------------------------------------------------------------------------
# convolutions
# ---------------------------------------------------------------------------------------------
class Conv3d(tnn.Conv3d):
    """nn.Conv3d parameter container; forward = explicit pad kernel (reflect/replicate) + conv kernel."""

    def _pads(self):
        if isinstance(self.padding, str):
            if self.padding == "valid":
                lo = hi = (0, 0, 0)
            else:  # 'same' (stride 1 only, as in torch)
                tot = [d * (k - 1) for d, k in zip(self.dilation, self.kernel_size)]
                lo = tuple(t // 2 for t in tot)
                hi = tuple(t - l for t, l in zip(tot, lo))
        else:
            lo = hi = tuple(self.padding)
        return lo, hi

    def fused(self, x, act=ACT_NONE, slope=0.0, want_stats=False, out=None):
        if tuple(self.dilation) != (1, 1, 1):
            raise NotImplementedError("raygun_b200.Conv3d: dilation is not used by raygun and not implemented")
        lo, hi = self._pads()
        pad = (0, 0, 0)
        if any(lo) or any(hi):
            if self.padding_mode == "zeros" and lo == hi:
                pad = lo
            elif (self.padding_mode in ("reflect", "replicate") and lo == hi and not torch.is_grad_enabled() and x.dim() == 5
                  and x.is_cuda and not want_stats and tuple(self.stride) == (1, 1, 1) and x.dtype == torch.bfloat16
                  and K.arena_of(x, lo, hi) is None and F._stem_eligible(x, self.weight, (1, 1, 1), int(self.groups))
                  and CV.stem_direct_ok(self.weight, False) and all(p_ < s_ for p_, s_ in zip(lo, x.shape[2:]))):
                # inference, the 1 -> ngf stem of a U-Net: the small-K kernel gathers its taps through mirrored / clamped
                # indices, so the explicit F.pad of the reference (UNet.py:61-69) costs no padded copy of the input
                with K.arena_scope(True):
                    return CV.stem_conv_fprop(x, self.weight, self.bias, lo, act, slope, out=out,
                                              pad_mode=F.PAD_MODES[self.padding_mode])
            elif self.padding_mode in ("zeros", "reflect", "replicate"):
                arena = K.arena_of(x, lo, hi) if (x.dim() == 5 and x.is_cuda) else None
                if arena is not None and not torch.is_grad_enabled():
                    # x is the interior of a pad arena: mirror the shell in place, read the padded view (no copy)
                    full, m = arena
                    K.border_fill(full, m, F.PAD_MODES[self.padding_mode])
                    d, h, w = x.shape[2:]
                    x = full[:, :, m - lo[0]: m + d + hi[0], m - lo[1]: m + h + hi[1], m - lo[2]: m + w + hi[2]]
                else:
                    # training: arena interiors take the same in-place route through autograd (F._PadArena); other
                    # inputs are copied, and the adjoint folds the conv's fresh data gradient in place unless the
                    # producer is a norm layer (dense-gradient kernels)
                    x = F.pad3d(x, lo, hi, self.padding_mode,
                                fold=F.PAD_FOLD and not getattr(x, "_rg_dense_grad", False))
            else:
                raise NotImplementedError(f"padding_mode {self.padding_mode!r} is not implemented")
        return F.conv3d(x, self.weight, self.bias, self.stride, pad, self.groups, act, slope, want_stats, out=out)

    def forward(self, x):
        return self.fused(x)


class ConvTranspose3d(tnn.ConvTranspose3d):
    def fused(self, x, act=ACT_NONE, slope=0.0, want_stats=False, out=None):
        if self.groups != 1 or tuple(self.dilation) != (1, 1, 1):
            raise NotImplementedError("raygun_b200.ConvTranspose3d: groups/dilation not implemented")
        return F.conv_transpose3d(x, self.weight, self.bias, self.stride, self.padding, self.output_padding, act,
                                  slope, want_stats, out=out)

    def forward(self, x, output_size=None):
        if output_size is not None:
            raise NotImplementedError("output_size is not supported")
        return self.fused(x)


# ---------------------------------------------------------------------------------------------
# normalisation
# ---------------------------------------------------------------------------------------------
class _NormMixin:
    _kind = NORM_INSTANCE

    def fused(self, x, stats=None, act=ACT_NONE, slope=0.0, residual=None):
        use_batch_stats = self.training or not self.track_running_stats or self.running_mean is None
        gamma = self.weight if self.affine else None
        beta = self.bias if self.affine else None
        n, c = x.shape[0], x.shape[1]
        if use_batch_stats:
            mean, rstd, var = F.batch_stats(x, self._kind, self.eps, stats)
            if self.training and self.track_running_stats and self.running_mean is not None:
                self._update_running(mean, var, x)
            training = True
        else:
            rm, rv = self.running_mean.float(), self.running_var.float()
            mean = rm if self._kind == NORM_BATCH else rm.repeat(n)
            rstd = torch.rsqrt(rv + self.eps)
            rstd = rstd if self._kind == NORM_BATCH else rstd.repeat(n)
            training = False
        return F.norm_act(x, self._kind, mean.contiguous(), rstd.contiguous(), gamma, beta, residual, act, slope,
                          training)

    @torch.no_grad()
    def _update_running(self, mean, var, x):
        # tiny [C]-sized updates; torch semantics: unbiased variance, momentum None = cumulative average
        n, c = x.shape[0], x.shape[1]
        cnt = x.shape[2] * x.shape[3] * x.shape[4] * (n if self._kind == NORM_BATCH else 1)
        if self._kind == NORM_INSTANCE:
            mean, var = mean.view(n, c).mean(0), var.view(n, c).mean(0)
        unbiased = var * (cnt / max(cnt - 1, 1))
        if self.num_batches_tracked is not None:
            self.num_batches_tracked += 1
        m = self.momentum if self.momentum is not None else 1.0 / float(self.num_batches_tracked)
        self.running_mean.mul_(1 - m).add_(mean.to(self.running_mean.dtype), alpha=m)
        self.running_var.mul_(1 - m).add_(unbiased.to(self.running_var.dtype), alpha=m)


class InstanceNorm3d(_NormMixin, tnn.InstanceNorm3d):
    _kind = NORM_INSTANCE

    def forward(self, x):
        return self.fused(x)


class BatchNorm3d(_NormMixin, tnn.BatchNorm3d):
    _kind = NORM_BATCH

    def forward(self, x):
        return self.fused(x)


_NORM_CLASSES = {tnn.InstanceNorm3d: InstanceNorm3d, tnn.BatchNorm3d: BatchNorm3d}


def make_norm(norm_layer, channels):
    """norm_layer: a class, a functools.partial of one, or any callable (BaseSystem.py:193-199,
    read_config '#partial(...)#' strings).  torch's 3-D norm classes map to kernel-backed subclasses."""
    if isinstance(norm_layer, functools.partial) and norm_layer.func in _NORM_CLASSES:
        return _NORM_CLASSES[norm_layer.func](channels, *norm_layer.args, **norm_layer.keywords)
    if isinstance(norm_layer, type) and norm_layer in _NORM_CLASSES:
        return _NORM_CLASSES[norm_layer](channels)
    return norm_layer(channels)


def norm_base_class(norm_layer):
    return norm_layer.func if isinstance(norm_layer, functools.partial) else norm_layer


# ---------------------------------------------------------------------------------------------
# padding / pooling / upsampling
# ---------------------------------------------------------------------------------------------
class ReflectionPad3d(tnn.ReflectionPad3d):
    def forward(self, x):
        p = self.padding  # (wl, wr, hl, hr, dl, dr)
        return F.pad3d(x, (p[4], p[2], p[0]), (p[5], p[3], p[1]), "reflect")


class ReplicationPad3d(tnn.ReplicationPad3d):
    def forward(self, x):
        p = self.padding
        return F.pad3d(x, (p[4], p[2], p[0]), (p[5], p[3], p[1]), "replicate")


class MaxPool3d(tnn.MaxPool3d):
    def forward(self, x):
        k, s = _t3(self.kernel_size), _t3(self.stride)
        if k != s or _t3(self.padding) != (0, 0, 0) or _t3(self.dilation) != (1, 1, 1) or self.ceil_mode:
            raise NotImplementedError("raygun_b200.MaxPool3d supports kernel == stride, no padding (as raygun uses it)")
        if any(i < f for i, f in zip(x.shape[2:], k)):
            raise RuntimeError(f"max_pool3d: input {tuple(x.shape)} smaller than the pooling window {k}")
        return F.max_pool3d(x, k)


class Upsample(tnn.Upsample):
    def forward(self, x):
        if self.mode != "nearest" or x.dim() != 5:
            return super().forward(x)
        f = self.scale_factor
        f = (f,) * 3 if not isinstance(f, (tuple, list)) else tuple(f)
        if any(float(v) != int(v) for v in f):
            return super().forward(x)
        return F.upsample_nearest(x, tuple(int(v) for v in f))


# ---------------------------------------------------------------------------------------------
# fused execution of a layer list
# ---------------------------------------------------------------------------------------------
def run_layers(layers, x, residual=None, out=None):
    """Run `layers` (an iterable of modules) on x with peephole fusion:
         conv -> [norm] -> [activation]   => conv epilogue (stats / activation) + one norm-apply pass
       `residual`, if given, is added right after the LAST layer, which must then be a norm or conv
       (ResnetBlock: x + conv_block(x)); it is fused into that norm-apply pass."""
    layers = list(layers)
    i, n = 0, len(layers)
    while i < n:
        m = layers[i]
        nxt = layers[i + 1] if i + 1 < n else None
        if isinstance(m, (Conv3d, ConvTranspose3d)):
            if isinstance(nxt, _NormMixin):
                y, st = m.fused(x, want_stats=True)
                a = layers[i + 2] if i + 2 < n else None
                ac = act_code(a) if a is not None else None
                # the residual is added right after the LAST layer: fused only when that layer is this norm (a list that
                # ends in norm -> activation keeps act(norm(x)) + residual: the trailing add below)
                last = (i + 2 >= n) and ac is None
                res = residual if last else None
                if ac is not None:
                    x = nxt.fused(y, st, ac[0], ac[1], residual=res)
                    i += 3
                else:
                    x = nxt.fused(y, st, residual=res)
                    i += 2
                if res is not None:
                    residual = None
                continue
            ac = act_code(nxt) if nxt is not None else None
            if ac is not None and ac[0] != ACT_NONE:
                x = m.fused(x, ac[0], ac[1], out=out if (i + 2 >= n and residual is None) else None)
                i += 2
                continue
            x = m.fused(x, out=out if (i + 1 >= n and residual is None) else None)
            i += 1
            continue
        x = m(x)
        i += 1
    if residual is not None:
        x = F.add_act(x, residual) if x.dim() == 5 else x + residual
    return x
