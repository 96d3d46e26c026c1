"""ResNet generator -- drop-in for ``raygun.torch.networks.ResNet`` (reference: networks/ResNet.py).

Same constructor kwargs, same ``self.model`` Sequential (indices, ``len``) and therefore the same
``state_dict`` keys.  The 3-D variant is built from the kernel-backed layers of ``raygun_b200.nn``
and executed with conv/norm/activation/residual fusion; the 2-D variant (outside the 3-D hot path,
SURVEY.md 8 a9) is assembled from stock ``torch.nn`` layers and simply delegated.
"""
from __future__ import annotations

import torch
import torch.nn as tnn

from .. import functional as F
from .. import nn as L
from .utils import NoiseBlock, ParameterizedNoiseBlock


class _Layers:
    """The layer vocabulary of one dimensionality."""

    def __init__(self, ndims):
        self.ndims = ndims
        if ndims == 3:
            self.Conv, self.ConvT = L.Conv3d, L.ConvTranspose3d
            self.ReflPad, self.ReplPad = L.ReflectionPad3d, L.ReplicationPad3d
            self.InstanceNorm = tnn.InstanceNorm3d
            self.Dropout = tnn.Dropout
        else:
            self.Conv, self.ConvT = tnn.Conv2d, tnn.ConvTranspose2d
            self.ReflPad, self.ReplPad = tnn.ReflectionPad2d, tnn.ReplicationPad2d
            self.InstanceNorm = tnn.InstanceNorm2d
            self.Dropout = tnn.Dropout

    def norm(self, norm_layer, ch):
        return L.make_norm(norm_layer, ch) if self.ndims == 3 else norm_layer(ch)

    def act(self, activation):
        return L.make_activation(activation) if self.ndims == 3 else activation()

    def padder(self, padding_type, width):
        """-> (list of explicit pad modules, conv padding argument); ResNet.py:264-279,388-401."""
        pt = padding_type.lower()
        if pt in ("reflect", "same"):
            return [self.ReflPad(width)], 0
        if pt == "replicate":
            return [self.ReplPad(width)], 0
        if pt == "zeros":
            return [], width
        if pt == "valid":
            return [], "valid"
        return None, None


def _uses_bias(norm_layer, instance_cls):
    # ResNet.py:259-262 -- bias only when the norm is exactly InstanceNorm (or a partial of it)
    return L.norm_base_class(norm_layer) == instance_cls


class _ResnetBlock(tnn.Module):
    """x + conv_block(x), cropping x for valid padding (ResNet.py:347-443)."""

    def __init__(self, ndims, dim, padding_type, norm_layer, use_dropout, use_bias, activation=tnn.ReLU):
        super().__init__()
        self._ls = _Layers(ndims)
        self.conv_block = self.build_conv_block(dim, padding_type, norm_layer, use_dropout, use_bias, activation)
        self.padding_type = padding_type

    def build_conv_block(self, dim, padding_type, norm_layer, use_dropout, use_bias, activation=tnn.ReLU):
        ls = self._ls
        pads, p = ls.padder(padding_type, 1)
        if pads is None:
            raise NotImplementedError("padding [%s] is not implemented" % padding_type)
        block = list(pads)
        block += [ls.Conv(dim, dim, kernel_size=3, padding=p, bias=use_bias), ls.norm(norm_layer, dim),
                  ls.act(activation)]
        if use_dropout:
            block += [ls.Dropout(0.2)]
        block += list(ls.padder(padding_type, 1)[0])
        block += [ls.Conv(dim, dim, kernel_size=3, padding=p, bias=use_bias), ls.norm(norm_layer, dim)]
        return tnn.Sequential(*block)

    def crop(self, x, shape):
        nd = self._ls.ndims
        sl = [slice(None)] * (x.dim() - nd)
        for have, want in zip(x.shape[-nd:], shape):
            o = (have - want) // 2
            sl.append(slice(o, o + want))
        return x[tuple(sl)]

    def forward(self, x):
        if self._ls.ndims != 3:
            res = self.conv_block(x)
            return (self.crop(x, res.shape[-2:]) if self.padding_type == "valid" else x) + res
        if self.padding_type == "valid":
            res = L.run_layers(self.conv_block, x)
            return F.add_act(res, self.crop(x, res.shape[-3:]))
        return L.run_layers(self.conv_block, x, residual=x)     # skip add fused into the last norm pass


class ResnetBlock3D(_ResnetBlock):
    def __init__(self, dim, padding_type, norm_layer, use_dropout, use_bias, activation=tnn.ReLU):
        super().__init__(3, dim, padding_type, norm_layer, use_dropout, use_bias, activation)


class ResnetBlock2D(_ResnetBlock):
    def __init__(self, dim, padding_type, norm_layer, use_dropout, use_bias, activation=tnn.ReLU):
        super().__init__(2, dim, padding_type, norm_layer, use_dropout, use_bias, activation)


def _build_generator(self, ndims, input_nc=1, output_nc=None, ngf=64, norm_layer=None, use_dropout=False, n_blocks=6,
                     padding_type="reflect", activation=tnn.ReLU, add_noise=False, n_downsampling=2):
    assert n_blocks >= 0
    tnn.Module.__init__(self)
    ls = _Layers(ndims)
    if norm_layer is None:
        norm_layer = ls.InstanceNorm
    use_bias = _uses_bias(norm_layer, ls.InstanceNorm)
    if output_nc is None:
        output_nc = ngf
    pads, p = ls.padder(padding_type, 3)
    if pads is None:            # unknown padding: the reference silently falls through to p = 0
        pads, p = [], 0
    updown_p = 0 if padding_type.lower() == "valid" else 1
    Block = ResnetBlock3D if ndims == 3 else ResnetBlock2D

    model = list(pads)
    model += [ls.Conv(input_nc, ngf, kernel_size=7, padding=p, bias=use_bias), ls.norm(norm_layer, ngf), ls.act(activation)]
    for i in range(n_downsampling):
        c = ngf * 2 ** i
        model += [ls.Conv(c, c * 2, kernel_size=3, stride=2, padding=updown_p, bias=use_bias),
                  ls.norm(norm_layer, c * 2), ls.act(activation)]
    c = ngf * 2 ** n_downsampling
    for _ in range(n_blocks):
        model.append(Block(c, padding_type=padding_type.lower(), norm_layer=norm_layer, use_dropout=use_dropout,
                           use_bias=use_bias, activation=activation))
    if add_noise == "param":
        model.append(ParameterizedNoiseBlock())
    elif add_noise:
        model.append(NoiseBlock())
    for i in range(n_downsampling):
        c = ngf * 2 ** (n_downsampling - i)
        extra = int(i == 0 and (add_noise is not False))
        model += [ls.ConvT(c + extra, int(c / 2), kernel_size=3, stride=2, padding=updown_p,
                           output_padding=updown_p, bias=use_bias),
                  ls.norm(norm_layer, int(c / 2)), ls.act(activation)]
    model += list(ls.padder(padding_type, 3)[0] or [])
    # reference quirk: the 2-D generator's last conv emits ngf channels, ignoring output_nc
    model.append(ls.Conv(ngf, output_nc if ndims == 3 else ngf, kernel_size=7, padding=p))
    self.model = tnn.Sequential(*model)
    self._ndims = ndims


class ResnetGenerator3D(tnn.Module):
    """ResNet.py:226-344."""

    def __init__(self, input_nc=1, output_nc=None, ngf=64, norm_layer=tnn.InstanceNorm3d, use_dropout=False,
                 n_blocks=6, padding_type="reflect", activation=tnn.ReLU, add_noise=False, n_downsampling=2):
        _build_generator(self, 3, input_nc, output_nc, ngf, norm_layer, use_dropout, n_blocks, padding_type,
                         activation, add_noise, n_downsampling)

    def forward(self, input):
        return _forward(self, input)


class ResnetGenerator2D(tnn.Module):
    """ResNet.py:6-124 (delegated to stock torch.nn)."""

    def __init__(self, input_nc=1, output_nc=None, ngf=64, norm_layer=tnn.InstanceNorm2d, use_dropout=False,
                 n_blocks=6, padding_type="reflect", activation=tnn.ReLU, add_noise=False, n_downsampling=2):
        _build_generator(self, 2, input_nc, output_nc, ngf, norm_layer, use_dropout, n_blocks, padding_type,
                         activation, add_noise, n_downsampling)

    def forward(self, input):
        return _forward(self, input)


def _forward(self, x):
    if self._ndims != 3:
        return self.model(x)
    y = L.run_layers(self.model, F.to_compute(x))
    return F.cast(y, torch.float32) if not getattr(self, "_keep_compute_dtype", False) else y


class ResNet(ResnetGenerator2D, ResnetGenerator3D):
    """Dispatcher with the reference's signature ``ResNet(ndims, **kwargs)`` (ResNet.py:446-473)."""

    def __init__(self, ndims, **kwargs):
        if ndims == 2:
            ResnetGenerator2D.__init__(self, **kwargs)
        elif ndims == 3:
            ResnetGenerator3D.__init__(self, **kwargs)
        else:
            raise ValueError(ndims, "Only 2D or 3D currently implemented. Feel free to contribute more!")

    def forward(self, input):
        return _forward(self, input)
