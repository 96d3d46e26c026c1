"""PatchGAN discriminator -- drop-in for ``raygun.torch.networks.NLayerDiscriminator``
(reference: networks/NLayerDiscriminator.py:5-202): same kwargs, same ``self.model`` Sequential
layout / ``state_dict`` keys, ``FOV`` property on the 2-D variant.  3-D runs on the sm_100a kernels
(stride-2 convs + LeakyReLU(0.2) fused in the epilogue, norm statistics from the conv epilogue);
2-D is delegated to stock torch.nn.
"""
from __future__ import annotations

import torch
import torch.nn as tnn

from .. import functional as F
from .. import nn as L


def _build(self, ndims, input_nc=1, ngf=64, n_layers=3, norm_layer=None, kw=4, downsampling_kw=None):
    tnn.Module.__init__(self)
    if ndims == 3:
        Conv, inst = L.Conv3d, tnn.InstanceNorm3d
        norm = lambda ch: L.make_norm(norm_layer, ch)
        lrelu = lambda: L.LeakyReLU(0.2, True)
    else:
        Conv, inst = tnn.Conv2d, tnn.InstanceNorm2d
        norm = lambda ch: norm_layer(ch)
        lrelu = lambda: tnn.LeakyReLU(0.2, True)
    # bias is redundant in front of an affine BatchNorm (NLayerDiscriminator.py:124-129)
    use_bias = L.norm_base_class(norm_layer) == inst
    ds_kw = kw if downsampling_kw is None else downsampling_kw
    pad = 1

    seq = [Conv(input_nc, ngf, kernel_size=ds_kw, stride=2, padding=pad), lrelu()]
    mult = 1
    for n in range(1, n_layers):
        prev, mult = mult, min(2 ** n, 8)
        seq += [Conv(ngf * prev, ngf * mult, kernel_size=ds_kw, stride=2, padding=pad, bias=use_bias),
                norm(ngf * mult), lrelu()]
    prev, mult = mult, min(2 ** n_layers, 8)
    seq += [Conv(ngf * prev, ngf * mult, kernel_size=kw, stride=1, padding=pad, bias=use_bias), norm(ngf * mult),
            lrelu()]
    seq += [Conv(ngf * mult, 1, kernel_size=kw, stride=1, padding=pad)]      # 1-channel patch prediction map
    self.model = tnn.Sequential(*seq)
    self._ndims = ndims


def _forward(self, x):
    if self._ndims != 3:
        return self.model(x)
    y = L.run_layers(self.model, F.to_compute(x))
    return F.cast(y, torch.float32)


class NLayerDiscriminator2D(tnn.Module):
    def __init__(self, input_nc=1, ngf=64, n_layers=3, norm_layer=tnn.BatchNorm2d, kw=4, downsampling_kw=None):
        _build(self, 2, input_nc, ngf, n_layers, norm_layer, kw, downsampling_kw)

    @property
    def FOV(self):
        """Receptive field of one output unit (NLayerDiscriminator.py:79-97)."""
        ks = [(m.kernel_size[-1], m.stride[-1]) for m in self.model if hasattr(m, "kernel_size")]
        r = 1
        for k, s in reversed(ks[1:]):
            r = s * r + (k - s)
        return r

    def forward(self, input):
        return _forward(self, input)


class NLayerDiscriminator3D(tnn.Module):
    def __init__(self, input_nc=1, ngf=64, n_layers=3, norm_layer=tnn.BatchNorm3d, kw=4, downsampling_kw=None):
        _build(self, 3, input_nc, ngf, n_layers, norm_layer, kw, downsampling_kw)

    def forward(self, input):
        return _forward(self, input)


class NLayerDiscriminator(NLayerDiscriminator2D, NLayerDiscriminator3D):
    """``NLayerDiscriminator(ndims, **kwargs)`` (NLayerDiscriminator.py:183-202)."""

    def __init__(self, ndims, **kwargs):
        if ndims == 2:
            NLayerDiscriminator2D.__init__(self, **kwargs)
        elif ndims == 3:
            NLayerDiscriminator3D.__init__(self, **kwargs)
        else:
            raise ValueError(ndims, "Only 2D or 3D currently implemented. Feel free to contribute more!")

    def forward(self, input):
        return _forward(self, input)
