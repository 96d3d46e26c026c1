"""Residual U-Net -- drop-in for ``raygun.torch.networks.ResidualUNet`` (reference:
networks/ResidualUNet.py:364-701): the U-Net of ``UNet.py`` whose last decoder ConvPass omits its
trailing activation (``final``, :93-97,:636) plus a global input skip ``act(crop(x) + y)`` (:689-701),
which here is one fused add+activation kernel per head."""
from __future__ import annotations

import torch
import torch.nn as tnn

from .. import functional as F
from .. import nn as L
from .UNet import ConvDownsample, ConvPass, MaxDownsample, Upsample, UNet, _activation, _build_unet, _center_crop

__all__ = ["ResidualUNet", "ConvPass", "ConvDownsample", "MaxDownsample", "Upsample"]


class ResidualUNet(UNet):
    def __init__(self, input_nc, ngf, fmap_inc_factor, downsample_factors, kernel_size_down=None, kernel_size_up=None,
                 activation="ReLU", output_nc=None, num_heads=1, constant_upsample=False, downsample_method="max",
                 padding_type="valid", residual=False, norm_layer=None, add_noise=False):
        tnn.Module.__init__(self)
        self.padding_type = padding_type
        self.activation = _activation(activation, len(downsample_factors[0]))
        _build_unet(self, input_nc, ngf, fmap_inc_factor, downsample_factors, kernel_size_down, kernel_size_up,
                    activation, output_nc, num_heads, constant_upsample, downsample_method, padding_type, residual,
                    norm_layer, add_noise, final_level0=True)

    def crop(self, x, shape):
        return _center_crop(x, tuple(shape))

    def forward(self, x):
        if self.ndims != 3:
            y = self.rec_forward(self.num_levels - 1, x)
            if self.padding_type.lower() == "valid":
                x = self.crop(x, y[0].shape[-self.ndims:])
            y = [self.activation(x + yi) for yi in y]
            return y[0] if self.num_heads == 1 else y
        xc = F.to_compute(x)
        ys = self.rec_forward(self.num_levels - 1, xc)
        skip = self.crop(xc, ys[0].shape[-3:]) if self.padding_type.lower() == "valid" else xc
        if skip.shape[1] != ys[0].shape[1]:
            if skip.shape[1] != 1:
                raise RuntimeError(f"ResidualUNet: input channels {skip.shape[1]} cannot broadcast to {ys[0].shape[1]}")
            skip = skip.expand(-1, ys[0].shape[1], -1, -1, -1)   # stride-0 view; the kernel is stride-generic
        ac = L.act_code(self.activation)
        outs = []
        for y in ys:
            if ac is not None:
                outs.append(F.add_act(y, skip, ac[0], ac[1]))
            else:
                outs.append(self.activation(F.add_act(y, skip)))
        return self._finish(outs)
