"""Model mirrors that sit next to the networks in the reference and reuse them unchanged.

``MTLSDModel`` (reference: ``models/MTLSDModel.py:12-63``): the affinity + LSD segmentation model is the SAME
``networks.UNet`` followed by two 1x1x1 ``ConvPass`` heads with a Sigmoid (``aff_head``: ngf -> num_affs, ``lsd_head``: ngf -> 10).
Both heads read the decoder output once each through the streaming 1x1x1 kernel (<= 4 channels, `rg_pw_conv_fprop`, bias + Sigmoid
fused) or the direct kernel (10 channels), with the activation in the convolution epilogue; the module tree and state_dict keys
(``unet.*``, ``aff_head.conv_pass.0.*``, ``lsd_head.conv_pass.0.*``) are the reference's, so its checkpoints load strictly.
The MTLSD / ACLSD *systems* (gunpowder pipelines, weighted-MSE training loop) stay out of scope."""
from __future__ import annotations

import torch

from . import functional as F
from .networks import UNet
from .networks.UNet import ConvPass

_DEFAULT_UNET = {"input_nc": 1, "ngf": 12, "fmap_inc_factor": 6, "downsample_factors": [(2, 2, 2), (2, 2, 2), (2, 2, 2)],
                 "constant_upsample": True}


class MTLSDModel(torch.nn.Module):
    def __init__(self, unet_kwargs=None, num_affs=3):
        super().__init__()
        unet_kwargs = dict(_DEFAULT_UNET if unet_kwargs is None else unet_kwargs)
        self.unet = UNet(**unet_kwargs)
        self.aff_head = ConvPass(unet_kwargs["ngf"], num_affs, [[1, 1, 1]], activation="Sigmoid")
        self.lsd_head = ConvPass(unet_kwargs["ngf"], 10, [[1, 1, 1]], activation="Sigmoid")
        self.output_arrays = ["pred_affs", "pred_lsds"]
        self.data_dict = {}

    def forward(self, raw):
        self.data_dict.update({"raw": raw.detach()})
        self.unet._keep_compute_dtype = True          # the heads consume the decoder output in the compute dtype
        try:
            z = self.unet(raw)
        finally:
            self.unet._keep_compute_dtype = False
        affs = F.cast(self.aff_head(z), torch.float32)
        lsds = F.cast(self.lsd_head(z), torch.float32)
        return affs, lsds
