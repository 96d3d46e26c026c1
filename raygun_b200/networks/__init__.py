"""Drop-in for ``raygun.torch.networks`` (reference: networks/__init__.py:4-8): the same exported
names, so ``BaseSystem.get_network``'s ``getattr(networks, net_type)`` seam resolves here."""
from .NLayerDiscriminator import NLayerDiscriminator
from .ResidualUNet import ResidualUNet
from .ResNet import ResNet
from .UNet import ConvDownsample, ConvPass, MaxDownsample, UNet, Upsample
from .utils import *  # noqa: F401,F403
