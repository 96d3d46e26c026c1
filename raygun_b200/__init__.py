"""raygun_b200 -- a B200-native (sm_100a) implementation of the dense 3-D convolution hot path of
htem/raygun, behind the unchanged ``raygun.torch.networks`` module API.

    from raygun_b200 import networks            # UNet, ResidualUNet, ResNet, NLayerDiscriminator ...
    from raygun_b200 import set_precision       # 'bf16' (tcgen05 tensor cores) | 'fp32' (parity path)

Host code is Python/PyTorch (memory, streams, autograd bookkeeping, torch.distributed); all arithmetic
on 5-D tensors runs in hand-written CUDA from ``libraygun_b200.so`` through a C ABI
(``include/raygun_b200.h``).  There is no CPU or cuDNN fallback.
"""
from . import networks  # noqa: F401
from . import nn  # noqa: F401
from .kernels import get_precision, precision, set_precision  # noqa: F401

__version__ = "0.1.0"
