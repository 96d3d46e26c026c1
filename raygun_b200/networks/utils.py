"""Helpers of ``raygun.torch.networks.utils`` (reference: networks/utils.py:7-105), same names and
behaviour, so ``init_weights`` / ``set_norm_mode`` / ``get_running_norm_stats`` keep working on the
kernel-backed modules (they match layers by class NAME: our norm classes keep "Norm" in their names,
our conv classes keep "Conv")."""
from __future__ import annotations

import torch
from torch.nn import init as _init

__all__ = ["get_norm_layers", "get_running_norm_stats", "set_norm_mode", "init_weights", "NoiseBlock",
           "ParameterizedNoiseBlock"]


def _is_norm(module) -> bool:
    return "norm" in type(module).__name__.lower()


def get_norm_layers(net):
    """utils.py:7-8."""
    return [m for m in net.modules() if _is_norm(m)]


def get_running_norm_stats(net):
    """Concatenated running means / variances of every norm layer; (None, None) if any layer
    does not track them (utils.py:11-28)."""
    layers = get_norm_layers(net)
    try:
        return (torch.cat([l.running_mean for l in layers]), torch.cat([l.running_var for l in layers]))
    except Exception:
        return None, None


def set_norm_mode(net, mode="train"):
    """'train' | 'eval' | 'fix_norms' (train mode with every norm layer frozen) -- utils.py:31-42."""
    if mode == "eval":
        net.eval()
    elif mode in ("train", "fix_norms"):
        net.train()
        if mode == "fix_norms":
            for m in net.modules():
                if _is_norm(m):
                    m.eval()


_INITIALISERS = {
    "normal": lambda w, gain: _init.normal_(w, 0.0, gain),
    "xavier": lambda w, gain: _init.xavier_normal_(w, gain=gain),
    "kaiming": lambda w, gain: _init.kaiming_normal_(w, a=0, mode="fan_in"),
    "orthogonal": lambda w, gain: _init.orthogonal_(w, gain=gain),
}


def init_weights(net, init_type="normal", init_gain=0.02, nonlinearity="relu"):
    """Re-initialise Conv*/Linear weights (biases -> 0); BatchNorm2d only among the norms -- the
    reference never touches 3-D norm layers (utils.py:45-81)."""
    def visit(m):
        name = type(m).__name__
        if hasattr(m, "weight") and ("Conv" in name or "Linear" in name):
            if init_type not in _INITIALISERS:
                raise NotImplementedError("initialization method [%s] is not implemented" % init_type)
            _INITIALISERS[init_type](m.weight.data, init_gain)
            if getattr(m, "bias", None) is not None:
                _init.constant_(m.bias.data, 0.0)
        elif "BatchNorm2d" in name:
            _init.normal_(m.weight.data, 1.0, init_gain)
            _init.constant_(m.bias.data, 0.0)

    print("initialize network with %s" % init_type)
    net.apply(visit)


def _append_channel(x, noise):
    """cat([x, noise], 1) with the product's copy kernels for 5-D CUDA tensors."""
    if x.dim() == 5 and x.is_cuda:
        from .. import functional as F
        return F.crop_cat(x, F.cast(noise, x.dtype), tuple(x.shape[2:]))
    return torch.cat([x, noise], 1)


class NoiseBlock(torch.nn.Module):
    """Appends one N(0,1) feature map (utils.py:84-94); the RNG stays PyTorch's."""

    def forward(self, x):
        shape = list(x.shape)
        shape[1] = 1
        noise = torch.empty(shape, device=x.device).normal_()
        return _append_channel(x, noise.requires_grad_())


class ParameterizedNoiseBlock(torch.nn.Module):
    """Appends noise ~ N(x[:,0], relu(x[:,1])) (utils.py:97-105)."""

    def forward(self, x):
        noise = torch.normal(x[:, 0, ...].float(), torch.relu(x[:, 1, ...].float())).unsqueeze(1)
        return _append_channel(x, noise.requires_grad_())
