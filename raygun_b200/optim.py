"""Optimiser of the training step: Adam over FLAT fp32 buckets, one fused kernel per step.

Replaces the ``torch.optim.Adam`` call sites of the reference (``losses/LinkCycleLoss.py:78,128``; wiring
``systems/CycleGAN.py:167-197`` through ``optimizers/utils.py:get_base_optimizer``), for the subset raygun configures
(``lr``, ``betas``, ``eps``; no amsgrad, no weight decay).

Layout in HBM: every parameter of the group becomes a VIEW of one flat fp32 buffer (``state_dict`` keys, shapes and
dtypes are untouched; ``load_state_dict`` copies in place), and so does every ``.grad``.  That buys
  * one ``rg_adam_flat`` launch per optimiser step (m, v, p update; 16 B read + 12 B written per parameter) instead of a
    multi-tensor foreach, with the step counter on the device so the step replays inside a CUDA graph;
  * weight-gradient kernels that write straight into the bucket (``conv.conv_wgrad(dw_out=..., accumulate=True)``): no
    per-parameter gradient tensors, no AccumulateGrad adds, one memset per ``zero_grad``;
  * a gradient bucket NCCL can all-reduce in place, chunk by chunk (``dist.GradSync``), with the 1/world of the averaging
    folded into the update (``grad_scale``);
  * the bf16 tensor-core packs of the updated weights re-packed in place right after the update (``conv.repack``), inside
    the same captured graph.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import conv as CV
from ._C import check, count_launch, lib, ptr, require_device, stream_ptr

ALIGN = 4      # fp32 elements: every view starts 16-byte aligned


class FlatAdam(torch.optim.Optimizer):
    """Adam (torch.optim.Adam semantics, bias-corrected, eps outside the square root) over a flat bucket.

    The parameters must already live on their CUDA device: construction re-homes ``p.data`` and ``p.grad`` into the
    buckets (a later ``module.to(...)`` would detach them again).  ``zero_grad`` clears the bucket and KEEPS the views,
    whatever ``set_to_none`` says -- the gradient kernels accumulate into them."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, grad_scale=1.0):
        if weight_decay or amsgrad:     # default_cycleGAN_conf.json passes weight_decay = 0 explicitly
            raise NotImplementedError("FlatAdam: weight_decay / amsgrad are not implemented (raygun's configs use neither)")
        params = list(params)
        if not params:
            raise ValueError("FlatAdam: empty parameter list")
        super().__init__(params, dict(lr=float(lr), betas=tuple(betas), eps=float(eps)))
        self.params = [p for g in self.param_groups for p in g["params"]]
        dev = self.params[0].device
        for p in self.params:
            if p.device != dev or p.dtype != torch.float32:
                raise ValueError("FlatAdam: parameters must be fp32 tensors on one device")
        require_device(self.params[0])
        self.offsets, total = [], 0
        for p in self.params:
            self.offsets.append(total)
            total += (p.numel() + ALIGN - 1) // ALIGN * ALIGN
        self.numel = total
        self.flat_p = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_m = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_v = torch.zeros(total, dtype=torch.float32, device=dev)
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        self.grad_scale = float(grad_scale)
        with torch.no_grad():
            for p, off in zip(self.params, self.offsets):
                n = p.numel()
                view = self.flat_p[off: off + n].view(p.shape)
                view.copy_(p.data)
                p.data = view
                g = self.flat_g[off: off + n].view(p.shape)
                p.grad = g
                p._rg_bucket_grad = g        # functional.* writes weight gradients here and returns None to autograd
        CV.invalidate_packs()                # the parameters moved: every derived pack is keyed on the data pointer

    def zero_grad(self, set_to_none: bool = True):
        self.flat_g.zero_()
        for p, off in zip(self.params, self.offsets):      # re-attach views a caller may have dropped (p.grad = None)
            if p.grad is None or p.grad.data_ptr() != p._rg_bucket_grad.data_ptr():
                p.grad = p._rg_bucket_grad

    @torch.no_grad()
    def step(self, closure=None):
        if closure is not None:
            raise NotImplementedError("FlatAdam.step: closures are not supported")
        g = self.param_groups[0]
        b1, b2 = g["betas"]
        check(lib().rg_adam_flat(ptr(self.flat_p), ptr(self.flat_g), ptr(self.flat_m), ptr(self.flat_v),
                                 C.c_int64(self.numel), C.c_float(g["lr"]), C.c_float(b1), C.c_float(b2), C.c_float(g["eps"]),
                                 ptr(self.step_dev), C.c_float(self.grad_scale), stream_ptr(self.flat_p.device)),
              "rg_adam_flat")
        count_launch(2)
        # the fp32 masters changed behind Python's version counters: refresh the bf16 tensor-core packs in place
        CV.repack(self.params)

    def state_dict(self):
        return {"flat_m": self.flat_m, "flat_v": self.flat_v, "step": self.step_dev,
                "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]}

    def load_state_dict(self, sd):
        self.flat_m.copy_(sd["flat_m"])
        self.flat_v.copy_(sd["flat_v"])
        self.step_dev.copy_(sd["step"])


def bucket_grad(w):
    """The gradient-bucket view of parameter `w` when it is live (FlatAdam owns `w` and `w.grad` still is that view)."""
    b = getattr(w, "_rg_bucket_grad", None)
    if b is None:
        return None
    g = w.grad
    return b if g is not None and g.data_ptr() == b.data_ptr() else None
