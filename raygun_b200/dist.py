"""Data-parallel plumbing: one process per GPU, torch.distributed (NCCL over NVLink 5 / NVSwitch; gloo on CPU).

The reference has no distributed code at all.  CycleGAN training shards over the batch: InstanceNorm
is per sample and every loss is a batch mean, so the average of the ranks' shard gradients IS the full-batch
gradient.  The exchange happens where the reference steps its optimizers -- inside LinkCycleLoss
(losses/LinkCycleLoss.py:78,128) -- through the ``grad_sync`` hook of ``raygun_b200.cyclegan.LinkCycleLoss``.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def init_from_env(backend: str | None = None):
    """Initialise the default process group from torchrun's environment; returns (rank, world, local_rank)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend, **kw)
    return rank, world, local


class GradSync:
    """Average the gradients of parameter groups across ranks with ONE all-reduce per group.

    groups: {"G": [params of G1 and G2], "D": [params of D1 and D2]}.  ``sync("G")`` flattens the existing
    ``.grad`` tensors into a persistent fp32 bucket, all-reduces it (sum), scales by 1/world and scatters the
    result back in place.  Parameters without a gradient contribute zeros (so every rank reduces the same bytes)."""

    def __init__(self, groups, group=None):
        self.groups = {k: [p for p in v] for k, v in groups.items()}
        self.pg = group
        self.buckets = {}
        self.bytes_reduced = 0

    def _bucket(self, name):
        if name not in self.buckets:
            ps = self.groups[name]
            n = sum(p.numel() for p in ps)
            self.buckets[name] = torch.zeros(n, dtype=torch.float32, device=ps[0].device)
        return self.buckets[name]

    def __call__(self, name):
        if not dist.is_initialized() or dist.get_world_size(self.pg) == 1:
            return
        flat = self._bucket(name)
        off = 0
        for p in self.groups[name]:
            n = p.numel()
            if p.grad is None:
                flat[off: off + n].zero_()
            else:
                flat[off: off + n].copy_(p.grad.reshape(-1))
            off += n
        dist.all_reduce(flat, group=self.pg)
        flat.div_(dist.get_world_size(self.pg))
        self.bytes_reduced += flat.numel() * 4
        off = 0
        for p in self.groups[name]:
            n = p.numel()
            if p.grad is not None:
                p.grad.copy_(flat[off: off + n].view_as(p.grad))
            off += n

    sync = __call__


def shard_batch(x: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """Equal contiguous shard of the batch dimension (global batch must divide by the world size)."""
    if x.shape[0] % world:
        raise ValueError(f"global batch {x.shape[0]} is not divisible by world size {world}")
    per = x.shape[0] // world
    return x[rank * per: (rank + 1) * per]
