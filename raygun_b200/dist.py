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
    """Average the gradients of parameter groups across ranks.

    groups: {"G": optimizer or [params of G1 and G2], "D": ...}.

    * A group given as an ``optim.FlatAdam`` is reduced IN PLACE in its flat gradient bucket (every ``.grad`` is a view of
      it): ``begin(name)`` issues one asynchronous all-reduce per ~25 MB chunk -- NCCL runs them on its own stream, after
      everything queued so far, concurrently with whatever the caller launches next (LinkCycleLoss launches the whole
      discriminator phase, which does not depend on the generator update: the exchange is hidden behind its forward and
      backward) -- and ``finish(name)`` makes the current stream wait for them.  The 1/world of the average is folded
      into the Adam update (``FlatAdam.grad_scale``): no scaling pass.  Both calls are capturable in a CUDA graph.
    * A group given as a parameter list (any other optimizer) is flattened into a persistent bucket, reduced with one
      blocking all-reduce, scaled and scattered back: ``sync(name)``.
    ``__call__(name)`` = begin + finish (or sync)."""

    CHUNK_BYTES = 25 << 20

    def __init__(self, groups, group=None, chunk_bytes=None):
        self.pg = group
        self.flat, self.groups = {}, {}
        for k, v in groups.items():
            if hasattr(v, "flat_g"):
                self.flat[k] = v
                if dist.is_initialized():
                    v.grad_scale = 1.0 / dist.get_world_size(self.pg)
            else:
                self.groups[k] = [p for p in v]
        self.chunk = int(chunk_bytes or self.CHUNK_BYTES) // 4
        self.buckets = {}
        self.pending = {}
        self.bytes_reduced = 0

    def _active(self):
        return dist.is_initialized() and dist.get_world_size(self.pg) > 1

    # -- flat buckets: asynchronous, chunked ----------------------------------------------------------------------
    def begin(self, name):
        if not self._active():
            return
        if name not in self.flat:
            return self.sync(name)
        flat = self.flat[name].flat_g
        works = []
        for off in range(0, flat.numel(), self.chunk):
            works.append(dist.all_reduce(flat[off: off + self.chunk], group=self.pg, async_op=True))
        self.pending[name] = works
        self.bytes_reduced += flat.numel() * 4

    def finish(self, name):
        for w in self.pending.pop(name, ()):
            w.wait()

    # -- parameter lists: one blocking bucket ---------------------------------------------------------------------
    def _bucket(self, name):
        if name not in self.buckets:
            ps = self.groups[name]
            n = sum(p.numel() for p in ps)
            self.buckets[name] = torch.zeros(n, dtype=torch.float32, device=ps[0].device)
        return self.buckets[name]

    def sync(self, name):
        if not self._active():
            return
        if name in self.flat:
            self.begin(name)
            return self.finish(name)
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

    __call__ = sync


def shard_batch(x: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """Equal contiguous shard of the batch dimension (global batch must divide by the world size)."""
    if x.shape[0] % world:
        raise ValueError(f"global batch {x.shape[0]} is not divisible by world size {world}")
    per = x.shape[0] // world
    return x[rank * per: (rank + 1) * per]


# ---------------------------------------------------------------------------------------------
# Spatial sharding building block (spatial model parallelism for networks whose single sample does not fit
# one GPU, e.g. Deep-UNet).  NOT wired into the networks and not measured on NVLink: the exchange primitive and its
# adjoint only, verified under gloo (tests/test_dist_cpu.py).
# ---------------------------------------------------------------------------------------------
def _exchange(send_prev, send_next, recv_prev_like, recv_next_like, rank, world, group):
    """Send `send_prev` to rank - 1 and `send_next` to rank + 1; receive the matching slabs.  Missing neighbours (the two ends
    of the chain) and empty slabs are skipped.  Returns (from_prev, from_next), None where nothing was received."""
    ops, from_prev, from_next = [], None, None
    if rank > 0 and recv_prev_like is not None:
        from_prev = torch.empty_like(recv_prev_like)
        ops.append(dist.P2POp(dist.irecv, from_prev, rank - 1, group))
    if rank + 1 < world and recv_next_like is not None:
        from_next = torch.empty_like(recv_next_like)
        ops.append(dist.P2POp(dist.irecv, from_next, rank + 1, group))
    if rank > 0 and send_prev is not None:
        ops.append(dist.P2POp(dist.isend, send_prev.contiguous(), rank - 1, group))
    if rank + 1 < world and send_next is not None:
        ops.append(dist.P2POp(dist.isend, send_next.contiguous(), rank + 1, group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return from_prev, from_next


class _HaloExchange(torch.autograd.Function):
    """y = x extended along `dim` by `lo` planes taken from the previous rank's END and `hi` planes from the next rank's START.
    The two ranks at the ends of the chain get zeros there (the caller applies the network's own padding at the volume faces).
    Adjoint: the gradient of a halo plane is sent back to the rank that owns the plane and added to it."""

    @staticmethod
    def forward(ctx, x, dim, lo, hi, group):
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        n = x.shape[dim]
        if max(lo, hi) > n:
            raise ValueError("halo_exchange: halo wider than the local shard")
        first = x.narrow(dim, 0, hi) if hi else None             # what the previous rank needs as its high halo
        last = x.narrow(dim, n - lo, lo) if lo else None          # what the next rank needs as its low halo
        from_prev, from_next = _exchange(first, last, last, first, rank, world, group)
        parts = []
        if lo:
            parts.append(from_prev if from_prev is not None else torch.zeros_like(last))
        parts.append(x)
        if hi:
            parts.append(from_next if from_next is not None else torch.zeros_like(first))
        ctx.meta = (dim, lo, hi, group, n)
        return torch.cat(parts, dim)

    @staticmethod
    def backward(ctx, g):
        dim, lo, hi, group, n = ctx.meta
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        g_lo = g.narrow(dim, 0, lo) if lo else None               # belongs to the previous rank's last `lo` planes
        g_hi = g.narrow(dim, lo + n, hi) if hi else None          # belongs to the next rank's first `hi` planes
        from_prev, from_next = _exchange(g_lo, g_hi, g_hi, g_lo, rank, world, group)
        gx = g.narrow(dim, lo, n).clone()
        if from_prev is not None:                                 # the previous rank's high halo = my first `hi` planes
            gx.narrow(dim, 0, hi).add_(from_prev)
        if from_next is not None:                                 # the next rank's low halo = my last `lo` planes
            gx.narrow(dim, n - lo, lo).add_(from_next)
        return gx, None, None, None, None


def halo_exchange(x, dim: int, lo: int, hi: int | None = None, group=None):
    """Extend the local shard `x` of a tensor that is split in equal chunks along `dim` over the ranks of `group` (rank order =
    position) by the neighbouring ranks' boundary planes: what a convolution with `lo` / `hi` taps before / after the centre
    needs in order to produce this rank's outputs exactly as the unsharded layer would.  Differentiable.  World size 1: the
    zero-extended tensor (identical semantics, no communication)."""
    hi = lo if hi is None else hi
    if lo == 0 and hi == 0:
        return x
    if not (dist.is_available() and dist.is_initialized()):
        pad = [0, 0] * x.dim()
        pad[2 * (x.dim() - 1 - dim)], pad[2 * (x.dim() - 1 - dim) + 1] = lo, hi
        return torch.nn.functional.pad(x, pad)
    return _HaloExchange.apply(x, dim, lo, hi, group)
