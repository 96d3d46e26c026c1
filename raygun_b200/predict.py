"""Blockwise volume prediction on the sm_100a path.

Restates the Daisy task geometry of the reference (``predict.py:73-108,154-164``: read roi at the origin, write roi
= read roi shrunk by ``context = (read - write) // 2``, blocks shifted by the write size while the read roi stays
inside the volume) and the worker arithmetic
(``torch/predict/worker.py:100-147``: uint8 -> float, ``/255``, ``*scale + shift``, model, centre crop, ``*255``, clamp,
TRUNCATING cast to uint8) over an in-memory volume.  Daisy / zarr are I/O and stay the reference's.

Blocks are independent given their halo, so ranks take a static round-robin share of the block list and never
communicate.  With fit="mirror" the blocks also cover the outer shell, the input being mirrored at the volume faces.
Unlike the reference worker (which builds an autograd graph per block, worker.py:116), inference runs under no_grad
and the block crosses PCIe as uint8, not float32.
"""
from __future__ import annotations

import torch

from . import functional as F
from . import kernels as K


def block_origins(vol_shape, write_shape, read_shape=None, fit="valid"):
    """Origins of the write blocks, z-major (the order Daisy hands them out for a dependency-free task).

    fit="valid" is the reference's behaviour (`predict.py:154-164` leaves daisy's default fit): the first write block sits
    at `context`, and only blocks whose READ roi lies inside the volume are processed -- the outer `context` shell of
    the output stays unwritten.  fit="mirror" (an extension) tiles the whole volume from the origin and mirrors the
    input at the volume faces; voxels nearer than the network's field of view to a face are then approximate."""
    if fit == "mirror" or read_shape is None:
        return [(z, y, x) for z in range(0, vol_shape[0], write_shape[0]) for y in range(0, vol_shape[1], write_shape[1])
                for x in range(0, vol_shape[2], write_shape[2])]
    if fit != "valid":
        raise NotImplementedError(f"fit [{fit}] is not implemented")
    ctx = [(r - w) // 2 for r, w in zip(read_shape, write_shape)]
    axes = [range(c, v - r + c + 1, w) for c, v, r, w in zip(ctx, vol_shape, read_shape, write_shape)]
    return [(z, y, x) for z in axes[0] for y in axes[1] for x in axes[2]]


def rank_blocks(origins, rank=0, world=1):
    return origins[rank::world]


def _block_slices(volume_shape, origin, read_shape, context):
    lo_pad, hi_pad, sl = [], [], []
    for o, r, c, v in zip(origin, read_shape, context, volume_shape):
        a, b = o - c, o - c + r
        lo, hi = max(0, -a), max(0, b - v)
        lo_pad.append(lo)
        hi_pad.append(hi)
        sl.append(slice(a + lo, b - hi))
    return tuple(sl), tuple(lo_pad), tuple(hi_pad)


def read_block(volume_u8: torch.Tensor, origin, read_shape, context, device, scale=2.0, shift=-1.0, dtype=None, stage=None):
    """Fetch one read block as a normalised (1,1,d,h,w) device tensor; parts outside the volume are mirrored.
    `stage`: optional pinned uint8 scratch (>= block size) so the strided host sub-volume crosses PCIe as ONE
    asynchronous pinned copy instead of a pageable one."""
    sl, lo_pad, hi_pad = _block_slices(volume_u8.shape, origin, read_shape, context)
    sub = volume_u8[sl]
    if not sub.is_cuda:
        if stage is not None:
            st = stage[: sub.numel()].view(sub.shape)
            st.copy_(sub)                                   # host gather (strided -> contiguous, pinned)
            sub = st.to(device, non_blocking=True)
        else:
            sub = sub.contiguous().to(device, non_blocking=True)
    x = K.u8_normalise(sub.contiguous(), scale, shift, dtype or torch.float32)[None, None]
    if any(lo_pad) or any(hi_pad):
        ok = all(p < s for p, s in zip(lo_pad, x.shape[2:])) and all(p < s for p, s in zip(hi_pad, x.shape[2:]))
        x = F.pad3d(x, tuple(lo_pad), tuple(hi_pad), "reflect" if ok else "replicate")
    return x


class _HostRing:
    """A few pinned staging buffers with one CUDA event each: the host may run ahead of the GPU by len(ring) blocks
    (gather block k+1 / scatter block k-1 while block k computes) without reusing a buffer that is still in flight."""

    def __init__(self, nbytes, depth, device):
        self.bufs = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(depth)]
        self.events = [None] * depth
        self.pending = [None] * depth
        self.i = 0
        self.device = device

    def acquire(self):
        k = self.i
        self.i = (self.i + 1) % len(self.bufs)
        self.retire(k)
        return k, self.bufs[k]

    def retire(self, k):
        if self.events[k] is not None:
            self.events[k].synchronize()
            self.events[k] = None
        if self.pending[k] is not None:          # deferred host scatter of a finished device->host copy
            dst, src = self.pending[k]
            dst.copy_(src)
            self.pending[k] = None

    def release_after_gpu(self, k, pending=None):
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.events[k] = ev
        self.pending[k] = pending

    def drain(self):
        for k in range(len(self.bufs)):
            self.retire(k)


@torch.no_grad()
def predict_volume(model, volume_u8: torch.Tensor, read_shape, write_shape, out: torch.Tensor | None = None, rank=0,
                   world=1, scale_shift=(2.0, -1.0), device=None, fit="valid", dtype=None):
    """Run `model` blockwise over a (D,H,W) uint8 volume; returns (out, n_blocks_done, voxels_written).

    `out` (uint8, same shape as the volume, CPU or CUDA) receives this rank's blocks; other voxels are untouched."""
    if volume_u8.dim() != 3 or volume_u8.dtype != torch.uint8:
        raise ValueError("volume must be a (D,H,W) uint8 tensor")
    read_shape, write_shape = tuple(read_shape), tuple(write_shape)
    if any((r - w) % 2 or r < w for r, w in zip(read_shape, write_shape)):
        raise ValueError("read_shape - write_shape must be even and non-negative in every dimension")
    context = tuple((r - w) // 2 for r, w in zip(read_shape, write_shape))
    device = device or torch.device("cuda", torch.cuda.current_device())
    if out is None:
        out = torch.zeros(volume_u8.shape, dtype=torch.uint8, device="cpu").pin_memory()
    was_training = model.training
    model.eval()
    done = vox = 0
    host_in = not volume_u8.is_cuda and volume_u8.is_pinned()
    host_out = not out.is_cuda
    nbytes = read_shape[0] * read_shape[1] * read_shape[2]
    ring_in = _HostRing(nbytes, 3, device) if host_in else None
    ring_out = _HostRing(write_shape[0] * write_shape[1] * write_shape[2], 3, device) if host_out else None
    for origin in rank_blocks(block_origins(volume_u8.shape, write_shape, read_shape, fit), rank, world):
        if ring_in is not None:
            k, stage = ring_in.acquire()
            x = read_block(volume_u8, origin, read_shape, context, device, *scale_shift, dtype=dtype, stage=stage)
            ring_in.release_after_gpu(k)
        else:
            x = read_block(volume_u8, origin, read_shape, context, device, *scale_shift, dtype=dtype)
        y = model(x)
        y = y[0] if isinstance(y, (list, tuple)) else y
        if tuple(y.shape[2:]) == read_shape:          # 'same' network: crop the halo (worker.py:124-129)
            y = y[:, :, context[0]: read_shape[0] - context[0], context[1]: read_shape[1] - context[1],
                  context[2]: read_shape[2] - context[2]]
        elif tuple(y.shape[2:]) != write_shape:
            raise RuntimeError(f"model output {tuple(y.shape[2:])} matches neither read {read_shape} nor write {write_shape}")
        ext = tuple(min(w, v - o) for w, v, o in zip(write_shape, volume_u8.shape, origin))
        y = y[:, :1, : ext[0], : ext[1], : ext[2]]
        blk = K.denorm_clamp_u8(y)[0, 0]               # *255, clamp, truncating cast -- on the device
        dst = out[origin[0]: origin[0] + ext[0], origin[1]: origin[1] + ext[1], origin[2]: origin[2] + ext[2]]
        if ring_out is not None:
            k, stage = ring_out.acquire()
            st = stage[: blk.numel()].view(blk.shape)
            st.copy_(blk, non_blocking=True)           # one pinned D2H; scattered into `out` when the buffer is reused
            ring_out.release_after_gpu(k, pending=(dst, st))
        else:
            dst.copy_(blk, non_blocking=True)
        done += 1
        vox += ext[0] * ext[1] * ext[2]
    if ring_out is not None:
        ring_out.drain()
    if ring_in is not None:
        ring_in.drain()
    torch.cuda.synchronize(device)
    model.train(was_training)
    return out, done, vox
