"""Blockwise volume prediction on the sm_100a path.

Restates the Daisy task geometry of the reference (``predict.py:73-108,154-164``: read roi at the origin, write roi
= read roi shrunk by ``context = (read - write) // 2``, blocks shifted by the write size while the read roi stays
inside the volume) and the worker arithmetic
(``torch/predict/worker.py:100-147``: uint8 -> float, ``/255``, ``*scale + shift``, model, centre crop, ``*255``, clamp,
TRUNCATING cast to uint8).  The volume is a (D,H,W) uint8 tensor in page-locked host memory (or resident on the device);
``blockio.py`` puts a chunked on-disk dataset behind the same loop.

Block fits
  * ``"valid"`` -- the reference's behaviour (`predict.py:154-164` leaves daisy's default fit): only blocks whose READ
    roi lies inside the volume are processed, the outer `context` shell of the output stays unwritten.
  * ``"clip"``  -- the whole volume is tiled from the origin; a read roi that would leave the volume is CLIPPED at the
    face instead, so the network's own reflect padding acts there exactly as it does in a whole-volume forward: with
    halo >= the receptive radius and origins = 0 mod the total down-sampling factor the stitched output equals the
    whole-volume output ("reflect-pad the volume boundary exactly as the whole-volume reference
    would").  Boundary blocks therefore have smaller read extents.
  * ``"mirror"`` -- full-size reads, the input mirrored at the volume faces (approximate near the faces).

Data path per block: ONE strided 3-D DMA from the pinned volume into a device uint8 block on a copy stream
(`rg_h2d_block`), `rg_u8_normalise`, the model, `rg_denorm_clamp_u8` on the cropped view, ONE strided 3-D DMA back
(`rg_d2h_block`) on a second copy stream; both block buffers are double-buffered with CUDA events, so the transfers of
blocks k+1 / k-1 overlap the convolutions of block k and the host only enqueues.  Blocks are independent given their
halo: ranks take a cost-balanced share of the block list and never communicate.
Unlike the reference worker (which builds an autograd graph per block, worker.py:116), inference runs under no_grad
and the block crosses PCIe as uint8, not float32.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import torch

from . import functional as F
from . import kernels as K
from ._C import check, count_launch, lib, ptr


@dataclass(frozen=True)
class Block:
    origin: tuple        # write roi origin (z, y, x) in the volume
    ext: tuple           # write roi extent (clipped at the far faces)
    read_lo: tuple       # read roi origin (may be negative for fit="mirror")
    read_ext: tuple      # read roi extent
    crop: tuple          # offset of the write roi inside the network output of the read roi

    @property
    def cost(self):
        return self.read_ext[0] * self.read_ext[1] * self.read_ext[2]

    @property
    def voxels(self):
        return self.ext[0] * self.ext[1] * self.ext[2]


def block_origins(vol_shape, write_shape, read_shape=None, fit="valid"):
    """Origins of the write blocks, z-major (the order Daisy hands them out for a dependency-free task)."""
    if fit in ("mirror", "clip") or read_shape is None:
        return [(z, y, x) for z in range(0, vol_shape[0], write_shape[0]) for y in range(0, vol_shape[1], write_shape[1])
                for x in range(0, vol_shape[2], write_shape[2])]
    if fit != "valid":
        raise NotImplementedError(f"fit [{fit}] is not implemented")
    ctx = [(r - w) // 2 for r, w in zip(read_shape, write_shape)]
    axes = [range(c, v - r + c + 1, w) for c, v, r, w in zip(ctx, vol_shape, read_shape, write_shape)]
    return [(z, y, x) for z in axes[0] for y in axes[1] for x in axes[2]]


def block_list(vol_shape, read_shape, write_shape, fit="valid"):
    """The task's blocks with their read / write rois (predict.py:92-108: context = (read - write) // 2)."""
    ctx = tuple((r - w) // 2 for r, w in zip(read_shape, write_shape))
    out = []
    for o in block_origins(vol_shape, write_shape, read_shape, fit):
        ext = tuple(min(w, v - a) for w, v, a in zip(write_shape, vol_shape, o))
        if fit == "clip":
            lo = tuple(max(0, a - c) for a, c in zip(o, ctx))
            hi = tuple(min(v, a + e + c) for v, a, e, c in zip(vol_shape, o, ext, ctx))
            rext = tuple(h - l for h, l in zip(hi, lo))
            crop = tuple(a - l for a, l in zip(o, lo))
        else:
            lo = tuple(a - c for a, c in zip(o, ctx))
            rext, crop = tuple(read_shape), ctx
        out.append(Block(tuple(o), ext, lo, rext, crop))
    return out


def rank_blocks(origins, rank=0, world=1):
    """Static round-robin share of a block list (equal-cost blocks)."""
    return origins[rank::world]


def partition_blocks(blocks, rank=0, world=1):
    """Cost-balanced static share of `blocks` (longest-processing-time greedy on the read-roi voxel count: boundary blocks of
    fit="clip" are cheaper than interior ones).  Deterministic; every rank computes the same assignment; each rank's share
    keeps the task order."""
    if world == 1:
        return list(blocks)
    order = sorted(range(len(blocks)), key=lambda i: (-blocks[i].cost, i))
    load, mine = [0] * world, []
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        load[r] += blocks[i].cost
        if r == rank:
            mine.append(i)
    return [blocks[i] for i in sorted(mine)]


def _copy_block(fn_name, *args):
    check(getattr(lib(), fn_name)(*args), fn_name)
    count_launch()


def _i64x3(v):
    return (C.c_int64 * 3)(*[int(a) for a in v])


def read_block(volume_u8: torch.Tensor, origin, read_shape, context, device, scale=2.0, shift=-1.0, dtype=None, stage=None):
    """Fetch one read block as a normalised (1,1,d,h,w) device tensor; parts outside the volume are mirrored
    (fit="mirror").  Synchronous helper for tests and tools: `predict_volume` pipelines the same steps."""
    lo = tuple(o - c for o, c in zip(origin, context))
    blk = Block(tuple(origin), tuple(read_shape), lo, tuple(read_shape), tuple(context))
    u8, lo_pad, hi_pad = _fetch(volume_u8, blk, device, None, None)
    return _normalise(u8, lo_pad, hi_pad, scale, shift, dtype)


def _fetch(volume_u8, blk, device, buf, stream):
    """uint8 sub-volume of the read roi that lies inside the volume -> dense device tensor (+ the mirror padding owed)."""
    lo_pad = tuple(max(0, -a) for a in blk.read_lo)
    hi_pad = tuple(max(0, a + r - v) for a, r, v in zip(blk.read_lo, blk.read_ext, volume_u8.shape))
    org = tuple(a + l for a, l in zip(blk.read_lo, lo_pad))
    ext = tuple(r - l - h for r, l, h in zip(blk.read_ext, lo_pad, hi_pad))
    n = ext[0] * ext[1] * ext[2]
    dst = (buf[:n] if buf is not None else torch.empty(n, dtype=torch.uint8, device=device)).view(ext)
    st = stream if stream is not None else torch.cuda.current_stream(device)
    _copy_block("rg_h2d_block", ptr(volume_u8), _i64x3(volume_u8.shape), _i64x3(org), _i64x3(ext), ptr(dst),
                C.c_void_p(st.cuda_stream))
    return dst, lo_pad, hi_pad


def _normalise(u8, lo_pad, hi_pad, scale, shift, dtype):
    x = K.u8_normalise(u8, scale, shift, dtype or torch.float32)[None, None]
    if any(lo_pad) or any(hi_pad):
        ok = all(p < s for p, s in zip(lo_pad, x.shape[2:])) and all(p < s for p, s in zip(hi_pad, x.shape[2:]))
        x = F.pad3d(x, tuple(lo_pad), tuple(hi_pad), "reflect" if ok else "replicate")
    return x


@torch.no_grad()
def predict_volume(model, volume_u8: torch.Tensor, read_shape, write_shape, out: torch.Tensor | None = None, rank=0,
                   world=1, scale_shift=(2.0, -1.0), device=None, fit="valid", dtype=None, blocks=None):
    """Run `model` blockwise over a (D,H,W) uint8 volume; returns (out, n_blocks_done, voxels_written).

    `volume_u8` / `out`: page-locked host tensors (asynchronous 3-D DMAs; pageable memory works but serialises) or device
    tensors.  `out` (uint8, same shape as the volume) receives this rank's blocks; other voxels are untouched.
    `blocks`: an explicit list of `Block`s to process instead of this rank's share of the whole task (bench / retries)."""
    if volume_u8.dim() != 3 or volume_u8.dtype != torch.uint8 or not volume_u8.is_contiguous():
        raise ValueError("volume must be a contiguous (D,H,W) uint8 tensor")
    read_shape, write_shape = tuple(read_shape), tuple(write_shape)
    if any((r - w) % 2 or r < w for r, w in zip(read_shape, write_shape)):
        raise ValueError("read_shape - write_shape must be even and non-negative in every dimension")
    device = device or torch.device("cuda", torch.cuda.current_device())
    if out is None:
        out = torch.zeros(volume_u8.shape, dtype=torch.uint8, device="cpu").pin_memory()
    if tuple(out.shape) != tuple(volume_u8.shape) or out.dtype != torch.uint8 or not out.is_contiguous():
        raise ValueError("out must be a contiguous uint8 tensor of the volume's shape")
    if blocks is None:
        blocks = partition_blocks(block_list(volume_u8.shape, read_shape, write_shape, fit), rank, world)
    was_training = model.training
    model.eval()
    # cyclegan.Generator: skip the final cast to fp32 (the uint8 quantisation reads the compute dtype directly)
    run = getattr(model, "forward_compute", None) or model
    main = torch.cuda.current_stream(device)
    s_in, s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
    max_in = max((b.cost for b in blocks), default=0)
    max_out = max((b.voxels for b in blocks), default=0)
    in_bufs = [torch.empty(max_in, dtype=torch.uint8, device=device) for _ in range(2)]
    out_bufs = [torch.empty(max_out, dtype=torch.uint8, device=device) for _ in range(2)]
    in_free = [None, None]        # recorded on `main` once the normalise kernel has consumed the buffer
    out_free = [None, None]       # recorded on `s_out` once the D2H copy has drained the buffer
    s_in.wait_stream(main)
    s_out.wait_stream(main)

    def fetch(k):
        b = blocks[k]
        if in_free[k & 1] is not None:
            s_in.wait_event(in_free[k & 1])
        u8, lo_pad, hi_pad = _fetch(volume_u8, b, device, in_bufs[k & 1], s_in)
        ev = torch.cuda.Event()
        ev.record(s_in)
        return u8, lo_pad, hi_pad, ev

    done = vox = 0
    nxt = fetch(0) if blocks else None
    for k, b in enumerate(blocks):
        u8, lo_pad, hi_pad, ready = nxt
        nxt = fetch(k + 1) if k + 1 < len(blocks) else None          # block k+1 crosses PCIe while block k computes
        main.wait_event(ready)
        x = _normalise(u8, lo_pad, hi_pad, *scale_shift, dtype)
        in_free[k & 1] = torch.cuda.Event()
        in_free[k & 1].record(main)
        for attempt in (0, 1):
            oom = False
            try:
                y = run(x)
            except torch.OutOfMemoryError:
                if attempt:
                    raise
                oom = True
            if not oom:
                break
            # block shapes differ (clipped face / edge / corner blocks): cached segments of the previous shape can leave
            # the allocator fragmented.  Hand them back and retry once (PYTORCH_CUDA_ALLOC_CONF=expandable_segments:True
            # avoids the situation altogether).
            torch.cuda.synchronize(device)
            torch.cuda.empty_cache()
        y = y[0] if isinstance(y, (list, tuple)) else y
        full = tuple(r for r in b.read_ext)
        if tuple(y.shape[2:]) == full:                # 'same' network: crop the halo (worker.py:124-129)
            c = b.crop
        elif all(s == r - 2 * c for s, r, c in zip(y.shape[2:], full, b.crop)):
            c = (0, 0, 0)                             # 'valid' network: the output already is the write roi
        else:
            raise RuntimeError(f"model output {tuple(y.shape[2:])} matches neither the read roi {full} nor the write roi")
        y = y[:, :1, c[0]: c[0] + b.ext[0], c[1]: c[1] + b.ext[1], c[2]: c[2] + b.ext[2]]
        if out_free[k & 1] is not None:
            main.wait_event(out_free[k & 1])
        blk = out_bufs[k & 1][: b.voxels].view(1, 1, *b.ext)
        K.denorm_clamp_u8(y, out=blk)                 # *255, clamp, truncating cast -- on the device
        wrote = torch.cuda.Event()
        wrote.record(main)
        s_out.wait_event(wrote)
        _copy_block("rg_d2h_block", ptr(blk), _i64x3(b.ext), _i64x3((0, 0, 0)), ptr(out), _i64x3(out.shape), _i64x3(b.origin),
                    _i64x3(b.ext), C.c_void_p(s_out.cuda_stream))
        out_free[k & 1] = torch.cuda.Event()
        out_free[k & 1].record(s_out)
        done += 1
        vox += b.voxels
    main.wait_stream(s_out)
    main.wait_stream(s_in)
    torch.cuda.synchronize(device)
    model.train(was_training)
    return out, done, vox
