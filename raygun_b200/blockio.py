"""On-disk block I/O for blockwise prediction: chunked uint8 datasets behind the worker loop.

The reference reads its source with ``daisy.open_ds(path, dataset)`` (zarr / N5), prepares the destination with
``daisy.prepare_ds(..., write_size=write_roi.shape)`` -- so one write block is one chunk -- and lets every worker do
``source.to_ndarray(block.read_roi)`` / ``destination[block.write_roi] = out`` (``predict.py:70-125``,
``torch/predict/worker.py:75-160``).  daisy and zarr are not available here and are I/O plumbing, not arithmetic; what
the hot path needs of them is restated: the two container formats' RAW (uncompressed) chunk layouts, a ROI reader that
gathers the chunks a read roi touches, and a chunk-aligned ROI writer.

  * zarr v2: ``<path>/<ds>/.zarray`` (JSON: shape, chunks, dtype "|u1", compressor null, order "C") and one file per chunk
    named ``i.j.k`` (or ``i/j/k`` with dimension_separator "/"), always a FULL chunk, edge chunks padded with fill_value;
    daisy's ``.zattrs`` carry ``offset`` / ``resolution`` (world units).
  * N5: ``<path>/<ds>/attributes.json`` (dimensions / blockSize in x,y,z order, dataType "uint8", compression raw) and
    block files ``x/y/z``: big-endian header (uint16 mode = 0, uint16 ndim, ndim x uint32 block extents) + the voxels with
    x fastest, i.e. the bytes of a C-ordered (z,y,x) array; edge blocks are stored at their true (clipped) extents.
Compressed chunks (blosc, gzip) raise NotImplementedError: decompression is host work outside this path.

`predict_dataset` is the worker loop over such datasets: a reader thread fills a page-locked staging block for block
k+1 while block k computes, a writer thread drains finished blocks to their chunks; the device side is the pipeline of
``predict.predict_volume`` (contiguous DMA, `rg_u8_normalise`, model, `rg_denorm_clamp_u8`, DMA back).
"""
from __future__ import annotations

import json
import os
import struct
import threading
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

from . import predict as P


class ChunkedU8:
    """A (D,H,W) uint8 array stored as raw chunks in zarr-v2 or N5 layout."""

    def __init__(self, root, fmt, shape, chunks, fill=0, sep=".", offset=(0, 0, 0), voxel_size=(1, 1, 1)):
        self.root, self.fmt = root, fmt
        self.shape, self.chunks = tuple(int(s) for s in shape), tuple(int(c) for c in chunks)
        self.fill, self.sep = int(fill), sep
        self.offset, self.voxel_size = tuple(offset), tuple(voxel_size)
        self.grid = tuple(-(-s // c) for s, c in zip(self.shape, self.chunks))
        self._rmw_lock = threading.Lock()

    # ---- chunk files -------------------------------------------------------------------------------------------
    def _chunk_path(self, idx):
        if self.fmt == "zarr":
            return os.path.join(self.root, self.sep.join(str(i) for i in idx)) if self.sep == "." else \
                os.path.join(self.root, *[str(i) for i in idx])
        return os.path.join(self.root, *[str(i) for i in reversed(idx)])          # N5: x/y/z

    def _chunk_extent(self, idx):
        return tuple(min(c, s - i * c) for i, c, s in zip(idx, self.chunks, self.shape))

    def read_chunk(self, idx):
        """The chunk as a (cz,cy,cx) uint8 array holding at least its in-volume extent (None: never written -> fill)."""
        path = self._chunk_path(idx)
        if not os.path.exists(path):
            return None
        if self.fmt == "zarr":
            a = np.fromfile(path, dtype=np.uint8)
            if a.size != self.chunks[0] * self.chunks[1] * self.chunks[2]:
                raise IOError(f"{path}: {a.size} bytes, expected a full raw chunk (compressed chunks are not supported)")
            return a.reshape(self.chunks)
        with open(path, "rb") as f:
            mode, nd = struct.unpack(">HH", f.read(4))
            dims = struct.unpack(">" + "I" * nd, f.read(4 * nd))
            if mode != 0 or nd != 3:
                raise IOError(f"{path}: unsupported N5 block header (mode {mode}, ndim {nd})")
            a = np.frombuffer(f.read(), dtype=np.uint8)
        ext = tuple(reversed(dims))
        if a.size != ext[0] * ext[1] * ext[2]:
            raise IOError(f"{path}: payload does not match its header (compressed blocks are not supported)")
        return a.reshape(ext)

    def write_chunk(self, idx, data):
        """`data`: the chunk's in-volume extent (clipped at the array's far faces)."""
        path = self._chunk_path(idx)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        tmp = f"{path}.tmp{os.getpid()}"
        if self.fmt == "zarr":
            full = data
            if tuple(data.shape) != self.chunks:
                full = np.full(self.chunks, self.fill, dtype=np.uint8)
                full[: data.shape[0], : data.shape[1], : data.shape[2]] = data
            with open(tmp, "wb") as f:
                f.write(np.ascontiguousarray(full).tobytes())
        else:
            with open(tmp, "wb") as f:
                f.write(struct.pack(">HHIII", 0, 3, *reversed(data.shape)))
                f.write(np.ascontiguousarray(data).tobytes())
        os.replace(tmp, path)

    # ---- ROI access (source.to_ndarray(roi) / destination[roi] = out, worker.py:100,160) ------------------------------
    def _touched(self, origin, ext):
        lo = [o // c for o, c in zip(origin, self.chunks)]
        hi = [(o + e - 1) // c for o, e, c in zip(origin, ext, self.chunks)]
        return [(z, y, x) for z in range(lo[0], hi[0] + 1) for y in range(lo[1], hi[1] + 1) for x in range(lo[2], hi[2] + 1)]

    def read_roi(self, origin, ext, out: np.ndarray, pool: ThreadPoolExecutor | None = None):
        """out[...] = array[origin : origin + ext] (out: a (ext) uint8 array, e.g. a view of page-locked memory)."""
        if any(o < 0 or o + e > s for o, e, s in zip(origin, ext, self.shape)):
            raise ValueError(f"read roi {origin}+{ext} outside the dataset {self.shape}")

        def one(idx):
            c0 = [i * c for i, c in zip(idx, self.chunks)]
            lo = [max(o, a) for o, a in zip(origin, c0)]
            hi = [min(o + e, a + c) for o, e, a, c in zip(origin, ext, c0, self.chunks)]
            dst = out[lo[0] - origin[0]: hi[0] - origin[0], lo[1] - origin[1]: hi[1] - origin[1], lo[2] - origin[2]: hi[2] - origin[2]]
            ch = self.read_chunk(idx)
            if ch is None:
                dst[...] = self.fill
            else:
                dst[...] = ch[lo[0] - c0[0]: hi[0] - c0[0], lo[1] - c0[1]: hi[1] - c0[1], lo[2] - c0[2]: hi[2] - c0[2]]

        idxs = self._touched(origin, ext)
        if pool is None or len(idxs) == 1:
            for i in idxs:
                one(i)
        else:
            list(pool.map(one, idxs))
        return out

    def write_roi(self, origin, data: np.ndarray, pool: ThreadPoolExecutor | None = None):
        """array[origin : origin + data.shape] = data.  Chunks the roi covers completely are written straight out; a chunk it
        covers partially is read, updated and rewritten -- safe within one process only, so ranks that write concurrently
        need chunk-aligned write rois (prepare_ds(write_size = write block) + fit="clip" gives exactly that; the
        reference relies on daisy's read_write_conflict bookkeeping instead)."""
        ext = data.shape
        if any(o < 0 or o + e > s for o, e, s in zip(origin, ext, self.shape)):
            raise ValueError(f"write roi {origin}+{tuple(ext)} outside the dataset {self.shape}")

        def one(idx):
            c0 = [i * c for i, c in zip(idx, self.chunks)]
            e = self._chunk_extent(idx)
            lo = [max(o, a) for o, a in zip(origin, c0)]
            hi = [min(o + x, a + b) for o, x, a, b in zip(origin, ext, c0, e)]
            src = data[lo[0] - origin[0]: hi[0] - origin[0], lo[1] - origin[1]: hi[1] - origin[1], lo[2] - origin[2]: hi[2] - origin[2]]
            if all(l == a and h == a + b for l, h, a, b in zip(lo, hi, c0, e)):
                self.write_chunk(idx, src)
                return
            with self._rmw_lock:
                old = self.read_chunk(idx)
                cur = np.full(e, self.fill, dtype=np.uint8)
                if old is not None:
                    cur[...] = old[: e[0], : e[1], : e[2]]
                cur[lo[0] - c0[0]: hi[0] - c0[0], lo[1] - c0[1]: hi[1] - c0[1], lo[2] - c0[2]: hi[2] - c0[2]] = src
                self.write_chunk(idx, cur)

        idxs = self._touched(origin, ext)
        if pool is None or len(idxs) == 1:
            for i in idxs:
                one(i)
        else:
            list(pool.map(one, idxs))

    def to_ndarray(self):
        return self.read_roi((0, 0, 0), self.shape, np.empty(self.shape, dtype=np.uint8))


def _fmt_of(path):
    p = path.rstrip("/")
    if p.endswith(".n5"):
        return "n5"
    if p.endswith(".zarr"):
        return "zarr"
    raise ValueError(f"{path}: container must end in .zarr or .n5")


def open_ds(path, ds_name):
    """daisy.open_ds(path, ds_name) for raw uint8 datasets (predict.py:70)."""
    root = os.path.join(path, ds_name)
    fmt = _fmt_of(path)
    if fmt == "zarr":
        with open(os.path.join(root, ".zarray")) as f:
            meta = json.load(f)
        if meta.get("compressor") is not None or meta.get("filters"):
            raise NotImplementedError(f"{root}: compressed / filtered chunks are outside this path (raw chunks only)")
        if np.dtype(meta["dtype"]) != np.uint8 or meta.get("order", "C") != "C" or len(meta["shape"]) != 3:
            raise NotImplementedError(f"{root}: only 3-D C-ordered uint8 arrays are supported")
        attrs = {}
        if os.path.exists(os.path.join(root, ".zattrs")):
            with open(os.path.join(root, ".zattrs")) as f:
                attrs = json.load(f)
        return ChunkedU8(root, "zarr", meta["shape"], meta["chunks"], meta.get("fill_value") or 0,
                         meta.get("dimension_separator", "."), attrs.get("offset", (0, 0, 0)), attrs.get("resolution", (1, 1, 1)))
    with open(os.path.join(root, "attributes.json")) as f:
        meta = json.load(f)
    if meta.get("compression", {}).get("type", "raw") != "raw" or meta.get("dataType") != "uint8" or len(meta["dimensions"]) != 3:
        raise NotImplementedError(f"{root}: only raw 3-D uint8 N5 datasets are supported")
    return ChunkedU8(root, "n5", list(reversed(meta["dimensions"])), list(reversed(meta["blockSize"])), 0, "/",
                     list(reversed(meta.get("offset", (0, 0, 0)))), list(reversed(meta.get("resolution", (1, 1, 1)))))


def prepare_ds(path, ds_name, shape, chunks, offset=(0, 0, 0), voxel_size=(1, 1, 1), delete=False):
    """daisy.prepare_ds(filename, ds_name, total_roi, voxel_size, dtype=uint8, write_size=chunks) (predict.py:111-125)."""
    import shutil
    root = os.path.join(path, ds_name)
    fmt = _fmt_of(path)
    if delete and os.path.isdir(root):
        shutil.rmtree(root)
    os.makedirs(root, exist_ok=True)
    if fmt == "zarr":
        if not os.path.exists(os.path.join(path, ".zgroup")):
            with open(os.path.join(path, ".zgroup"), "w") as f:
                json.dump({"zarr_format": 2}, f)
        with open(os.path.join(root, ".zarray"), "w") as f:
            json.dump({"zarr_format": 2, "shape": list(shape), "chunks": list(chunks), "dtype": "|u1", "compressor": None,
                       "fill_value": 0, "filters": None, "order": "C", "dimension_separator": "."}, f)
        with open(os.path.join(root, ".zattrs"), "w") as f:
            json.dump({"offset": list(offset), "resolution": list(voxel_size)}, f)
    else:
        if not os.path.exists(os.path.join(path, "attributes.json")):
            with open(os.path.join(path, "attributes.json"), "w") as f:
                json.dump({"n5": "2.0.0"}, f)
        with open(os.path.join(root, "attributes.json"), "w") as f:
            json.dump({"dimensions": list(reversed(shape)), "blockSize": list(reversed(chunks)), "dataType": "uint8",
                       "compression": {"type": "raw"}, "offset": list(reversed(offset)), "resolution": list(reversed(voxel_size))}, f)
    return open_ds(path, ds_name)


@torch.no_grad()
def predict_dataset(model, source: ChunkedU8, destination: ChunkedU8, read_shape, write_shape, rank=0, world=1,
                    scale_shift=(2.0, -1.0), device=None, fit="valid", dtype=None, io_threads=8):
    """The worker loop (worker.py:94-161) over on-disk datasets; returns (n_blocks_done, voxels_written).

    Geometry as `predict.block_list` (fit "valid" = the reference's Daisy task, "clip" = whole volume with clipped read
    rois).  The destination's chunks must divide the write blocks (prepare_ds(write_size=...) makes them equal)."""
    from . import kernels as K
    read_shape, write_shape = tuple(read_shape), tuple(write_shape)
    if fit == "mirror":
        raise NotImplementedError("predict_dataset: fit='mirror' is not implemented (use 'valid' or 'clip')")
    device = device or torch.device("cuda", torch.cuda.current_device())
    blocks = P.partition_blocks(P.block_list(source.shape, read_shape, write_shape, fit), rank, world)
    if not blocks:
        return 0, 0
    was_training = model.training
    model.eval()
    max_in, max_out = max(b.cost for b in blocks), max(b.voxels for b in blocks)
    h_in = [torch.empty(max_in, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
    h_out = [torch.empty(max_out, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
    d_in = [torch.empty(max_in, dtype=torch.uint8, device=device) for _ in range(2)]
    d_out = [torch.empty(max_out, dtype=torch.uint8, device=device) for _ in range(2)]
    main = torch.cuda.current_stream(device)
    s_in, s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
    s_in.wait_stream(main)
    s_out.wait_stream(main)
    pool = ThreadPoolExecutor(max_workers=io_threads)
    h_in_free = [threading.Event(), threading.Event()]      # set once the H2D copy that read the staging block completed
    for e in h_in_free:
        e.set()
    in_free, out_free, writes = [None, None], [None, None], [None, None]

    def load(k):                                            # reader thread: disk -> page-locked staging block
        b = blocks[k]
        h_in_free[k & 1].wait()
        h_in_free[k & 1].clear()
        view = h_in[k & 1][: b.cost].view(b.read_ext).numpy()
        source.read_roi(b.read_lo, b.read_ext, view, pool)
        return view

    def store(k, ev):                                       # writer thread: staging block -> chunks
        b = blocks[k]
        ev.synchronize()
        destination.write_roi(b.origin, h_out[k & 1][: b.voxels].view(b.ext).numpy(), pool)

    reader = ThreadPoolExecutor(max_workers=1)
    writer = ThreadPoolExecutor(max_workers=1)
    pending = reader.submit(load, 0)
    done = vox = 0
    for k, b in enumerate(blocks):
        pending.result()
        nxt = reader.submit(load, k + 1) if k + 1 < len(blocks) else None
        if in_free[k & 1] is not None:
            s_in.wait_event(in_free[k & 1])
        u8 = d_in[k & 1][: b.cost].view(b.read_ext)
        with torch.cuda.stream(s_in):
            u8.copy_(h_in[k & 1][: b.cost].view(b.read_ext), non_blocking=True)
            copied = torch.cuda.Event()
            copied.record(s_in)
        # release the staging block for the reader once the DMA has drained it
        threading.Thread(target=lambda ev=copied, e=h_in_free[k & 1]: (ev.synchronize(), e.set()), daemon=True).start()
        main.wait_event(copied)
        x = K.u8_normalise(u8, *scale_shift, dtype or torch.float32)[None, None]
        in_free[k & 1] = torch.cuda.Event()
        in_free[k & 1].record(main)
        y = model(x)
        y = y[0] if isinstance(y, (list, tuple)) else y
        c = b.crop if tuple(y.shape[2:]) == tuple(b.read_ext) else (0, 0, 0)
        y = y[:, :1, c[0]: c[0] + b.ext[0], c[1]: c[1] + b.ext[1], c[2]: c[2] + b.ext[2]]
        if tuple(y.shape[2:]) != tuple(b.ext):
            raise RuntimeError(f"model output {tuple(y.shape[2:])} does not cover the write roi {b.ext}")
        if writes[k & 1] is not None:
            writes[k & 1].result()                          # the staging block of block k-2 has been written out
        if out_free[k & 1] is not None:
            main.wait_event(out_free[k & 1])
        blk = d_out[k & 1][: b.voxels].view(1, 1, *b.ext)
        K.denorm_clamp_u8(y, out=blk)
        wrote = torch.cuda.Event()
        wrote.record(main)
        s_out.wait_event(wrote)
        with torch.cuda.stream(s_out):
            h_out[k & 1][: b.voxels].view(b.ext).copy_(blk[0, 0], non_blocking=True)
            out_free[k & 1] = torch.cuda.Event()
            out_free[k & 1].record(s_out)
        writes[k & 1] = writer.submit(store, k, out_free[k & 1])
        pending = nxt
        done += 1
        vox += b.voxels
    for w in writes:
        if w is not None:
            w.result()
    main.wait_stream(s_in)
    main.wait_stream(s_out)
    torch.cuda.synchronize(device)
    for ex in (reader, writer, pool):
        ex.shutdown(wait=True)
    model.train(was_training)
    return done, vox
