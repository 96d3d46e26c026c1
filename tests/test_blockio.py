"""On-disk block I/O of blockwise prediction (raygun_b200/blockio.py): raw zarr-v2 / N5 chunk layouts, ROI reads and writes
(`source.to_ndarray(block.read_roi)` / `destination[block.write_roi] = out`, torch/predict/worker.py:100,160) on the CPU;
the GPU test runs the whole worker loop from one dataset into another and compares with the in-memory predictor."""
import json
import os
import struct

import numpy as np
import pytest
import torch

from raygun_b200 import blockio as B


@pytest.mark.parametrize("ext", ["zarr", "n5"])
def test_chunk_layouts_and_roi_round_trip(tmp_path, ext):
    rng = np.random.default_rng(0)
    vol = rng.integers(0, 256, size=(37, 50, 41), dtype=np.uint8)
    path = str(tmp_path / f"vol.{ext}")
    ds = B.prepare_ds(path, "raw", vol.shape, (16, 24, 16), offset=(0, 0, 0), voxel_size=(4, 4, 4))
    assert ds.grid == (3, 3, 3)
    ds.write_roi((0, 0, 0), vol)
    # container metadata and chunk files are the formats' own
    if ext == "zarr":
        meta = json.load(open(os.path.join(path, "raw", ".zarray")))
        assert meta["compressor"] is None and meta["dtype"] == "|u1" and meta["chunks"] == [16, 24, 16] and meta["order"] == "C"
        raw = np.fromfile(os.path.join(path, "raw", "2.2.2"), dtype=np.uint8).reshape(16, 24, 16)     # padded edge chunk
        assert np.array_equal(raw[:5, :2, :9], vol[32:, 48:, 32:]) and int(raw[5:].max()) == 0
        assert json.load(open(os.path.join(path, "raw", ".zattrs")))["resolution"] == [4, 4, 4]
    else:
        meta = json.load(open(os.path.join(path, "raw", "attributes.json")))
        assert meta["dimensions"] == [41, 50, 37] and meta["blockSize"] == [16, 24, 16] and meta["compression"] == {"type": "raw"}
        with open(os.path.join(path, "raw", "2", "2", "2"), "rb") as f:                                # x/y/z, clipped block
            mode, nd, bx, by, bz = struct.unpack(">HHIII", f.read(16))
            payload = np.frombuffer(f.read(), dtype=np.uint8)
        assert (mode, nd, bx, by, bz) == (0, 3, 9, 2, 5) and np.array_equal(payload.reshape(5, 2, 9), vol[32:, 48:, 32:])
    again = B.open_ds(path, "raw")
    assert again.shape == vol.shape and again.chunks == (16, 24, 16)
    assert np.array_equal(again.to_ndarray(), vol)
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(4) as pool:
        for org, e in (((3, 5, 7), (20, 30, 19)), ((16, 24, 16), (16, 24, 16)), ((30, 40, 30), (7, 10, 11)), ((0, 0, 0), (1, 1, 1))):
            out = np.empty(e, dtype=np.uint8)
            again.read_roi(org, e, out, pool)
            assert np.array_equal(out, vol[org[0]: org[0] + e[0], org[1]: org[1] + e[1], org[2]: org[2] + e[2]])
    # partial-chunk write (read-modify-write) and never-written chunks (fill value)
    patch = rng.integers(0, 256, size=(10, 11, 12), dtype=np.uint8)
    again.write_roi((5, 6, 7), patch)
    vol[5:15, 6:17, 7:19] = patch
    assert np.array_equal(again.to_ndarray(), vol)
    empty = B.prepare_ds(path, "out", vol.shape, (16, 24, 16), delete=True)
    assert int(empty.to_ndarray().max()) == 0
    with pytest.raises(ValueError):
        again.read_roi((30, 40, 30), (8, 10, 11), np.empty((8, 10, 11), dtype=np.uint8))


def test_compressed_datasets_are_refused(tmp_path):
    path = str(tmp_path / "c.zarr")
    B.prepare_ds(path, "raw", (8, 8, 8), (4, 4, 4))
    meta_p = os.path.join(path, "raw", ".zarray")
    meta = json.load(open(meta_p))
    meta["compressor"] = {"id": "blosc"}
    json.dump(meta, open(meta_p, "w"))
    with pytest.raises(NotImplementedError):
        B.open_ds(path, "raw")


@pytest.mark.gpu
@pytest.mark.parametrize("ext", ["zarr", "n5"])
def test_predict_dataset_equals_in_memory_prediction(tmp_path, ext):
    """The worker loop over on-disk datasets (reader thread -> pinned staging -> DMA -> model -> DMA -> writer thread ->
    chunks) writes exactly what `predict_volume` writes for the same geometry, for the reference's "valid" fit and for
    the whole-volume "clip" fit."""
    import contextlib
    import sys
    import raygun_b200 as rb
    from raygun_b200 import predict as P
    from raygun_b200.cyclegan import make_generator
    from raygun_b200.synth import em_like_u8
    torch.manual_seed(3)
    with contextlib.redirect_stdout(sys.stderr):
        g = make_generator("unet", dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3, padding_type="same",
                                        final_activation="Tanh")).to("cuda").eval()
    vol = em_like_u8((96, 120, 144), seed=11)
    src_path, dst_path = str(tmp_path / f"src.{ext}"), str(tmp_path / f"dst.{ext}")
    src = B.prepare_ds(src_path, "volumes/raw", vol.shape, (32, 40, 48))
    src.write_roi((0, 0, 0), vol)
    with rb.precision("bf16"):
        for fit in ("clip", "valid"):
            dst = B.prepare_ds(dst_path, f"raw_{fit}", vol.shape, (48, 48, 48), delete=True)
            nb, vox = 0, 0
            for rank in range(2):
                n, v = B.predict_dataset(g, B.open_ds(src_path, "volumes/raw"), dst, (96,) * 3, (48,) * 3, rank=rank, world=2,
                                         fit=fit, dtype=torch.bfloat16, io_threads=4)
                nb, vox = nb + n, vox + v
            want, nb_w, vox_w = P.predict_volume(g, torch.from_numpy(vol), (96,) * 3, (48,) * 3, fit=fit, dtype=torch.bfloat16)
            assert (nb, vox) == (nb_w, vox_w) and nb > 0
            got = B.open_ds(dst_path, f"raw_{fit}").to_ndarray()
            assert np.array_equal(got, want.numpy()), fit
            assert float(got.std()) > 1.0
