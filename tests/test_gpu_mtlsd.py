"""MTLSDModel mirror (reference models/MTLSDModel.py:12-63: networks.UNet + two 1x1x1 Sigmoid ConvPass heads) against the
fixture generated from the REFERENCE (tests/golden/mtlsd.npz, oracle/make_golden_mtlsd.py): fp32 parity mode max-abs <= 1e-4 on
both outputs and every parameter gradient; bf16 mode within 2e-2."""
import json
import os

import numpy as np
import pytest
import torch

from tests._golden import GOLDEN


def _load():
    z = np.load(os.path.join(GOLDEN, "mtlsd.npz"))
    spec = json.loads(bytes(z["spec"]).decode())
    kw = dict(spec["kw"], downsample_factors=[tuple(f) for f in spec["kw"]["downsample_factors"]])
    sd = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd/")}
    grads = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("grad/")}
    return kw, spec["num_affs"], sd, grads, {k: torch.from_numpy(z[k]) for k in ("x", "affs", "lsds", "ga", "gl")}


def test_mtlsd_state_dict_matches_the_reference_fixture():
    from raygun_b200.models import MTLSDModel
    kw, na, sd, grads, _ = _load()
    m = MTLSDModel(unet_kwargs=kw, num_affs=na)
    assert list(m.state_dict()) == list(sd) and set(grads) == {k for k, _ in m.named_parameters()}
    m.load_state_dict(sd)                      # strict
    assert m.output_arrays == ["pred_affs", "pred_lsds"]


@pytest.mark.gpu
def test_mtlsd_forward_backward_vs_reference_fixture():
    import raygun_b200 as rb
    from raygun_b200.models import MTLSDModel
    kw, na, sd, grads, t = _load()
    for mode, tol in (("fp32", 1e-4), ("bf16", 2e-2)):
        m = MTLSDModel(unet_kwargs=kw, num_affs=na)
        m.load_state_dict(sd)
        m = m.to("cuda")
        with rb.precision(mode):
            affs, lsds = m(t["x"].to("cuda"))
            ((affs * t["ga"].to("cuda")).sum() + (lsds * t["gl"].to("cuda")).sum()).backward()
        assert affs.dtype == torch.float32 and tuple(affs.shape) == tuple(t["affs"].shape) and tuple(lsds.shape) == tuple(t["lsds"].shape)
        assert float((affs.cpu() - t["affs"]).abs().max()) <= tol, mode
        assert float((lsds.cpu() - t["lsds"]).abs().max()) <= tol, mode
        for k, p in m.named_parameters():
            ref = grads[k]
            scale = max(1.0, float(ref.abs().max()))
            err = float((p.grad.cpu() - ref).abs().max())
            if mode == "fp32":
                assert err <= 1e-4 * scale, (k, err)
            else:       # bf16: direction and size of every gradient (cosine >= 0.98; tiny-norm gradients: equally tiny)
                a, b = p.grad.cpu().double().flatten(), ref.double().flatten()
                if float(b.norm()) > 1e-3:
                    assert float(torch.dot(a, b) / (a.norm() * b.norm())) >= 0.98, k
