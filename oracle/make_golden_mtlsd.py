"""TEST INFRASTRUCTURE ONLY.  Generate tests/golden/mtlsd.npz from the REFERENCE's MTLSDModel (models/MTLSDModel.py:12-63,
imported verbatim through oracle/ref_loader.py; build container only): weights, one input, both head outputs and every
parameter gradient of sum(affs * ga) + sum(lsds * gl) for fixed random ga / gl.

    python -m oracle.make_golden_mtlsd
"""
import importlib
import json
import os

import numpy as np
import torch

from oracle import ref_loader

KW = {"input_nc": 1, "ngf": 8, "fmap_inc_factor": 2, "downsample_factors": [[2, 2, 2], [2, 2, 2]], "constant_upsample": True}


def main():
    ref_loader.load()
    M = importlib.import_module("raygun.torch.models.MTLSDModel").MTLSDModel
    torch.manual_seed(77)
    kw = dict(KW, downsample_factors=[tuple(f) for f in KW["downsample_factors"]])
    m = M(unet_kwargs=kw, num_affs=3)
    x = torch.randn(1, 1, 48, 48, 48).clamp(-1, 1)
    affs, lsds = m(x)
    ga, gl = torch.randn_like(affs), torch.randn_like(lsds)
    ((affs * ga).sum() + (lsds * gl).sum()).backward()
    out = {"spec": np.frombuffer(json.dumps({"kw": KW, "num_affs": 3}).encode(), dtype=np.uint8), "x": x.numpy(),
           "affs": affs.detach().numpy(), "lsds": lsds.detach().numpy(), "ga": ga.numpy(), "gl": gl.numpy()}
    for k, v in m.state_dict().items():
        out["sd/" + k] = v.numpy()
    for k, p in m.named_parameters():
        out["grad/" + k] = p.grad.numpy()
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "mtlsd.npz")
    np.savez_compressed(path, **out)
    print(path, tuple(affs.shape), tuple(lsds.shape), os.path.getsize(path))


if __name__ == "__main__":
    main()
