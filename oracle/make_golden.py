"""TEST INFRASTRUCTURE ONLY.  Generate tests/golden/*.npz from the REFERENCE.

Run in the build container (needs /root/reference):

    python -m oracle.make_golden

Each fixture holds: the constructor spec (JSON), the reference module's
state_dict, a seeded input, the reference output, and -- for loss =
sum(out * r) with a seeded r -- the reference gradients w.r.t. the input and
every parameter.  The fixtures are small (ngf 4..8) so they can be committed.
"""
from __future__ import annotations

import functools
import itertools
import json
import os

import numpy as np
import torch

from . import ref_loader
from . import restate as O

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

IN3 = dict(kind="instance", affine=False)
BN3 = dict(kind="batch", affine=True)


def _norm_layer(norm):
    if norm is None:
        return None
    if norm["kind"] == "instance":
        return functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
    return torch.nn.BatchNorm3d


CASES = {
    # name: (kind, ctor kwargs (JSON-able), input shape)
    "unet_same": ("unet", dict(input_nc=1, ngf=8, fmap_inc_factor=2,
                               downsample_factors=[[2, 2, 2]] * 2, padding_type="same"), (1, 1, 16, 16, 16)),
    "unet_valid": ("unet", dict(input_nc=1, ngf=4, fmap_inc_factor=2,
                                downsample_factors=[[2, 2, 2]] * 2, padding_type="valid"), (1, 1, 44, 44, 44)),
    "unet_residual_in": ("unet", dict(input_nc=2, ngf=4, fmap_inc_factor=3,
                                      downsample_factors=[[2, 2, 2]], padding_type="same",
                                      residual=True, norm=IN3, output_nc=2), (2, 2, 12, 12, 12)),
    "unet_convdown": ("unet", dict(input_nc=1, ngf=4, fmap_inc_factor=2,
                                   downsample_factors=[[2, 2, 2]], padding_type="valid",
                                   downsample_method="convolve", constant_upsample=True,
                                   activation="LeakyReLU"), (1, 1, 24, 24, 24)),
    "residualunet_same": ("residualunet", dict(input_nc=1, ngf=4, fmap_inc_factor=2,
                                               downsample_factors=[[2, 2, 2]], padding_type="same",
                                               residual=True, output_nc=1), (1, 1, 16, 16, 16)),
    "resnet_same": ("resnet", dict(input_nc=1, ngf=8, n_blocks=2, padding_type="same"), (2, 1, 16, 16, 16)),
    "resnet_zeros": ("resnet", dict(input_nc=1, ngf=4, n_blocks=1, padding_type="zeros"), (1, 1, 12, 12, 12)),
    "resnet_replicate": ("resnet", dict(input_nc=1, ngf=4, n_blocks=1, padding_type="replicate"), (1, 1, 12, 12, 12)),
    "resnet_valid": ("resnet", dict(input_nc=1, ngf=4, n_blocks=1, padding_type="valid"), (1, 1, 40, 40, 40)),
    "nld_instance": ("nld", dict(input_nc=1, ngf=8, n_layers=3, kw=3, downsampling_kw=2, norm=IN3), (2, 1, 32, 32, 32)),
    "nld_batch": ("nld", dict(input_nc=1, ngf=8, n_layers=3, norm=BN3), (2, 1, 32, 32, 32)),
}


def build_reference(R, kind, kw):
    kw = dict(kw)
    norm = kw.pop("norm", "absent")
    if kind in ("unet", "residualunet"):
        if norm != "absent":
            kw["norm_layer"] = _norm_layer(norm)
        kw["downsample_factors"] = [tuple(f) for f in kw["downsample_factors"]]
        return (R.UNet if kind == "unet" else R.ResidualUNet)(**kw)
    if kind == "resnet":
        if norm != "absent":
            kw["norm_layer"] = _norm_layer(norm)
        return R.ResNet(3, **kw)
    if kind == "nld":
        if norm != "absent":
            kw["norm_layer"] = _norm_layer(norm)
        return R.NLayerDiscriminator(3, **kw)
    raise ValueError(kind)


def oracle_forward(kind, kw, sd, x):
    kw = dict(kw)
    if kind in ("unet", "residualunet"):
        spec = O.unet_spec(residual_unet=(kind == "residualunet"), **kw)
        return O.unet_forward(sd, spec, x)
    if kind == "resnet":
        kw.setdefault("norm", IN3)
        return O.resnet3d_forward(sd, O.resnet_spec(**kw), x)
    if kind == "nld":
        kw.setdefault("norm", BN3)
        return O.nld3d_forward(sd, O.nld_spec(**kw), x)
    raise ValueError(kind)


def make_case(R, name, kind, kw, shape):
    torch.manual_seed(1234)
    net = build_reference(R, kind, kw)
    R.init_weights(net, "kaiming")
    # non-trivial biases so bias handling is exercised
    with torch.no_grad():
        for k, p in net.named_parameters():
            if k.endswith("bias"):
                p.normal_(0, 0.1)
    g = torch.Generator().manual_seed(99)
    x = torch.randn(shape, generator=g).clamp(-1, 1).requires_grad_(True)
    y = net(x)
    ys = y if isinstance(y, (list, tuple)) else [y]
    r = [torch.randn(t.shape, generator=g) for t in ys]
    loss = sum((a * b).sum() for a, b in zip(ys, r))
    params = dict(net.named_parameters())
    grads = torch.autograd.grad(loss, [x] + list(params.values()), allow_unused=True)
    out = {"spec": np.frombuffer(json.dumps(dict(kind=kind, kw=kw)).encode(), dtype=np.uint8),
           "x": x.detach().numpy(), "dx": grads[0].numpy()}
    for i, (t, rr) in enumerate(zip(ys, r)):
        out[f"y{i}"] = t.detach().numpy()
        out[f"r{i}"] = rr.numpy()
    for k, v in net.state_dict().items():
        out["sd/" + k] = v.detach().numpy()
    for (k, _), gr in zip(params.items(), grads[1:]):
        if gr is not None:
            out["grad/" + k] = gr.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    return sum(v.nbytes for v in out.values())


def make_cycle(R):
    """Three LinkCycleLoss steps of a tiny 3D CycleGAN (class-default lambdas)."""
    torch.manual_seed(42)
    gkw = dict(input_nc=1, ngf=4, n_blocks=1, padding_type="same")
    dkw = dict(input_nc=1, ngf=4, n_layers=2, kw=3, downsampling_kw=2)
    mk_g = lambda: torch.nn.Sequential(R.ResNet(3, **gkw), torch.nn.Conv3d(4, 1, 1), torch.nn.Tanh())
    mk_d = lambda: R.NLayerDiscriminator(3, norm_layer=_norm_layer(IN3), **dkw)
    nets = [mk_g(), mk_g(), mk_d(), mk_d()]
    for n in nets:
        R.init_weights(n, "kaiming")
    G1, G2, D1, D2 = nets
    out = {"spec": np.frombuffer(json.dumps(dict(gkw=gkw, dkw=dkw, lr=1e-3)).encode(), dtype=np.uint8)}
    for nm, n in zip(("g1", "g2", "d1", "d2"), nets):
        for k, v in n.state_dict().items():
            out[f"sd/{nm}/{k}"] = v.detach().numpy().copy()
    oG = torch.optim.Adam(itertools.chain(G1.parameters(), G2.parameters()), lr=1e-3)
    oD = torch.optim.Adam(itertools.chain(D1.parameters(), D2.parameters()), lr=1e-3)
    model = R.CycleModel(G1, G2)
    loss = R.LinkCycleLoss(D1, G1, D2, G2, oG, oD, 3)
    g = torch.Generator().manual_seed(7)
    losses = []
    for it in range(3):
        a = torch.randn((2, 1, 16, 16, 16), generator=g).clamp(-1, 1)
        b = torch.randn((2, 1, 16, 16, 16), generator=g).clamp(-1, 1)
        out[f"a{it}"], out[f"b{it}"] = a.numpy(), b.numpy()
        fb, cb, fa, ca = model(a, b)
        loss(a, fa, ca, b, fb, cb)
        losses.append([loss.loss_dict[f"Total_Loss/{k}"] for k in ("G1", "G2", "D1", "D2")])
        if it == 0:
            out["fake_B0"] = fb.detach().numpy()
    out["losses"] = np.asarray(losses, dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "cycle_step.npz"), **out)


def make_split_cycle(R):
    """Three SplitCycleLoss steps (CycleModel(split=True), one optimizer per generator) from the SAME initial weights
    and batches as cycle_step.npz: only the reference losses are stored (split_cycle_step.npz)."""
    z = np.load(os.path.join(OUT, "cycle_step.npz"))
    spec = json.loads(bytes(z["spec"]).decode())
    gkw, dkw = spec["gkw"], spec["dkw"]
    mk_g = lambda: torch.nn.Sequential(R.ResNet(3, **gkw), torch.nn.Conv3d(4, 1, 1), torch.nn.Tanh())
    mk_d = lambda: R.NLayerDiscriminator(3, norm_layer=_norm_layer(IN3), **dkw)
    nets = [mk_g(), mk_g(), mk_d(), mk_d()]
    for nm, n in zip(("g1", "g2", "d1", "d2"), nets):
        pre = f"sd/{nm}/"
        n.load_state_dict({k[len(pre):]: torch.from_numpy(z[k]).clone() for k in z.files if k.startswith(pre)})
    G1, G2, D1, D2 = nets
    oG1 = torch.optim.Adam(G1.parameters(), lr=1e-3)
    oG2 = torch.optim.Adam(G2.parameters(), lr=1e-3)
    oD = torch.optim.Adam(itertools.chain(D1.parameters(), D2.parameters()), lr=1e-3)
    model = R.CycleModel(G1, G2, split=True)
    loss = R.SplitCycleLoss(D1, G1, D2, G2, oG1, oG2, oD, 3)
    losses = []
    for it in range(3):
        a, b = torch.from_numpy(z[f"a{it}"]), torch.from_numpy(z[f"b{it}"])
        fb, cb, fa, ca = model(a, b)
        loss(a, fa, ca, b, fb, cb)
        losses.append([loss.loss_dict[f"Total_Loss/{k}"] for k in ("G1", "G2", "D1", "D2")])
    np.savez_compressed(os.path.join(OUT, "split_cycle_step.npz"), losses=np.asarray(losses, dtype=np.float64))
    return losses


def main():
    os.makedirs(OUT, exist_ok=True)
    R = ref_loader.load()
    import sys
    if "--only-split" in sys.argv:          # add the SplitCycleLoss fixture without regenerating the others
        print(make_split_cycle(R))
        return
    for name, (kind, kw, shape) in CASES.items():
        nb = make_case(R, name, kind, kw, shape)
        print(f"{name}: {nb / 1e6:.2f} MB raw")
    make_cycle(R)
    make_split_cycle(R)
    # structural constants the reference's own tests assert
    d = R.NLayerDiscriminator(3)
    consts = dict(
        nld3d_default_params=sum(p.numel() for p in d.parameters()),  # nlayerdiscriminator_test.py:62-66
        nld3d_default_len=len(d.model),
        resnet3d_default_len=len(R.ResNet(3).model),
    )
    with open(os.path.join(OUT, "structural.json"), "w") as f:
        json.dump(consts, f, indent=1)
    print(consts)


if __name__ == "__main__":
    main()
