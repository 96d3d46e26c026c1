"""200-step loss curves of the CUDA path (bf16 tensor-core mode and fp32 parity mode) against the ORACLE
(oracle/restate.py: the reference's CycleModel + LinkCycleLoss + Adam step restated functionally; here evaluated with
stock torch fp32 ops on the GPU for speed, TF32 off -- test infrastructure, never the product path).
north_star: "loss curves within 1% over 200 steps".  Prints one JSON object per workload with the curves and their
deviations; the bounds are codified in tests/test_gpu_oracle.py.

    python -m tests.loss_curves [--steps 200] [--what gen,cycle]
"""
import argparse
import contextlib
import functools
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def em_like(n, side, seed):
    """Synthetic EM-like patches in [-1, 1]: smoothed noise + membranes (same generator family as bench.synth_batch)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(n, 1, side, side, side, generator=g)
    k = torch.ones(1, 1, 3, 3, 3) / 27.0
    for _ in range(2):
        x = torch.nn.functional.conv3d(x, k, padding=1)
    x = (x - x.amin()) / (x.amax() - x.amin())
    return (x * 2 - 1).contiguous()


def generator_curve(mode, steps, dev, side=32, pool=4, lr=2e-5):
    """ResNet generator (ngf 64, 3 blocks, 'same') regressing a fixed smooth target with SmoothL1 + Adam."""
    import raygun_b200 as rb
    from raygun_b200 import functional as F
    from raygun_b200.cyclegan import make_generator
    torch.manual_seed(7)
    with contextlib.redirect_stdout(sys.stderr):
        g = make_generator("resnet", dict(input_nc=1, output_nc=1, ngf=64, n_blocks=3, padding_type="same",
                                          final_activation="Tanh")).to(dev)
    xs = em_like(pool, side, 100).to(dev)
    ts = (-xs).roll(1, dims=2).contiguous()            # contrast inversion + one-voxel shift: learnable, not identity
    opt = torch.optim.Adam(g.parameters(), lr=lr, betas=(0.9, 0.999))
    out = []
    with rb.precision(mode):
        for it in range(steps):
            i = it % pool
            opt.zero_grad(set_to_none=True)
            loss = F.smooth_l1_loss(ts[i:i + 1], g(xs[i:i + 1]))
            loss.backward()
            opt.step()
            out.append(float(loss.detach()))
    return out


def cyclegan_curve(mode, steps, dev, side=32, pool=4, lr=1e-5):
    """LinkCycleLoss step (2 G-UNet ngf 64 + 2 NLayerDiscriminator ngf 64, LSGAN + SmoothL1 cycle), Adam lr 1e-5
    (default_cycleGAN_conf.json)."""
    import raygun_b200 as rb
    from raygun_b200 import networks as N
    from raygun_b200.cyclegan import CycleGANStep, make_generator
    G_KW = dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3, padding_type="same",
                final_activation="Tanh")
    D_KW = dict(input_nc=1, ngf=64, n_layers=3, kw=3, downsampling_kw=2)
    torch.manual_seed(11)
    IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
    with contextlib.redirect_stdout(sys.stderr):          # initialised on the CPU (the RNG stream the oracle twin uses)
        g1, g2 = (make_generator("unet", dict(G_KW)) for _ in range(2))
        d1, d2 = (N.NLayerDiscriminator(3, norm_layer=IN, **D_KW) for _ in range(2))
        for d in (d1, d2):
            N.init_weights(d, "kaiming")
    g1, g2, d1, d2 = (m.to(dev) for m in (g1, g2, d1, d2))
    a = em_like(pool, side, 200).to(dev)
    b = (-em_like(pool, side, 201)).to(dev)
    out = {k: [] for k in ("G1", "G2", "D1", "D2")}
    with rb.precision(mode):
        step = CycleGANStep(g1, g2, d1, d2, lr=lr)
        for it in range(steps):
            i = it % pool
            r = step.step(a[i:i + 1], b[i:i + 1])
            for k in out:
                out[k].append(float(r[k]))
    return out


def _oracle_backend(variant):
    """fp32 arithmetic for the oracle on the GPU: no TF32; variant 0 = deterministic cuDNN algorithms, variant 1 = the
    reference's own setting (`cudnn.benchmark = True`, systems/CycleGAN.py:11): a second, equally valid fp32 evaluation
    of the same algorithm -- the yardstick for how far two correct implementations drift apart."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.deterministic = variant == 0
    torch.backends.cudnn.benchmark = variant != 0


def _amp(variant):
    """variant 2: the reference's own reduced-precision mode, torch.autocast(bfloat16) -- the yardstick for a bf16 path."""
    return torch.autocast("cuda", dtype=torch.bfloat16, enabled=variant == 2)


def oracle_generator_curve(steps, dev, side=32, pool=4, lr=2e-5, variant=0):
    """The workload of `generator_curve` through the oracle: same initial weights, batches, loss and Adam."""
    from oracle import restate as O
    from raygun_b200.cyclegan import make_generator
    torch.manual_seed(7)
    with contextlib.redirect_stdout(sys.stderr):
        g = make_generator("resnet", dict(input_nc=1, output_nc=1, ngf=64, n_blocks=3, padding_type="same",
                                          final_activation="Tanh"))
    sd = {k: v.detach().clone().to(dev) for k, v in g.state_dict().items()}
    spec = O.resnet_spec(input_nc=1, ngf=64, n_blocks=3, padding_type="same", norm=dict(kind="instance", affine=False))
    xs = em_like(pool, side, 100).to(dev)
    ts = (-xs).roll(1, dims=2).contiguous()
    _oracle_backend(variant)
    opt = O.AdamState({k: v for k, v in sd.items() if v.is_floating_point()}, lr, (0.9, 0.999))
    out = []
    for it in range(steps):
        i = it % pool
        names = list(opt.p)
        for k in names:
            opt.p[k].requires_grad_(True)
        with _amp(variant):
            y = O.generator_forward(sd, "resnet", spec, xs[i:i + 1], final_activation="Tanh")
        loss = O.smooth_l1(ts[i:i + 1], y.float())
        grads = torch.autograd.grad(loss, [opt.p[k] for k in names], allow_unused=True)
        for k in names:
            opt.p[k].requires_grad_(False)
        opt.step(dict(zip(names, grads)))
        out.append(float(loss.detach()))
    return out


def oracle_cyclegan_curve(steps, dev, side=32, pool=4, lr=1e-5, variant=0):
    """The workload of `cyclegan_curve` through oracle.restate.CycleStep (LinkCycleLoss.py:137-176 semantics)."""
    from oracle import restate as O
    from raygun_b200 import networks as N
    from raygun_b200.cyclegan import make_generator
    G_KW = dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3, padding_type="same",
                final_activation="Tanh")
    D_KW = dict(input_nc=1, ngf=64, n_layers=3, kw=3, downsampling_kw=2)
    torch.manual_seed(11)
    IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
    with contextlib.redirect_stdout(sys.stderr):
        g1, g2 = (make_generator("unet", dict(G_KW)) for _ in range(2))
        d1, d2 = (N.NLayerDiscriminator(3, norm_layer=IN, **D_KW) for _ in range(2))
        for d in (d1, d2):
            N.init_weights(d, "kaiming")
    sds = [{k: v.detach().clone().to(dev) for k, v in m.state_dict().items()} for m in (g1, g2, d1, d2)]
    uspec = O.unet_spec(input_nc=1, ngf=64, fmap_inc_factor=2, downsample_factors=[(2, 2, 2)] * 2, padding_type="same")
    dspec = O.nld_spec(input_nc=1, ngf=64, n_layers=3, norm=dict(kind="instance", affine=False), kw=3, downsampling_kw=2)
    _oracle_backend(variant)
    def g_fwd(sd, x):
        with _amp(variant):
            return O.generator_forward(sd, "unet", uspec, x, final_activation="Tanh").float()

    def d_fwd(sd, x):
        with _amp(variant):
            return O.nld3d_forward(sd, dspec, x).float()

    cs = O.CycleStep(*sds, g_fwd=g_fwd, d_fwd=d_fwd, lr=lr)
    a = em_like(pool, side, 200).to(dev)
    b = (-em_like(pool, side, 201)).to(dev)
    out = {k: [] for k in ("G1", "G2", "D1", "D2")}
    for it in range(steps):
        i = it % pool
        r = cs.step(a[i:i + 1], b[i:i + 1])
        for k in out:
            out[k].append(float(r[k]))
    return out


def deviation(ref, got, window=10):
    """max over steps of |got - ref| / |ref|, per step and for the curves smoothed over `window` steps."""
    r = torch.tensor(ref, dtype=torch.float64)
    g = torch.tensor(got, dtype=torch.float64)
    per_step = float(((g - r).abs() / r.abs().clamp_min(1e-12)).max())
    k = torch.ones(1, 1, window, dtype=torch.float64) / window
    rs = torch.nn.functional.conv1d(r[None, None], k)[0, 0]
    gs = torch.nn.functional.conv1d(g[None, None], k)[0, 0]
    smooth = float(((gs - rs).abs() / rs.abs().clamp_min(1e-12)).max())
    mean = float((g.mean() - r.mean()).abs() / r.mean().abs())
    return dict(per_step_max_rel=per_step, smoothed_max_rel=smooth, window=window, mean_rel=mean,
                first=[ref[0], got[0]], last=[ref[-1], got[-1]])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--what", default="gen,cycle")
    ap.add_argument("--side", type=int, default=32)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    what = args.what.split(",")
    if "gen" in what:
        o = oracle_generator_curve(args.steps, dev, side=args.side)
        o2 = oracle_generator_curve(args.steps, dev, side=args.side, variant=1)
        o3 = oracle_generator_curve(args.steps, dev, side=args.side, variant=2)
        f = generator_curve("fp32", args.steps, dev, side=args.side)
        h = generator_curve("bf16", args.steps, dev, side=args.side)
        f2 = generator_curve("fp32", args.steps, dev, side=args.side)
        h2 = generator_curve("bf16", args.steps, dev, side=args.side)
        print(json.dumps(dict(workload="generator", steps=args.steps, fp32_vs_oracle=deviation(o, f),
                              bf16_vs_oracle=deviation(o, h), oracle_vs_oracle_other_cudnn_algos=deviation(o, o2),
                              oracle_autocast_bf16_vs_oracle=deviation(o, o3),
                              fp32_rerun_identical=f == f2, bf16_rerun_identical=h == h2,
                              first_step_over_1pct=dict(fp32=first_over(o, f), bf16=first_over(o, h), oracle2=first_over(o, o2)),
                              oracle=o[::20], fp32=f[::20], bf16=h[::20])), flush=True)
    if "cycle" in what:
        o = oracle_cyclegan_curve(args.steps, dev, side=args.side)
        o2 = oracle_cyclegan_curve(args.steps, dev, side=args.side, variant=1)
        o3 = oracle_cyclegan_curve(args.steps, dev, side=args.side, variant=2)
        f = cyclegan_curve("fp32", args.steps, dev, side=args.side)
        h = cyclegan_curve("bf16", args.steps, dev, side=args.side)
        f2 = cyclegan_curve("fp32", args.steps, dev, side=args.side)
        h2 = cyclegan_curve("bf16", args.steps, dev, side=args.side)
        print(json.dumps(dict(workload="cyclegan", steps=args.steps,
                              fp32_vs_oracle={k: deviation(o[k], f[k]) for k in f},
                              bf16_vs_oracle={k: deviation(o[k], h[k]) for k in f},
                              oracle_vs_oracle_other_cudnn_algos={k: deviation(o[k], o2[k]) for k in f},
                              oracle_autocast_bf16_vs_oracle={k: deviation(o[k], o3[k]) for k in f},
                              fp32_rerun_identical=f == f2, bf16_rerun_identical=h == h2,
                              first_step_over_1pct={k: dict(fp32=first_over(o[k], f[k]), bf16=first_over(o[k], h[k]),
                                                            oracle2=first_over(o[k], o2[k]), oracle_autocast=first_over(o[k], o3[k]))
                                                    for k in f},
                              oracle={k: v[::20] for k, v in o.items()}, fp32={k: v[::20] for k, v in f.items()},
                              bf16={k: v[::20] for k, v in h.items()})), flush=True)


def first_over(ref, got, tol=1e-2):
    """First step whose relative deviation exceeds `tol` (None: never)."""
    for i, (r, g) in enumerate(zip(ref, got)):
        if abs(g - r) > tol * max(abs(r), 1e-12):
            return i
    return None


if __name__ == "__main__":
    main()
