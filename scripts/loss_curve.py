"""200-step loss curves, bf16 tensor-core path vs the fp32 path of the same kernels library (the fp32 path is pinned
to the reference at 1e-4 per layer / 2e-3 per CycleGAN loss by tests/test_gpu_networks.py).  north_star: "loss curves
within 1% over 200 steps".  Prints one JSON object per workload with the per-step curves and their deviation; the
numbers are codified in tests/test_gpu_networks.py::test_bf16_loss_curves_follow_fp32_over_200_steps.

    python scripts/loss_curve.py [--steps 200] [--what gen,cycle]
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
    with contextlib.redirect_stdout(sys.stderr):
        g1, g2 = (make_generator("unet", dict(G_KW)).to(dev) for _ in range(2))
        d1, d2 = (N.NLayerDiscriminator(3, norm_layer=IN, **D_KW).to(dev) for _ in range(2))
        for d in (d1, d2):
            N.init_weights(d, "kaiming")
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
        f = generator_curve("fp32", args.steps, dev, side=args.side)
        h = generator_curve("bf16", args.steps, dev, side=args.side)
        f2 = generator_curve("fp32", args.steps, dev, side=args.side)
        print(json.dumps(dict(workload="generator", steps=args.steps, bf16_vs_fp32=deviation(f, h),
                              fp32_vs_fp32_rerun=deviation(f, f2), fp32=f[::20], bf16=h[::20])))
    if "cycle" in what:
        f = cyclegan_curve("fp32", args.steps, dev, side=args.side)
        h = cyclegan_curve("bf16", args.steps, dev, side=args.side)
        f2 = cyclegan_curve("fp32", args.steps, dev, side=args.side)
        print(json.dumps(dict(workload="cyclegan", steps=args.steps,
                              bf16_vs_fp32={k: deviation(f[k], h[k]) for k in f},
                              fp32_vs_fp32_rerun={k: deviation(f[k], f2[k]) for k in f},
                              fp32={k: v[::20] for k, v in f.items()}, bf16={k: v[::20] for k, v in h.items()})))


if __name__ == "__main__":
    main()
