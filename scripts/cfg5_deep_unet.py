"""cfg5 (BASELINE.json configs[4]): deep 3-D U-Net generator (4 down-sampling levels, fmap_inc_factor 3, ngf 64)
forward + backward on one 1x1xS^3 patch, bf16 tensor-core path; prints time per step, peak HBM and which conv kernels
ran.  `--check` first compares the bf16 forward at 32^3 with the fp32 path of the same library (2e-2 relative).

    python scripts/cfg5_deep_unet.py [--sides 64,128,256] [--steps 3] [--check]
"""
import argparse
import contextlib
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

DEEP_KW = dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=3, num_downs=5, padding_type="same",
               final_activation="Tanh")


def deep_unet_check(g, dev, side=32):
    """bf16 tensor-core path vs the fp32 path of the same library on one 1x1x32^3 patch: output (north_star: 2e-2
    relative) and the gradient of every parameter (cosine)."""
    import time
    import raygun_b200 as rb
    from raygun_b200 import functional as F
    torch.manual_seed(6)
    x = torch.rand(1, 1, side, side, side, device=dev) * 2 - 1
    t = torch.rand(1, 1, side, side, side, device=dev) * 2 - 1
    params = list(g.parameters())
    out = {}
    for mode in ("fp32", "bf16"):
        for p in params:
            p.grad = None
        t0 = time.perf_counter()
        with rb.precision(mode):
            y = g(x)
            F.smooth_l1_loss(t, y).backward()
        torch.cuda.synchronize()
        out[mode] = (y.detach().float(), [p.grad.detach().float().clone() for p in params], time.perf_counter() - t0)
    yf, gf, tf = out["fp32"]
    yh, gh, th = out["bf16"]
    for p in params:
        p.grad = None
    cos = [float((a * b).sum() / (a.norm() * b.norm() + 1e-30)) for a, b in zip(gf, gh)]
    rel = float((yh - yf).norm() / yf.norm())
    gn = float(torch.sqrt(sum((b ** 2).sum() for b in gh)) / torch.sqrt(sum((a ** 2).sum() for a in gf)))
    return dict(check=f"bf16 vs fp32 path, forward+backward at {side}^3", rel_rms=rel, min_grad_cosine=min(cos),
                mean_grad_cosine=sum(cos) / len(cos), grad_norm_ratio=gn, seconds_fp32=tf, seconds_bf16=th,
                ok=rel <= 2e-2 and min(cos) >= 0.98)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sides", default="64,128,256")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--profile", action="store_true", help="per-kernel device-time table of the last side")
    args = ap.parse_args()
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    from raygun_b200 import functional as F
    from raygun_b200.cyclegan import make_generator
    dev = torch.device("cuda", 0)
    torch.manual_seed(5)
    with contextlib.redirect_stdout(sys.stderr):
        g = make_generator("unet", dict(DEEP_KW)).to(dev)
    params = list(g.parameters())
    n_par = sum(p.numel() for p in params)
    print(json.dumps(dict(parameters=n_par, weights_gb=n_par * 4 / 1e9,
                          hbm_after_build_gb=torch.cuda.memory_allocated() / 1e9)), flush=True)
    if args.check:
        print(json.dumps(deep_unet_check(g, dev)), flush=True)
    rb.set_precision("bf16")
    for side in [int(s) for s in args.sides.split(",")]:
        x = torch.rand(1, 1, side, side, side, device=dev) * 2 - 1
        t = torch.rand(1, 1, side, side, side, device=dev) * 2 - 1

        def step():
            for p in params:
                p.grad = None
            loss = F.smooth_l1_loss(t, g(x))
            loss.backward()
            return loss

        torch.cuda.reset_peak_memory_stats()
        for k in CV.stats:
            CV.stats[k] = 0
        loss = step()
        torch.cuda.synchronize()
        stats = {k: v for k, v in CV.stats.items() if v}
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        gnorm = float(torch.sqrt(sum((p.grad.float() ** 2).sum() for p in params)))
        print(json.dumps(dict(side=side, ms_per_step=ms, voxels_per_s=side ** 3 / (ms * 1e-3), loss=float(loss.detach()),
                              grad_norm=gnorm, finite=bool(torch.isfinite(torch.tensor(gnorm))),
                              peak_hbm_gb=torch.cuda.max_memory_allocated() / 1e9, conv_paths=stats)), flush=True)
    if args.profile:
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            step()
            torch.cuda.synchronize()
        ev = [e for e in prof.key_averages() if e.device_time_total > 0]
        tot = sum(e.device_time_total for e in ev)
        print(f"total device time {tot / 1e3:.2f} ms over {sum(e.count for e in ev)} kernels")
        for e in sorted(ev, key=lambda e: -e.device_time_total)[:24]:
            print(f"{e.device_time_total / 1e3:9.3f} ms {100 * e.device_time_total / tot:5.1f}%  x{e.count:4d}  {e.key[:110]}")
        CV.profile = []
        step()
        torch.cuda.synchronize()
        prof2, CV.profile = CV.profile, None
        rows = [(a.elapsed_time(b), fl, kind, kname) for kind, fl, a, b, kname in prof2]
        tf = sum(r[1] for r in rows)
        tm = sum(r[0] for r in rows)
        print(f"conv launches: {len(rows)}, {tf / 1e12:.1f} TFLOP in {tm:.1f} ms = {tf / tm / 1e9:.0f} TFLOP/s")
        for dt, fl, kind, kname in sorted(rows, key=lambda r: -r[0])[:40]:
            print(f"{dt:8.3f} ms  {fl / 1e12:7.3f} TFLOP  {fl / dt / 1e9:7.0f} TF/s  {kind:6s} {kname}")


if __name__ == "__main__":
    main()
