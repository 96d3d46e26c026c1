"""Time single convolution layers (fprop / dgrad / wgrad) of the BASELINE configs with CUDA events.
usage: python scripts/bench_conv_layer.py [head|unet|res|all] [--ring-max N]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import raygun_b200 as rb
from raygun_b200 import conv as CV

LAYERS = {
    # name: (n, cin, cout, kernel, pad, in spatial)
    "head7": (2, 64, 64, (7, 7, 7), (0, 0, 0), (134, 134, 134)),
    "stem7x1x1": (2, 64, 64, (7, 1, 1), (0, 0, 0), (134, 128, 128)),
    "unet64_64": (1, 64, 64, (3, 3, 3), (1, 1, 1), (64, 64, 64)),
    "unet192_64": (1, 192, 64, (3, 3, 3), (1, 1, 1), (64, 64, 64)),
    "unet128_128": (1, 128, 128, (3, 3, 3), (1, 1, 1), (32, 32, 32)),
    "unet384_128": (1, 384, 128, (3, 3, 3), (1, 1, 1), (32, 32, 32)),
    "res256": (2, 256, 256, (3, 3, 3), (0, 0, 0), (34, 34, 34)),
    "unet64_64_256": (1, 64, 64, (3, 3, 3), (1, 1, 1), (256, 256, 256)),
    # the G-UNet layers at the cfg4 block size (read 256^3)
    "inf192_64": (1, 192, 64, (3, 3, 3), (1, 1, 1), (256, 256, 256)),
    "inf64_128": (1, 64, 128, (3, 3, 3), (1, 1, 1), (128, 128, 128)),
    "inf128_128": (1, 128, 128, (3, 3, 3), (1, 1, 1), (128, 128, 128)),
    "inf384_128": (1, 384, 128, (3, 3, 3), (1, 1, 1), (128, 128, 128)),
    "inf128_256": (1, 128, 256, (3, 3, 3), (1, 1, 1), (64, 64, 64)),
    "inf256_256": (1, 256, 256, (3, 3, 3), (1, 1, 1), (64, 64, 64)),
}


def timeit(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if "--lib" in sys.argv:      # A/B against an older build of the library (pair with --no-halo if it predates it)
        import ctypes
        from raygun_b200 import _C
        l = ctypes.CDLL(sys.argv[sys.argv.index("--lib") + 1])
        l.rg_last_error.restype = ctypes.c_char_p
        _C._lib = l
    if "--no-halo" in sys.argv:
        CV.HALO_WGRAD = False
    if "--ring-max" in sys.argv:
        CV.RING_MAX_PRODUCED = int(sys.argv[sys.argv.index("--ring-max") + 1])
    names = [k for k in LAYERS if which == "all" or k.startswith(which)]
    dev = torch.device("cuda:0")
    for name in names:
        n, cin, cout, k, p, sp = LAYERS[name]
        od = [i + 2 * pp - kk + 1 for i, kk, pp in zip(sp, k, p)]
        x = torch.randn(n, cin, *sp, device=dev).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        w = torch.randn(cout, cin, *k, device=dev) * 0.05
        dy = torch.randn(n, cout, *od, device=dev).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        flops = 2.0 * n * od[0] * od[1] * od[2] * cout * cin * k[0] * k[1] * k[2]
        with rb.precision("bf16"):
            r0 = CV.stats.get("ring", 0)
            t_f = timeit(lambda: CV.conv_fprop(x, w, None, (1, 1, 1), p, 1, 0, 0.0, want_stats=True))
            ring_f = CV.stats.get("ring", 0) > r0
            if "--fprop-only" in sys.argv:
                print(f"{name:16s} {flops / 1e9:9.1f} GF | fprop {t_f:8.3f} ms {flops / t_f / 1e9:7.1f} TF/s {'ring' if ring_f else 'igemm'}", flush=True)
                continue
            r0 = CV.stats.get("ring", 0)
            t_d = timeit(lambda: CV.conv_dgrad(dy, w, tuple(x.shape), (1, 1, 1), p))
            ring_d = CV.stats.get("ring", 0) > r0
            r0 = CV.stats.get("halo_wgrad", 0)
            t_w = timeit(lambda: CV.conv_wgrad(x, dy, tuple(w.shape), (1, 1, 1), p))
            halo_w = CV.stats.get("halo_wgrad", 0) > r0
        print(f"{name:16s} {flops / 1e9:9.1f} GF | fprop {t_f:8.3f} ms {flops / t_f / 1e9:7.1f} TF/s {'ring' if ring_f else 'igemm'}"
              f" | dgrad {t_d:8.3f} ms {flops / t_d / 1e9:7.1f} TF/s {'ring' if ring_d else 'igemm'}"
              f" | wgrad {t_w:8.3f} ms {flops / t_w / 1e9:7.1f} TF/s {'halo' if halo_w else 'igemm'}", flush=True)


if __name__ == "__main__":
    main()
