"""Times the HBM-bound G-UNet inference kernels alone at a cfg4-class block (default 516 + 2 x 24 = 564^3 read block):
the 1 -> 64 stem, both ConvTranspose3d(k = s = 2) launches into the channel slice of the skip concat buffer, the 2^3 max
pool.  Prints ms per launch and algorithmic GB/s (bytes that must cross HBM once).

    python scripts/bench_tail_kernels.py [side]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import raygun_b200 as rb
from raygun_b200 import conv as CV
from raygun_b200 import kernels as K

side = int(sys.argv[1]) if len(sys.argv) > 1 else 564
side -= side % 4
rb.set_precision("bf16")
dev = torch.device("cuda:0")
torch.manual_seed(0)


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(reps))
    return ts[len(ts) // 2]


with torch.no_grad():
    vox = side ** 3
    # stem 1 -> 64, 3^3, zero padding
    x1 = torch.randn(1, 1, side, side, side, device=dev).bfloat16()
    w = torch.randn(64, 1, 3, 3, 3, device=dev) * 0.1
    b = torch.randn(64, device=dev) * 0.1
    y = K.empty5((1, 64, side, side, side), torch.bfloat16, dev)
    ms = timed(lambda: CV.stem_conv_fprop(x1, w, b, (1, 1, 1), 1, 0.0, out=y))
    print(f"stem 1->64 {side}^3: {ms:.3f} ms  {vox * 130 / ms / 1e6:.0f} GB/s algorithmic")
    # max pool of the 64-channel skip tensor
    ms = timed(lambda: K.maxpool_fwd(y, (2, 2, 2)))
    print(f"maxpool 64ch {side}^3: {ms:.3f} ms  {vox * 128 * 1.125 / ms / 1e6:.0f} GB/s algorithmic")
    del x1
    for name, cin, cout, s_in in (("upconv L0 128->64", 128, 64, side // 2), ("upconv L1 256->128", 256, 128, side // 4)):
        x = K.empty5((1, cin, s_in, s_in, s_in), torch.bfloat16, dev).normal_()
        wt = torch.randn(cin, cout, 2, 2, 2, device=dev) * 0.05
        bt = torch.randn(cout, device=dev) * 0.1
        cat = K.empty5((1, 2 * cout, 2 * s_in, 2 * s_in, 2 * s_in), torch.bfloat16, dev)
        right = cat[:, cout:]
        out = CV.upconv2_fprop(x, wt, bt, out=right)
        assert out.data_ptr() == right.data_ptr()
        ms = timed(lambda: CV.upconv2_fprop(x, wt, bt, out=right))
        byts = x.numel() * 2 + 8 * s_in ** 3 * cout * 2
        print(f"{name} in {s_in}^3: {ms:.3f} ms  {byts / ms / 1e6:.0f} GB/s algorithmic  {2.0 * x.numel() * cout * 8 / ms / 1e9:.0f} TFLOP/s")
        # spot check against the fp32 formula on a corner
        xs = x[:, :, :4, :4, :4].float()
        ref = torch.nn.functional.conv_transpose3d(xs, wt.bfloat16().float(), bt, stride=2)
        got = right[:, :, :8, :8, :8].float()
        print("   max rel err", float((got - ref).abs().max() / ref.abs().max()))
        del x, cat, right, out
