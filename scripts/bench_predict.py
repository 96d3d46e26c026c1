"""cfg4 in the small: blockwise G-UNet prediction (read 256^3 -> write 208^3, halo 24) over a synthetic uint8 volume
held in pinned host memory; reports voxels/s, algorithmic TFLOP/s (1.850 MFLOP per written voxel) and a
per-kernel device-time table.  usage: python scripts/bench_predict.py [blocks_per_dim=2] [--table]"""
import contextlib
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import raygun_b200 as rb
from raygun_b200 import predict as P
from raygun_b200.cyclegan import make_generator
from raygun_b200.synth import em_like_u8

nb = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 2
R = int(os.environ.get("PREDICT_READ", "256"))
W = R - 48
side = nb * W + (R - W)
dev = torch.device("cuda:0")
rb.set_precision("bf16")
torch.manual_seed(42)
with contextlib.redirect_stdout(sys.stderr):
    g = make_generator("unet", dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3, padding_type="same",
                                    final_activation="Tanh")).to(dev).eval()
vol = torch.from_numpy(em_like_u8((side, side, side), seed=42)).pin_memory()
out = torch.zeros_like(vol).pin_memory()


def run():
    return P.predict_volume(g, vol, (R, R, R), (W, W, W), out=out, dtype=torch.bfloat16)


run()
torch.cuda.synchronize()
t0 = time.perf_counter()
_, n, vox = run()
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print(f"volume {side}^3, {n} blocks: {dt * 1e3:.1f} ms, {vox / dt / 1e6:.1f} Mvoxel/s written, "
      f"{vox * 1.850e6 / dt / 1e12:.1f} algorithmic TFLOP/s ({n * R ** 3 * 1.850e6 / dt / 1e12:.1f} executed incl. halo)")
print("peak memory", torch.cuda.max_memory_allocated() / 2 ** 30, "GiB")
if "--table" in sys.argv:
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        run()
        torch.cuda.synchronize()
    ev = [e for e in prof.key_averages() if e.device_time_total > 0]
    tot = sum(e.device_time_total for e in ev)
    print(f"total device time {tot / 1e3:.2f} ms")
    for e in sorted(ev, key=lambda e: -e.device_time_total)[:25]:
        print(f"{e.device_time_total / 1e3:9.3f} ms {100 * e.device_time_total / tot:5.1f}%  x{e.count:4d}  {e.key[:110]}")
