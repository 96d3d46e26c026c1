"""Per-kernel device-time table of one CycleGAN training step (cfg3) on one GPU.
usage: python scripts/profile_cyclegan.py [batch=8] [single|reference]"""
import contextlib
import functools
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import raygun_b200 as rb
from raygun_b200 import kernels as K, networks as N
from raygun_b200.cyclegan import CycleGANStep, make_generator

rb.set_precision("bf16")
dev = torch.device("cuda:0")
torch.manual_seed(42)
IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
with contextlib.redirect_stdout(sys.stderr):
    g1, g2 = (make_generator("unet", dict(bench.G_KW)).to(dev) for _ in range(2))
    d1, d2 = (N.NLayerDiscriminator(3, norm_layer=IN, **bench.D_KW).to(dev) for _ in range(2))
BATCH = int(sys.argv[1]) if len(sys.argv) > 1 else 8
SINGLE = (sys.argv[2] if len(sys.argv) > 2 else "single") == "single"
step = CycleGANStep(g1, g2, d1, d2, lr=1e-5, single_backward=SINGLE)
a = K.u8_normalise(bench.synth_batch(BATCH, 64, 42).to(dev), dtype=torch.bfloat16)
b = K.u8_normalise(bench.synth_batch(BATCH, 64, 43).to(dev), dtype=torch.bfloat16)
for _ in range(3):
    step.step(a, b)
torch.cuda.synchronize()
t0 = time.perf_counter()
step.step(a, b)
torch.cuda.synchronize()
print(f"wall {1e3 * (time.perf_counter() - t0):.1f} ms per step")
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step.step(a, b)
    torch.cuda.synchronize()
ev = [e for e in prof.key_averages() if e.device_time_total > 0]
tot = sum(e.device_time_total for e in ev)
print(f"total device time {tot / 1e3:.2f} ms over {sum(e.count for e in ev)} kernels")
for e in sorted(ev, key=lambda e: -e.device_time_total)[:48]:
    print(f"{e.device_time_total / 1e3:9.3f} ms {100 * e.device_time_total / tot:5.1f}%  x{e.count:4d}  {e.key[:110]}")
