"""Per-kernel device-time table of one bench step (torch.profiler / CUPTI)."""
import os, sys, contextlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import raygun_b200 as rb
from raygun_b200 import functional as F, kernels as K
from raygun_b200.cyclegan import make_generator
import bench

wl = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
batch, side, n_blocks, ngf = bench.WORKLOADS[wl]
dev = torch.device("cuda:0")
rb.set_precision("bf16")
torch.manual_seed(42)
with contextlib.redirect_stdout(sys.stderr):
    g = make_generator("resnet", dict(input_nc=1, output_nc=1, ngf=ngf, n_blocks=n_blocks, padding_type="same", final_activation="Tanh")).to(dev)
x = K.u8_normalise(bench.synth_batch(batch, side, 42).to(dev), dtype=torch.float32)
t = K.u8_normalise(bench.synth_batch(batch, side, 43).to(dev), dtype=torch.float32)
params = list(g.parameters())
def step():
    for p in params: p.grad = None
    y = g(x); loss = F.smooth_l1_loss(t, y); loss.backward()
for _ in range(3): step()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step(); torch.cuda.synchronize()
ev = [e for e in prof.key_averages() if e.device_time_total > 0 and e.device_type.name == "CUDA"]
tot = sum(e.device_time_total for e in ev)
print(f"total device time {tot/1e3:.2f} ms")
for e in sorted(ev, key=lambda e: -e.device_time_total)[:30]:
    print(f"{e.device_time_total/1e3:9.3f} ms {100*e.device_time_total/tot:5.1f}%  x{e.count:4d}  {e.key[:100]}")
