"""Per-launch table of the tensor-core convolution launches of one cfg2 step (conv.profile hooks), grouped by
(kind, kernel, algorithmic FLOPs) — the FLOP count identifies the layer class."""
import os, sys, contextlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import raygun_b200 as rb
from raygun_b200 import functional as F, kernels as K, conv as CV
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
CV.profile = []
step(); torch.cuda.synchronize()
prof, CV.profile = CV.profile, None
groups = {}
for kind, fl, a, b, kname in prof:
    d = groups.setdefault((kind, kname, round(fl / 1e9, 1)), [0.0, 0])
    d[0] += a.elapsed_time(b); d[1] += 1
tot = sum(v[0] for v in groups.values())
print(f"conv launches {len(prof)}, {tot:.2f} ms")
for (kind, kname, gf), (ms, n) in sorted(groups.items(), key=lambda kv: -kv[1][0]):
    print(f"{ms:8.3f} ms  x{n:3d}  {kind:6s} {kname:20s} {gf:9.1f} GF/launch  {gf * n / ms:7.0f} TF/s")
