"""Small, deterministic launch sequences for `ncu --set full` captures (profiles/r02_ncu_*):

    python scripts/ncu_targets.py cfg4block [side]   two G-UNet inference blocks (564^3-class read block by default 384)
    python scripts/ncu_targets.py cfg2step           two G-ResNet9 forward+backward steps on 2x1x128^3
    python scripts/ncu_targets.py cfg3step [batch]   two CycleGAN steps (eager), global batch 8 of 64^3

The first pass warms caches / allocator; capture the second with `-s <launches of one pass>`."""
import contextlib
import functools
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import raygun_b200 as rb
from raygun_b200 import functional as F
from raygun_b200 import kernels as K
from raygun_b200 import networks as N
from raygun_b200.cyclegan import CycleGANStep, make_generator

what = sys.argv[1]
rb.set_precision("bf16")
dev = torch.device("cuda:0")
torch.manual_seed(42)
with contextlib.redirect_stdout(sys.stderr):
    if what == "cfg4block":
        side = int(sys.argv[2]) if len(sys.argv) > 2 else 384
        g = make_generator("unet", dict(bench.G_KW)).to(dev).eval()
        u8 = bench.build_volume((side, side, side), pinned=False).to(dev)
        with torch.no_grad():
            for _ in range(2):
                y = g(K.u8_normalise(u8, dtype=torch.bfloat16)[None, None])
                out = K.denorm_clamp_u8(y)
    elif what == "cfg2step":
        g = make_generator("resnet", dict(input_nc=1, output_nc=1, ngf=64, n_blocks=9, padding_type="same", final_activation="Tanh")).to(dev)
        x = K.u8_normalise(bench.synth_batch(2, 128, 42).to(dev), dtype=torch.float32)
        t = K.u8_normalise(bench.synth_batch(2, 128, 43).to(dev), dtype=torch.float32)
        for _ in range(2):
            for p in g.parameters():
                p.grad = None
            F.smooth_l1_loss(t, g(x)).backward()
    elif what == "cfg3step":
        batch = int(sys.argv[2]) if len(sys.argv) > 2 else 8
        IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
        g1, g2 = (make_generator("unet", dict(bench.G_KW)).to(dev) for _ in range(2))
        d1, d2 = (N.NLayerDiscriminator(3, norm_layer=IN, **bench.D_KW).to(dev) for _ in range(2))
        step = CycleGANStep(g1, g2, d1, d2, lr=1e-5, single_backward=True)
        a = K.u8_normalise(bench.synth_batch(batch, 64, 42).to(dev), dtype=torch.bfloat16)
        b = K.u8_normalise(bench.synth_batch(batch, 64, 43).to(dev), dtype=torch.bfloat16)
        for _ in range(2):
            step.step(a, b)
    else:
        raise SystemExit(__doc__)
torch.cuda.synchronize()
print("ok", what)
