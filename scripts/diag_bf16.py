"""Diagnostics (GPU): bf16 gradient error sources + first igemm throughput numbers."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import raygun_b200 as rb
from raygun_b200 import conv as CV, kernels as K, networks as N
from oracle import restate as O

DEV = "cuda:0"

def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).abs().max()) / max(1e-6, float(b.abs().max()))

def grads_case():
    torch.manual_seed(13)
    net = N.ResNet(3, input_nc=1, ngf=64, n_blocks=2, padding_type="same")
    N.init_weights(net, "kaiming")
    sd = {k: v.detach().clone() for k, v in net.state_dict().items()}
    x = torch.randn(1, 1, 24, 24, 24).clamp(-1, 1)
    spec = O.resnet_spec(input_nc=1, ngf=64, n_blocks=2, padding_type="same", norm=dict(kind="instance", affine=False))
    sdr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    ref = O.resnet3d_forward(sdr, spec, x)
    r = torch.randn_like(ref)
    (ref * r).sum().backward()
    # torch autocast comparator on GPU (test-only): same net via restatement under bf16 autocast
    sdc = {k: v.clone().to(DEV).requires_grad_(True) for k, v in sd.items()}
    with torch.autocast("cuda", dtype=torch.bfloat16):
        yc = O.resnet3d_forward(sdc, spec, x.to(DEV))
    (yc.float() * r.to(DEV)).sum().backward()
    keys = ["model.1.weight", "model.4.weight", "model.10.conv_block.1.weight", "model.11.conv_block.5.weight", "model.12.weight", "model.19.weight"]
    print("torch autocast bf16 vs fp32 oracle: y", rel(yc, ref), {k: round(rel(sdc[k].grad, sdr[k].grad), 4) for k in keys})
    net = net.to(DEV)
    for mode, ig in (("fp32", "1"), ("bf16", "0"), ("bf16", "1")):
        os.environ["RAYGUN_B200_IGEMM"] = ig
        net.zero_grad()
        with rb.precision(mode):
            y = net(x.to(DEV))
            (y * r.to(DEV)).sum().backward()
        p = dict(net.named_parameters())
        print(mode, "igemm=" + ig, "y", round(rel(y, ref), 5), {k: round(rel(p[k].grad, sdr[k].grad), 4) for k in keys})
    os.environ["RAYGUN_B200_IGEMM"] = "1"

def bench_conv(n, cin, cout, k, s, p, sp, iters=5):
    x = torch.randn(n, cin, *sp, device=DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
    w = torch.randn(cout, cin, k, k, k, device=DEV) * 0.05
    b = torch.zeros(cout, device=DEV)
    with rb.precision("bf16"):
        for _ in range(2):
            y, _, _ = CV.conv_fprop(x, w, b, (s,) * 3, (p,) * 3, want_stats=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            y, _, _ = CV.conv_fprop(x, w, b, (s,) * 3, (p,) * 3, want_stats=True)
        e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    fl = 2.0 * y.numel() * cin * k ** 3
    print(f"fprop n{n} {cin}->{cout} k{k} s{s} p{p} in{sp} out{tuple(y.shape[2:])}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s")

if __name__ == "__main__":
    grads_case()
    bench_conv(2, 256, 256, 3, 1, 0, (34, 34, 34))
    bench_conv(2, 64, 128, 3, 2, 1, (128, 128, 128))
    bench_conv(2, 128, 256, 3, 2, 1, (64, 64, 64))
    bench_conv(1, 64, 64, 7, 1, 0, (70, 70, 70))
    bench_conv(1, 64, 64, 3, 1, 1, (64, 64, 64))
    bench_conv(1, 192, 64, 3, 1, 1, (64, 64, 64))
    bench_conv(1, 128, 128, 3, 1, 1, (32, 32, 32))
    bench_conv(2, 512, 512, 3, 1, 1, (32, 32, 32))
