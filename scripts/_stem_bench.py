import sys, torch
sys.path.insert(0, '/root/repo')
import raygun_b200 as rb
from raygun_b200 import conv as CV
rb.set_precision("bf16")
dev = "cuda:0"
x = torch.randn(1, 1, 258, 258, 258, device=dev).bfloat16()
w = torch.randn(64, 1, 3, 3, 3, device=dev) * 0.1
b = torch.randn(64, device=dev)
with torch.no_grad():
    for _ in range(3):
        y = CV.stem_conv_fprop(x, w, b, (0, 0, 0), 1, 0.0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        y = CV.stem_conv_fprop(x, w, b, (0, 0, 0), 1, 0.0)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
print(f"stem 1->64 3^3 at 256^3: {ms:.3f} ms, {y.numel() * 2 / ms / 1e6:.1f} GB/s written")
