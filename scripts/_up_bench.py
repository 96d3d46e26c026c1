import sys, torch
sys.path.insert(0, '/root/repo')
import raygun_b200 as rb
from raygun_b200 import conv as CV
rb.set_precision("bf16")
dev = "cuda:0"
x = torch.randn(1, 128, 128, 128, 128, device=dev).bfloat16().contiguous(memory_format=torch.channels_last_3d)
w = torch.randn(128, 128, 2, 2, 2, device=dev) * 0.05
b = torch.randn(128, device=dev)
with torch.no_grad(), rb.kernels.arena_scope(False):
    for _ in range(3):
        y = CV.upconv2_fprop(x, w, b, 0, 0.0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        y = CV.upconv2_fprop(x, w, b, 0, 0.0)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
print(f"upconv2 128->128 at 128^3: {ms:.3f} ms, {y.numel() * 2 / ms / 1e6:.1f} GB/s written")
