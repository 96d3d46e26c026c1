"""Developer tool: where does the ring kernel's MMA-issuing warp wait?  Needs build/libdbg.so
(scripts/build_ring_dbg.sh).  usage: python scripts/ring_timing.py <layer name from bench_conv_layer.LAYERS> [--stats]"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch

import raygun_b200 as rb
from raygun_b200 import _C
from raygun_b200 import conv as CV
from bench_conv_layer import LAYERS

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = ctypes.CDLL(os.path.join(root, "build", os.environ.get("RING_DBG_LIB", "libdbg.so")))
lib.rg_last_error.restype = ctypes.c_char_p
_C._lib = lib
name = sys.argv[1]
n, cin, cout, k, p, sp = LAYERS[name]
dev = torch.device("cuda:0")
x = torch.randn(n, cin, *sp, device=dev).bfloat16().contiguous(memory_format=torch.channels_last_3d)
w = torch.randn(cout, cin, *k, device=dev) * 0.05
with rb.precision("bf16"):
    for _ in range(3):
        CV.conv_fprop(x, w, None, (1, 1, 1), p, 1, 0, 0.0, want_stats="--stats" in sys.argv)
    torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * (148 * 8))()
assert lib.rg_ring_debug_read(buf) == 0
t = torch.tensor(list(buf), dtype=torch.float64).view(148, 8)
tot = t[:, 4]
names = ["table build", "tempty wait", "afull wait", "bfull wait"]
print(f"{name}: MMA warp total {tot.mean():.0f} clk (min {tot.min():.0f}, max {tot.max():.0f})")
for i, nm in enumerate(names):
    print(f"  {nm:12s} {100 * (t[:, i] / tot).mean():5.1f} %  (max CTA {100 * (t[:, i] / tot).max():5.1f} %)")
