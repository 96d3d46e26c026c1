"""Developer diagnostic: 1x1x1 conv + Sigmoid head (MTLSD), ours (fp32 mode) vs stock torch, forward and gradients."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import raygun_b200 as rb
from raygun_b200.networks.UNet import ConvPass
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
torch.manual_seed(0)
dev = "cuda"
for shape, cout, act in (((1, 8, 8, 8, 8), 3, "Sigmoid"), ((1, 8, 8, 8, 8), 3, "ReLU"), ((1, 8, 8, 8, 8), 10, "Sigmoid"), ((2, 8, 9, 7, 5), 3, "Sigmoid"),
                         ((1, 8, 16, 16, 16), 3, "Sigmoid")):
    head = ConvPass(8, cout, [[1, 1, 1]], activation=act).to(dev)
    ref = torch.nn.Sequential(torch.nn.Conv3d(8, cout, 1), getattr(torch.nn, act)()).to(dev)
    ref[0].load_state_dict(head.conv_pass[0].state_dict())
    x = torch.randn(*shape, device=dev)
    g = torch.randn(shape[0], cout, *shape[2:], device=dev)
    xo, xr = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    with rb.precision("fp32"):
        yo = head(xo)
        (yo * g).sum().backward()
    yr = ref(xr)
    (yr * g).sum().backward()
    wo, wr = head.conv_pass[0].weight.grad, ref[0].weight.grad
    bo, br = head.conv_pass[0].bias.grad, ref[0].bias.grad
    print(shape, cout, act, "y %.2e  dx %.2e  dw %.2e (max %.2f)  db %.2e" % (float((yo - yr).abs().max()), float((xo.grad - xr.grad).abs().max()),
          float((wo - wr).abs().max()), float(wr.abs().max()), float((bo - br).abs().max())))
    if float((wo - wr).abs().max()) > 1e-4:
        print("   ours ", wo.flatten()[:8].tolist()); print("   torch", wr.flatten()[:8].tolist())
