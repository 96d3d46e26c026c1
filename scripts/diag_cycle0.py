"""Developer diagnostic: step-0 forward of the cfg3 networks, ours vs stock torch ops on the same device."""
import sys, os, functools, contextlib, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raygun_b200 as rb
from raygun_b200 import networks as N
from raygun_b200.cyclegan import make_generator
from tests.loss_curves import em_like
dev = torch.device("cuda", 0)
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
G_KW = dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3, padding_type="same", final_activation="Tanh")
D_KW = dict(input_nc=1, ngf=64, n_layers=3, kw=3, downsampling_kw=2)
torch.manual_seed(11)
IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
with contextlib.redirect_stdout(sys.stderr):
    g1 = make_generator("unet", dict(G_KW)).to(dev)
    d1 = N.NLayerDiscriminator(3, norm_layer=IN, **D_KW).to(dev)
    N.init_weights(d1, "kaiming")
B = (-em_like(4, 32, 201))[0:1].to(dev)
# stock-torch twin of the discriminator (same layer list)
import torch.nn as tnn
ref = tnn.Sequential(*[tnn.Conv3d(m.in_channels, m.out_channels, m.kernel_size, m.stride, m.padding) if isinstance(m, tnn.Conv3d)
                       else (tnn.InstanceNorm3d(m.num_features) if isinstance(m, tnn.InstanceNorm3d) else tnn.LeakyReLU(0.2))
                       for m in d1.model]).to(dev)
ref.load_state_dict(d1.model.state_dict())
for mode in ("fp32", "bf16"):
    with rb.precision(mode), torch.no_grad():
        for k in range(1, len(d1.model) + 1):
            from raygun_b200 import nn as L
            yo = L.run_layers(list(d1.model)[:k], B)
            yr = ref[:k](B)
            print(mode, "D layers[:%d]" % k, type(d1.model[k - 1]).__name__, tuple(yo.shape), tuple(yr.shape),
                  "max diff %.3e  ref max %.3e" % (float((yo.float() - yr).abs().max()), float(yr.abs().max())))
        print(mode, "whole D", float((d1(B).float() - ref(B)).abs().max()))
