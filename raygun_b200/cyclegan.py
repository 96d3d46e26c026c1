"""Host-side mirror of the CycleGAN model / loss glue of the reference, driving the sm_100a kernels.

Mirrors (same names, argument meaning, ordering of side effects):
  * ``BaseSystem.get_network`` wrapper      systems/BaseSystem.py:163-242  -> ``make_generator`` / ``make_discriminator``
  * ``GANLoss``                             losses/GANLoss.py:3-64
  * ``CycleModel.forward``                  models/CycleModel.py:37-89
  * ``LinkCycleLoss``                       losses/LinkCycleLoss.py:9-176 (+ BaseCompetentLoss.py:8-48)
  * the gunpowder ``Train`` node's step     train/BaseTrain.py:63-75       -> ``CycleGANStep``
The loss reductions run as fused forward+gradient warp-shuffle kernels; everything else is control flow.
"""
from __future__ import annotations

import functools
import itertools

import torch
import torch.nn as tnn

from . import conv as CV
from . import functional as F
from . import networks as N
from . import nn as L
from .optim import FlatAdam


# ---------------------------------------------------------------------------------------------
# get_network
# ---------------------------------------------------------------------------------------------
class Generator(tnn.Sequential):
    """Sequential(net, Conv3d(ngf -> output_nc, 1x1x1, padding=padding_type), final_activation)
    (BaseSystem.py:209-224) executed without leaving the compute dtype between the three stages."""

    def forward(self, x):
        net = self[0]
        if getattr(net, "ndims", getattr(net, "_ndims", 3)) != 3:
            return super().forward(x)
        return F.cast(self.forward_compute(x), torch.float32)

    def forward_compute(self, x):
        """forward() without the final cast to fp32: the result stays in the compute dtype (bf16 storage in bf16 mode).
        Blockwise prediction quantises it to uint8 directly (predict.predict_volume) -- same values, one pass less."""
        net = self[0]
        if getattr(net, "ndims", getattr(net, "_ndims", 3)) != 3:
            return super().forward(x)
        tail = list(self)[1:]
        req = self._head_request(net, tail) if not torch.is_grad_enabled() else None
        net._keep_compute_dtype = True
        try:
            if req is not None:
                with CV.offer_head(req):
                    y = net(x)
            else:
                y = net(x)
        finally:
            net._keep_compute_dtype = False
        if req is None or not req.done:
            y = L.run_layers(tail, y)
        return y

    @staticmethod
    def _head_request(net, tail):
        """Inference: offer the 1x1x1 head (+ activation) to the network's last convolution (conv.HeadRequest) when the
        network ends in conv -> activation with nothing behind it (a plain U-Net: the top r_conv ConvPass)."""
        if not (1 <= len(tail) <= 2 and isinstance(tail[0], L.Conv3d) and tuple(tail[0].kernel_size) == (1, 1, 1)
                and tail[0].groups == 1 and not any(tail[0]._pads()[0])):
            return None
        ac = (L.ACT_NONE, 0.0)
        if len(tail) == 2:
            ac = L.act_code(tail[1])
            if ac is None:
                return None
        last = getattr(net, "last_conv_before_head", None)
        conv = last() if callable(last) else None
        if conv is None:
            return None
        return CV.HeadRequest(conv.weight, tail[0].weight, tail[0].bias, ac[0], ac[1])


def _downsample_factors(kw, ndims):
    # BaseSystem.get_downsample_factors (BaseSystem.py:145-161)
    if "downsample_factors" not in kw:
        f = kw.pop("down_factor", 2)
        n = kw.pop("num_downs", 3)
        kw["downsample_factors"] = [(f,) * ndims] * (n - 1)
    return kw


def make_generator(net_type, net_kwargs, output_nc=None, final_activation=None, ndims=3, init="kaiming"):
    """What ``BaseSystem.get_network(net_type, net_kwargs)`` builds for a generator
    (BaseSystem.py:163-242): pops ``final_activation`` / ``output_nc`` from the kwargs, builds the net,
    appends Conv{n}d(ngf -> output_nc, 1x..x1, padding=padding_type) [+ activation], kaiming init."""
    kw = dict(net_kwargs)
    final_activation = kw.pop("final_activation", final_activation)
    output_nc = kw.pop("output_nc", output_nc if output_nc is not None else kw.get("input_nc", 1))
    t = net_type.lower()
    if t == "unet":
        net = N.UNet(**_downsample_factors(kw, ndims))
    elif t == "residualunet":
        net = N.ResidualUNet(**_downsample_factors(kw, ndims))
    elif t == "resnet":
        net = N.ResNet(ndims, **kw)
    elif hasattr(N, net_type):
        net = getattr(N, net_type)(**kw)
    else:
        raise ValueError(f"Unknown network type requested: {net_type}")
    conv = L.Conv3d if ndims == 3 else tnn.Conv2d
    layers = [net, conv(kw.get("ngf", 64), output_nc, (1,) * ndims, padding=kw.get("padding_type", "valid")
                        if kw.get("padding_type", "valid") in ("same", "valid") else 0)]
    if isinstance(final_activation, tnn.Module):
        layers.append(final_activation)               # an instance, e.g. from a "#torch.nn.Tanh()#" config expression
    elif final_activation is not None:
        layers.append(L.make_activation(final_activation) if ndims == 3 else
                      (getattr(tnn, final_activation) if isinstance(final_activation, str) else final_activation)())
    g = Generator(*layers)
    if init:
        N.init_weights(g, init, nonlinearity="relu")
    return g


def make_discriminator(net_kwargs, ndims=3, init="kaiming"):
    """get_network for 'patchgan'/'nlayerdiscriminator': norm forced to InstanceNorm without affine or
    running stats (BaseSystem.py:192-200)."""
    norm = {3: tnn.InstanceNorm3d, 2: tnn.InstanceNorm2d}[ndims]
    kw = dict(net_kwargs)
    kw["norm_layer"] = functools.partial(norm, affine=False, track_running_stats=False)
    d = N.NLayerDiscriminator(ndims, **kw)
    if init:
        N.init_weights(d, init, nonlinearity="leaky_relu")
    return d


# ---------------------------------------------------------------------------------------------
# losses
# ---------------------------------------------------------------------------------------------
class GANLoss(tnn.Module):
    """losses/GANLoss.py: lsgan = MSE against a constant label (one fused reduce+grad kernel, no label
    tensor is ever materialised); wgangp = -/+ mean; vanilla = BCE-with-logits (stock torch op)."""

    def __init__(self, gan_mode, target_real_label=1.0, target_fake_label=0.0):
        super().__init__()
        self.register_buffer("real_label", torch.tensor(target_real_label))
        self.register_buffer("fake_label", torch.tensor(target_fake_label))
        self._labels = (float(target_real_label), float(target_fake_label))
        self.gan_mode = gan_mode
        if gan_mode == "vanilla":
            self.loss = tnn.BCEWithLogitsLoss()
        elif gan_mode in ("lsgan", "wgangp"):
            self.loss = None
        else:
            raise NotImplementedError("gan mode %s not implemented" % gan_mode)

    def get_target_tensor(self, prediction, target_is_real):
        return (self.real_label if target_is_real else self.fake_label).expand_as(prediction)

    def __call__(self, prediction, target_is_real):
        if self.gan_mode == "lsgan":
            return F.mse_const_loss(prediction, self._labels[0] if target_is_real else self._labels[1])
        if self.gan_mode == "wgangp":
            m = F.mean_loss(prediction)
            return -m if target_is_real else m
        return self.loss(prediction, self.get_target_tensor(prediction, target_is_real))


def _cycle_loss_fn(l1_loss):
    """Map the reference's ``l1_loss`` module argument onto the fused kernels."""
    if l1_loss is None or isinstance(l1_loss, tnn.SmoothL1Loss):
        beta = 1.0 if l1_loss is None else float(l1_loss.beta)
        if l1_loss is None or l1_loss.reduction == "mean":
            return lambda real, pred: F.smooth_l1_loss(real, pred, beta)
    if isinstance(l1_loss, tnn.L1Loss) and l1_loss.reduction == "mean":
        return lambda real, pred: F.l1_loss(real, pred)
    if isinstance(l1_loss, str):
        return {"smoothl1": lambda r, p: F.smooth_l1_loss(r, p, 1.0), "l1": F.l1_loss}[l1_loss.lower()]
    return l1_loss          # any other callable runs as given


class CycleModel(tnn.Module):
    """models/CycleModel.py:7-89; the trilinear sampling bottleneck of 3-D models runs csrc/resize.cu (2-D: torch)."""

    def __init__(self, netG1, netG2, scale_factor_A=None, scale_factor_B=None, split=False, freeze_norms_at=None,
                 **kwargs):
        super().__init__()
        self.netG1, self.netG2 = netG1, netG2
        self.nets = [netG1, netG2]
        self.output_arrays = ["fake_B", "cycled_B", "fake_A", "cycled_A"]
        self.scale_factor_A, self.scale_factor_B, self.split = scale_factor_A, scale_factor_B, split
        self.freeze_norms_at = freeze_norms_at
        self.cycle = True
        self.crop_pad = None
        for k, v in kwargs.items():
            setattr(self, k, v)

    def sampling_bottleneck(self, array, scale_factor):
        if array.dim() == 5 and array.is_cuda and len(scale_factor) == 3:
            return F.sampling_bottleneck(array, scale_factor)      # rg_resize_trilinear_fwd / _bwd (csrc/resize.cu)
        size = array.shape[-len(scale_factor):]
        mode = {2: "bilinear", 3: "trilinear"}[len(size)]
        down = tnn.functional.interpolate(array, scale_factor=scale_factor, mode=mode, align_corners=True)
        return tnn.functional.interpolate(down, size=size, mode=mode, align_corners=True)

    def set_crop_pad(self, crop_pad, ndims):
        self.crop_pad = (slice(None),) * 2 + (slice(crop_pad, -crop_pad),) * ndims

    def set_norm_modes(self, mode="train"):
        for net in self.nets:
            N.set_norm_mode(net, mode)

    def update_status(self, step):
        if self.freeze_norms_at is not None and step >= self.freeze_norms_at:
            self.set_norm_modes(mode="fix_norms")

    def _half(self, first, second, real, scale):
        fake = first(real)
        if self.crop_pad is not None:
            fake = fake[self.crop_pad]
        if scale:
            fake = self.sampling_bottleneck(fake, scale)
        cycled = None
        if self.cycle:
            cycled = second(fake.detach() if self.split else fake)
            if self.crop_pad is not None:
                cycled = cycled[self.crop_pad]
        return fake, cycled

    def forward(self, real_A=None, real_B=None):
        assert real_A is not None or real_B is not None, "Must have some real input to generate outputs)"
        fake_B = cycled_A = fake_A = cycled_B = None
        if real_A is not None:
            fake_B, cycled_A = self._half(self.netG1, self.netG2, real_A, self.scale_factor_B)
        if real_B is not None:
            fake_A, cycled_B = self._half(self.netG2, self.netG1, real_B, self.scale_factor_A)
        return fake_B, cycled_B, fake_A, cycled_A


_DEFAULT_G_LAMBDA = {s: {"l1_loss": {"cycled": 10, "identity": 0}, "gan_loss": {"fake": 1, "cycled": 0}} for s in "AB"}
_DEFAULT_D_LAMBDA = {s: {"real": 1, "fake": 1, "cycled": 0} for s in "AB"}


class LinkCycleLoss(tnn.Module):
    """losses/LinkCycleLoss.py.  ``forward`` performs the whole optimisation step, in the reference's
    order: G losses -> backward (per side, retain_graph) -> optimizer_G.step -> D losses on the
    PRE-step fakes (detached) -> backward -> optimizer_D.step; returns the total whose ``backward`` is a
    no-op.  Extras: ``grad_sync(phase)`` is called right before each optimizer step (data-parallel
    gradient all-reduce hooks in there); ``single_backward`` sums the two generator losses and
    back-propagates once (same gradients, 4 instead of 6 generator traversals)."""

    def __init__(self, netD1, netG1, netD2, netG2, optimizer_G, optimizer_D, dims, l1_loss=None, g_lambda_dict=None,
                 d_lambda_dict=None, gan_mode="lsgan", grad_sync=None, single_backward=False, **kwargs):
        super().__init__()
        self.netD1, self.netG1, self.netD2, self.netG2 = netD1, netG1, netD2, netG2
        self.optimizer_G, self.optimizer_D, self.dims = optimizer_G, optimizer_D, dims
        self._l1_arg = l1_loss if l1_loss is not None else tnn.SmoothL1Loss()
        self.l1_loss = _cycle_loss_fn(l1_loss)
        self.g_lambda_dict = g_lambda_dict or _DEFAULT_G_LAMBDA
        self.d_lambda_dict = d_lambda_dict or _DEFAULT_D_LAMBDA
        self.gan_mode = gan_mode
        self.gan_loss = GANLoss(gan_mode=gan_mode)
        self.grad_sync, self.single_backward = grad_sync, single_backward
        self.loss_dict, self.data_dict = {}, {}
        for k, v in kwargs.items():
            setattr(self, k, v)

    # -- BaseCompetentLoss helpers ---------------------------------------------------------------
    def set_requires_grad(self, nets, requires_grad=False):
        for net in nets if isinstance(nets, list) else [nets]:
            if net is not None:
                for p in net.parameters():
                    p.requires_grad = requires_grad

    def crop(self, x, shape):
        """Centre crop as a strided view (BaseCompetentLoss.py:33-42); the kernels are stride-generic."""
        nd = len(shape)
        sl = [slice(None)] * (x.dim() - nd)
        for have, want in zip(x.shape[-nd:], shape):
            o = (have - want) // 2
            sl.append(slice(o, o + want))
        return x[tuple(sl)]

    def clamp_weights(self, net, min=-0.01, max=0.01):
        for module in net.model:
            if hasattr(module, "weight") and hasattr(module.weight, "data"):
                module.weight.data = module.weight.data.clamp(min, max)

    # -- the step ----------------------------------------------------------------------------------
    def _d_loss(self, side, dnet, data):
        loss = 0
        for key, lam in self.d_lambda_dict[side].items():
            if lam != 0:
                this = self.gan_loss(dnet(data[key].detach()), key == "real")
                self.loss_dict[f"Discriminator_{side}/{key}"] = this
                loss = loss + lam * this
        return loss

    def backward_D(self, side, dnet, data_dict):
        loss = self._d_loss(side, dnet, data_dict)
        loss.backward()
        return loss

    def backward_Ds(self, data_dict, n_loop=5):
        self.set_requires_grad([self.netG1, self.netG2], False)
        self.optimizer_D.zero_grad(set_to_none=True)      # once, BEFORE the wgangp loop, as LinkCycleLoss.py:66 does
        loops = n_loop if self.gan_mode.lower() == "wgangp" else 1
        for _ in range(loops):
            loss_D1 = self.backward_D("B", self.netD1, data_dict["B"])
            loss_D2 = self.backward_D("A", self.netD2, data_dict["A"])
            if self.grad_sync is not None:
                self.grad_sync("D")
            self.optimizer_D.step()
            if loops > 1:
                self.clamp_weights(self.netD1)
                self.clamp_weights(self.netD2)
        self.set_requires_grad([self.netG1, self.netG2], True)
        return loss_D1, loss_D2

    def _g_loss(self, side, gnet, dnet, data):
        loss = 0
        real = data["real"]
        for fcn_name, lambdas in self.g_lambda_dict[side].items():
            for key, lam in lambdas.items():
                if lam == 0:
                    continue
                if key == "identity" and key not in data:
                    data["identity"] = gnet(real)
                pred = data[key]
                if fcn_name == "l1_loss":
                    r = real
                    if tuple(real.shape[-self.dims:]) != tuple(pred.shape[-self.dims:]):
                        r = self.crop(real, pred.shape[-self.dims:])
                    this = self.l1_loss(r, pred)
                else:
                    this = self.gan_loss(dnet(pred), True)
                self.loss_dict[f"{fcn_name}/{key}_{side}"] = this
                loss = loss + lam * this
        return loss

    def backward_G(self, side, gnet, dnet, data_dict):
        loss = self._g_loss(side, gnet, dnet, data_dict)
        loss.backward(retain_graph=True)
        return loss

    def backward_Gs(self, data_dict, defer_step=False):
        """LinkCycleLoss.py:116-135.  defer_step (data-parallel overlap): the generator gradients' all-reduce is only
        STARTED here; `finish_Gs` waits for it and steps the optimizer after the discriminator phase, which reads neither
        the generator weights nor their gradients (it works on the detached, pre-step fakes) -- same results, with the
        exchange hidden behind the discriminators' forward and backward."""
        self.set_requires_grad([self.netD1, self.netD2], False)
        self.optimizer_G.zero_grad(set_to_none=True)
        if self.single_backward:
            loss_G1 = self._g_loss("B", self.netG1, self.netD1, data_dict["B"])
            loss_G2 = self._g_loss("A", self.netG2, self.netD2, data_dict["A"])
            (loss_G1 + loss_G2).backward()
        else:
            loss_G1 = self.backward_G("B", self.netG1, self.netD1, data_dict["B"])
            loss_G2 = self.backward_G("A", self.netG2, self.netD2, data_dict["A"])
        if defer_step:
            self.grad_sync.begin("G")
        else:
            if self.grad_sync is not None:
                self.grad_sync("G")
            self.optimizer_G.step()
        self.set_requires_grad([self.netD1, self.netD2], True)
        return loss_G1, loss_G2

    def finish_Gs(self):
        self.grad_sync.finish("G")
        self.optimizer_G.step()

    def forward(self, real_A, fake_A, cycled_A, real_B, fake_B, cycled_B):
        self.data_dict.update(real_A=real_A, fake_A=fake_A, cycled_A=cycled_A, real_B=real_B, fake_B=fake_B,
                              cycled_B=cycled_B)
        if tuple(real_A.shape[-self.dims:]) != tuple(fake_B.shape[-self.dims:]):
            real_A = self.crop(real_A, fake_A.shape[-self.dims:])
            real_B = self.crop(real_B, fake_B.shape[-self.dims:])
        data = {"A": {"real": real_A, "fake": fake_A, "cycled": cycled_A},
                "B": {"real": real_B, "fake": fake_B, "cycled": cycled_B}}
        overlap = (self.grad_sync is not None and hasattr(self.grad_sync, "begin") and type(self) is LinkCycleLoss
                   and "G" in getattr(self.grad_sync, "flat", {}))
        loss_G1, loss_G2 = self.backward_Gs(data, defer_step=True) if overlap else self.backward_Gs(data)
        loss_D1, loss_D2 = self.backward_Ds(data)
        if overlap:
            self.finish_Gs()
        # one host sync for the four logged scalars (the reference does four float() calls); under CUDA-graph capture
        # (CycleGANStep(graph=True)) the read is deferred to after the replay: log_totals()
        self.loss_vec = torch.stack([loss_D1.detach(), loss_D2.detach(), loss_G1.detach(), loss_G2.detach()])
        if not getattr(self, "defer_host_sync", False):
            self.log_totals()
        total = (loss_G1 + loss_G2 + loss_D1 + loss_D2).detach()
        total.backward = lambda: None
        return total

    def log_totals(self):
        vals = self.loss_vec.tolist()
        self.loss_dict.update({"Total_Loss/D1": vals[0], "Total_Loss/D2": vals[1], "Total_Loss/G1": vals[2],
                               "Total_Loss/G2": vals[3]})
        return vals


class SplitCycleLoss(LinkCycleLoss):
    """losses/SplitCycleLoss.py:10-194: the two generators are optimised one after the other with their own optimizers
    (used with ``CycleModel(split=True)``, whose second traversal sees the detached fake, CycleModel.py:52-57,75-80):
    G1 loss -> backward -> optimizer_G1.step; G2 loss -> backward -> optimizer_G2.step; then the discriminators as in
    LinkCycleLoss.  ``grad_sync`` is called with "G1", "G2" and "D"."""

    def __init__(self, netD1, netG1, netD2, netG2, optimizer_G1, optimizer_G2, optimizer_D, dims, l1_loss=None,
                 g_lambda_dict=None, d_lambda_dict=None, gan_mode="lsgan", grad_sync=None, **kwargs):
        super().__init__(netD1, netG1, netD2, netG2, None, optimizer_D, dims, l1_loss=l1_loss, g_lambda_dict=g_lambda_dict,
                         d_lambda_dict=d_lambda_dict, gan_mode=gan_mode, grad_sync=grad_sync, **kwargs)
        self.optimizer_G1, self.optimizer_G2 = optimizer_G1, optimizer_G2

    def backward_G(self, side, gnet, dnet, data_dict):
        loss = self._g_loss(side, gnet, dnet, data_dict)
        loss.backward()                                           # SplitCycleLoss.py:116 (no retain_graph)
        return loss

    def backward_Gs(self, data_dict):
        self.set_requires_grad([self.netD1, self.netD2], False)
        # G1 first (SplitCycleLoss.py:125-137)
        self.set_requires_grad([self.netG1], True)
        self.set_requires_grad([self.netG2], False)
        self.optimizer_G1.zero_grad(set_to_none=True)
        loss_G1 = self.backward_G("B", self.netG1, self.netD1, data_dict["B"])
        if self.grad_sync is not None:
            self.grad_sync("G1")
        self.optimizer_G1.step()
        # then G2 (SplitCycleLoss.py:139-151)
        self.set_requires_grad([self.netG2], True)
        self.set_requires_grad([self.netG1], False)
        self.optimizer_G2.zero_grad(set_to_none=True)
        loss_G2 = self.backward_G("A", self.netG2, self.netD2, data_dict["A"])
        if self.grad_sync is not None:
            self.grad_sync("G2")
        self.optimizer_G2.step()
        self.set_requires_grad([self.netG1, self.netD1, self.netD2], True)
        return loss_G1, loss_G2


# ---------------------------------------------------------------------------------------------
# the training step harness (what gp.torch.Train does around model + loss)
# ---------------------------------------------------------------------------------------------
class CycleGANStep:
    """model(real_A, real_B) -> loss(real_A, fake_A, cycled_A, real_B, fake_B, cycled_B); Adam for G1+G2 and
    D1+D2 as wired by systems/CycleGAN.py:167-197 (lr 1e-5, betas (0.9, 0.999) from default_cycleGAN_conf.json).

    graph=True: after `graph_warmup` eager steps the whole step -- four generator forwards, the six loss/backward
    traversals, the gradient all-reduces of `grad_sync` (NCCL collectives are capturable) and both Adam updates
    (capturable=True: step counters on the device) -- is captured ONCE into a CUDA graph over static input buffers and
    replayed afterwards: one launch per step instead of ~1500.  At one 64^3 sample per GPU (cfg3 on 8 GPUs) the eager
    step is bound by the host's launch rate, not by the device.  Input shapes must not change after the capture."""

    def __init__(self, g1, g2, d1, d2, lr=1e-5, betas=(0.9, 0.999), l1_loss=None, g_lambda_dict=None,
                 d_lambda_dict=None, gan_mode="lsgan", grad_sync=None, single_backward=False, dims=3,
                 optimizer_cls=None, graph=False, graph_warmup=3):
        """optimizer_cls: None = `optim.FlatAdam` (flat fp32 buckets, one fused Adam kernel, tensor-core packs re-packed
        in place; parameters on a CUDA device) -- "torch" = `torch.optim.Adam` (capturable under `graph`), or any
        optimizer class.  grad_sync: a `dist.GradSync`, or "auto" to build one over the two optimizers when
        torch.distributed is initialised."""
        on_gpu = next(g1.parameters()).is_cuda
        if optimizer_cls is None and on_gpu:
            opt = FlatAdam
        elif optimizer_cls in (None, "torch"):
            opt = functools.partial(torch.optim.Adam, capturable=True) if graph else torch.optim.Adam
        else:
            opt = optimizer_cls
        self.nets = (g1, g2, d1, d2)
        self.optimizer_G = opt(itertools.chain(g1.parameters(), g2.parameters()), lr=lr, betas=betas)
        self.optimizer_D = opt(itertools.chain(d1.parameters(), d2.parameters()), lr=lr, betas=betas)
        self.managed_packs = isinstance(self.optimizer_G, FlatAdam) and isinstance(self.optimizer_D, FlatAdam)
        if grad_sync == "auto":
            from .dist import GradSync
            grad_sync = GradSync({"G": self.optimizer_G if isinstance(self.optimizer_G, FlatAdam) else
                                  list(g1.parameters()) + list(g2.parameters()),
                                  "D": self.optimizer_D if isinstance(self.optimizer_D, FlatAdam) else
                                  list(d1.parameters()) + list(d2.parameters())}) if torch.distributed.is_initialized() else None
        self.model = CycleModel(g1, g2)
        self.loss = LinkCycleLoss(d1, g1, d2, g2, self.optimizer_G, self.optimizer_D, dims, l1_loss=l1_loss,
                                  g_lambda_dict=g_lambda_dict, d_lambda_dict=d_lambda_dict, gan_mode=gan_mode,
                                  grad_sync=grad_sync, single_backward=single_backward)
        self.want_graph, self.graph_warmup = bool(graph), int(graph_warmup)
        self.graph = None                 # torch.cuda.CUDAGraph once captured
        self.graph_error = None           # why the capture was abandoned (the step then stays eager)
        self.graph_launches = 0           # C-ABI launches recorded into the graph
        self._eager_steps = 0

    def _run(self, real_A, real_B):
        fake_B, cycled_B, fake_A, cycled_A = self.model(real_A, real_B)
        self.loss(real_A, fake_A, cycled_A, real_B, fake_B, cycled_B)

    def _totals(self):
        d = self.loss.loss_dict
        return {"G1": d["Total_Loss/G1"], "G2": d["Total_Loss/G2"], "D1": d["Total_Loss/D1"], "D2": d["Total_Loss/D2"]}

    def _capture(self, real_A, real_B):
        import gc
        from . import _C
        # the eager steps' autograd graphs (kept alive by the logged loss tensors) own AccumulateGrad nodes bound to the
        # eager stream; a gradient accumulated through one of them would leave the capturing stream and invalidate the
        # capture.  Drop every reference so the capture builds its own nodes.
        self.loss.loss_dict.clear()
        self.loss.data_dict.clear()
        self.loss.loss_vec = None
        gc.collect()
        CV.invalidate_packs()             # every weight pack is (re)built inside the capture: no pack kernel is skipped
        torch.cuda.synchronize()
        self._in_a, self._in_b = real_A.clone(), real_B.clone()
        graph = torch.cuda.CUDAGraph()
        self.loss.defer_host_sync = True
        l0 = _C.launch_count
        try:
            with torch.cuda.graph(graph):
                self._run(self._in_a, self._in_b)
        except Exception as exc:          # noqa: BLE001 -- any capture failure: stay eager, say why
            self.loss.defer_host_sync = False
            self.want_graph, self.graph_error = False, repr(exc)
            for net in self.nets:         # the step toggles requires_grad around its phases: undo a half-run step
                for p in net.parameters():
                    p.requires_grad = True
            torch.cuda.synchronize()
            return False
        self.graph, self.graph_launches = graph, _C.launch_count - l0
        return True

    def step(self, real_A, real_B):
        if self.want_graph and self.graph is None and self._eager_steps >= self.graph_warmup:
            self._capture(real_A, real_B)
        if self.graph is not None:
            if tuple(real_A.shape) != tuple(self._in_a.shape) or tuple(real_B.shape) != tuple(self._in_b.shape):
                raise ValueError("CycleGANStep(graph=True): input shapes are frozen by the captured graph")
            self._in_a.copy_(real_A, non_blocking=True)
            self._in_b.copy_(real_B, non_blocking=True)
            self.graph.replay()
            if not self.managed_packs:
                # a torch optimizer updated the weights inside the graph without touching Python's version counters:
                # eager forwards after this replay must not trust their cached packs (FlatAdam re-packs in place instead)
                CV.invalidate_packs()
            self.loss.log_totals()
            return self._totals()
        self._eager_steps += 1
        self._run(real_A, real_B)
        return self._totals()
