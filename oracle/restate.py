"""TEST INFRASTRUCTURE ONLY.  Functional CPU restatement of raygun's hot path.

Every function takes a ``state_dict``-like mapping (``name -> tensor``) and a
small spec, and evaluates the reference's algorithm with plain
``torch.nn.functional`` calls on the CPU (fp32 or fp64).  Nothing here is an
``nn.Module`` and nothing imports the product package, so a bug in the
product's module classes cannot hide in the checker.

Parity status: PINNED -- ``tests/test_oracle_pinned.py`` checks every function
against the reference modules imported from ``/root/reference`` (when present)
and against the committed ``tests/golden/*.npz`` generated from them.

Reference citations are relative to ``/root/reference/src/raygun/torch``.
"""
from __future__ import annotations

import math
from typing import Mapping, Sequence

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# --------------------------------------------------------------------------
# small helpers
# --------------------------------------------------------------------------
def center_crop(x: Tensor, spatial: Sequence[int]) -> Tensor:
    """networks/UNet.py:97-106 -- centre crop of the trailing spatial dims."""
    nd = len(spatial)
    sl = [slice(None)] * (x.dim() - nd)
    for have, want in zip(x.shape[-nd:], spatial):
        o = (have - want) // 2
        sl.append(slice(o, o + want))
    return x[tuple(sl)]


def activation_fn(act):
    """networks/UNet.py:41-47 -- name in torch.nn, a class, or None."""
    if act is None:
        return lambda t: t
    if isinstance(act, str):
        return getattr(torch.nn, act)()
    if isinstance(act, type):
        return act()
    return act


def norm_apply(sd: Mapping[str, Tensor], prefix: str, x: Tensor, norm: dict | None,
               training: bool = True) -> Tensor:
    """InstanceNorm3d / BatchNorm3d as torch defines them:
    biased variance for normalisation, eps inside the sqrt."""
    if norm is None:
        return x
    eps = norm.get("eps", 1e-5)
    w = sd.get(prefix + "weight") if norm.get("affine", norm["kind"] == "batch") else None
    b = sd.get(prefix + "bias") if w is not None else None
    rm, rv = sd.get(prefix + "running_mean"), sd.get(prefix + "running_var")
    red = (2, 3, 4) if norm["kind"] == "instance" else (0, 2, 3, 4)
    use_running = (rm is not None) and not training
    if use_running:
        mean = rm.view(1, -1, 1, 1, 1)
        var = rv.view(1, -1, 1, 1, 1)
    else:
        mean = x.mean(dim=red, keepdim=True)
        var = x.var(dim=red, unbiased=False, keepdim=True)
    y = (x - mean) / torch.sqrt(var + eps)
    if w is not None:
        y = y * w.view(1, -1, 1, 1, 1) + b.view(1, -1, 1, 1, 1)
    return y


def conv3d(x, w, b=None, stride=1, pad=0, pad_mode="zeros"):
    """nn.Conv3d semantics; pad may be an int, 'same' or 'valid'."""
    k = w.shape[-3:]
    if pad == "valid":
        pad = 0
    if pad == "same":
        tot = [kk - 1 for kk in k]
        lo = [t // 2 for t in tot]
        hi = [t - l for t, l in zip(tot, lo)]
    else:
        lo = hi = [pad] * 3 if isinstance(pad, int) else list(pad)
    if any(lo) or any(hi):
        pads = (lo[2], hi[2], lo[1], hi[1], lo[0], hi[0])
        mode = {"zeros": "constant", "reflect": "reflect", "replicate": "replicate",
                "circular": "circular"}[pad_mode]
        x = F.pad(x, pads, mode=mode)
    return F.conv3d(x, w, b, stride=stride)


# --------------------------------------------------------------------------
# U-Net family  (networks/UNet.py, networks/ResidualUNet.py)
# --------------------------------------------------------------------------
def _as_k3(k):
    return (k,) * 3 if isinstance(k, int) else tuple(k)


def conv_pass(sd, prefix, x, kernel_sizes, padding, act, residual=False, norm=None,
              final=False, training=True):
    """ConvPass.forward, networks/UNet.py:108-117 (+ResidualUNet.py:93-126).

    Sequential indices follow the constructor (:54-95): per kernel one conv,
    then the norm (if any), then the activation unless this is the last conv of
    a residual (or ``final``) pass.
    """
    f = activation_fn(act)
    idx = 0
    y = x
    nk = len(kernel_sizes)
    for i, _k in enumerate(kernel_sizes):
        w = sd[f"{prefix}conv_pass.{idx}.weight"]
        b = sd.get(f"{prefix}conv_pass.{idx}.bias")
        y = conv3d(y, w, b, pad=padding, pad_mode="reflect")
        idx += 1
        if norm is not None:
            y = norm_apply(sd, f"{prefix}conv_pass.{idx}.", y, norm, training)
            idx += 1
        last = i == nk - 1
        if not ((residual and last) or (final and last)):
            y = f(y)
            idx += 1
    if not residual:
        return y
    wi = sd[f"{prefix}x_init_map.weight"]
    groups = min(wi.shape[0], x.shape[1])
    init_x = F.conv3d(x, wi, None, groups=groups)
    if padding.lower() == "valid":
        init_x = center_crop(init_x, y.shape[-3:])
    return init_x + y if final else f(init_x + y)


def crop_to_factor(x, factor, kernel_sizes):
    """Upsample.crop_to_factor, networks/UNet.py:275-325."""
    sp = x.shape[-3:]
    cc = [sum(_as_k3(k)[d] - 1 for k in kernel_sizes) for d in range(3)]
    ns = [int(math.floor(float(s - c) / f)) for s, c, f in zip(sp, cc, factor)]
    tgt = tuple(n * f + c for n, c, f in zip(ns, cc, factor))
    if tgt != tuple(sp):
        assert all(t > c for t, c in zip(tgt, cc)), "feature map too small"
        return center_crop(x, tgt)
    return x


def unet_spec(input_nc, ngf, fmap_inc_factor, downsample_factors, kernel_size_down=None,
              kernel_size_up=None, activation="ReLU", output_nc=None, num_heads=1,
              constant_upsample=False, downsample_method="max", padding_type="valid",
              residual=False, norm=None, residual_unet=False):
    nl = len(downsample_factors) + 1
    if kernel_size_down is None:
        kernel_size_down = [[(3, 3, 3), (3, 3, 3)]] * nl
    if kernel_size_up is None:
        kernel_size_up = [[(3, 3, 3), (3, 3, 3)]] * (nl - 1)
    return dict(input_nc=input_nc, ngf=ngf, inc=fmap_inc_factor,
                factors=[tuple(f) for f in downsample_factors], ks_down=kernel_size_down,
                ks_up=kernel_size_up, act=activation, output_nc=output_nc,
                heads=num_heads, const_up=constant_upsample, down=downsample_method.lower(),
                padding=padding_type, residual=residual, norm=norm,
                residual_unet=residual_unet, levels=nl)


def unet_forward(sd, spec, x, prefix="", training=True):
    """UNet.rec_forward / forward, networks/UNet.py:626-666;
    ResidualUNet.forward global skip, networks/ResidualUNet.py:689-701."""
    nl = spec["levels"]
    pad = spec["padding"]
    valid = pad.lower() == "valid"
    # crop factors, networks/UNet.py:506-522
    crop_factors, prod = [], None
    for fac in spec["factors"][::-1]:
        if valid:
            prod = list(fac) if prod is None else [a * b for a, b in zip(fac, prod)]
        else:
            prod = None
        crop_factors.append(prod)
    crop_factors = crop_factors[::-1]
    f_act = activation_fn(spec["act"])

    def rec(level, f_in):
        i = nl - level - 1
        f_left = conv_pass(sd, f"{prefix}l_conv.{i}.", f_in, spec["ks_down"][i], pad,
                           spec["act"], spec["residual"], spec["norm"], training=training)
        if level == 0:
            return [f_left] * spec["heads"]
        fac = spec["factors"][i]
        if spec["down"] == "max":
            g_in = F.max_pool3d(f_left, fac, stride=fac)
        else:
            w = sd[f"{prefix}l_down.{i}.conv_pass.0.weight"]
            b = sd.get(f"{prefix}l_down.{i}.conv_pass.0.bias")
            g_in = F.conv3d(f_left, w, b, stride=fac)
            if spec["norm"] is not None:
                g_in = norm_apply(sd, f"{prefix}l_down.{i}.conv_pass.1.", g_in, spec["norm"], training)
            g_in = f_act(g_in)
        gs = rec(level - 1, g_in)
        outs = []
        for h in range(spec["heads"]):
            if spec["const_up"]:
                g_up = F.interpolate(gs[h], scale_factor=tuple(float(v) for v in fac), mode="nearest")
            else:
                g_up = F.conv_transpose3d(gs[h], sd[f"{prefix}r_up.{h}.{i}.up.weight"],
                                          sd.get(f"{prefix}r_up.{h}.{i}.up.bias"), stride=fac)
            if crop_factors[i] is not None:
                g_up = crop_to_factor(g_up, crop_factors[i], spec["ks_up"][i])
            f_c = center_crop(f_left, g_up.shape[-3:])
            cat = torch.cat([f_c, g_up], dim=1)
            final = spec["residual_unet"] and i == 0
            outs.append(conv_pass(sd, f"{prefix}r_conv.{h}.{i}.", cat, spec["ks_up"][i], pad,
                                  spec["act"], spec["residual"], spec["norm"], final=final,
                                  training=training))
        return outs

    ys = rec(nl - 1, x)
    if spec["residual_unet"]:
        # ResidualUNet.py:689-701: act(crop(x) + y) per head
        ys = [f_act(center_crop(x, y.shape[-3:]) + y) if valid else f_act(x + y) for y in ys]
    return ys[0] if spec["heads"] == 1 else ys


# --------------------------------------------------------------------------
# ResNet generator (networks/ResNet.py:226-443)
# --------------------------------------------------------------------------
def resnet_spec(input_nc=1, output_nc=None, ngf=64, norm=dict(kind="instance", affine=False),
                use_dropout=False, n_blocks=6, padding_type="reflect", activation="ReLU",
                n_downsampling=2):
    return dict(input_nc=input_nc, output_nc=output_nc if output_nc is not None else ngf,
                ngf=ngf, norm=norm, n_blocks=n_blocks, padding=padding_type.lower(),
                act=activation, n_down=n_downsampling, use_dropout=use_dropout)


def _explicit_pad(x, p, padding):
    """ReflectionPad3d / ReplicationPad3d layers, networks/ResNet.py:269-279,391-401."""
    if padding in ("reflect", "same"):
        return F.pad(x, (p,) * 6, mode="reflect"), 0
    if padding == "replicate":
        return F.pad(x, (p,) * 6, mode="replicate"), 0
    if padding == "zeros":
        return x, p
    if padding == "valid":
        return x, 0
    raise NotImplementedError(padding)


def resnet3d_forward(sd, spec, x, prefix="", training=True):
    """ResnetGenerator3D.forward, networks/ResNet.py:342-344, unrolled.

    Sequential indices follow the constructor (:281-340): optional padder
    modules occupy an index each; dropout (use_dropout) is not restated
    (stochastic; the benchmark configs leave it off)."""
    assert not spec["use_dropout"], "dropout is stochastic; not restated"
    pad = spec["padding"]
    has_padder = pad in ("reflect", "same", "replicate")
    updown_p = 0 if pad == "valid" else 1
    f = activation_fn(spec["act"])
    norm = spec["norm"]
    m = f"{prefix}model."
    idx = 0

    def nrm(t, i):
        return norm_apply(sd, f"{m}{i}.", t, norm, training)

    # stem: [padder] conv7 norm act
    y, p = _explicit_pad(x, 3, pad)
    idx += 1 if has_padder else 0
    y = F.conv3d(y, sd[f"{m}{idx}.weight"], sd.get(f"{m}{idx}.bias"), padding=p)
    y = f(nrm(y, idx + 1))
    idx += 3
    for _ in range(spec["n_down"]):
        y = F.conv3d(y, sd[f"{m}{idx}.weight"], sd.get(f"{m}{idx}.bias"), stride=2,
                     padding=updown_p)
        y = f(nrm(y, idx + 1))
        idx += 3
    for _ in range(spec["n_blocks"]):
        cb = f"{m}{idx}.conv_block."
        j = 0
        r, p = _explicit_pad(y, 1, pad)
        j += 1 if has_padder else 0
        r = F.conv3d(r, sd[f"{cb}{j}.weight"], sd.get(f"{cb}{j}.bias"), padding=p)
        r = f(norm_apply(sd, f"{cb}{j + 1}.", r, norm, training))
        j += 3
        r, p = _explicit_pad(r, 1, pad)
        j += 1 if has_padder else 0
        r = F.conv3d(r, sd[f"{cb}{j}.weight"], sd.get(f"{cb}{j}.bias"), padding=p)
        r = norm_apply(sd, f"{cb}{j + 1}.", r, norm, training)
        y = (center_crop(y, r.shape[-3:]) if pad == "valid" else y) + r
        idx += 1
    for _ in range(spec["n_down"]):
        y = F.conv_transpose3d(y, sd[f"{m}{idx}.weight"], sd.get(f"{m}{idx}.bias"), stride=2,
                               padding=updown_p, output_padding=updown_p)
        y = f(nrm(y, idx + 1))
        idx += 3
    y, p = _explicit_pad(y, 3, pad)
    idx += 1 if has_padder else 0
    return F.conv3d(y, sd[f"{m}{idx}.weight"], sd.get(f"{m}{idx}.bias"), padding=p)


# --------------------------------------------------------------------------
# PatchGAN discriminator (networks/NLayerDiscriminator.py:104-180)
# --------------------------------------------------------------------------
def nld_spec(input_nc=1, ngf=64, n_layers=3, norm=dict(kind="batch", affine=True), kw=4,
             downsampling_kw=None):
    return dict(input_nc=input_nc, ngf=ngf, n_layers=n_layers, norm=norm, kw=kw,
                ds_kw=downsampling_kw if downsampling_kw is not None else kw)


def nld3d_forward(sd, spec, x, prefix="", training=True):
    m = f"{prefix}model."
    norm = spec["norm"]
    lrelu = lambda t: F.leaky_relu(t, 0.2)
    y = lrelu(F.conv3d(x, sd[f"{m}0.weight"], sd.get(f"{m}0.bias"), stride=2, padding=1))
    idx = 2
    for _ in range(1, spec["n_layers"]):
        y = F.conv3d(y, sd[f"{m}{idx}.weight"], sd.get(f"{m}{idx}.bias"), stride=2, padding=1)
        y = lrelu(norm_apply(sd, f"{m}{idx + 1}.", y, norm, training))
        idx += 3
    y = F.conv3d(y, sd[f"{m}{idx}.weight"], sd.get(f"{m}{idx}.bias"), stride=1, padding=1)
    y = lrelu(norm_apply(sd, f"{m}{idx + 1}.", y, norm, training))
    idx += 3
    return F.conv3d(y, sd[f"{m}{idx}.weight"], sd.get(f"{m}{idx}.bias"), stride=1, padding=1)


# --------------------------------------------------------------------------
# get_network wrapper (systems/BaseSystem.py:209-224): net -> 1x1x1 conv -> act
# --------------------------------------------------------------------------
def generator_forward(sd, kind, spec, x, final_activation="Tanh", padding_type="valid",
                      training=True):
    """Sequential(net, Conv3d(ngf, output_nc, 1, padding=padding_type), act)."""
    if kind == "unet":
        y = unet_forward(sd, spec, x, prefix="0.", training=training)
    elif kind == "resnet":
        y = resnet3d_forward(sd, spec, x, prefix="0.", training=training)
    else:
        raise ValueError(kind)
    y = F.conv3d(y, sd["1.weight"], sd.get("1.bias"))
    return activation_fn(final_activation)(y)


# --------------------------------------------------------------------------
# losses (losses/GANLoss.py:48-64, losses/LinkCycleLoss.py:22,87-114)
# --------------------------------------------------------------------------
def gan_loss(pred, is_real, mode="lsgan"):
    if mode == "lsgan":
        return ((pred - (1.0 if is_real else 0.0)) ** 2).mean()
    if mode == "wgangp":
        return -pred.mean() if is_real else pred.mean()
    if mode == "vanilla":
        t = torch.full_like(pred, 1.0 if is_real else 0.0)
        return F.binary_cross_entropy_with_logits(pred, t)
    raise NotImplementedError(mode)


def smooth_l1(a, b, beta=1.0):
    d = (a - b).abs()
    return torch.where(d < beta, 0.5 * d * d / beta, d - 0.5 * beta).mean()


def l1(a, b):
    return (a - b).abs().mean()


# --------------------------------------------------------------------------
# blockwise prediction arithmetic (predict/worker.py:100-147, predict.py:92-108)
# --------------------------------------------------------------------------
def u8_normalise(vol_u8, scale=2.0, shift=-1.0):
    """worker.py:101-114: float(data); -= iinfo.min; /= iinfo.max; *scale; +shift."""
    x = vol_u8.to(torch.float32)
    x = x - 0.0
    x = x / 255.0
    return x * scale + shift


def denorm_clamp_u8(out):
    """worker.py:133-147: out*255, clamp [0,255], truncating cast to uint8."""
    o = out * 255.0
    o = torch.clamp(o, 0, 255)
    return o.to(torch.uint8)  # truncation toward zero, like numpy astype


def block_grid(vol_shape, read, write):
    """Read/write ROIs of a Daisy-style blockwise task (predict.py:92-108,154-164):
    write blocks tile the volume from the origin; context = (read-write)//2.
    Yields (write_origin, write_shape) with edge blocks clipped to the volume."""
    ctx = tuple((r - w) // 2 for r, w in zip(read, write))
    out = []
    for z in range(0, vol_shape[0], write[0]):
        for y in range(0, vol_shape[1], write[1]):
            for x in range(0, vol_shape[2], write[2]):
                o = (z, y, x)
                s = tuple(min(w, v - oo) for w, v, oo in zip(write, vol_shape, o))
                out.append((o, s))
    return ctx, out


# --------------------------------------------------------------------------
# CycleGAN training step (models/CycleModel.py:37-89 + losses/LinkCycleLoss.py:43-176
# + torch.optim.Adam as wired by systems/CycleGAN.py:167-197)
# --------------------------------------------------------------------------
class AdamState:
    """torch.optim.Adam (no amsgrad, weight_decay 0) restated for a dict of tensors."""

    def __init__(self, params: dict, lr=1e-5, betas=(0.9, 0.999), eps=1e-8):
        self.p, self.lr, self.b1, self.b2, self.eps = params, lr, betas[0], betas[1], eps
        self.m = {k: torch.zeros_like(v) for k, v in params.items()}
        self.v = {k: torch.zeros_like(v) for k, v in params.items()}
        self.t = 0

    def step(self, grads: dict):
        self.t += 1
        c1 = 1 - self.b1 ** self.t
        c2 = 1 - self.b2 ** self.t
        with torch.no_grad():
            for k, g in grads.items():
                if g is None:
                    continue
                self.m[k].mul_(self.b1).add_(g, alpha=1 - self.b1)
                self.v[k].mul_(self.b2).addcmul_(g, g, value=1 - self.b2)
                denom = (self.v[k].sqrt() / math.sqrt(c2)).add_(self.eps)
                self.p[k].addcdiv_(self.m[k], denom, value=-self.lr / c1)


class CycleStep:
    """One LinkCycleLoss.forward over (real_A, real_B), functional.

    ``g_fwd(sd, x)`` / ``d_fwd(sd, x)`` evaluate a generator / discriminator from
    a parameter dict.  Parameter dicts are updated in place, in the reference's
    order: G step first (optimizer_G over G1+G2), then the D step using the
    fakes computed BEFORE the G step (LinkCycleLoss.py:53 ``.detach()``)."""

    def __init__(self, g1, g2, d1, d2, g_fwd, d_fwd, lr=1e-5, betas=(0.9, 0.999),
                 lambda_cycle=10.0, lambda_identity=0.0, lambda_gan=1.0, cycle_loss="smoothl1"):
        self.g1, self.g2, self.d1, self.d2 = g1, g2, d1, d2
        self.g_fwd, self.d_fwd = g_fwd, d_fwd
        self.lc, self.li, self.lg = lambda_cycle, lambda_identity, lambda_gan
        self.cyc = smooth_l1 if cycle_loss == "smoothl1" else l1
        gp = {f"g1.{k}": v for k, v in g1.items()}
        gp.update({f"g2.{k}": v for k, v in g2.items()})
        dp = {f"d1.{k}": v for k, v in d1.items()}
        dp.update({f"d2.{k}": v for k, v in d2.items()})
        self.opt_g = AdamState(gp, lr, betas)
        self.opt_d = AdamState(dp, lr, betas)

    @staticmethod
    def _leaf(sd, on):
        for v in sd.values():
            if v.is_floating_point():
                v.requires_grad_(on)

    def step(self, real_A, real_B):
        g1, g2, d1, d2 = self.g1, self.g2, self.d1, self.d2
        for sd in (g1, g2):
            self._leaf(sd, True)
        for sd in (d1, d2):
            self._leaf(sd, False)
        fake_B = self.g_fwd(g1, real_A)
        cycled_A = self.g_fwd(g2, fake_B)
        fake_A = self.g_fwd(g2, real_B)
        cycled_B = self.g_fwd(g1, fake_A)
        rA = center_crop(real_A, fake_A.shape[-3:])
        rB = center_crop(real_B, fake_B.shape[-3:])

        def g_side(real, cycled, fake, gnet, dnet):
            loss = self.lc * self.cyc(center_crop(real, cycled.shape[-3:]), cycled)
            if self.li != 0:
                ident = self.g_fwd(gnet, real)
                loss = loss + self.li * self.cyc(center_crop(real, ident.shape[-3:]), ident)
            return loss + self.lg * gan_loss(self.d_fwd(dnet, fake), True)

        loss_G1 = g_side(rB, cycled_B, fake_B, g1, d1)
        loss_G2 = g_side(rA, cycled_A, fake_A, g2, d2)
        names = [k for k, v in self.opt_g.p.items() if v.requires_grad]
        leaves = [self.opt_g.p[k] for k in names]
        ga = torch.autograd.grad(loss_G1, leaves, retain_graph=True, allow_unused=True)
        gb = torch.autograd.grad(loss_G2, leaves, allow_unused=True)
        grads = {}
        for k, a, b in zip(names, ga, gb):
            grads[k] = a if b is None else (b if a is None else a + b)
        for sd in (g1, g2):
            self._leaf(sd, False)
        self.opt_g.step(grads)

        for sd in (d1, d2):
            self._leaf(sd, True)
        fB, fA = fake_B.detach(), fake_A.detach()
        loss_D1 = gan_loss(self.d_fwd(d1, rB), True) + gan_loss(self.d_fwd(d1, fB), False)
        loss_D2 = gan_loss(self.d_fwd(d2, rA), True) + gan_loss(self.d_fwd(d2, fA), False)
        names = [k for k, v in self.opt_d.p.items() if v.requires_grad]
        leaves = [self.opt_d.p[k] for k in names]
        gd = torch.autograd.grad(loss_D1 + loss_D2, leaves, allow_unused=True)
        for sd in (d1, d2):
            self._leaf(sd, False)
        self.opt_d.step(dict(zip(names, gd)))
        return dict(G1=float(loss_G1), G2=float(loss_G2), D1=float(loss_D1), D2=float(loss_D2))
