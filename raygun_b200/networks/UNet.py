"""U-Net -- drop-in for ``raygun.torch.networks.UNet`` (reference: networks/UNet.py:12-666).

Same classes (ConvPass, ConvDownsample, MaxDownsample, Upsample, UNet), constructor kwargs,
attributes (``l_conv, l_down, r_up, r_conv, ndims, num_levels, num_heads, dims ...``) and
``state_dict`` keys.  3-D networks run on the sm_100a kernels: conv(+bias+act) epilogues, the skip
connection's crop + concat as two strided copies into one NDHWC buffer, max-pool / nearest
up-sampling / k=s transposed convolution kernels.  2-D kernels build stock torch.nn layers.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn as tnn

from .. import functional as F
from .. import kernels as K
from .. import nn as L
from .._C import ACT_NONE
from .utils import NoiseBlock, ParameterizedNoiseBlock


def _conv_cls(dims):
    if dims == 3:
        return L.Conv3d
    if dims == 2:
        return tnn.Conv2d
    raise RuntimeError("%dD convolution not implemented" % dims)


def _activation(activation, dims):
    if dims == 3:
        return L.make_activation(activation)
    if activation is None:
        return tnn.Identity()
    return getattr(tnn, activation)() if isinstance(activation, str) else activation()


def _center_crop(x, shape):
    """Centre crop of the trailing len(shape) dims, as a view (UNet.py:97-106)."""
    nd = len(shape)
    sl = [slice(None)] * (x.dim() - nd)
    for have, want in zip(x.shape[-nd:], shape):
        o = (have - want) // 2
        sl.append(slice(o, o + want))
    return x[tuple(sl)]


class ConvPass(tnn.Module):
    """k x (conv [+norm] +activation), optional residual via a grouped 1x1 map (UNet.py:12-117);
    ``final`` (ResidualUNet.py:23,93-97,123-126) drops the trailing activation."""

    def __init__(self, input_nc, output_nc, kernel_sizes, activation, padding="valid", residual=False,
                 padding_mode="reflect", norm_layer=None, final=False):
        super().__init__()
        self.dims = len(kernel_sizes[0])
        conv = _conv_cls(self.dims)
        self.activation = _activation(activation, self.dims)
        self.residual = residual
        self.padding = padding
        self.final = final
        layers = []
        last = len(kernel_sizes) - 1
        for i, ks in enumerate(kernel_sizes):
            self.dims = len(ks)
            layers.append(conv(input_nc, output_nc, ks, padding=padding, padding_mode=padding_mode))
            if residual and i == 0:
                self.x_init_map = conv(input_nc, output_nc, tuple(int(v) for v in np.ones(self.dims, dtype=int)),
                                       padding=padding, padding_mode=padding_mode, bias=False,
                                       groups=min(input_nc, output_nc))
            if norm_layer is not None:
                layers.append(L.make_norm(norm_layer, output_nc) if self.dims == 3 else norm_layer(output_nc))
            if not ((residual or final) and i == last):
                layers.append(self.activation)
            input_nc = output_nc
        self.conv_pass = tnn.Sequential(*layers)

    def crop(self, x, shape):
        return _center_crop(x, tuple(shape))

    def forward(self, x, out=None):
        """`out` (3-D inference only): where the last convolution should write (a slice of the skip concat buffer)."""
        if self.dims != 3:
            if not self.residual:
                return self.conv_pass(x)
            res = self.conv_pass(x)
            init_x = self.x_init_map(x)
            if self.padding.lower() == "valid":
                init_x = self.crop(init_x, res.shape[-self.dims:])
            return init_x + res if self.final else self.activation(init_x + res)
        if not self.residual:
            return L.run_layers(self.conv_pass, x, out=out)
        res = L.run_layers(self.conv_pass, x)
        init_x = self.x_init_map(x)
        if self.padding.lower() == "valid":
            init_x = self.crop(init_x, res.shape[-3:])
        if self.final:
            return F.add_act(res, init_x)
        ac = L.act_code(self.activation)
        if ac is not None:
            return F.add_act(res, init_x, ac[0], ac[1])
        return self.activation(F.add_act(res, init_x))


class ConvDownsample(tnn.Module):
    """Strided valid convolution [+norm] + activation (UNet.py:120-188)."""

    def __init__(self, input_nc, output_nc, kernel_sizes, downsample_factor, activation, padding="valid",
                 padding_mode="reflect", norm_layer=None):
        super().__init__()
        self.dims = len(kernel_sizes)
        conv = _conv_cls(self.dims)
        self.activation = _activation(activation, self.dims)
        self.padding = padding
        layers = [conv(input_nc, output_nc, kernel_sizes, stride=downsample_factor, padding="valid",
                       padding_mode=padding_mode)]
        if norm_layer is not None:
            layers.append(L.make_norm(norm_layer, output_nc) if self.dims == 3 else norm_layer(output_nc))
        layers.append(self.activation)
        self.conv_pass = tnn.Sequential(*layers)

    def forward(self, x):
        return L.run_layers(self.conv_pass, x) if self.dims == 3 else self.conv_pass(x)


class MaxDownsample(tnn.Module):
    """MaxPool(f, stride=f); ``flexible`` lets the pool drop the ragged border (UNet.py:191-238)."""

    def __init__(self, downsample_factor, flexible=True):
        super().__init__()
        self.dims = len(downsample_factor)
        self.downsample_factor = downsample_factor
        self.flexible = flexible
        pool = {2: tnn.MaxPool2d, 3: L.MaxPool3d, 4: L.MaxPool3d}[self.dims]
        self.down = pool(downsample_factor, stride=downsample_factor)

    def forward(self, x):
        if not self.flexible:
            self.check_mismatch(x.size())
            return self.down(x)
        try:
            return self.down(x)
        except Exception:
            self.check_mismatch(x.size())
            raise

    def check_mismatch(self, size):
        for d in range(1, self.dims + 1):
            if size[-d] % self.downsample_factor[-d] != 0:
                raise RuntimeError("Can not downsample shape %s with factor %s, mismatch in spatial dimension %d"
                                   % (size, self.downsample_factor, self.dims - d))


class Upsample(tnn.Module):
    """Up-sample g_out, crop for translation equivariance, crop the skip, concatenate
    (UNet.py:241-351)."""

    def __init__(self, scale_factor, mode=None, input_nc=None, output_nc=None, crop_factor=None,
                 next_conv_kernel_sizes=None):
        super().__init__()
        if crop_factor is not None:
            assert next_conv_kernel_sizes is not None, "crop_factor and next_conv_kernel_sizes have to be given together"
        self.crop_factor = crop_factor
        self.next_conv_kernel_sizes = next_conv_kernel_sizes
        self.dims = len(scale_factor)
        if mode == "transposed_conv":
            up = {2: tnn.ConvTranspose2d, 3: L.ConvTranspose3d}[self.dims]
            self.up = up(input_nc, output_nc, kernel_size=scale_factor, stride=scale_factor)
        else:
            up = L.Upsample if self.dims == 3 else tnn.Upsample
            self.up = up(scale_factor=tuple(scale_factor), mode=mode)

    def _factor_target(self, spatial, factor, kernel_sizes):
        # largest s' = n*f + c <= s with c the crop the following valid convs will take (UNet.py:283-312)
        crop = tuple(sum(ks[d] - 1 for ks in kernel_sizes) for d in range(self.dims))
        ns = (int(math.floor(float(s - c) / f)) for s, c, f in zip(spatial, crop, factor))
        return tuple(n * f + c for n, c, f in zip(ns, crop, factor)), crop

    def crop_to_factor(self, x, factor, kernel_sizes):
        spatial = tuple(x.shape[-self.dims:])
        target, crop = self._factor_target(spatial, factor, kernel_sizes)
        if target != spatial:
            assert all(t > c for t, c in zip(target, crop)), (
                "Feature map with shape %s is too small to ensure translation equivariance with factor %s and "
                "following convolutions %s" % (tuple(x.shape), factor, kernel_sizes))
            return self.crop(x, target)
        return x

    def crop(self, x, shape):
        return _center_crop(x, tuple(shape))

    def forward(self, f_left, g_out):
        g_up = self.up(g_out)
        g_cropped = self.crop_to_factor(g_up, self.crop_factor, self.next_conv_kernel_sizes) \
            if self.crop_factor is not None else g_up
        target = tuple(g_cropped.shape[-self.dims:])
        if self.dims == 3:
            return F.crop_cat(f_left, g_up, target)      # both centre crops + concat: two strided copies
        return torch.cat([self.crop(f_left, target), g_cropped], dim=1)


def _build_unet(self, input_nc, ngf, fmap_inc_factor, downsample_factors, kernel_size_down=None, kernel_size_up=None,
                activation="ReLU", output_nc=None, num_heads=1, constant_upsample=False, downsample_method="max",
                padding_type="valid", residual=False, norm_layer=None, add_noise=False, final_level0=False):
    self.ndims = len(downsample_factors[0])
    self.num_levels = len(downsample_factors) + 1
    self.num_heads = num_heads
    self.input_nc = input_nc
    self.output_nc = output_nc if output_nc else ngf
    self.residual = residual
    if add_noise == "param":
        self.noise_layer = ParameterizedNoiseBlock()
    elif add_noise:
        self.noise_layer = NoiseBlock()
    else:
        self.noise_layer = None
    nl, inc = self.num_levels, fmap_inc_factor
    if kernel_size_down is None:
        kernel_size_down = [[(3,) * self.ndims, (3,) * self.ndims]] * nl
    if kernel_size_up is None:
        kernel_size_up = [[(3,) * self.ndims, (3,) * self.ndims]] * (nl - 1)

    # crop factors for translation equivariance under valid padding (UNet.py:506-522)
    crop_factors, prod = [], None
    for factor in downsample_factors[::-1]:
        pt = padding_type.lower()
        if pt == "valid":
            prod = list(factor) if prod is None else [a * b for a, b in zip(factor, prod)]
        elif pt == "same":
            prod = None
        else:
            raise RuntimeError(f"Invalid padding_type option: {padding_type}")
        crop_factors.append(prod)
    crop_factors = crop_factors[::-1]

    is_max = downsample_method.lower() == "max"
    self.l_conv = tnn.ModuleList([
        ConvPass(input_nc if lvl == 0 else ngf * inc ** (lvl - is_max), ngf * inc ** lvl, kernel_size_down[lvl],
                 activation=activation, padding=padding_type, residual=residual, norm_layer=norm_layer)
        for lvl in range(nl)])
    self.dims = self.l_conv[0].dims
    if is_max:
        self.l_down = tnn.ModuleList([MaxDownsample(downsample_factors[lvl]) for lvl in range(nl - 1)])
    elif downsample_method.lower() == "convolve":
        self.l_down = tnn.ModuleList([
            ConvDownsample(ngf * inc ** lvl, ngf * inc ** (lvl + 1), kernel_size_down[lvl][0], downsample_factors[lvl],
                           activation=activation, padding=padding_type, norm_layer=norm_layer)
            for lvl in range(nl - 1)])
    else:
        raise RuntimeError(f'Unknown downsampling method {downsample_method}. Use "max" or "convolve" instead.')
    noise = add_noise is not False
    self.r_up = tnn.ModuleList([
        tnn.ModuleList([
            Upsample(downsample_factors[lvl], mode="nearest" if constant_upsample else "transposed_conv",
                     input_nc=ngf * inc ** (lvl + 1) + int(lvl == 1 and noise), output_nc=ngf * inc ** (lvl + 1),
                     crop_factor=crop_factors[lvl], next_conv_kernel_sizes=kernel_size_up[lvl])
            for lvl in range(nl - 1)])
        for _ in range(num_heads)])
    self.r_conv = tnn.ModuleList([
        tnn.ModuleList([
            ConvPass(ngf * inc ** lvl + ngf * inc ** (lvl + 1),
                     ngf * inc ** lvl if output_nc is None or lvl != 0 else output_nc, kernel_size_up[lvl],
                     activation=activation, padding=padding_type, residual=residual, norm_layer=norm_layer,
                     **({"final": lvl == 0} if final_level0 else {}))
            for lvl in range(nl - 1)])
        for _ in range(num_heads)])


class UNet(tnn.Module):
    """UNet.py:354-666."""

    def __init__(self, input_nc, ngf, fmap_inc_factor, downsample_factors, kernel_size_down=None, kernel_size_up=None,
                 activation="ReLU", output_nc=None, num_heads=1, constant_upsample=False, downsample_method="max",
                 padding_type="valid", residual=False, norm_layer=None, add_noise=False):
        super().__init__()
        _build_unet(self, input_nc, ngf, fmap_inc_factor, downsample_factors, kernel_size_down, kernel_size_up,
                    activation, output_nc, num_heads, constant_upsample, downsample_method, padding_type, residual,
                    norm_layer, add_noise)

    def _cat_channels(self, i):
        """(skip channels, up-sampled channels) of level i when the fused concat path applies, else None."""
        up = self.r_up[0][i].up
        last = [m for m in self.l_conv[i].conv_pass if isinstance(m, (L.Conv3d,))]
        if not (self.num_heads == 1 and self.ndims == 3 and not self.residual and last
                and isinstance(up, L.ConvTranspose3d) and self.r_up[0][i].crop_factor is None
                and str(self.l_conv[i].padding).lower() == "same"):
            return None
        return last[-1].out_channels, up.out_channels

    def last_conv_before_head(self):
        """The convolution whose activated output IS the network output (the generator head may be folded into its
        epilogue, cyclegan.Generator): the last conv of the top r_conv ConvPass of a plain single-head U-Net."""
        if self.ndims != 3 or self.num_heads != 1 or self.residual or self.noise_layer is not None or self.num_levels < 2:
            return None
        layers = list(self.r_conv[0][0].conv_pass)
        if len(layers) >= 2 and isinstance(layers[-2], L.Conv3d) and L.act_code(layers[-1]) is not None:
            return layers[-2]
        return None

    def _train_cat_forward(self, level, i, f_in):
        """Training twin of the in-place skip concat below: the left ConvPass and the transposed conv write their outputs
        into the two channel slices of ONE pad-arena concat buffer (functional.mark_train_out) and `F.cat_join` ties the
        buffer to both producers for autograd -- the two strided copies of `F.crop_cat` per level and traversal leave the
        step.  Returns None (nothing executed) when the level's shape does not qualify."""
        ch = self._cat_channels(i)
        if ch is None or (ch[0] + ch[1]) % 8 or ch[0] % 8:
            return None
        lp = [m for m in self.l_conv[i].conv_pass]
        if not (len(lp) >= 2 and isinstance(lp[-2], L.Conv3d) and L.act_code(lp[-1]) is not None):
            return None                                   # a norm layer behind the last conv: its output is not the conv's
        with torch.no_grad():          # plain buffer views, like the ones the kernels' autograd Functions create themselves
            cat = K.empty5_arena((f_in.shape[0], ch[0] + ch[1], *f_in.shape[2:]), torch.bfloat16, f_in.device)
            left, right = F.mark_train_out(cat[:, : ch[0]]), F.mark_train_out(cat[:, ch[0]:])
        f_left = self.l_conv[i](f_in, out=left)
        if f_left.shape != left.shape or f_left.data_ptr() != left.data_ptr():
            raise RuntimeError("UNet: the skip ConvPass did not write into the concat buffer (unexpected kernel route)")
        g_in = self.l_down[i](f_left)
        gs_out = self.rec_forward(level - 1, g_in)
        g_up = self.r_up[0][i].up.fused(gs_out[0], out=right)
        if g_up.shape != right.shape or g_up.data_ptr() != right.data_ptr():
            # the up-sampled tensor came out elsewhere (e.g. a shape the fused kernel does not take): classic concat
            return [self.r_conv[0][i](F.crop_cat(f_left, g_up, tuple(g_up.shape[2:])))]
        return [self.r_conv[0][i](F.cat_join(f_left, g_up, cat))]

    def rec_forward(self, level, f_in):
        i = self.num_levels - level - 1          # index into the module lists
        cat = None
        train_cat = (level > 0 and f_in.dim() == 5 and f_in.is_cuda and torch.is_grad_enabled() and F.TRAIN_ARENA
                     and F.TRAIN_CAT and K.get_precision() == "bf16" and K.use_tensor_cores())
        if train_cat:
            out = self._train_cat_forward(level, i, f_in)
            if out is not None:
                return out
        if level > 0 and f_in.dim() == 5 and f_in.is_cuda and K.arena_enabled():
            ch = self._cat_channels(i)
            if ch is not None:
                # inference: the skip tensor and the up-sampled tensor are written straight into the two channel
                # slices of ONE padded concat buffer (no crop: 'same' padding), so `torch.cat` (UNet.py:349-351) and
                # the reflect pad of the following conv cost no copy at all
                cat = K.empty5_arena((f_in.shape[0], ch[0] + ch[1], *f_in.shape[2:]), torch.bfloat16, f_in.device)
        if cat is not None:
            left = cat[:, : ch[0]]
            f_left = self.l_conv[i](f_in, out=left)
            if f_left.data_ptr() != left.data_ptr() or f_left.shape != left.shape:
                if f_left.shape == left.shape:
                    K.copy_into(left, f_left)
                    f_left = left
                else:
                    cat = None
        else:
            f_left = self.l_conv[i](f_in)
        if level == 0:
            if self.noise_layer is not None:
                f_left = self.noise_layer(f_left)
            return [f_left] * self.num_heads
        g_in = self.l_down[i](f_left)
        gs_out = self.rec_forward(level - 1, g_in)
        if cat is not None:
            right = cat[:, ch[0]:]
            g_up = self.r_up[0][i].up.fused(gs_out[0], out=right)
            if g_up.shape == right.shape:
                if g_up.data_ptr() != right.data_ptr():
                    K.copy_into(right, g_up)
                return [self.r_conv[0][i](cat)]
            return [self.r_conv[0][i](F.crop_cat(f_left, g_up, tuple(g_up.shape[2:])))]
        return [self.r_conv[h][i](self.r_up[h][i](f_left, gs_out[h])) for h in range(self.num_heads)]

    def _finish(self, ys):
        if self.ndims == 3 and not getattr(self, "_keep_compute_dtype", False):
            ys = [F.cast(y, torch.float32) for y in ys]
        return ys[0] if self.num_heads == 1 else ys

    def forward(self, x):
        if self.ndims == 3:
            x = F.to_compute(x)
        return self._finish(self.rec_forward(self.num_levels - 1, x))
