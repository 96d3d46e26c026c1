"""GPU parity of the small-K stem convolution (stem_conv_sm100.cu: im2col rows gathered straight into the tensor-core
operand tile) against fp32 convolution of the same bf16-rounded operands: the unrolled 1 -> 64 3^3 U-Net stem, the
generic tap loops (K = 9 .. 64, several K steps), ragged tiles, zero padding, strided outputs and inputs."""
import pytest
import torch
import torch.nn.functional as TF

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _q(t):
    return t.bfloat16().float()


def _rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    assert a.shape == b.shape
    return float((a - b).abs().max()) / max(1e-6, float(b.abs().max()))


CASES = [
    # cin, cout, kernel, pad, spatial, batch
    (1, 64, (3, 3, 3), (1, 1, 1), (9, 17, 11), 2),        # the U-Net stem (unrolled gather), ragged tiles, zero padding
    (1, 64, (3, 3, 3), (0, 0, 0), (10, 18, 10), 1),       # valid
    (1, 32, (1, 3, 3), (0, 1, 1), (4, 16, 8), 1),         # K = 9: one K step
    (2, 64, (2, 3, 3), (1, 1, 1), (5, 9, 12), 1),         # K = 36: three K steps, two input channels
    (1, 128, (4, 4, 4), (2, 2, 2), (6, 10, 9), 1),        # K = 64: four K steps, 128 produced channels
    (3, 256, (1, 1, 1), (0, 0, 0), (3, 16, 16), 2),       # K = 3, widest output
]


@pytest.mark.parametrize("case", CASES)
def test_stem_conv_matches_fp32_reference(case):
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    cin, cout, k, p, sp, n = case
    torch.manual_seed(31)
    x = _q(torch.randn(n, cin, *sp))
    conv = torch.nn.Conv3d(cin, cout, k, padding=p)
    w, b = _q(conv.weight.detach()), conv.bias.detach()
    ref = TF.relu(TF.conv3d(x, w, b, padding=p))
    with rb.precision("bf16"), torch.no_grad():
        assert CV.stem_direct_ok(w, False)
        before = CV.stats.get("stem_direct", 0)
        y = CV.stem_conv_fprop(x.to(DEV).bfloat16(), w.to(DEV), b.to(DEV), p, 1, 0.0)
        assert CV.stats.get("stem_direct", 0) == before + 1
    assert y.dtype == torch.bfloat16 and tuple(y.shape) == tuple(ref.shape)
    assert _rel(y, ref) <= 1e-2, _rel(y, ref)


def test_stem_conv_many_tiles_strided_input_and_channel_slice_output():
    """More tiles than SMs x stages (persistent loop, both builder groups, barrier phases wrap), the input a cropped
    view and the output a channel slice of a wider NDHWC buffer; no bias, no activation."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    torch.manual_seed(32)
    big = _q(torch.randn(1, 1, 44, 70, 52))
    x = big[:, :, 2:42, 3:67, 1:49]
    w = _q(torch.randn(64, 1, 3, 3, 3) * 0.2)
    ref = TF.conv3d(x, w, None, padding=1)
    with rb.precision("bf16"), torch.no_grad():
        xg = big.to(DEV).bfloat16()[:, :, 2:42, 3:67, 1:49]
        full = torch.zeros(1, 192, 40, 64, 48, device=DEV, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last_3d)
        out = full[:, 64:128]
        y = CV.stem_conv_fprop(xg, w.to(DEV), None, (1, 1, 1), 0, 0.0, out=out)
        assert y.data_ptr() == out.data_ptr()
    assert _rel(y, ref) <= 1e-2
    assert float(full[:, :64].abs().max()) == 0.0 and float(full[:, 128:].abs().max()) == 0.0


def test_unet_first_conv_takes_the_direct_stem_kernel_and_trains():
    """Through the module API: forward uses the direct kernel, backward still produces the weight gradient."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV, nn as L
    torch.manual_seed(33)
    conv = L.Conv3d(1, 64, 3, padding=1).to(DEV)
    x = _q(torch.randn(1, 1, 12, 16, 16)).to(DEV)
    ref_conv = torch.nn.Conv3d(1, 64, 3, padding=1)
    ref_conv.load_state_dict({k: v.cpu() for k, v in conv.state_dict().items()})
    yr = ref_conv(x.cpu())
    yr.sum().backward()
    with rb.precision("bf16"):
        before = CV.stats.get("stem_direct", 0)
        y = conv(x)
        assert CV.stats.get("stem_direct", 0) == before + 1
        y.float().sum().backward()
    assert _rel(y, yr) <= 2e-2
    assert _rel(conv.weight.grad, ref_conv.weight.grad) <= 2e-2


PADDED = [
    # cin, cout, kernel, spatial, batch, padding_mode
    (1, 64, (3, 3, 3), (9, 17, 11), 2, "reflect"),        # the G-UNet stem with 'same' reflect padding (UNet.py:61-69)
    (1, 64, (3, 3, 3), (37, 6, 70), 1, "reflect"),        # several 4 x 32 tiles per row, overhang on both in-plane axes
    (1, 64, (3, 3, 3), (5, 9, 33), 1, "replicate"),
    (2, 64, (3, 3, 3), (6, 10, 9), 1, "reflect"),         # generic builder (two input channels: K = 54)
]


@pytest.mark.parametrize("case", PADDED)
def test_stem_conv_reads_reflect_and_replicate_padding_through_mirrored_indices(case):
    """Inference: nn.Conv3d(padding='same', padding_mode='reflect') on the 1 -> ngf stem runs the small-K kernel on the
    UNPADDED input (taps outside it gathered through mirrored / clamped indices) -- no padded copy -- and equals
    F.pad + conv3d of the same bf16-rounded operands; a float input and the training path keep their routes."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    from raygun_b200 import nn as L
    cin, cout, k, sp, n, mode = case
    torch.manual_seed(35)
    x = _q(torch.randn(n, cin, *sp))
    conv = L.Conv3d(cin, cout, k, padding="same", padding_mode=mode)
    w, b = _q(conv.weight.detach()), conv.bias.detach()
    ref = TF.relu(TF.conv3d(TF.pad(x, (1,) * 6, mode=mode), w, b))
    conv = conv.to(DEV)
    with rb.precision("bf16"):
        xg = x.to(DEV).bfloat16()
        if cin > 1:
            xg = xg.contiguous(memory_format=torch.channels_last_3d)
        before = dict(CV.stats)
        with torch.no_grad():
            from raygun_b200 import _C
            launches = _C.launch_count
            y = conv.fused(xg, 1, 0.0)
        assert CV.stats.get("stem_direct", 0) == before.get("stem_direct", 0) + 1
        assert _C.launch_count == launches + 1, "a padded copy was made"
        assert y.shape == (n, cout, *sp) and y.dtype == torch.bfloat16
        assert _rel(y, ref) <= 1e-2, _rel(y, ref)
        # training keeps the explicit pad (its adjoint folds the data gradient) and agrees with the inference route
        yt = conv.fused(xg, 1, 0.0)
        assert _rel(yt, y.float()) <= 1e-2


SMALLK = [
    # cin, cout, kernel, stride, pad, spatial, batch
    (1, 64, (3, 3, 3), (1, 1, 1), (1, 1, 1), (9, 17, 11), 2),       # U-Net stem, zero 'same' padding, ragged row tiles
    (1, 64, (3, 3, 3), (1, 1, 1), (0, 0, 0), (12, 10, 70), 1),      # valid conv on a pre-padded tensor, two tiles per row
    (1, 64, (3, 3, 3), (2, 2, 2), (1, 1, 1), (16, 16, 16), 3),      # NLayerDiscriminator's first conv (stride 2)
    (2, 128, (2, 3, 3), (1, 1, 1), (1, 1, 1), (5, 9, 12), 1),       # K = 36 (64-wide accumulators), two channel tiles
    (1, 64, (4, 4, 4), (2, 2, 2), (1, 1, 1), (12, 12, 12), 1),      # K = 64, the classic PatchGAN kw = 4
]


@pytest.mark.parametrize("case", SMALLK)
def test_smallk_wgrad_matches_autograd_and_accumulates_into_a_bucket(case):
    """rg_smallk_wgrad (the FMA-bound weight gradient of Cin * taps <= 64 layers: UNet.py:61-69 stem,
    NLayerDiscriminator.py:137-145 first conv) vs torch autograd on the same bf16-rounded operands; the bias gradient; the
    accumulate flag (gradient buckets); bit-reproducible across two calls."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    cin, cout, k, s, p, sp, n = case
    torch.manual_seed(37)
    x = _q(torch.randn(n, cin, *sp))
    w = torch.randn(cout, cin, *k, requires_grad=True)
    b = torch.zeros(cout, requires_grad=True)
    y = TF.conv3d(x, w, b, stride=s, padding=p)
    dy = _q(torch.randn_like(y))
    y.backward(dy)
    with rb.precision("bf16"):
        xg = x.to(DEV).bfloat16()
        if cin > 1:
            xg = xg.contiguous(memory_format=torch.channels_last_3d)
        dyg = dy.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        before = CV.stats.get("smallk_wgrad", 0)
        dw, db = CV.conv_wgrad(xg, dyg, tuple(w.shape), s, p, 1, want_bias=True)
        assert CV.stats.get("smallk_wgrad", 0) == before + 1, "small-K kernel was not taken"
        dw2, db2 = CV.conv_wgrad(xg, dyg, tuple(w.shape), s, p, 1, want_bias=True)
        bucket = torch.full(tuple(w.shape), 0.5, device=DEV)
        CV.conv_wgrad(xg, dyg, tuple(w.shape), s, p, 1, dw_out=bucket, accumulate=True)
    assert torch.equal(dw, dw2) and torch.equal(db, db2)
    assert _rel(dw, w.grad) <= 2e-5, _rel(dw, w.grad)
    assert _rel(db, b.grad) <= 2e-5
    assert float((bucket - 0.5 - dw).abs().max()) <= 1e-5 * max(1.0, float(dw.abs().max()))
