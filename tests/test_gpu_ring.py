"""GPU parity of the slice-ring convolution kernel (ring_conv_sm100.cu: halo-resident slices, row-shifted UMMA
descriptors, TMEM accumulator ring) against fp32 convolution of the SAME bf16-rounded operands on the CPU.
Covers: 7^3 / 3^3 / anisotropic kernels, several reduction chunks and produced tiles, ragged spatial extents
(tiles and depth segments that do not divide), zero padding through TMA out-of-bounds fill, the fused bias /
activation / statistics epilogue, stride-1 data gradients, and a multi-wave problem (more units than SMs)."""
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


def _t3(v):
    return (v,) * 3 if isinstance(v, int) else tuple(v)


FPROP = [
    # cin, cout, kernel, pad, spatial (d,h,w), batch
    (64, 64, 7, 0, (14, 22, 14), 1),          # the ResNet head shape class (valid 7^3 on a padded tensor)
    (64, 64, 7, 3, (9, 9, 9), 1),
    (64, 64, 3, 1, (8, 8, 8), 1),
    (64, 64, 3, 1, (19, 17, 11), 2),          # ragged tiles, several groups, depth not a multiple of S
    (192, 64, 3, 1, (7, 16, 8), 1),           # three reduction chunks (U-Net decoder 192 -> 64)
    (128, 64, 3, 0, (10, 9, 12), 2),
    (64, 64, (7, 1, 1), (3, 0, 0), (20, 16, 8), 1),   # the folded stem: taps along depth only
    (64, 64, (1, 3, 3), (0, 1, 1), (5, 16, 16), 1),
    (64, 64, 5, 2, (11, 10, 9), 1),           # KD = 5: lag-2 configuration with S = 2
    (64, 64, 1, 0, (9, 8, 8), 2),
    (64, 64, (2, 4, 3), (1, 2, 1), (9, 10, 11), 1),
]


@pytest.mark.parametrize("case", FPROP)
def test_ring_fprop_with_bias_activation_and_stats(case):
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    cin, cout, k, p, sp, n = case
    torch.manual_seed(21)
    x = _q(torch.randn(n, cin, *sp))
    conv = torch.nn.Conv3d(cin, cout, k, padding=p)
    w, b = _q(conv.weight.detach()), conv.bias.detach()
    ref = TF.leaky_relu(TF.conv3d(x, w, b, padding=p), 0.2)
    before = CV.stats.get("ring", 0)
    with rb.precision("bf16"):
        xg = x.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        y, ssum, ssq = CV.conv_fprop(xg, w.to(DEV), b.to(DEV), (1, 1, 1), _t3(p), 1, 2, 0.2, want_stats=True)
    assert CV.stats.get("ring", 0) == before + 1, "slice-ring kernel was not taken"
    assert y.dtype == torch.bfloat16
    assert _rel(y, ref) <= 1e-2, _rel(y, ref)
    yq = y.float().cpu()
    assert _rel(ssum.view(n, cout), yq.sum(dim=(2, 3, 4))) <= 2e-3
    assert _rel(ssq.view(n, cout), (yq * yq).sum(dim=(2, 3, 4))) <= 2e-3


def test_ring_fprop_many_units_and_channel_slice_output(monkeypatch):
    """More units than SMs (persistent loop, several units per CTA, stats flushed per (n, tile)); two produced
    tiles when the 64-channel limit is lifted; output written into a channel slice of a wider NDHWC buffer."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    monkeypatch.setattr(CV, "RING_MAX_PRODUCED", 128)
    torch.manual_seed(22)
    n, cin, cout, sp = 2, 64, 128, (24, 40, 72)
    x = _q(torch.randn(n, cin, *sp))
    conv = torch.nn.Conv3d(cin, cout, 3, padding=1)
    w = _q(conv.weight.detach())
    ref = TF.conv3d(x, w, None, padding=1)
    before = CV.stats.get("ring", 0)
    with rb.precision("bf16"):
        xg = x.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        y, ssum, ssq = CV.conv_fprop(xg, w.to(DEV), None, (1, 1, 1), (1, 1, 1), 1, 0, 0.0, want_stats=True)
    assert CV.stats.get("ring", 0) == before + 1
    assert _rel(y, ref) <= 1e-2
    yq = y.float().cpu()
    assert _rel(ssum.view(n, cout), yq.sum(dim=(2, 3, 4))) <= 2e-3
    assert _rel(ssq.view(n, cout), (yq * yq).sum(dim=(2, 3, 4))) <= 2e-3


DGRAD = [
    # cin, cout, kernel, pad, input spatial
    (64, 64, 7, 0, (14, 22, 14)),
    (64, 64, 3, 1, (9, 10, 11)),
    (64, 128, 3, 1, (8, 16, 8)),      # dgrad reduces over cout = 128 (two chunks), produces cin = 64
    (64, 64, (7, 1, 1), (3, 0, 0), (12, 8, 8)),
    (64, 64, 5, 1, (9, 9, 9)),
]


@pytest.mark.parametrize("case", DGRAD)
def test_ring_dgrad_matches_autograd(case):
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    cin, cout, k, p, sp = case
    torch.manual_seed(23)
    x = torch.randn(2, cin, *sp, requires_grad=True)
    conv = torch.nn.Conv3d(cin, cout, k, padding=p, bias=False)
    w = _q(conv.weight.detach())
    y = TF.conv3d(x, w, None, padding=p)
    dy = _q(torch.randn_like(y))
    (ref,) = torch.autograd.grad(y, x, dy)
    before = CV.stats.get("ring", 0)
    with rb.precision("bf16"):
        dyg = dy.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        dx, _, _ = CV.conv_dgrad(dyg, w.to(DEV), tuple(x.shape), (1, 1, 1), _t3(p))
    assert CV.stats.get("ring", 0) == before + 1, "slice-ring kernel was not taken"
    assert _rel(dx, ref) <= 1e-2, _rel(dx, ref)


WGRAD = [
    # cin, cout, kernel, pad, input spatial, batch
    (64, 64, 7, 0, (12, 22, 14), 1),          # head shape class: 25 tap pairs in 4 groups, unpaired last tap
    (64, 64, 3, 1, (9, 10, 11), 2),           # zero padding: X planes outside the volume are skipped, ragged tiles
    (128, 64, 3, 1, (6, 16, 8), 1),           # two ci chunks
    (64, 128, 3, 0, (7, 18, 10), 1),          # BN = 128
    (64, 256, (1, 3, 3), (0, 1, 1), (3, 16, 16), 1),   # BN = 256: two accumulators per CTA
    (64, 64, (7, 1, 1), (3, 0, 0), (12, 16, 8), 1),
    (64, 64, 1, 0, (5, 8, 8), 2),
    (64, 64, (2, 4, 3), (1, 2, 1), (9, 10, 11), 1),
    (64, 64, (4, 3, 3), (2, 1, 1), (9, 10, 11), 2),   # even KD >= 4: every depth tap handled by the depth-paired N = 128 variant
    (128, 64, (5, 1, 3), (2, 0, 1), (7, 9, 16), 1),   # odd KD: paired taps + the last tap alone, two ci chunks
]


@pytest.mark.parametrize("case", WGRAD)
def test_halo_wgrad_matches_autograd(case):
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    cin, cout, k, p, sp, n = case
    torch.manual_seed(24)
    x = _q(torch.randn(n, cin, *sp))
    w = torch.nn.Conv3d(cin, cout, k, padding=p, bias=False).weight.detach().requires_grad_(True)
    y = TF.conv3d(x, w, None, padding=p)
    dy = _q(torch.randn_like(y))
    (ref,) = torch.autograd.grad(y, w, dy)
    before = CV.stats.get("halo_wgrad", 0)
    with rb.precision("bf16"):
        xg = x.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        dyg = dy.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        dw, db = CV.conv_wgrad(xg, dyg, tuple(w.shape), (1, 1, 1), _t3(p), want_bias=True)
    assert CV.stats.get("halo_wgrad", 0) == before + 1, "halo wgrad kernel was not taken"
    assert dw.dtype == torch.float32 and tuple(dw.shape) == tuple(w.shape)
    assert _rel(dw, ref) <= 2e-3, _rel(dw, ref)           # fp32 accumulation of exact bf16 products
    assert _rel(db, dy.sum(dim=(0, 2, 3, 4))) <= 2e-3


def test_halo_wgrad_many_splits_is_linear_in_dy():
    """A multi-wave problem (more CTAs than SMs, several voxel splits) plus a size-independent property:
    dW(x, a*dy1 + dy2) == a*dW(x, dy1) + dW(x, dy2) up to bf16 rounding of the combined dy."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    torch.manual_seed(25)
    n, c, sp = 2, 64, (20, 48, 40)
    x = torch.randn(n, c, *sp, device=DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
    dy1 = torch.randn(n, c, *sp, device=DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
    dy2 = torch.randn(n, c, *sp, device=DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
    comb = (2.0 * dy1.float() + dy2.float()).bfloat16().contiguous(memory_format=torch.channels_last_3d)
    with rb.precision("bf16"):
        f = lambda g: CV.conv_wgrad(x, g, (c, c, 3, 3, 3), (1, 1, 1), (1, 1, 1))[0]
        a, b, ab = f(dy1), f(dy2), f(comb)
    ref = TF.conv3d(x.float().transpose(0, 1), comb.float().transpose(0, 1), padding=1).transpose(0, 1)   # [co][ci][k..]
    assert _rel(ab, ref) <= 5e-3
    assert _rel(ab, 2.0 * a + b) <= 1e-2


HEAD = [
    # cin, kernel, pad, spatial, batch, head channels, head activation
    (64, 3, 1, (19, 17, 11), 2, 1, "tanh"),      # the G-UNet tail: 3^3 64 -> 64 ReLU, then 64 -> 1 + Tanh
    (128, 3, 1, (9, 20, 13), 1, 3, "none"),      # two reduction chunks (the concat buffer), 3 head channels
    (64, 7, 3, (9, 9, 9), 1, 4, "sigmoid"),
]


@pytest.mark.parametrize("case", HEAD)
def test_ring_fused_head_equals_the_two_layer_result(case):
    """rg_ring_run_head (BaseSystem.py:209-224 folded into the last ConvPass layer, UNet.py:52-87): the head's output
    equals 1x1x1 conv + activation applied to the bf16 tensor the unfused layer stores -- compared with an fp32 CPU
    evaluation of the same bf16-rounded operands and with the library's own two-kernel path."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    cin, k, p, sp, n, hn, hact = case
    torch.manual_seed(33)
    x = _q(torch.randn(n, cin, *sp))
    conv = torch.nn.Conv3d(cin, 64, k, padding=p)
    head = torch.nn.Conv3d(64, hn, 1)
    w, b = _q(conv.weight.detach()), conv.bias.detach()
    hw, hb = head.weight.detach(), head.bias.detach()
    mid = _q(TF.relu(TF.conv3d(x, w, b, padding=p)))
    ref = TF.conv3d(mid, hw, hb)
    code = {"none": 0, "tanh": 3, "sigmoid": 4}[hact]
    ref = {"none": lambda t: t, "tanh": torch.tanh, "sigmoid": torch.sigmoid}[hact](ref)
    with rb.precision("bf16"), torch.no_grad():
        xg = x.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        wg, bg, hwg, hbg = w.to(DEV), b.to(DEV), hw.to(DEV), hb.to(DEV)
        before = CV.stats.get("ring_head", 0)
        with CV.offer_head(CV.HeadRequest(wg, hwg, hbg, code, 0.0)) as req:
            y, _, _ = CV.conv_fprop(xg, wg, bg, (1, 1, 1), _t3(p), 1, 1, 0.0)
        assert req.done and CV.stats.get("ring_head", 0) == before + 1, "fused head was not taken"
        assert y.shape == (n, hn, *sp) and y.dtype == torch.bfloat16
        y2, _, _ = CV.conv_fprop(xg, wg, bg, (1, 1, 1), _t3(p), 1, 1, 0.0)
        z2, _, _ = CV.conv_fprop(y2, hwg, hbg, (1, 1, 1), (0, 0, 0), 1, code, 0.0)
    scale = max(1.0, float(ref.abs().max()))
    assert float((y.float().cpu() - ref).abs().max()) <= 1e-2 * scale
    # against the two-kernel path: same operands, same bf16 intermediate; only the fp32 summation order of 64 products and
    # the final bf16 rounding can differ (one bf16 ulp = 2^-8 relative)
    assert float((y.float() - z2.float()).abs().max()) <= 2 ** -7 * scale
    assert float((y.float() - z2.float()).abs().mean()) <= 2e-4 * scale


def test_generator_inference_folds_the_head_into_the_last_convolution():
    """cyclegan.Generator under no_grad: Sequential(UNet, Conv3d(64 -> 1, 1), Tanh) (BaseSystem.py:209-224) runs the head
    inside the last ring launch; the result matches the unfused execution to one bf16 ulp, and training (grad enabled)
    never takes the fused path."""
    import contextlib
    import sys
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    from raygun_b200.cyclegan import make_generator
    torch.manual_seed(8)
    with contextlib.redirect_stdout(sys.stderr):
        g = make_generator("unet", dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3,
                                        padding_type="same", final_activation="Tanh")).to(DEV).eval()
    x = torch.rand(1, 1, 40, 48, 56).mul(2).sub(1).to(DEV)
    with rb.precision("bf16"):
        with torch.no_grad():
            before = CV.stats.get("ring_head", 0), CV.stats.get("pointwise", 0)
            y = g(x)
            assert CV.stats.get("ring_head", 0) == before[0] + 1 and CV.stats.get("pointwise", 0) == before[1]
            old, CV.FUSE_HEAD = CV.FUSE_HEAD, False
            try:
                z = g(x)
            finally:
                CV.FUSE_HEAD = old
            assert CV.stats.get("ring_head", 0) == before[0] + 1 and CV.stats.get("pointwise", 0) == before[1] + 1
        before = CV.stats.get("ring_head", 0)
        g(x).sum().backward()
        assert CV.stats.get("ring_head", 0) == before
    assert y.shape == z.shape == x.shape and y.dtype == torch.float32
    assert float((y - z).abs().max()) <= 2 ** -7 and float((y - z).abs().mean()) <= 2e-4
    assert float(y.std()) > 1e-3
