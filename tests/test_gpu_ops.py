"""GPU parity, operator level: every C-ABI kernel vs the CPU oracle (plain torch.nn.functional
fp32/fp64 on the same seeded inputs).  fp32 path tolerance: max-abs <= 1e-4 (north_star)."""
import pytest
import torch
import torch.nn.functional as TF

pytestmark = pytest.mark.gpu

TOL = 1e-4


def _dev():
    return torch.device("cuda:0")


def _close(a, b, tol=TOL, what=""):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    err = float((a - b).abs().max()) if a.numel() else 0.0
    assert err <= tol * max(1.0, float(b.abs().max()) if b.numel() else 1.0), (what, err)


CONV_CASES = [
    # cin, cout, k, stride, pad, groups, spatial
    (1, 8, 3, 1, 0, 1, (9, 10, 11)),
    (5, 7, 3, 1, 1, 1, (8, 8, 8)),
    (4, 6, 7, 1, 3, 1, (9, 9, 9)),
    (3, 5, 3, 2, 1, 1, (9, 10, 8)),
    (6, 4, 2, 2, 1, 1, (9, 9, 9)),
    (4, 8, 4, 2, 1, 1, (10, 10, 10)),
    (4, 12, 1, 1, 0, 4, (5, 6, 7)),
    (8, 1, 4, 1, 1, 1, (6, 6, 6)),
    (3, 70, (1, 3, 2), (1, 2, 1), (0, 1, 0), 1, (4, 9, 7)),
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_direct_conv3d_forward_backward_fp32(case):
    import raygun_b200 as rb
    from raygun_b200 import functional as F
    cin, cout, k, s, p, g, sp = case
    torch.manual_seed(1)
    x = torch.randn(2, cin, *sp, requires_grad=True)
    conv = torch.nn.Conv3d(cin, cout, k, stride=s, padding=p, groups=g)
    y_ref = conv(x)
    r = torch.randn_like(y_ref)
    gx, gw, gb = torch.autograd.grad((y_ref * r).sum(), [x, conv.weight, conv.bias])
    with rb.precision("fp32"):
        xg = x.detach().to(_dev()).requires_grad_(True)
        w = conv.weight.detach().to(_dev()).requires_grad_(True)
        b = conv.bias.detach().to(_dev()).requires_grad_(True)
        y = F.conv3d(xg, w, b, s, p, g)
        y.backward(r.to(_dev()))
    _close(y, y_ref, what="y")
    _close(xg.grad, gx, what="dx")
    _close(w.grad, gw, 2e-4, what="dw")
    _close(b.grad, gb, 2e-4, what="db")


@pytest.mark.parametrize("case", [(6, 4, 2, 2, 0, 0, (4, 5, 6)), (5, 3, 3, 2, 1, 1, (4, 4, 5)), (4, 4, 3, 1, 1, 0, (5, 5, 5)),
                                  (3, 6, 4, 2, 1, 0, (4, 4, 4))])
def test_direct_conv_transpose3d_forward_backward_fp32(case):
    import raygun_b200 as rb
    from raygun_b200 import functional as F
    cin, cout, k, s, p, op, sp = case
    torch.manual_seed(2)
    x = torch.randn(2, cin, *sp, requires_grad=True)
    ct = torch.nn.ConvTranspose3d(cin, cout, k, stride=s, padding=p, output_padding=op)
    y_ref = ct(x)
    r = torch.randn_like(y_ref)
    gx, gw, gb = torch.autograd.grad((y_ref * r).sum(), [x, ct.weight, ct.bias])
    with rb.precision("fp32"):
        xg = x.detach().to(_dev()).requires_grad_(True)
        w = ct.weight.detach().to(_dev()).requires_grad_(True)
        b = ct.bias.detach().to(_dev()).requires_grad_(True)
        y = F.conv_transpose3d(xg, w, b, s, p, op)
        y.backward(r.to(_dev()))
    _close(y, y_ref, what="y")
    _close(xg.grad, gx, what="dx")
    _close(w.grad, gw, 2e-4, what="dw")
    _close(b.grad, gb, 2e-4, what="db")


@pytest.mark.parametrize("mode", ["reflect", "replicate", "zeros"])
@pytest.mark.parametrize("lo,hi", [((1, 1, 1), (1, 1, 1)), ((3, 3, 3), (3, 3, 3)), ((0, 1, 2), (1, 0, 2))])
def test_pad3d_and_its_adjoint(mode, lo, hi):
    from raygun_b200 import functional as F
    torch.manual_seed(3)
    x = torch.randn(2, 3, 5, 6, 7, requires_grad=True)
    pads = (lo[2], hi[2], lo[1], hi[1], lo[0], hi[0])
    y_ref = TF.pad(x, pads, mode={"zeros": "constant"}.get(mode, mode))
    r = torch.randn_like(y_ref)
    (gx,) = torch.autograd.grad((y_ref * r).sum(), [x])
    xg = x.detach().to(_dev()).requires_grad_(True)
    y = F.pad3d(xg, lo, hi, mode)
    y.backward(r.to(_dev()))
    _close(y, y_ref, 0.0, "pad")
    _close(xg.grad, gx, 1e-6, "pad adjoint")


@pytest.mark.parametrize("kind", ["instance", "batch"])
@pytest.mark.parametrize("act", ["none", "relu", "leaky"])
@pytest.mark.parametrize("layout", ["ncdhw", "ndhwc"])
def test_norm_act_forward_backward(kind, act, layout):
    from raygun_b200 import functional as F
    from raygun_b200 import nn as L
    torch.manual_seed(4)
    x = (torch.randn(2, 6, 5, 6, 7) * 2 + 0.5).requires_grad_(True)
    res = torch.randn(2, 6, 5, 6, 7, requires_grad=True)
    ref_norm = torch.nn.InstanceNorm3d(6) if kind == "instance" else torch.nn.BatchNorm3d(6)
    if kind == "batch":
        with torch.no_grad():
            ref_norm.weight.uniform_(0.5, 1.5)
            ref_norm.bias.uniform_(-0.5, 0.5)
    f = {"none": lambda t: t, "relu": torch.relu, "leaky": lambda t: TF.leaky_relu(t, 0.2)}[act]
    y_ref = f(ref_norm(x) + res)
    r = torch.randn_like(y_ref)
    params = [x, res] + (list(ref_norm.parameters()) if kind == "batch" else [])
    grads = torch.autograd.grad((y_ref * r).sum(), params)
    ours = (L.InstanceNorm3d(6) if kind == "instance" else L.BatchNorm3d(6)).to(_dev())
    if kind == "batch":
        with torch.no_grad():
            ours.weight.copy_(ref_norm.weight)
            ours.bias.copy_(ref_norm.bias)
    code = {"none": (0, 0.0), "relu": (1, 0.0), "leaky": (2, 0.2)}[act]
    xg = x.detach().to(_dev())
    rg = res.detach().to(_dev())
    if layout == "ndhwc":
        xg = xg.contiguous(memory_format=torch.channels_last_3d)
    xg.requires_grad_(True)
    rg.requires_grad_(True)
    y = ours.fused(xg, None, code[0], code[1], residual=rg)
    y.backward(r.to(_dev()))
    _close(y, y_ref, what="y")
    _close(xg.grad, grads[0], 2e-4, "dx")
    _close(rg.grad, grads[1], what="dres")
    if kind == "batch":
        _close(ours.weight.grad, grads[2], 2e-4, "dgamma")
        _close(ours.bias.grad, grads[3], 2e-4, "dbeta")
        _close(ours.running_mean, ref_norm.running_mean, what="running_mean")
        _close(ours.running_var, ref_norm.running_var, what="running_var")
        ours.eval(), ref_norm.eval()
        _close(ours(x.detach().to(_dev())), ref_norm(x.detach()), what="eval")


def test_maxpool_upsample_activation_cropcat():
    from raygun_b200 import functional as F
    torch.manual_seed(5)
    x = torch.randn(2, 3, 7, 8, 6, requires_grad=True)
    for f in [(2, 2, 2), (1, 2, 3)]:
        y_ref = TF.max_pool3d(x, f, stride=f)
        r = torch.randn_like(y_ref)
        (gx,) = torch.autograd.grad((y_ref * r).sum(), [x])
        xg = x.detach().to(_dev()).requires_grad_(True)
        y = F.max_pool3d(xg, f)
        y.backward(r.to(_dev()))
        _close(y, y_ref, 0.0, "pool")
        _close(xg.grad, gx, 0.0, "pool bwd")
        y_ref = TF.interpolate(x, scale_factor=tuple(float(v) for v in f), mode="nearest")
        r = torch.randn_like(y_ref)
        (gx,) = torch.autograd.grad((y_ref * r).sum(), [x])
        xg = x.detach().to(_dev()).requires_grad_(True)
        y = F.upsample_nearest(xg, f)
        y.backward(r.to(_dev()))
        _close(y, y_ref, 0.0, "up")
        _close(xg.grad, gx, 1e-6, "up bwd")
    for code, fn in [(1, torch.relu), (2, lambda t: TF.leaky_relu(t, 0.2)), (3, torch.tanh), (4, torch.sigmoid)]:
        y_ref = fn(x)
        r = torch.randn_like(y_ref)
        (gx,) = torch.autograd.grad((y_ref * r).sum(), [x])
        xg = x.detach().to(_dev()).requires_grad_(True)
        y = F.activation(xg, code, 0.2)
        y.backward(r.to(_dev()))
        _close(y, y_ref, 1e-6, "act")
        _close(xg.grad, gx, 1e-5, "act bwd")
    a = torch.randn(2, 3, 9, 10, 8, requires_grad=True)
    b = torch.randn(2, 2, 7, 6, 6, requires_grad=True)
    from oracle.restate import center_crop
    y_ref = torch.cat([center_crop(a, (5, 6, 4)), center_crop(b, (5, 6, 4))], 1)
    r = torch.randn_like(y_ref)
    ga, gb = torch.autograd.grad((y_ref * r).sum(), [a, b])
    ag, bg = a.detach().to(_dev()).requires_grad_(True), b.detach().to(_dev()).requires_grad_(True)
    y = F.crop_cat(ag, bg, (5, 6, 4))
    y.backward(r.to(_dev()))
    _close(y, y_ref, 0.0, "cropcat")
    _close(ag.grad, ga, 0.0, "cropcat da")
    _close(bg.grad, gb, 0.0, "cropcat db")


def test_losses_match_oracle():
    from oracle import restate as O
    from raygun_b200 import functional as F
    torch.manual_seed(6)
    a = (torch.randn(2, 1, 9, 9, 9) * 1.5).requires_grad_(True)
    b = torch.randn(2, 1, 9, 9, 9, requires_grad=True)
    cases = [("mse1", lambda u, v: O.gan_loss(u, True), lambda u, v: F.mse_const_loss(u, 1.0)),
             ("mse0", lambda u, v: O.gan_loss(u, False), lambda u, v: F.mse_const_loss(u, 0.0)),
             ("smoothl1", O.smooth_l1, F.smooth_l1_loss), ("l1", O.l1, F.l1_loss)]
    for name, ref_fn, our_fn in cases:
        l_ref = ref_fn(a, b) * 3.0
        ga, gb = torch.autograd.grad(l_ref, [a, b], allow_unused=True)
        ag, bg = a.detach().to(_dev()).requires_grad_(True), b.detach().to(_dev()).requires_grad_(True)
        l = our_fn(ag, bg) * 3.0
        l.backward()
        assert abs(float(l) - float(l_ref)) <= 1e-5 * max(1.0, abs(float(l_ref))), name
        _close(ag.grad, ga, 1e-6, name + " da")
        if gb is not None:
            _close(bg.grad, gb, 1e-6, name + " db")


def test_predict_arithmetic_bit_exact():
    from oracle import restate as O
    from raygun_b200 import kernels as K
    u = torch.arange(256, dtype=torch.uint8).repeat(40)
    x = K.u8_normalise(u.to(_dev()), dtype=torch.float32)
    assert torch.equal(x.cpu(), O.u8_normalise(u))
    v = torch.linspace(-0.2, 1.2, 30000).view(1, 1, 30, 20, 50)
    out = K.denorm_clamp_u8(v.to(_dev())[:, :, 2:-2, 3:-3, 4:-4])
    assert torch.equal(out.cpu(), O.denorm_clamp_u8(v[:, :, 2:-2, 3:-3, 4:-4]))


def test_adam_matches_oracle_restatement():
    from oracle.restate import AdamState
    from raygun_b200 import kernels as K
    torch.manual_seed(7)
    p = torch.randn(10007)
    ref = {"p": p.clone()}
    st = AdamState(ref, lr=1e-3)
    pg = p.to(_dev())
    m, v = torch.zeros_like(pg), torch.zeros_like(pg)
    for step in range(1, 4):
        g = torch.randn(10007)
        st.step({"p": g})
        K.adam_step(pg, g.to(_dev()), m, v, 1e-3, 0.9, 0.999, 1e-8, step)
    _close(pg, ref["p"], 1e-6, "adam")


# ---- vectorised bf16 NDHWC fast paths (8 channels / thread) vs the oracle on bf16-rounded operands ----
def _bf(t):
    return t.bfloat16().float()


def _ndhwc(t):
    return t.to(_dev()).bfloat16().contiguous(memory_format=torch.channels_last_3d)


@pytest.mark.parametrize("mode", ["reflect", "replicate", "zeros"])
@pytest.mark.parametrize("lo,hi", [((1, 1, 1), (1, 1, 1)), ((3, 3, 3), (3, 3, 3)), ((0, 2, 1), (2, 0, 1))])
@pytest.mark.parametrize("c", [16, 64])          # 64 channels x 9 voxels per row: the row-based adjoint kernel
@pytest.mark.parametrize("fold", [False, True])   # True: rg_pad3d_fold, the in-place O(surface) adjoint
def test_vec_pad_and_adjoint_bf16_ndhwc(mode, lo, hi, c, fold):
    from raygun_b200 import functional as F
    torch.manual_seed(21)
    x = _bf(torch.randn(2, c, 5, 6, 9 if c == 64 else 7)).requires_grad_(True)
    pads = (lo[2], hi[2], lo[1], hi[1], lo[0], hi[0])
    y_ref = TF.pad(x, pads, mode={"zeros": "constant"}.get(mode, mode))
    r = _bf(torch.randn_like(y_ref))
    (gx,) = torch.autograd.grad((y_ref * r).sum(), [x])
    xg = _ndhwc(x.detach()).requires_grad_(True)
    y = F.pad3d(xg, lo, hi, mode, fold=fold)
    y.backward(_ndhwc(r).clone())
    assert y.dtype == torch.bfloat16
    _close(y, y_ref, 0.0, "vec pad")
    _close(xg.grad, gx, 1e-2, "vec pad adjoint")
    if fold:        # the folded interior view equals the copying adjoint bit for bit (same fp32 gather order)
        from raygun_b200 import kernels as K
        m = F.PAD_MODES[mode]
        a = K.pad3d_bwd(_ndhwc(r), tuple(x.shape), lo, m)
        b = K.pad3d_fold(_ndhwc(r).clone(), lo, hi, m)
        assert tuple(b.shape) == tuple(x.shape) and torch.equal(a, b)


@pytest.mark.parametrize("kind", ["instance", "batch"])
@pytest.mark.parametrize("c", [16, 192, 320])
def test_vec_norm_act_bf16_ndhwc(kind, c):
    from raygun_b200 import nn as L
    torch.manual_seed(22)
    x = _bf(torch.randn(2, c, 5, 6, 7) * 2 + 0.5).requires_grad_(True)
    res = _bf(torch.randn(2, c, 5, 6, 7)).requires_grad_(True)
    ref_norm = torch.nn.InstanceNorm3d(c) if kind == "instance" else torch.nn.BatchNorm3d(c)
    y_ref = TF.leaky_relu(ref_norm(x) + res, 0.2)
    r = _bf(torch.randn_like(y_ref))
    gx, gr = torch.autograd.grad((y_ref * r).sum(), [x, res])
    ours = (L.InstanceNorm3d(c) if kind == "instance" else L.BatchNorm3d(c)).to(_dev())
    xg, rg = _ndhwc(x.detach()).requires_grad_(True), _ndhwc(res.detach()).requires_grad_(True)
    y = ours.fused(xg, None, 2, 0.2, residual=rg)
    y.backward(_ndhwc(r))
    _close(y, y_ref, 1e-2, "vec norm y")
    _close(xg.grad, gx, 2e-2, "vec norm dx")
    _close(rg.grad, gr, 1e-2, "vec norm dres")


def test_vec_act_copy_cropcat_bf16_ndhwc():
    from oracle.restate import center_crop
    from raygun_b200 import functional as F
    torch.manual_seed(23)
    x = _bf(torch.randn(2, 24, 6, 5, 7)).requires_grad_(True)
    y_ref = torch.tanh(x)
    r = _bf(torch.randn_like(y_ref))
    (gx,) = torch.autograd.grad((y_ref * r).sum(), [x])
    xg = _ndhwc(x.detach()).requires_grad_(True)
    y = F.activation(xg, 3)
    y.backward(_ndhwc(r))
    _close(y, y_ref, 1e-2, "vec act")
    _close(xg.grad, gx, 1e-2, "vec act bwd")
    a = _bf(torch.randn(2, 16, 9, 10, 8)).requires_grad_(True)
    b = _bf(torch.randn(2, 8, 7, 6, 6)).requires_grad_(True)
    y_ref = torch.cat([center_crop(a, (5, 6, 4)), center_crop(b, (5, 6, 4))], 1)
    r = _bf(torch.randn_like(y_ref))
    ga, gb = torch.autograd.grad((y_ref * r).sum(), [a, b])
    ag, bg = _ndhwc(a.detach()).requires_grad_(True), _ndhwc(b.detach()).requires_grad_(True)
    y = F.crop_cat(ag, bg, (5, 6, 4))
    y.backward(_ndhwc(r))
    _close(y, y_ref, 0.0, "vec cropcat")
    _close(ag.grad, ga, 0.0, "vec cropcat da")
    _close(bg.grad, gb, 0.0, "vec cropcat db")


DEV = "cuda:0"


@pytest.mark.parametrize("mode", ["reflect", "replicate", "zeros"])
def test_border_fill_equals_explicit_padding(mode):
    """rg_pad_border_fill on the interior-written arena == F.pad of the interior (bit-exact copy of bf16 values)."""
    from raygun_b200 import functional as F
    from raygun_b200 import kernels as K
    torch.manual_seed(5)
    x = torch.randn(2, 16, 5, 7, 9).bfloat16()
    inner = K.empty5_arena(tuple(x.shape), torch.bfloat16, torch.device(DEV), margin=2)
    inner.copy_(x.to(DEV))
    full = inner._base
    full_before = full.clone()
    K.border_fill(full, 2, F.PAD_MODES[mode])
    tmode = {"zeros": "constant"}.get(mode, mode)
    ref = torch.nn.functional.pad(x.float(), (2,) * 6, mode=tmode).bfloat16()
    assert torch.equal(full.cpu(), ref)
    assert torch.equal(full[:, :, 2:-2, 2:-2, 2:-2].cpu(), full_before[:, :, 2:-2, 2:-2, 2:-2].cpu())
    assert K.arena_of(inner, (1, 1, 1), (2, 2, 2)) is not None and K.arena_of(inner, (3, 0, 0), (0, 0, 0)) is None
    assert K.arena_of(inner[:, :8], (1, 1, 1), (1, 1, 1)) is None          # a channel slice is not the arena interior


def test_vectorised_maxpool_matches_torch_on_strided_ndhwc_views():
    from raygun_b200 import kernels as K
    torch.manual_seed(6)
    x = torch.randn(2, 24, 9, 10, 12).bfloat16()
    xg = x.to(DEV).contiguous(memory_format=torch.channels_last_3d)
    for f in ((2, 2, 2), (1, 2, 3), (3, 2, 1)):
        ref = torch.nn.functional.max_pool3d(x.float(), f, f)
        assert torch.equal(K.maxpool_fwd(xg, f).float().cpu(), ref)
        sl = xg[:, 8:24]                                                     # channel slice: voxel stride 24, 16 channels
        assert torch.equal(K.maxpool_fwd(sl, f).float().cpu(), ref[:, 8:24])


def test_vectorised_maxpool_backward_matches_autograd_including_ties():
    """Row-based NDHWC bf16 backward: the first maximal element in (d, h, w) order gets the gradient (ties are common
    after ReLU), remainder voxels outside any window get zero."""
    from raygun_b200 import kernels as K
    torch.manual_seed(7)
    x = torch.relu(torch.randn(2, 24, 9, 10, 12)).bfloat16()             # many exact zeros -> ties
    xg = x.to(DEV).contiguous(memory_format=torch.channels_last_3d)
    for f in ((2, 2, 2), (1, 2, 3), (3, 2, 1)):
        xr = x.float().requires_grad_(True)
        y = torch.nn.functional.max_pool3d(xr, f, f)
        g = torch.randn_like(y).bfloat16().float()
        (ref,) = torch.autograd.grad(y, xr, g)
        gg = g.bfloat16().to(DEV).contiguous(memory_format=torch.channels_last_3d)
        dx = K.maxpool_bwd(xg, gg, f)
        assert torch.equal(dx.float().cpu(), ref), f


def test_training_conv_outputs_stay_dense_and_inference_uses_arenas():
    """autograd.Function.forward always runs with grad mode off: the caller's grad mode has to be handed down, or
    training silently gets the margin-allocated inference layout (strided outputs -> slow norm kernels)."""
    import raygun_b200 as rb
    from raygun_b200 import nn as L
    torch.manual_seed(5)
    conv = L.Conv3d(64, 64, 3, padding=1).to(DEV)
    x = torch.randn(1, 64, 8, 8, 8, device=DEV)
    with rb.precision("bf16"):
        y = conv(x.clone().requires_grad_(True))
        assert y.requires_grad and y.is_contiguous(memory_format=torch.channels_last_3d)
        with torch.no_grad():
            z = conv(x)
        assert getattr(z._base, "_rg_arena", None) == 1 and not z.is_contiguous(memory_format=torch.channels_last_3d)
    assert torch.equal(y.detach().float(), z.float())


def test_strided_channel_inner_views_are_conv_operands_in_place():
    """Pad-arena interiors / channel slices feed the TMA and pointwise kernels without a densifying copy."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV, kernels as K
    torch.manual_seed(6)
    full = torch.randn(1, 192, 12, 12, 12, device=DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
    x = full[:, 64:192, 1:11, 1:11, 1:11]                       # channel slice of an interior view
    assert K.channels_inner_bf16(x) and K.conv_operand(x) is x
    xd = x.contiguous(memory_format=torch.channels_last_3d)
    with rb.precision("bf16"), torch.no_grad():
        wt = torch.randn(128, 128, 2, 2, 2, device=DEV) * 0.05     # ConvTranspose3d k = s = 2 (UNet.py:265-269)
        a, _, _ = CV.conv_dgrad(x, wt, (1, 128, 20, 20, 20), (2, 2, 2), (0, 0, 0))
        b, _, _ = CV.conv_dgrad(xd, wt, (1, 128, 20, 20, 20), (2, 2, 2), (0, 0, 0))
        assert torch.equal(a.float(), b.float())
        w1 = torch.randn(1, 128, 1, 1, 1, device=DEV) * 0.1        # 1x1x1 head (BaseSystem.py:209-224)
        a, _, _ = CV.conv_fprop(x, w1, None, (1, 1, 1), (0, 0, 0), 1, 3, 0.0)
        b, _, _ = CV.conv_fprop(xd, w1, None, (1, 1, 1), (0, 0, 0), 1, 3, 0.0)
        assert torch.equal(a.float(), b.float())
        w3 = torch.randn(256, 128, 3, 3, 3, device=DEV) * 0.05     # igemm path (256 produced channels)
        a, _, _ = CV.conv_fprop(x, w3, None, (1, 1, 1), (1, 1, 1), 1, 0, 0.0)
        b, _, _ = CV.conv_fprop(xd, w3, None, (1, 1, 1), (1, 1, 1), 1, 0, 0.0)
        assert torch.equal(a.float(), b.float())


def test_discriminator_edge_layers_leave_the_cuda_core_gemm():
    """NLayerDiscriminator edge layers in bf16 mode: the 1 -> 64 k2 s2 p1 stem's data gradient (streaming small-Cin
    kernel) and the 512 -> 1 3^3 head (tensor cores with the produced channel zero-padded to 64) against autograd."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV, functional as F
    torch.manual_seed(8)
    q = lambda t: t.bfloat16().float()
    # stem: conv 1 -> 64, kernel 2, stride 2, padding 1 on 16^3 (-> 9^3); gradient w.r.t. the image
    x = q(torch.randn(2, 1, 16, 16, 16)).requires_grad_(True)
    w = q(torch.randn(64, 1, 2, 2, 2) * 0.3)
    y = TF.conv3d(x, w, None, stride=2, padding=1)
    dy = q(torch.randn_like(y))
    (ref,) = torch.autograd.grad(y, x, dy)
    with rb.precision("bf16"):
        before = CV.stats.get("smallcin_dgrad", 0)
        dyg = dy.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        dx, _, _ = CV.conv_dgrad(dyg, w.to(DEV), (2, 1, 16, 16, 16), (2, 2, 2), (1, 1, 1))
        assert CV.stats.get("smallcin_dgrad", 0) == before + 1
    assert float((dx.float().cpu() - ref).abs().max()) <= 1e-2 * float(ref.abs().max())
    # head: conv 512 -> 1, 3^3, padding 1 on 9^3, forward + all three gradients through the functional API
    conv = torch.nn.Conv3d(512, 1, 3, padding=1)
    with torch.no_grad():
        conv.weight.copy_(q(conv.weight))
    xh = q(torch.randn(2, 512, 9, 9, 9)).requires_grad_(True)
    yh = conv(xh)
    g = q(torch.randn_like(yh))
    rx, rw, rb_ = torch.autograd.grad(yh, [xh, conv.weight, conv.bias], g)
    with rb.precision("bf16"):
        d0 = CV.stats["direct"]
        xg = xh.detach().to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d).requires_grad_(True)   # compute storage
        wg = conv.weight.detach().to(DEV).requires_grad_(True)
        bg = conv.bias.detach().to(DEV).requires_grad_(True)
        yg = F.conv3d(xg, wg, bg, 1, 1)
        yg.backward(g.to(DEV).to(yg.dtype))
        assert CV.stats["direct"] == d0, "the CUDA-core GEMM ran"
    rel = lambda a, b: float((a.float().cpu() - b).abs().max()) / max(1e-6, float(b.abs().max()))
    assert tuple(yg.shape) == tuple(yh.shape) and rel(yg, yh) <= 1e-2
    assert rel(xg.grad, rx) <= 1e-2 and rel(wg.grad, rw) <= 1e-2 and rel(bg.grad, rb_) <= 1e-2


@pytest.mark.parametrize("act", [1, 2, 3])
def test_fused_activation_backward_and_bias_gradient(act):
    """rg_act_bwd_bias: dx identical to rg_act_bwd, bias sum equal to the fp32 sum of the bf16 dx; and a
    Conv3d + ReLU layer's bias gradient through the functional API matches autograd."""
    from raygun_b200 import kernels as K
    torch.manual_seed(9)
    y = torch.randn(2, 64, 9, 10, 11, device=DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
    if act == 3:
        y = torch.tanh(y.float()).bfloat16().contiguous(memory_format=torch.channels_last_3d)
    dy = torch.randn(2, 64, 9, 10, 11, device=DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
    ref = K.act_bwd(y, dy, act, 0.2)
    dx, db = K.act_bwd_bias(y, dy, act, 0.2)
    assert db is not None and torch.equal(dx, ref)
    want = ref.float().sum(dim=(0, 2, 3, 4))
    assert float((db - want).abs().max()) <= 2e-3 * float(want.abs().max())
    # row-strided operands (interiors of padded buffers: pad-arena outputs, gradients folded in place) take the row kernel
    yp = torch.nn.functional.pad(y, (1, 1, 1, 1, 1, 1)).contiguous(memory_format=torch.channels_last_3d)
    dp = torch.nn.functional.pad(dy, (2, 0, 1, 2, 0, 1)).contiguous(memory_format=torch.channels_last_3d)
    yi, di = yp[:, :, 1:-1, 1:-1, 1:-1], dp[:, :, 0:-1, 1:-2, 2:]
    assert not yi.is_contiguous(memory_format=torch.channels_last_3d) and torch.equal(yi, y) and torch.equal(di, dy)
    dx2, db2 = K.act_bwd_bias(yi, di, act, 0.2)
    assert db2 is not None and torch.equal(dx2, ref)
    assert float((db2 - want).abs().max()) <= 2e-3 * float(want.abs().max())
    # a channel slice is row-strided too
    dx3, db3 = K.act_bwd_bias(y[:, :32], dy[:, :32], act, 0.2)
    assert db3 is not None and torch.equal(dx3, ref[:, :32])
    # fp32 operands fall back to the two-kernel path
    dx4, db4 = K.act_bwd_bias(y.float(), dy.float(), act, 0.2)
    assert db4 is None


def test_conv_relu_bias_gradient_through_the_fused_pass():
    import raygun_b200 as rb
    from raygun_b200 import functional as F
    torch.manual_seed(10)
    q = lambda t: t.bfloat16().float()
    conv = torch.nn.Conv3d(64, 64, 3, padding=1)
    with torch.no_grad():
        conv.weight.copy_(q(conv.weight))
    x = q(torch.randn(2, 64, 8, 8, 8))
    yr = TF.relu(conv(x))
    g = q(torch.randn_like(yr))
    rw, rb_ = torch.autograd.grad(yr, [conv.weight, conv.bias], g)
    with rb.precision("bf16"):
        xg = x.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        wg = conv.weight.detach().to(DEV).requires_grad_(True)
        bg = conv.bias.detach().to(DEV).requires_grad_(True)
        yg = F.conv3d(xg, wg, bg, 1, 1, act=1)
        yg.backward(g.to(DEV).to(yg.dtype))
    rel = lambda a, b: float((a.float().cpu() - b).abs().max()) / max(1e-6, float(b.abs().max()))
    assert rel(bg.grad, rb_) <= 1e-2 and rel(wg.grad, rw) <= 1e-2


def test_training_pad_arenas_and_in_place_fold_match_the_copying_path():
    """Two reflect-'same' Conv3d + ReLU layers (a G-UNet ConvPass, UNet.py:61-69) in training: with pad arenas and the
    in-place pad adjoint the output is bit-identical to the copying path and every gradient agrees; both match torch."""
    import raygun_b200 as rb
    from raygun_b200 import functional as F, nn as L
    torch.manual_seed(31)
    q = lambda t: t.bfloat16().float()
    c1 = L.Conv3d(64, 64, 3, padding="same", padding_mode="reflect").to(DEV)
    c2 = L.Conv3d(64, 128, 3, padding="same", padding_mode="reflect").to(DEV)
    with torch.no_grad():
        for c in (c1, c2):
            c.weight.copy_(q(c.weight))
    layers = [c1, L.ReLU(), c2, L.ReLU()]
    x0 = q(torch.randn(2, 64, 10, 12, 16, device=DEV))
    g0 = q(torch.randn(2, 128, 10, 12, 16, device=DEV))
    params = [c1.weight, c1.bias, c2.weight, c2.bias]

    def run(arena, fold):
        F.TRAIN_ARENA, F.PAD_FOLD = arena, fold
        try:
            with rb.precision("bf16"):
                x = x0.bfloat16().contiguous(memory_format=torch.channels_last_3d).requires_grad_(True)
                # a max-pool in front: its output is neither an arena interior nor a norm output -> copying pad + fold
                y = L.run_layers(layers, F.max_pool3d(F.upsample_nearest(x, (2, 2, 2)), (2, 2, 2)))
                grads = torch.autograd.grad(y, [x] + params, g0.to(y.dtype))
            return y.detach().float(), [g.float() for g in grads], y
        finally:
            F.TRAIN_ARENA, F.PAD_FOLD = True, True

    y_a, g_a, yt = run(True, True)
    assert getattr(yt._base, "_rg_arena", None) == 1          # conv + ReLU outputs are arena interiors in training
    y_b, g_b, yt_b = run(False, False)
    assert yt_b.is_contiguous(memory_format=torch.channels_last_3d)
    assert torch.equal(y_a, y_b)
    for a, b in zip(g_a, g_b):
        assert float((a - b).abs().max()) <= 2e-3 * max(1e-6, float(b.abs().max()))
    # torch reference (fp32 math on the bf16-rounded parameters)
    xr = x0.clone().requires_grad_(True)
    h = TF.relu(TF.conv3d(TF.pad(xr, (1,) * 6, mode="reflect"), c1.weight, c1.bias))
    yr = TF.relu(TF.conv3d(TF.pad(h, (1,) * 6, mode="reflect"), c2.weight, c2.bias))
    gr = torch.autograd.grad(yr, [xr] + params, g0)
    rel = lambda a, b: float((a - b.float()).abs().max()) / max(1e-6, float(b.abs().max()))
    assert rel(y_a, yr.detach()) <= 2e-2
    for a, b in zip(g_a, gr):
        assert rel(a, b) <= 1e-1          # bf16 activations through two layers (the copying path shows the same)


@pytest.mark.parametrize("shape,scale", [((2, 1, 16, 16, 16), (0.5, 0.5, 0.5)), ((1, 1, 20, 13, 17), (0.5, 0.75, 0.3)),
                                         ((1, 3, 9, 8, 7), (1.0, 0.5, 0.25)), ((1, 1, 5, 6, 4), (0.2, 0.2, 0.3))])
def test_sampling_bottleneck_trilinear_kernels_vs_interpolate(shape, scale):
    """CycleModel.sampling_bottleneck (CycleModel.py:24-30): down by scale_factor and back up, trilinear with
    align_corners=True, forward and adjoint (deterministic gather) vs F.interpolate / its autograd; and the CycleModel
    mirror routes 3-D CUDA tensors through the kernels."""
    from raygun_b200 import functional as F
    from raygun_b200.cyclegan import CycleModel
    dev = _dev()
    torch.manual_seed(21)
    x = torch.randn(*shape, device=dev, requires_grad=True)
    xr = x.detach().clone().requires_grad_(True)
    y = F.sampling_bottleneck(x, scale)
    down = TF.interpolate(xr, scale_factor=scale, mode="trilinear", align_corners=True)
    yr = TF.interpolate(down, size=shape[2:], mode="trilinear", align_corners=True)
    _close(y, yr, 1e-5, "bottleneck fwd")
    g = torch.randn_like(yr)
    y.backward(g)
    yr.backward(g)
    _close(x.grad, xr.grad, 1e-5, "bottleneck adjoint")
    y2 = F.sampling_bottleneck(x.detach(), scale)
    assert torch.equal(y2, y.detach())                       # bit-reproducible
    m = CycleModel(torch.nn.Identity(), torch.nn.Identity(), scale_factor_A=scale, scale_factor_B=scale)
    fake_B, cycled_B, fake_A, cycled_A = m(x.detach(), x.detach())
    assert torch.equal(fake_B, y.detach()) and torch.equal(cycled_A, fake_B)


@pytest.mark.parametrize("cin,cout", [(8, 3), (64, 1), (64, 64)])
def test_fp32_mode_reaches_the_backward_thread(cin, cout):
    """autograd runs backward on its own worker thread: the precision mode must be visible there.  Regression: with a
    thread-local mode, `with precision("fp32")` still dispatched the weight gradient of pointwise-eligible (<= 4 output channels)
    and tensor-core-eligible (64-multiple channels) layers to the bf16 kernels -- an error of ~2e-3 relative."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    from raygun_b200.networks.UNet import ConvPass
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = _dev()
    torch.manual_seed(4)
    k = [[1, 1, 1]] if cout < 64 else [[3, 3, 3]]
    head = ConvPass(cin, cout, k, activation="Sigmoid").to(dev)
    ref = torch.nn.Sequential(torch.nn.Conv3d(cin, cout, k[0]), torch.nn.Sigmoid()).to(dev)
    ref[0].load_state_dict(head.conv_pass[0].state_dict())
    x = torch.randn(1, cin, 10, 10, 10, device=dev)
    xo, xr = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    before = dict(CV.stats)
    with rb.precision("fp32"):
        yo = head(xo)
        g = torch.randn_like(yo)
        (yo * g).sum().backward()
    (ref(xr) * g).sum().backward()
    assert CV.stats.get("igemm", 0) == before.get("igemm", 0) and CV.stats.get("pointwise", 0) == before.get("pointwise", 0)
    _close(head.conv_pass[0].weight.grad, ref[0].weight.grad, TOL, "dw")
    _close(head.conv_pass[0].bias.grad, ref[0].bias.grad, TOL, "db")
    _close(xo.grad, xr.grad, TOL, "dx")
