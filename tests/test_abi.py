"""CPU-side checks: the C-ABI library builds, loads and exports every symbol of include/raygun_b200.h;
the drop-in classes reproduce the reference's module trees / state_dict keys; host logic that needs
no GPU.  No compute call is made here."""
import json
import os
import re

import pytest
import torch

import raygun_b200 as rb
from raygun_b200 import _C, _build
from raygun_b200 import networks as N
from tests._golden import GOLDEN, build_ours, case_names, load_case

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_builds_and_exports_every_declared_symbol():
    _build.build()
    lib = _C.lib()
    with open(os.path.join(ROOT, "include", "raygun_b200.h")) as f:
        header = f.read()
    declared = set(re.findall(r"^(?:int|size_t|long long|const char\*)\s+(rg_[a-z0-9_]+)\s*\(", header, flags=re.M))
    assert declared == set(_C.SYMBOLS), declared ^ set(_C.SYMBOLS)
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.rg_version() == 201


def test_sass_contains_blackwell_tensor_core_and_tma_instructions():
    import shutil
    import subprocess
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    out = subprocess.run(["cuobjdump", "-sass", _C.LIB_PATH], capture_output=True, text=True).stdout
    for mnemonic in ("UTCHMMA", "UTMALDG", "UTMASTG", "LDTM"):
        assert mnemonic in out, mnemonic
    assert "HMMA." not in out.replace("UTCHMMA", "")   # no legacy mma.sync path


@pytest.mark.parametrize("name", case_names())
def test_state_dict_keys_and_shapes_match_reference_fixtures(name):
    c = load_case(name)
    net = build_ours(c["kind"], c["kw"])
    ours = {k: tuple(v.shape) for k, v in net.state_dict().items()}
    ref = {k: tuple(v.shape) for k, v in c["sd"].items()}
    assert ours == ref
    net.load_state_dict(c["sd"])     # strict


def test_structural_constants_of_the_reference_tests():
    with open(os.path.join(GOLDEN, "structural.json")) as f:
        s = json.load(f)
    d = N.NLayerDiscriminator(3)
    assert sum(p.numel() for p in d.parameters()) == s["nld3d_default_params"] == 11048769
    assert len(d.model) == s["nld3d_default_len"] == 12
    assert len(N.ResNet(3).model) == s["resnet3d_default_len"]
    assert len(N.ResNet(3, n_blocks=4, n_downsampling=3).model) == 28      # resnet_test.py:17-22
    assert len(N.NLayerDiscriminator(2, n_layers=5).model) == 18            # nlayerdiscriminator_test.py:80
    assert 30 <= N.NLayerDiscriminator(2).FOV <= 110                        # :24-37
    with pytest.raises(ValueError):
        N.ResNet(4)
    with pytest.raises(ValueError):
        N.NLayerDiscriminator(4)


def test_unet_attributes():
    u = N.UNet(input_nc=3, ngf=16, fmap_inc_factor=2, downsample_factors=[(2, 2, 2)] * 2, padding_type="same")
    assert (u.ndims, u.num_levels, u.num_heads, u.input_nc, u.output_nc, u.dims) == (3, 3, 1, 3, 16, 3)
    assert len(u.l_conv) == 3 and len(u.l_down) == 2 and len(u.r_up[0]) == 2 and len(u.r_conv[0]) == 2
    md = N.MaxDownsample((2, 2, 2), flexible=False)
    with pytest.raises(RuntimeError):
        md(torch.zeros(1, 1, 5, 4, 4))


def test_two_d_networks_are_delegated_to_stock_torch():
    y = N.ResNet(2, ngf=4, n_blocks=1)(torch.randn(1, 1, 32, 32))
    assert y.shape == (1, 4, 32, 32)
    d = N.NLayerDiscriminator(2, ngf=4)(torch.randn(1, 1, 64, 64))
    assert d.shape[:2] == (1, 1)


def test_no_cpu_fallback_for_the_3d_path():
    u = N.UNet(input_nc=1, ngf=4, fmap_inc_factor=2, downsample_factors=[(2, 2, 2)], padding_type="same")
    with pytest.raises(_C.RaygunB200Error):
        u(torch.zeros(1, 1, 8, 8, 8))


def test_norm_helpers_work_on_the_kernel_backed_modules():
    net = N.ResNet(3, ngf=4, n_blocks=1, norm_layer=torch.nn.BatchNorm3d)
    assert len(N.get_norm_layers(net)) == 7
    N.set_norm_mode(net, "fix_norms")
    assert net.training and not any(m.training for m in N.get_norm_layers(net))
    m, v = N.get_running_norm_stats(net)
    assert m is not None and m.numel() == v.numel()
    N.init_weights(net, "kaiming")


def test_precision_switch():
    assert rb.get_precision() in ("bf16", "fp32")
    with rb.precision("fp32"):
        assert rb.get_precision() == "fp32"
    with pytest.raises(ValueError):
        rb.set_precision("fp16")


def test_ring_config_reported_by_the_library():
    """The host-side configuration logic (slices per weight stage, drain lag, smem budget) through the C ABI."""
    import ctypes as C
    from raygun_b200 import _C
    d = _C.IgemmDesc()
    d.mode, d.n, d.cin, d.cout = 0, 2, 64, 64
    d.in_dhw[:] = [134, 134, 134]
    d.out_dhw[:] = [128, 128, 128]
    d.kernel[:] = [7, 7, 7]
    d.stride[:] = [1, 1, 1]
    d.pad[:] = [0, 0, 0]
    d.x_strides[:] = [134 ** 3 * 64, 134 * 134 * 64, 134 * 64, 64]
    d.y_strides[:] = [128 ** 3 * 64, 128 * 128 * 64, 128 * 64, 64]
    out = (C.c_int * 8)()
    _C.check(_C.lib().rg_ring_config(C.byref(d), out), "rg_ring_config")
    S, lag, na, nb, seg_len, segs, smem, pitch = list(out)
    assert S == 2 and lag == 1 and na >= 3 and nb == 3
    assert S + 7 - 1 <= 8 and smem <= 227 * 1024 and pitch % 1024 == 0
    assert segs == 0 or seg_len * segs >= 128      # segs == 0: balanced flattened partition (the default)
    d.stride[:] = [2, 2, 2]
    assert _C.lib().rg_ring_supported(C.byref(d)) == 0


def test_arena_scope_hands_the_callers_grad_mode_down():
    import torch
    import raygun_b200 as rb
    from raygun_b200 import kernels as K
    with rb.precision("bf16"):
        assert not K.arena_enabled()                      # grad mode on: training layout
        with torch.no_grad():
            assert K.arena_enabled()
            with K.arena_scope(False):                    # e.g. inside Function.forward of a training step
                assert not K.arena_enabled()
            assert K.arena_enabled()
        with K.arena_scope(True):
            assert K.arena_enabled()
        assert not K.arena_enabled()
    with rb.precision("fp32"), torch.no_grad():
        assert not K.arena_enabled()


def test_split_cycle_loss_surface_mirrors_the_reference():
    """SplitCycleLoss.py:13-41: constructor order (netD1, netG1, netD2, netG2, optimizer_G1, optimizer_G2, optimizer_D,
    dims), class-default lambdas, and the three gradient-sync phases a data-parallel run hooks into."""
    import inspect
    from raygun_b200.cyclegan import LinkCycleLoss, SplitCycleLoss
    names = list(inspect.signature(SplitCycleLoss.__init__).parameters)[1:9]
    assert names == ["netD1", "netG1", "netD2", "netG2", "optimizer_G1", "optimizer_G2", "optimizer_D", "dims"]
    assert issubclass(SplitCycleLoss, LinkCycleLoss)
    import torch
    lin = lambda: torch.nn.Linear(1, 1)
    nets = [lin() for _ in range(4)]
    opt = lambda *ms: torch.optim.SGD([p for m in ms for p in m.parameters()], lr=0.1)
    loss = SplitCycleLoss(nets[0], nets[1], nets[2], nets[3], opt(nets[1]), opt(nets[3]), opt(nets[0], nets[2]), 3)
    assert loss.g_lambda_dict["A"]["l1_loss"] == {"cycled": 10, "identity": 0} and loss.d_lambda_dict["B"]["fake"] == 1
    assert loss.optimizer_G is None and loss.gan_mode == "lsgan"


def test_generator_offers_its_head_only_to_networks_that_end_in_conv_activation():
    """Host logic of the fused network head (cyclegan.Generator._head_request / UNet.last_conv_before_head;
    BaseSystem.py:209-224): a plain single-head U-Net offers Conv3d(ngf -> output_nc, 1) + final activation to its last
    convolution; a U-Net with norm layers, a residual U-Net, a ResNet and a head that is not 1x1x1 do not."""
    import contextlib
    import functools
    import io
    from raygun_b200 import conv as CV
    from raygun_b200 import nn as L
    from raygun_b200.cyclegan import Generator, make_generator
    kw = dict(input_nc=1, output_nc=1, ngf=8, fmap_inc_factor=2, num_downs=3, padding_type="same")
    with contextlib.redirect_stdout(io.StringIO()):
        plain = make_generator("unet", dict(kw, final_activation="Tanh"))
        no_act = make_generator("unet", dict(kw))
        IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
        normed = make_generator("unet", dict(kw, final_activation="Tanh", norm_layer=IN))
        resid = make_generator("residualunet", dict(kw, final_activation="Tanh"))
        resnet = make_generator("resnet", dict(input_nc=1, output_nc=1, ngf=8, n_blocks=2, padding_type="same",
                                               final_activation="Tanh"))
    assert isinstance(plain, Generator)
    req = Generator._head_request(plain[0], list(plain)[1:])
    last = plain[0].last_conv_before_head()
    assert isinstance(req, CV.HeadRequest) and isinstance(last, L.Conv3d)
    assert req.weight is last.weight and req.head_w is plain[1].weight and req.act == rb._C.ACT_TANH and not req.done
    assert last is list(plain[0].r_conv[0][0].conv_pass)[-2]
    req = Generator._head_request(no_act[0], list(no_act)[1:])
    assert req is not None and req.act == rb._C.ACT_NONE
    for g in (normed, resid, resnet):
        assert Generator._head_request(g[0], list(g)[1:]) is None
    # a head with a spatial kernel is left alone
    odd = torch.nn.Sequential(plain[0], L.Conv3d(8, 1, 3, padding=1))
    assert Generator._head_request(odd[0], list(odd)[1:]) is None
    # the offer is scoped: nothing is pending outside the context manager
    with CV.offer_head(CV.HeadRequest(last.weight, plain[1].weight, plain[1].bias)) as r:
        assert CV._head_request is r
    assert CV._head_request is None
