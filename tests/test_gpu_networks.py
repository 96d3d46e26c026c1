"""GPU parity, network level: the drop-in classes (fp32 kernel path) vs the committed golden vectors
generated from the reference modules (tests/golden, oracle/make_golden.py) -- outputs, input gradients and
every parameter gradient -- and the reference's own shape tests (tests/torch/networks/*_test.py) run
against the new classes."""
import functools

import pytest
import torch

from tests._golden import build_ours, case_names, load_case, load_cycle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).abs().max()) / max(1.0, float(b.abs().max()))


@pytest.mark.parametrize("name", case_names())
def test_fp32_forward_and_all_gradients_match_reference_golden(name):
    import raygun_b200 as rb
    c = load_case(name)
    net = build_ours(c["kind"], c["kw"])
    net.load_state_dict(c["sd"])
    net = net.to(DEV).train()
    with rb.precision("fp32"):
        x = c["x"].to(DEV).requires_grad_(True)
        y = net(x)
        ys = y if isinstance(y, (list, tuple)) else [y]
        assert len(ys) == len(c["ys"])
        for a, b in zip(ys, c["ys"]):
            assert a.dtype == torch.float32 and tuple(a.shape) == tuple(b.shape)
            assert _rel(a, b) <= 1e-4, ("forward", _rel(a, b))     # north_star: fp32 max-abs <= 1e-4 per layer
        loss = sum((a * r.to(DEV)).sum() for a, r in zip(ys, c["rs"]))
        loss.backward()
    assert _rel(x.grad, c["dx"]) <= 2e-4, ("dx", _rel(x.grad, c["dx"]))
    params = dict(net.named_parameters())
    for k, ref in c["grads"].items():
        g = params[k].grad
        if g is None:
            assert float(ref.abs().max()) < 1e-4, k
            continue
        # the bias in front of a norm layer has a mathematically zero gradient: rounding noise only
        tol = 3e-3 if k.endswith("bias") else 3e-4
        assert _rel(g, ref) <= tol, (k, _rel(g, ref))


def test_reference_shape_tests_on_the_new_classes():
    """tests/torch/networks/{unet,resnet,nlayerdiscriminator}_test.py shape assertions."""
    import raygun_b200 as rb
    from raygun_b200 import networks as N
    with rb.precision("fp32"), torch.no_grad():
        cp = N.ConvPass(1, 4, [(3, 3, 3)], "ReLU").to(DEV)                    # unet_test.py:11-27
        assert cp(torch.randn(1, 1, 10, 10, 10, device=DEV)).shape == (1, 4, 8, 8, 8)
        u = N.UNet(input_nc=3, ngf=16, fmap_inc_factor=2, downsample_factors=[(2, 2, 2)] * 2,
                   padding_type="same").to(DEV)                                # unet_test.py:172-196
        assert u(torch.randn(1, 3, 32, 32, 32, device=DEV)).shape == (1, 16, 32, 32, 32)
        g = N.ResNet(3, ngf=8).to(DEV)                                         # resnet_test.py:73-78
        assert g(torch.randn(1, 1, 32, 32, 32, device=DEV)).shape == (1, 8, 32, 32, 32)
        gv = N.ResNet(3, ngf=4, padding_type="valid", n_blocks=2).to(DEV)      # resnet_test.py:80-87 (scaled)
        out = gv(torch.randn(1, 1, 64, 64, 64, device=DEV))
        assert out.shape[2] == out.shape[3] == out.shape[4] and out.shape[1] == 4
        gn = N.ResNet(3, ngf=4, n_blocks=1, add_noise=True).to(DEV)            # resnet_test.py:98-106
        xin = torch.randn(1, 1, 16, 16, 16, device=DEV)
        assert not torch.equal(gn(xin), gn(xin))
        d = N.NLayerDiscriminator(3, ngf=8).to(DEV)                            # nlayerdiscriminator_test.py:42-50
        assert d(torch.randn(2, 1, 32, 32, 32, device=DEV)).shape == (2, 1, 2, 2, 2)
        md = N.MaxDownsample((2, 2, 2))
        assert md(torch.randn(1, 2, 8, 8, 8, device=DEV)).shape == (1, 2, 4, 4, 4)


def test_cyclegan_training_steps_match_reference_losses():
    """Three LinkCycleLoss steps (fp32 path) vs the losses the reference produced (golden cycle_step.npz)."""
    import raygun_b200 as rb
    from raygun_b200.cyclegan import CycleGANStep, make_generator
    from raygun_b200 import networks as N
    c = load_cycle()
    IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
    g1 = make_generator("resnet", dict(c["spec"]["gkw"]), output_nc=1, final_activation="Tanh").to(DEV)
    g2 = make_generator("resnet", dict(c["spec"]["gkw"]), output_nc=1, final_activation="Tanh").to(DEV)
    d1 = N.NLayerDiscriminator(3, norm_layer=IN, **c["spec"]["dkw"]).to(DEV)
    d2 = N.NLayerDiscriminator(3, norm_layer=IN, **c["spec"]["dkw"]).to(DEV)
    for net, nm in ((g1, "g1"), (g2, "g2"), (d1, "d1"), (d2, "d2")):
        net.load_state_dict(c["sds"][nm])
    with rb.precision("fp32"):
        step = CycleGANStep(g1, g2, d1, d2, lr=c["spec"]["lr"])
        with torch.no_grad():
            fb = g1(c["batches"][0][0].to(DEV))
        assert _rel(fb, c["fake_B0"]) <= 1e-4
        for it, (a, b) in enumerate(c["batches"]):
            out = step.step(a.to(DEV), b.to(DEV))
            got = [float(out[k]) for k in ("G1", "G2", "D1", "D2")]
            for gv, rv in zip(got, c["losses"][it]):
                # 2e-3: fp32 atomics make the weight-gradient summation order run-dependent and three Adam steps
                # amplify it (observed spread up to 1.0e-3 between identical runs)
                assert abs(gv - rv) <= 2e-3 * max(1.0, abs(rv)), (it, got, c["losses"][it])


def test_blockwise_prediction_equals_whole_volume_and_oracle_arithmetic():
    """cfg4 in miniature: G-UNet ('same', 2 downsamplings) over a uint8 volume in halo-24 blocks with origins = 0 mod 4
    reproduces the whole-volume forward (exact iff crop >= 22 and aligned origins), and the worker's
    normalise / denormalise arithmetic (truncating cast) matches the oracle restatement."""
    import raygun_b200 as rb
    from oracle import restate as O
    from raygun_b200 import kernels as K
    from raygun_b200 import predict as P
    from raygun_b200.cyclegan import make_generator
    from raygun_b200.synth import em_like_u8
    torch.manual_seed(3)
    g = make_generator("unet", dict(input_nc=1, output_nc=1, ngf=4, fmap_inc_factor=2, num_downs=3, padding_type="same",
                                    final_activation="Tanh")).to(DEV).eval()
    vol = torch.from_numpy(em_like_u8((80, 72, 76), seed=42))
    with rb.precision("fp32"), torch.no_grad():
        whole = g(K.u8_normalise(vol.to(DEV), dtype=torch.float32)[None, None])
        ref_u8 = K.denorm_clamp_u8(whole)[0, 0].cpu()
        outs, outs_m = [], []
        for rank in range(2):       # two "ranks" fill disjoint block sets of one output volume
            out, n, vox = P.predict_volume(g, vol, (64, 64, 64), (16, 16, 16), out=outs[0] if outs else None,
                                           rank=rank, world=2)
            outs.append(out)
            outm, nm, voxm = P.predict_volume(g, vol, (64, 64, 64), (16, 16, 16), out=outs_m[0] if outs_m else None,
                                              rank=rank, world=2, fit="mirror")
            outs_m.append(outm)
    out, outm = outs[0], outs_m[0]
    # reference geometry (daisy fit="valid"): blocks cover [24, 24 + 16k) per axis, the context shell stays unwritten
    inner = (slice(24, 56), slice(24, 40), slice(24, 40))
    assert len(P.block_origins(vol.shape, (16,) * 3, (64,) * 3)) == 2 * 1 * 1
    diff = (out[inner].int() - ref_u8[inner].int()).abs()
    assert int(diff.max()) <= 1 and float((diff > 0).float().mean()) < 1e-3      # fp32 rounding at truncation edges only
    mask = torch.ones_like(out, dtype=torch.bool)
    mask[inner] = False
    assert int(out[mask].max()) == 0
    # mirror fit covers everything; farther than the halo from every face it equals the whole-volume forward
    core = (slice(24, 56), slice(24, 48), slice(24, 52))
    diffm = (outm[core].int() - ref_u8[core].int()).abs()
    assert int(diffm.max()) <= 1 and float((diffm > 0).float().mean()) < 1e-3
    assert float((outm.int() - ref_u8.int()).abs().float().mean()) < 8.0          # shell: approximate, not garbage
    # oracle arithmetic on the same fp32 network output
    assert torch.equal(ref_u8, O.denorm_clamp_u8(whole.cpu())[0, 0])
    assert torch.equal(K.u8_normalise(vol.to(DEV), dtype=torch.float32).cpu(), O.u8_normalise(vol))


def test_unet_inference_arena_path_equals_copy_path(monkeypatch):
    """no_grad bf16 U-Net: conv outputs written into pad arenas, reflect padding as an in-place border fill and the
    skip concat written in place must give exactly the tensor the explicit pad / crop+cat copies give."""
    import raygun_b200 as rb
    from raygun_b200 import kernels as K
    from raygun_b200.cyclegan import make_generator
    torch.manual_seed(9)
    g = make_generator("unet", dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3, padding_type="same",
                                    final_activation="Tanh")).to(DEV).eval()
    from raygun_b200.synth import em_like_u8
    x = K.u8_normalise(torch.from_numpy(em_like_u8((24, 40, 32), seed=7)).to(DEV), dtype=torch.float32)[None, None]
    with rb.precision("bf16"), torch.no_grad():
        monkeypatch.setenv("RAYGUN_B200_ARENA", "0")
        y_copy = g(x).float().cpu()
        monkeypatch.setenv("RAYGUN_B200_ARENA", "1")
        y_arena = g(x).float().cpu()
    assert torch.equal(y_copy, y_arena)
    # and both follow the fp32 reference network within the bf16 tolerance of north_star (2e-2 relative)
    with rb.precision("fp32"), torch.no_grad():
        y32 = g(x).float().cpu()
    rel_rms = float((y_arena - y32).norm() / y32.norm())
    assert rel_rms <= 2e-2, rel_rms
    assert float((y_arena - y32).abs().max()) <= 4e-2 * max(1.0, float(y32.abs().max()))


def test_blockwise_prediction_is_independent_of_the_block_size_at_large_blocks():
    """Full-size property for cfg4: with halo 24 (>= 22) and block origins = 0 mod 4 the written voxels do not depend on
    the block geometry, so one 384^3 read block (64-bit offsets, > 2^31-element tensors, many waves of tiles in every
    kernel) must reproduce what eight 192^3 read blocks write on the region both cover."""
    import contextlib
    import sys
    import raygun_b200 as rb
    from raygun_b200 import predict as P
    from raygun_b200.cyclegan import make_generator
    from raygun_b200.synth import em_like_u8
    torch.manual_seed(42)
    with rb.precision("bf16"):
        with contextlib.redirect_stdout(sys.stderr):
            g = make_generator("unet", dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3,
                                            padding_type="same", final_activation="Tanh")).to(DEV).eval()
        vol = torch.from_numpy(em_like_u8((384, 384, 384), seed=7)).to(DEV)
        big, nb_big, _ = P.predict_volume(g, vol, (384,) * 3, (336,) * 3, out=torch.zeros_like(vol), dtype=torch.bfloat16)
        small, nb_small, _ = P.predict_volume(g, vol, (192,) * 3, (144,) * 3, out=torch.zeros_like(vol), dtype=torch.bfloat16)
    assert nb_big == 1 and nb_small == 8
    a = big[24:312, 24:312, 24:312]
    b = small[24:312, 24:312, 24:312]
    # Bit-identical: the slice-ring kernel anchors its slice groups on absolute plane indices (ring_conv_sm100.cu RingWalk),
    # so the fp32 accumulation order of a voxel no longer depends on the block's depth partition (r01: 12 % of the voxels
    # differed by one uint8 level between these two geometries).
    assert torch.equal(a, b), int((a.int() - b.int()).abs().max())
    assert float(a.float().std()) > 1.0                               # not a constant image


def test_split_cyclegan_training_steps_match_reference_losses():
    """Three SplitCycleLoss steps (CycleModel(split=True), one optimizer per generator, SplitCycleLoss.py:120-157) in the
    fp32 parity mode vs the losses the REFERENCE produced from the same weights and batches (split_cycle_step.npz,
    generated by oracle/make_golden.py --only-split)."""
    import itertools
    import os
    import numpy as np
    import raygun_b200 as rb
    from raygun_b200 import networks as N
    from raygun_b200.cyclegan import CycleModel, SplitCycleLoss, make_generator
    from tests._golden import GOLDEN
    c = load_cycle()
    ref = np.load(os.path.join(GOLDEN, "split_cycle_step.npz"))["losses"]
    IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
    g1 = make_generator("resnet", dict(c["spec"]["gkw"]), output_nc=1, final_activation="Tanh").to(DEV)
    g2 = make_generator("resnet", dict(c["spec"]["gkw"]), output_nc=1, final_activation="Tanh").to(DEV)
    d1 = N.NLayerDiscriminator(3, norm_layer=IN, **c["spec"]["dkw"]).to(DEV)
    d2 = N.NLayerDiscriminator(3, norm_layer=IN, **c["spec"]["dkw"]).to(DEV)
    for net, nm in ((g1, "g1"), (g2, "g2"), (d1, "d1"), (d2, "d2")):
        net.load_state_dict(c["sds"][nm])
    lr = c["spec"]["lr"]
    with rb.precision("fp32"):
        model = CycleModel(g1, g2, split=True)
        loss = SplitCycleLoss(d1, g1, d2, g2, torch.optim.Adam(g1.parameters(), lr=lr), torch.optim.Adam(g2.parameters(), lr=lr),
                              torch.optim.Adam(itertools.chain(d1.parameters(), d2.parameters()), lr=lr), 3)
        for it, (a, b) in enumerate(c["batches"]):
            a, b = a.to(DEV), b.to(DEV)
            fb, cb, fa, ca = model(a, b)
            loss(a, fa, ca, b, fb, cb)
            got = [loss.loss_dict[f"Total_Loss/{k}"] for k in ("G1", "G2", "D1", "D2")]
            for gv, rv in zip(got, ref[it]):
                assert abs(gv - rv) <= 2e-3 * max(1.0, abs(rv)), (it, got, ref[it].tolist())


def test_deep_unet_cfg5_bf16_path_follows_the_fp32_path():
    """cfg5 architecture (4 down-sampling levels, fmap_inc_factor 3, ngf 64: 64..5184 channels, 1.78 B parameters) on
    one 32^3 patch: bf16 tensor-core forward within 2e-2 of the fp32 path (north_star), every parameter gradient
    aligned with its fp32 counterpart, and every conv on a tensor-core kernel (192-column tiles for the x3 widths)."""
    import importlib.util
    import os
    from raygun_b200 import conv as CV
    from raygun_b200.cyclegan import make_generator
    spec = importlib.util.spec_from_file_location(
        "cfg5_deep_unet", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts", "cfg5_deep_unet.py"))
    M = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(M)
    torch.manual_seed(5)
    g = make_generator("unet", dict(M.DEEP_KW)).to(DEV)
    assert sum(p.numel() for p in g.parameters()) == 1783620993
    r = M.deep_unet_check(g, torch.device(DEV))
    assert r["rel_rms"] <= 2e-2 and r["min_grad_cosine"] >= 0.98 and abs(r["grad_norm_ratio"] - 1) <= 2e-2, r
    del g
    torch.cuda.empty_cache()


def test_cyclegan_step_replayed_as_one_cuda_graph_matches_the_eager_step():
    """CycleGANStep(graph=True): three eager steps, one capture, then replays.  Same seeds, same batches: the losses of
    every step and the final weights agree with the eager harness (wgrad atomics leave ~1e-3 run-to-run spread)."""
    import copy
    import raygun_b200 as rb
    from raygun_b200.cyclegan import CycleGANStep, make_generator
    from raygun_b200 import networks as N
    IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
    torch.manual_seed(17)
    gkw = dict(input_nc=1, output_nc=1, ngf=8, fmap_inc_factor=2, num_downs=3, padding_type="same", final_activation="Tanh")
    nets = [make_generator("unet", dict(gkw)).to(DEV) for _ in range(2)]
    nets += [N.NLayerDiscriminator(3, norm_layer=IN, input_nc=1, ngf=8, n_layers=2, kw=3, downsampling_kw=2).to(DEV)
             for _ in range(2)]
    twins = copy.deepcopy(nets)
    batches = [(torch.randn(2, 1, 32, 32, 32, device=DEV).tanh(), torch.randn(2, 1, 32, 32, 32, device=DEV).tanh())
               for _ in range(6)]
    with rb.precision("bf16"):
        eager = CycleGANStep(*nets, lr=1e-4)
        graphed = CycleGANStep(*twins, lr=1e-4, graph=True, graph_warmup=3)
        le = [eager.step(a, b) for a, b in batches]
        lg = [graphed.step(a, b) for a, b in batches]
    assert graphed.graph is not None and graphed.graph_error is None and graphed.graph_launches > 100
    for i, (x, y) in enumerate(zip(le, lg)):
        for k in ("G1", "G2", "D1", "D2"):
            assert abs(float(x[k]) - float(y[k])) <= 5e-3 * max(1.0, abs(float(x[k]))), (i, k, x, y)
    for ne, ng in zip(nets, twins):
        for (name, pe), (_, pg) in zip(ne.named_parameters(), ng.named_parameters()):
            # Adam moves a weight by ~lr per step whatever the gradient's size: near-zero gradients (zero-initialised
            # biases) follow the run-to-run noise of the fp32 atomics, so allow steps x lr absolute
            tol = max(5e-3 * float(pe.abs().max()), 6e-4)
            assert float((pe - pg).abs().max()) <= tol, name
    with pytest.raises(ValueError):
        graphed.step(batches[0][0][:1], batches[0][1][:1])
