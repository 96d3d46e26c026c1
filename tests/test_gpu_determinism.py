"""Bit-reproducibility of the CUDA path, the flat-bucket optimiser and the captured step.

The reference's CPU path and cuDNN's deterministic algorithms give the same bits run after run; so must this library
(the fp32 `red.global.add` finishes of the weight-gradient kernels and the statistics atomics made two runs
of the SAME kernels drift 5-6 % apart over 200 steps).  Every reduction is now two-stage with a fixed order
(csrc/reduce_rows.cuh, csrc/wgrad_finish.cuh), and the slice-ring kernel's accumulation order per output voxel does
not depend on the depth partition (ring_conv_sm100.cu: RingWalk)."""
import contextlib
import copy
import functools
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
IN = functools.partial(torch.nn.InstanceNorm3d, affine=False, track_running_stats=False)
G_KW = dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3, padding_type="same", final_activation="Tanh")
D_KW = dict(input_nc=1, ngf=64, n_layers=3, kw=3, downsampling_kw=2)


def _quiet():
    return contextlib.redirect_stdout(sys.stderr)


def _grads_of(make, x, t, mode):
    import raygun_b200 as rb
    from raygun_b200 import functional as F
    torch.manual_seed(5)
    with _quiet():
        g = make().to(DEV)
    with rb.precision(mode):
        y = g(x)
        loss = F.smooth_l1_loss(t, y)
        loss.backward()
    return y.detach().clone(), loss.detach().clone(), [p.grad.clone() for p in g.parameters()]


@pytest.mark.parametrize("mode", ["bf16", "fp32"])
@pytest.mark.parametrize("net", ["resnet", "unet"])
def test_generator_forward_backward_is_bit_reproducible(mode, net):
    """Same weights, same batch, two runs: outputs, loss and EVERY parameter gradient are bit-identical (tensor-core
    wgrad slabs, epilogue statistics rows, norm / bias / loss reducers: no atomics anywhere)."""
    from raygun_b200.cyclegan import make_generator
    ngf = 64 if mode == "bf16" else 8
    if net == "resnet":
        make = lambda: make_generator("resnet", dict(input_nc=1, output_nc=1, ngf=ngf, n_blocks=2, padding_type="same",
                                                     final_activation="Tanh"))
    else:
        make = lambda: make_generator("unet", dict(G_KW, ngf=ngf))
    gen = torch.Generator().manual_seed(1)
    x = torch.rand(2, 1, 32, 32, 32, generator=gen).mul(2).sub(1).to(DEV)
    t = torch.rand(2, 1, 32, 32, 32, generator=gen).mul(2).sub(1).to(DEV)
    y0, l0, g0 = _grads_of(make, x, t, mode)
    y1, l1, g1 = _grads_of(make, x, t, mode)
    assert torch.equal(y0, y1) and torch.equal(l0, l1)
    for a, b in zip(g0, g1):
        assert torch.equal(a, b)


def _cycle_nets(ngf=64, seed=17):
    from raygun_b200 import networks as N
    from raygun_b200.cyclegan import make_generator
    torch.manual_seed(seed)
    with _quiet():
        nets = [make_generator("unet", dict(G_KW, ngf=ngf)).to(DEV) for _ in range(2)]
        nets += [N.NLayerDiscriminator(3, norm_layer=IN, **dict(D_KW, ngf=ngf)).to(DEV) for _ in range(2)]
        for d in nets[2:]:
            N.init_weights(d, "kaiming")
    return nets


def _batches(n, side=32, batch=1, seed=3):
    gen = torch.Generator().manual_seed(seed)
    return [(torch.rand(batch, 1, side, side, side, generator=gen).mul(2).sub(1).to(DEV),
             torch.rand(batch, 1, side, side, side, generator=gen).mul(2).sub(1).to(DEV)) for _ in range(n)]


def test_cyclegan_training_is_bit_reproducible_and_graph_equals_eager_on_the_tensor_core_path():
    """ngf 64 (ring / igemm / halo-wgrad / stem / upconv2 kernels and their weight packs under capture and replay, ADVICE
    r01): an eager harness, a second eager harness and a CUDA-graph harness, same seeds and batches -> the four losses of
    every step and the final weights are IDENTICAL.  An eager forward between replays uses the packs the replay
    refreshed in place (no stale tensor-core weights)."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    from raygun_b200.cyclegan import CycleGANStep
    nets = _cycle_nets()
    twins, third = copy.deepcopy(nets), copy.deepcopy(nets)
    batches = _batches(7)
    probe = batches[0][0]
    with rb.precision("bf16"):
        before = dict(CV.stats)
        eager = CycleGANStep(*nets, lr=1e-4)
        eager2 = CycleGANStep(*third, lr=1e-4)
        graphed = CycleGANStep(*twins, lr=1e-4, graph=True, graph_warmup=2)
        le, lg, l2, probes_e, probes_g = [], [], [], [], []
        for a, b in batches:
            le.append(eager.step(a, b))
            l2.append(eager2.step(a, b))
            lg.append(graphed.step(a, b))
            with torch.no_grad():           # an eager forward between replays (validation / prediction / snapshots)
                probes_e.append(nets[0](probe).clone())
                probes_g.append(twins[0](probe).clone())
    assert graphed.graph is not None and graphed.graph_error is None and graphed.graph_launches > 100
    assert CV.stats.get("ring", 0) > before.get("ring", 0) and CV.stats.get("halo_wgrad", 0) > before.get("halo_wgrad", 0)
    for i in range(len(batches)):
        for k in ("G1", "G2", "D1", "D2"):
            assert float(le[i][k]) == float(l2[i][k]), ("eager vs eager", i, k, le[i], l2[i])
            assert float(le[i][k]) == float(lg[i][k]), ("eager vs graph", i, k, le[i], lg[i])
        assert torch.equal(probes_e[i], probes_g[i]), f"stale packs after replay {i}"
    for ne, ng in zip(nets, twins):
        for (name, pe), (_, pg) in zip(ne.named_parameters(), ng.named_parameters()):
            assert torch.equal(pe, pg), name
    with pytest.raises(ValueError):
        graphed.step(batches[0][0][:, :, :16], batches[0][1][:, :, :16])


def test_flat_adam_follows_torch_adam_and_single_backward_equals_the_reference_schedule():
    """optim.FlatAdam (flat buckets, device step counter, wgrad kernels writing into the bucket) vs torch.optim.Adam on
    per-parameter gradients, and LinkCycleLoss(single_backward=True) (4 generator traversals) vs the reference's two
    retain_graph backward calls (6): same losses, weights within float rounding of the update."""
    import raygun_b200 as rb
    from raygun_b200.cyclegan import CycleGANStep
    nets = _cycle_nets(ngf=8)
    twins, third = copy.deepcopy(nets), copy.deepcopy(nets)
    batches = _batches(3, batch=2)
    with rb.precision("fp32"):
        flat = CycleGANStep(*nets, lr=1e-4)
        stock = CycleGANStep(*twins, lr=1e-4, optimizer_cls="torch")
        single = CycleGANStep(*third, lr=1e-4, single_backward=True)
        from raygun_b200.optim import FlatAdam
        assert isinstance(flat.optimizer_G, FlatAdam) and isinstance(stock.optimizer_G, torch.optim.Adam)
        for a, b in batches:
            lf, ls, l1 = flat.step(a, b), stock.step(a, b), single.step(a, b)
            for k in ("G1", "G2", "D1", "D2"):
                assert abs(float(lf[k]) - float(ls[k])) <= 2e-5 * max(1.0, abs(float(ls[k]))), (k, lf, ls)
                assert abs(float(l1[k]) - float(ls[k])) <= 2e-5 * max(1.0, abs(float(ls[k]))), (k, l1, ls)
    for nf, ns, n1 in zip(nets, twins, third):
        for (name, pf), (_, ps), (_, p1) in zip(nf.named_parameters(), ns.named_parameters(), n1.named_parameters()):
            # Adam moves a weight by ~lr per step whatever the gradient's size: biases in front of an InstanceNorm have a
            # mathematically zero gradient, so their updates follow rounding noise -> allow steps x lr absolute
            assert float((pf.detach() - ps.detach()).abs().max()) <= 3.2e-4, name
            assert float((p1.detach() - ps.detach()).abs().max()) <= 3.2e-4, name
    # state_dict stays checkpoint-compatible: parameters are views of the flat bucket, not replaced objects
    sd = nets[0].state_dict()
    assert all(v.dtype == torch.float32 for v in sd.values())
    copy.deepcopy(twins[0]).load_state_dict(sd)


def test_blockwise_prediction_is_bitwise_independent_of_block_geometry_and_equals_the_whole_volume():
    """cfg4 property: halo 24 >= 22, write sizes (hence read origins) multiples of 12 -> the uint8 output does
    not depend on the read-block geometry, and with fit="clip" (read rois clipped at the volume faces) the stitched
    volume EQUALS the whole-volume forward, face voxels included -- on the bf16 tensor-core path, bit for bit."""
    import raygun_b200 as rb
    from raygun_b200 import kernels as K
    from raygun_b200 import predict as P
    from raygun_b200.cyclegan import make_generator
    from raygun_b200.synth import em_like_u8
    torch.manual_seed(42)
    with rb.precision("bf16"), torch.no_grad():
        with _quiet():
            g = make_generator("unet", dict(G_KW)).to(DEV).eval()
        vol = torch.from_numpy(em_like_u8((192, 168, 192), seed=7)).to(DEV)
        whole = K.denorm_clamp_u8(g(K.u8_normalise(vol, dtype=torch.bfloat16)[None, None]))[0, 0]
        outs = []
        for read, write in (((120,) * 3, (72,) * 3), ((96, 144, 84), (48, 96, 36)), ((180, 168, 192), (132, 120, 144))):
            out, nb, vox = P.predict_volume(g, vol, read, write, out=torch.zeros_like(vol), dtype=torch.bfloat16, fit="clip")
            assert vox == vol.numel() and nb == len(P.block_list(vol.shape, read, write, "clip"))
            outs.append(out)
        for o in outs:
            assert torch.equal(o, whole)
        # two "ranks": cost-balanced disjoint shares fill one output volume
        out2 = torch.zeros_like(vol)
        for r in range(2):
            P.predict_volume(g, vol, (120,) * 3, (72,) * 3, out=out2, rank=r, world=2, dtype=torch.bfloat16, fit="clip")
        assert torch.equal(out2, whole)
        assert float(whole.float().std()) > 1.0


def test_training_skip_concat_in_place_equals_the_copying_concat_bit_for_bit():
    """U-Net training forward + backward with the producers writing straight into the channel slices of the concat
    buffer (UNet._train_cat_forward / functional.cat_join; UNet.py:338-351) vs the two strided copies of F.crop_cat:
    same kernels on the same values -> identical outputs and identical parameter gradients, 2 copies per level fewer."""
    import raygun_b200 as rb
    from raygun_b200 import _C
    from raygun_b200 import functional as F
    from raygun_b200.cyclegan import make_generator
    gen = torch.Generator().manual_seed(11)
    x = torch.rand(2, 1, 32, 40, 48, generator=gen).mul(2).sub(1).to(DEV)
    t = torch.rand(2, 1, 32, 40, 48, generator=gen).mul(2).sub(1).to(DEV)
    make = lambda: make_generator("unet", dict(G_KW))
    res = {}
    for flag in (True, False):
        old, F.TRAIN_CAT = F.TRAIN_CAT, flag
        try:
            n0 = _C.launch_count
            res[flag] = _grads_of(make, x, t, "bf16") + (_C.launch_count - n0,)
        finally:
            F.TRAIN_CAT = old
    (y1, l1, g1, n1), (y0, l0, g0, n0) = res[True], res[False]
    assert torch.equal(y1, y0) and torch.equal(l1, l0)
    for a, b in zip(g1, g0):
        assert torch.equal(a, b)
    assert n1 == n0 - 4, (n1, n0)          # two levels x two copies of the forward concat
