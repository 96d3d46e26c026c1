"""Host-side multi-rank logic on CPU (gloo, world size 2): gradient bucket all-reduce and block partition."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from raygun_b200 import predict as P


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from raygun_b200.dist import GradSync, init_from_env, shard_batch
    r, w, _ = init_from_env("gloo")
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.Linear(3, 2))
    x = torch.arange(32, dtype=torch.float32).view(8, 4) / 10
    xs = shard_batch(x, r, w)
    net(xs).pow(2).mean().backward()
    net[1].bias.grad = None                      # a parameter without gradient on every rank
    sync = GradSync({"G": list(net.parameters())})
    sync("G")
    full = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.Linear(3, 2))
    full.load_state_dict(net.state_dict())
    full(x).pow(2).mean().backward()
    err = max(float((a.grad - b.grad).abs().max()) for a, b in zip(net.parameters(), full.parameters())
              if a.grad is not None)
    q.put((r, err, sync.bytes_reduced))
    dist.destroy_process_group()


def test_gradsync_average_equals_full_batch_gradient_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for r, err, nbytes in res:
        assert err < 1e-6, (r, err)
        assert nbytes == 4 * (12 + 3 + 6 + 2)


def test_block_partition_is_a_disjoint_cover():
    origins = P.block_origins((100, 64, 70), (32, 32, 32))
    assert len(origins) == 4 * 2 * 3 and origins[0] == (0, 0, 0) and origins[-1] == (96, 32, 64)
    parts = [P.rank_blocks(origins, r, 3) for r in range(3)]
    assert sorted(sum(parts, [])) == sorted(origins)
    assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_valid_fit_matches_daisy_geometry():
    """predict.py:154-164: read roi at the origin, shifted by the write size while it stays inside the volume."""
    o = P.block_origins((100, 64, 70), (16, 16, 16), (64, 64, 64), fit="valid")
    assert o[0] == (24, 24, 24) and len(o) == 3 * 1 * 1 and o[-1] == (56, 24, 24)
    assert P.block_origins((63, 64, 64), (16,) * 3, (64,) * 3) == []


class _FlatStub:
    """What dist.GradSync needs of an optim.FlatAdam (which itself needs a CUDA device): the flat gradient bucket."""

    def __init__(self, n):
        self.flat_g = torch.zeros(n)
        self.grad_scale = 1.0


def _flat_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from raygun_b200.dist import GradSync, init_from_env
    init_from_env("gloo")
    opt = _FlatStub(1000)
    sync = GradSync({"G": opt}, chunk_bytes=4 * 300)           # 4 chunks of <= 300 floats
    opt.flat_g.copy_(torch.arange(1000, dtype=torch.float32) * (rank + 1))
    sync.begin("G")                                            # asynchronous: the caller keeps launching work here
    busy = torch.ones(8).sum()
    sync.finish("G")
    want = torch.arange(1000, dtype=torch.float32) * sum(r + 1 for r in range(world))
    q.put((rank, float((opt.flat_g - want).abs().max()), opt.grad_scale, len(sync.pending), float(busy), sync.bytes_reduced))
    dist.destroy_process_group()


def test_gradsync_flat_bucket_chunked_async_allreduce_gloo_world2():
    """The data-parallel exchange of the CycleGAN step (LinkCycleLoss.py:78,128 call sites): the flat gradient bucket is
    summed in place by chunked asynchronous all-reduces started at `begin` and joined at `finish`; the 1/world of the
    average is handed to the optimiser's update instead of a scaling pass."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_flat_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for r, err, scale, pending, busy, nbytes in res:
        assert err == 0.0 and scale == 0.5 and pending == 0 and busy == 8.0 and nbytes == 4000


def test_clip_fit_geometry_and_cost_balanced_partition_of_the_cfg4_grid():
    """cfg4: the 2048 x 2048 x 1024 volume in 516^3 write blocks with halo 24, read rois clipped at the volume faces
    (predict.py:92-108 geometry; whole-volume semantics at the faces): 4 x 4 x 2 = 32 blocks that cover the volume exactly
    once, read origins multiples of 12, and a longest-processing-time partition that balances 8 ranks within 3 %."""
    vol, W, R = (1024, 2048, 2048), 516, 564
    blocks = P.block_list(vol, (R,) * 3, (W,) * 3, "clip")
    assert len(blocks) == 2 * 4 * 4
    assert sum(b.voxels for b in blocks) == vol[0] * vol[1] * vol[2]
    cover = torch.zeros(vol[0] // 4, vol[1] // 4, vol[2] // 4, dtype=torch.uint8)        # 4-voxel granularity suffices
    for b in blocks:
        assert all(o % 12 == 0 and e % 4 == 0 for o, e in zip(b.read_lo, b.read_ext))      # pooling + ring-group alignment
        assert all(lo >= 0 and lo + e <= v for lo, e, v in zip(b.read_lo, b.read_ext, vol))
        assert all(c in (0, 24) for c in b.crop)
        assert all(c + e <= r for c, e, r in zip(b.crop, b.ext, b.read_ext))
        z, y, x = (o // 4 for o in b.origin)
        cover[z: z + b.ext[0] // 4, y: y + b.ext[1] // 4, x: x + b.ext[2] // 4] += 1
    assert int(cover.min()) == 1 and int(cover.max()) == 1
    for world in (2, 4, 8):
        parts = [P.partition_blocks(blocks, r, world) for r in range(world)]
        assert sorted(b.origin for p in parts for b in p) == sorted(b.origin for b in blocks)
        loads = [sum(b.cost for b in p) for p in parts]
        assert max(loads) <= 1.03 * (sum(loads) / world), (world, loads)
    # equal-cost blocks (the reference's "valid" fit): the partition degenerates to the round robin of rank_blocks
    valid = P.block_list((100, 64, 70), (64,) * 3, (16,) * 3, "valid")
    assert [b.origin for b in P.partition_blocks(valid, 1, 2)] == P.rank_blocks([b.origin for b in valid], 1, 2)


def _halo_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch.nn.functional as TF
    from raygun_b200.dist import halo_exchange, init_from_env
    r, w, _ = init_from_env("gloo")
    torch.manual_seed(3)
    x = torch.randn(2, 3, 8 * w, 6, 5, dtype=torch.float64)          # full tensor, identical on every rank
    w1 = torch.randn(4, 3, 3, 3, 3, dtype=torch.float64, requires_grad=True)
    w2 = torch.randn(2, 4, 5, 1, 1, dtype=torch.float64, requires_grad=True)
    t = torch.randn(2, 2, 8 * w, 6, 5, dtype=torch.float64)

    def net(xin, sharded):
        # two 'same' convolutions (zero padding): depth taps come from the neighbours when sharded, from F.pad when not
        def ext(v, k):
            return halo_exchange(v, 2, k) if sharded else TF.pad(v, (0, 0, 0, 0, k, k))
        y = TF.relu(TF.conv3d(ext(xin, 1), w1, padding=(0, 1, 1)))
        return TF.conv3d(ext(y, 2), w2)

    n = x.shape[2] // w
    xs = x[:, :, r * n: (r + 1) * n].clone().requires_grad_(True)
    ys = net(xs, True)
    loss = ((ys - t[:, :, r * n: (r + 1) * n]) ** 2).sum()
    loss.backward()
    g1, g2 = w1.grad.clone(), w2.grad.clone()
    dist.all_reduce(g1)                                                # weight gradients: sum of the shards' partial sums
    dist.all_reduce(g2)
    w1.grad = w2.grad = None
    xf = x.clone().requires_grad_(True)
    yf = net(xf, False)
    ((yf - t) ** 2).sum().backward()
    err = max(float((ys - yf[:, :, r * n: (r + 1) * n]).abs().max()),
              float((xs.grad - xf.grad[:, :, r * n: (r + 1) * n]).abs().max()),
              float((g1 - w1.grad).abs().max()), float((g2 - w2.grad).abs().max()))
    q.put((r, err))
    dist.destroy_process_group()


def test_halo_exchange_makes_a_depth_sharded_conv_stack_equal_the_unsharded_one_gloo_world2():
    """dist.halo_exchange (the building block of spatial model parallelism): two stacked convolutions on a tensor
    split along depth over 2 ranks, halos of 1 and 2 planes -> outputs, input gradients and (all-reduced) weight gradients
    equal the unsharded computation in float64."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_halo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for r, err in res:
        assert err < 1e-10, (r, err)
