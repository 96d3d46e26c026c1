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
