"""CPU, world_size 2, gloo: the sharding / collective logic of nerf_shared_b200/dist.py.

The GPU kernels cannot run here, so the per-rank compute is the CPU oracle (tests may use it as the
checker); what is under test is that N ranks on equal ray shards + one SUM all-reduce scaled by
1/world reproduce the single-process large-batch gradient, and that row-sharded tiles gather into
the single-process image."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORLD = 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(fn, *args):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_entry, args=(fn, r, port, q) + args) for r in range(WORLD)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    out = {}
    while not q.empty():
        r, v = q.get()
        out[r] = v
    assert set(out) == set(range(WORLD))
    return out


def _entry(fn, rank, port, q, *args):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(WORLD),
                      LOCAL_RANK=str(rank))
    torch.set_num_threads(2)
    from nerf_shared_b200 import dist as nd
    r, w, _ = nd.init_from_env("gloo")
    assert (r, w) == (rank, WORLD)
    try:
        q.put((rank, fn(rank)))
    finally:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------
def test_shard_bounds_partition():
    from nerf_shared_b200.dist import shard_bounds
    for n in (0, 1, 7, 8, 100, 800, 1023, 65536):
        for world in (1, 2, 3, 4, 8):
            cover = []
            for r in range(world):
                b, e = shard_bounds(n, r, world)
                assert 0 <= b <= e <= n
                cover += list(range(b, e))
            assert cover == list(range(n))
            sizes = [shard_bounds(n, r, world)[1] - shard_bounds(n, r, world)[0] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def _w_gather(rank):
    from nerf_shared_b200 import dist as nd
    out = {}
    for H, W in ((8, 5), (7, 3), (1, 4), (101, 2)):      # even, ragged, fewer rows than ranks, odd
        full = torch.arange(H * W * 5, dtype=torch.float32).reshape(H, W, 5)
        b, e = nd.shard_bounds(H, rank, WORLD)
        got = nd.gather_rows(full[b:e].clone(), H)
        out[(H, W)] = bool(torch.equal(got, full))
    return out


def test_gather_rows_world2():
    res = _run(_w_gather)
    for r in range(WORLD):
        assert all(res[r].values()), res


def _w_grad(rank):
    """each rank: oracle gradients of its half of a 16-ray batch into a flat buffer, all-reduce, 1/world"""
    from nerf_shared_b200 import dist as nd
    from oracle import nerf_oracle as O

    class Opt:       # the two attributes allreduce_gradients touches on optim.FlatAdam
        grad_scale = 1.0
        flat_g = None

    N = 16
    g = torch.Generator().manual_seed(0)
    K = O.blender_intrinsics(20, 20)
    ro, rd = O.get_rays(20, 20, K, torch.from_numpy(O.pose_spherical(10.0, -30.0, 4.0))[:3, :4])
    idx = torch.randperm(400, generator=g)[:N]
    batch = torch.stack([ro.reshape(-1, 3)[idx], rd.reshape(-1, 3)[idx]], 0)
    target = torch.rand(N, 3, generator=g)
    pc, pf = O.make_params(11), O.make_params(12)
    kw = dict(n_samples=16, n_importance=16)

    def flat(gc, gf):
        return torch.cat([gc[k].reshape(-1) for k in sorted(gc)] + [gf[k].reshape(-1) for k in sorted(gf)])

    rays_s, tgt_s = nd.shard_ray_batch(batch, target)
    assert rays_s.shape == (2, N // WORLD, 3) and tgt_s.shape == (N // WORLD, 3)
    loss_s, gc, gf, _ = O.train_step_grads(pc, pf, rays_s[0], rays_s[1], tgt_s, 2.0, 6.0, **kw)
    opt = Opt()
    opt.flat_g = flat(gc, gf)
    nd.allreduce_gradients(opt)
    avg = opt.flat_g * opt.grad_scale
    loss_f, gc, gf, _ = O.train_step_grads(pc, pf, batch[0], batch[1], target, 2.0, 6.0, **kw)
    full = flat(gc, gf)
    return {"rel": float((avg - full).norm() / full.norm()), "scale": opt.grad_scale, "n": int(full.numel())}


def test_sharded_gradient_equals_large_batch_world2():
    res = _run(_w_grad)
    for r in range(WORLD):
        assert res[r]["scale"] == 0.5 and res[r]["n"] == 2 * 595844
        assert res[r]["rel"] < 2e-5, res
    assert res[0]["rel"] == res[1]["rel"]      # both ranks hold the same reduced buffer


def test_uneven_batch_is_refused():
    from nerf_shared_b200 import dist as nd
    assert nd.world_info() == (0, 1)
    rays, tgt = torch.zeros(2, 10, 3), torch.zeros(10, 3)
    a, b = nd.shard_ray_batch(rays, tgt)
    assert a.shape[1] == 10 and b.shape[0] == 10


def _w_uneven(rank):
    from nerf_shared_b200 import dist as nd
    try:
        nd.shard_ray_batch(torch.zeros(2, 7, 3), torch.zeros(7, 3))
    except ValueError:
        return True
    return False


def test_uneven_batch_is_refused_world2():
    res = _run(_w_uneven)
    assert res[0] and res[1]
