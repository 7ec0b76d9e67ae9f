"""Multi-GPU plumbing: one process per GPU, torch.distributed (NCCL over NVLink) -- SURVEY.md section 8e.

The reference is single-process / single-device (utils.py:18); rays are independent, so the path
shards with NO exchange inside it:

* training (BASELINE.json configs[2]): every rank renders its own contiguous slice of the ray batch
  with replicated models; the weight-gradient kernels accumulate into FlatAdam's flat fp32 buffer,
  ONE all-reduce(SUM) of that 4.77 MB buffer follows, and Adam applies 1/world (equal shards, so this
  is the gradient of the global-batch mean loss);
* rendering (configs[3]): image rows are split across ranks, each rank renders its rows, the
  [rows, W, 5] (rgb, disp, acc) tiles are all-gathered.

Everything here is index arithmetic and collectives on whatever tensors it is given, so the
world_size-2 `gloo` tests (tests/test_dist_cpu.py) exercise the same code on CPU tensors.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def world_info(group=None):
    """(rank, world) -- (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def init_from_env(backend: str = "nccl"):
    """init_process_group from torchrun's RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*; binds the CUDA
    device for nccl.  Returns (rank, world, local_rank)."""
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if backend == "nccl":
        torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        kw = {"device_id": torch.device("cuda", local)} if backend == "nccl" else {}
        dist.init_process_group(backend, **kw)
    return rank, world, local


def shard_bounds(n: int, rank: int, world: int):
    """Contiguous [begin, end) of `n` units for `rank`: sizes differ by at most one, earlier ranks larger."""
    base, rem = divmod(int(n), int(world))
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def shard_ray_batch(batch_rays: torch.Tensor, target: torch.Tensor, group=None):
    """This rank's slice of a global batch (batch_rays [2, N, 3], target [N, 3]).  N must divide evenly:
    with equal shards the SUM all-reduce scaled by 1/world is exactly the global-mean gradient."""
    rank, world = world_info(group)
    n = batch_rays.shape[1]
    if n % world != 0:
        raise ValueError("global ray batch of %d does not split evenly over %d ranks" % (n, world))
    b, e = shard_bounds(n, rank, world)
    return batch_rays[:, b:e], target[b:e]


def allreduce_gradients(optimizer, group=None):
    """One SUM all-reduce of the flat gradient buffer; FlatAdam.grad_scale = 1/world does the averaging
    inside the Adam kernel (no extra pass over the buffer)."""
    rank, world = world_info(group)
    optimizer.grad_scale = 1.0 / world
    if world > 1:
        dist.all_reduce(optimizer.flat_g, op=dist.ReduceOp.SUM, group=group)
    return optimizer.flat_g


def step_with_allreduce(optimizer, group=None):
    """Gradient exchange + optimizer step of a data-parallel main.py loop (the line pair a user adds at main.py:103-104).
    With the optimizer's symmetric gradient buffer (NCCL, default group) this is ONE kernel per rank: two-shot all-reduce
    over NVLink peer memory fused with Adam (FlatAdam.step_allreduce); otherwise NCCL all-reduce + optimizer.step()."""
    rank, world = world_info(group)
    optimizer.grad_scale = 1.0 / world
    if world > 1 and group is None and getattr(optimizer, "p2p", None) is not None:
        optimizer.push_hyper()
        optimizer.step_allreduce()
        optimizer.bump_versions()
        return
    if world > 1:
        dist.all_reduce(optimizer.flat_g, op=dist.ReduceOp.SUM, group=group)
    optimizer.step()


def gather_rows(local: torch.Tensor, H: int, group=None) -> torch.Tensor:
    """local: this rank's rows [r1 - r0, ...] (shard_bounds(H, rank, world)) -> the full [H, ...] tensor on
    every rank.  Rows are padded to the largest shard so one all_gather_into_tensor moves everything."""
    rank, world = world_info(group)
    if world == 1:
        return local
    b, e = shard_bounds(H, rank, world)
    assert local.shape[0] == e - b, (local.shape, b, e)
    rows_max = (H + world - 1) // world
    tail = tuple(local.shape[1:])
    send = local.contiguous()
    if e - b < rows_max:
        send = torch.cat([send, send.new_zeros((rows_max - (e - b),) + tail)], 0)
    recv = send.new_empty((world * rows_max,) + tail)
    dist.all_gather_into_tensor(recv, send, group=group)
    if H % world == 0:
        return recv
    parts = []
    for r in range(world):
        rb, re_ = shard_bounds(H, r, world)
        parts.append(recv[r * rows_max:r * rows_max + (re_ - rb)])
    return torch.cat(parts, 0)


def render_sharded(renderer, H, W, K, chunk, c2w, coarse_model, fine_model, group=None):
    """Renderer.render_from_pose with the image rows split across ranks.
    Every rank returns the full rgb [H, W, 3], disp [H, W], acc [H, W]."""
    from . import utils
    rank, world = world_info(group)
    r0, r1 = shard_bounds(H, rank, world)
    c2w = utils._on_gpu(c2w)
    pix = torch.arange(r0 * W, r1 * W, device=c2w.device, dtype=torch.int64)
    rays_o, rays_d = utils.get_rays_at(H, W, K, c2w[:3, :4], pix)
    rgb, disp, acc, _ = renderer.render_from_rays(H, W, K, chunk, torch.stack([rays_o, rays_d], 0), coarse_model,
                                                  fine_model, retraw=False)
    tile = torch.cat([rgb, disp[:, None], acc[:, None]], -1).reshape(r1 - r0, W, 5)
    full = gather_rows(tile, H, group)
    return full[..., :3], full[..., 3], full[..., 4]
