"""Host profile of main.py's loop call by call (utils.sample_random_ray_batch -> render_from_rays -> img2mse x2 ->
backward -> optimizer.step -> lr write), cProfile over 300 steps: where the host time of the eager loop goes.
    python tools/prof_eager.py > gpurun_out/prof_eager.txt"""
import cProfile
import os
import pstats
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from nerf_shared_b200 import utils

dev = torch.device("cuda:0")
utils.device = dev
args = bench.Args()
K, poses, images = bench.synth_scene()
torch.manual_seed(0)
coarse, fine = utils.create_nerf_models(args)
opt = utils.get_optimizer(coarse, fine, args)
renderer = utils.get_renderer(args, {"near": 2.0, "far": 6.0})
renderer.train()
poses_d = torch.tensor(poses, device=dev)
images_pinned = torch.from_numpy(images).pin_memory()
hwf = [bench.H, bench.W, float(K[0][0])]
i_train = np.arange(bench.N_IMAGES)


def one(i):
    batch_rays, target_s, _, _ = utils.sample_random_ray_batch(args, images_pinned, poses_d, None, 1024, False, None, i_train,
                                                                hwf, K, 0, i + 1)
    rgb, _, _, extras = renderer.render_from_rays(bench.H, bench.W, K, chunk=args.chunk, rays=batch_rays, coarse_model=coarse,
                                                  fine_model=fine, retraw=True)
    opt.zero_grad()
    img_loss = utils.img2mse(rgb, target_s)
    trans = extras['raw'][..., -1]
    loss = img_loss
    psnr = utils.mse2psnr(img_loss)
    img_loss0 = utils.img2mse(extras['rgb0'], target_s)
    loss = loss + img_loss0
    psnr0 = utils.mse2psnr(img_loss0)
    loss.backward()
    opt.step()
    for g in opt.param_groups:
        g['lr'] = 5e-4 * (0.1 ** (i / 250000))
    return loss


for i in range(20):
    one(i)
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(300):
    one(i)
t_host = time.perf_counter() - t0
torch.cuda.synchronize()
t_all = time.perf_counter() - t0
print("300 steps: host issue time %.3f ms/step, wall %.3f ms/step" % (t_host / 0.3, t_all / 0.3))
pr = cProfile.Profile()
pr.enable()
for i in range(300):
    one(i)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr, stream=sys.stdout).sort_stats("tottime").print_stats(35)
pstats.Stats(pr, stream=sys.stdout).sort_stats("cumtime").print_stats(45)
# host time per phase without the profiler (perf_counter around each call, no syncs)
import collections
acc = collections.defaultdict(float)
def timed_one(i):
    t = time.perf_counter
    a = t()
    batch_rays, target_s, _, _ = utils.sample_random_ray_batch(args, images_pinned, poses_d, None, 1024, False, None, i_train,
                                                                hwf, K, 0, i + 1)
    b = t(); acc["sample"] += b - a
    rgb, _, _, extras = renderer.render_from_rays(bench.H, bench.W, K, chunk=args.chunk, rays=batch_rays, coarse_model=coarse,
                                                  fine_model=fine, retraw=True)
    c = t(); acc["render"] += c - b
    opt.zero_grad()
    d = t(); acc["zero_grad"] += d - c
    img_loss = utils.img2mse(rgb, target_s)
    trans = extras['raw'][..., -1]
    loss = img_loss
    psnr = utils.mse2psnr(img_loss)
    img_loss0 = utils.img2mse(extras['rgb0'], target_s)
    loss = loss + img_loss0
    psnr0 = utils.mse2psnr(img_loss0)
    e = t(); acc["losses"] += e - d
    loss.backward()
    f = t(); acc["backward"] += f - e
    opt.step()
    g = t(); acc["opt.step"] += g - f
    for gr in opt.param_groups:
        gr['lr'] = 5e-4 * (0.1 ** (i / 250000))
    acc["lr"] += t() - g
_orig_stage = utils._pinned_stage
def _timed_stage(n):
    a = time.perf_counter()
    r = _orig_stage(n)
    acc["(stage wait inside sample)"] += time.perf_counter() - a
    return r
utils._pinned_stage = _timed_stage
for i in range(300):
    timed_one(i)
torch.cuda.synchronize()
print("host ms/step by phase (no profiler):", {k: round(v / 0.3, 4) for k, v in acc.items()}, "sum", round(sum(acc.values()) / 0.3, 4))
