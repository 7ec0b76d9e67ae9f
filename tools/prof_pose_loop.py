"""Host profile of the pose demo's loop (demo_est_rel_pose.py:78-97) against this package."""
import cProfile
import os
import pstats
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from nerf_shared_b200 import datasets, render_utils, utils

dev = torch.device("cuda:0")
utils.device = dev
Hf, Wf = 378, 504
images, poses, _, Kf = datasets.synthetic_llff(Hf, Wf, 2)


class A(bench.Args):
    N_importance = 64


torch.manual_seed(0)
mc, mf = utils.create_nerf_models(A())
for m in (mc, mf):
    for p in m.parameters():
        p.requires_grad_(False)
r = render_utils.Renderer(perturb=0.0, N_importance=64, N_samples=64, use_viewdirs=True, white_bkgd=False,
                          raw_noise_std=0.0, ndc=True, lindisp=False, near=0.0, far=1.0)
pose = torch.tensor(poses[0, :3, :4], device=dev).clone().requires_grad_(True)
batches = [np.random.choice(Hf * Wf, size=512, replace=False) for _ in range(16)]
k = [0]


def step():
    sel = batches[k[0] % 16]
    k[0] += 1
    ys, xs = sel // Wf, sel % Wf
    target_s = torch.Tensor(images[1][ys, xs]).to(dev)
    rays_o, rays_d = utils.get_rays(Hf, Wf, Kf, pose)
    ro, rd = rays_o[ys, xs], rays_d[ys, xs]
    rgb = r.render_from_rays(Hf, Wf, Kf, chunk=32768, rays=torch.stack([ro, rd], 0), coarse_model=mc, fine_model=mf, retraw=True)[0]
    loss = utils.img2mse(rgb, target_s)
    pose.grad = None
    loss.backward()
    return loss


for _ in range(10):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(100):
    step()
torch.cuda.synchronize()
print("pose loop: %.3f ms/step" % ((time.perf_counter() - t0) * 10))
import gc
gc.disable()
t0 = time.perf_counter()
for _ in range(100):
    step()
torch.cuda.synchronize()
print("pose loop, gc disabled: %.3f ms/step" % ((time.perf_counter() - t0) * 10))
gc.enable()
seg = [0.0] * 6
for _ in range(100):
    a = time.perf_counter()
    sel = batches[k[0] % 16]; k[0] += 1
    ys, xs = sel // Wf, sel % Wf
    target_s = torch.Tensor(images[1][ys, xs]).to(dev)
    b = time.perf_counter()
    rays_o, rays_d = utils.get_rays(Hf, Wf, Kf, pose)
    ro, rd = rays_o[ys, xs], rays_d[ys, xs]
    c = time.perf_counter()
    st = torch.stack([ro, rd], 0)
    rgb = r.render_from_rays(Hf, Wf, Kf, chunk=32768, rays=st, coarse_model=mc, fine_model=mf, retraw=True)[0]
    d = time.perf_counter()
    loss = utils.img2mse(rgb, target_s)
    pose.grad = None
    e = time.perf_counter()
    loss.backward()
    f = time.perf_counter()
    for i, (x, y) in enumerate(((a, b), (b, c), (c, d), (d, e), (e, f))):
        seg[i] += y - x
torch.cuda.synchronize()
print("segments ms/step: batch+target %.3f  get_rays+index %.3f  stack+render %.3f  mse %.3f  backward %.3f" % tuple(1e3 * s_ / 100 for s_ in seg[:5]))
pr = cProfile.Profile()
pr.enable()
for _ in range(100):
    step()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr, stream=sys.stdout).sort_stats("cumulative").print_stats(30)
