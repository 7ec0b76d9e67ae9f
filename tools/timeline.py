"""Kernel timeline of one graphed training step (torch.profiler / CUPTI): start, duration and stream of every kernel
of the last of a few replays -- shows what runs beside what in the two-branch graph.
    python tools/timeline.py [n_rays] > profiles/<round>_step_timeline.txt"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from nerf_shared_b200 import train, utils

n_rays = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dev = torch.device("cuda:0")
utils.device = dev
args = bench.Args()
args.N_rand = n_rays
K, poses, images = bench.synth_scene()
torch.manual_seed(0)
coarse, fine = utils.create_nerf_models(args)
opt = utils.get_optimizer(coarse, fine, args)
renderer = utils.get_renderer(args, {"near": 2.0, "far": 6.0})
renderer.train()
poses_d = torch.tensor(poses, device=dev)
step = train.GraphedTrainStep(renderer, coarse, fine, opt, bench.H, bench.W, K, n_rays, chunk=args.chunk)
pix = torch.from_numpy(np.random.choice(bench.H * bench.W, n_rays, replace=False)).to(dev)
ro, rd = utils.get_rays_at(bench.H, bench.W, K, poses_d[0, :3, :4], pix)
tgt = torch.from_numpy(images[0].reshape(-1, 3)[pix.cpu().numpy()]).to(dev)
rays = torch.stack([ro, rd], 0)
for _ in range(5):
    step(rays, tgt)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        step(rays, tgt)
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
marks = [i for i, e in enumerate(evs) if "step_prologue" in e.name]      # one per call, in front of the replay
if len(marks) >= 2:
    evs = evs[marks[-2]:marks[-1]]
t0 = evs[0].time_range.start
print("# one replay of the %d-ray step: start us, duration us, stream, kernel (mode=%s)" % (n_rays, step.mode))
for e in evs:
    print("%9.1f %8.1f  s%-3s %s" % (e.time_range.start - t0, e.time_range.end - e.time_range.start,
                                    getattr(e, "device_resource_id", "?"), e.name[:90]))
print("# span %.1f us" % (evs[-1].time_range.end - t0))
