"""GPU idle time inside the unchanged main.py loop (tools/prof_eager.py's loop): kernel / memcpy intervals of 40 steps
from torch.profiler (CUPTI), union of busy time against the span, and the largest gaps with their neighbours.
    python tools/gaps_eager.py > gpurun_out/gaps_eager.txt"""
import os
import sys

import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile

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
    loss = img_loss
    img_loss0 = utils.img2mse(extras['rgb0'], target_s)
    loss = loss + img_loss0
    loss.backward()
    opt.step()
    return loss


for i in range(30):
    one(i)
torch.cuda.synchronize()
N = 40
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(N):
        one(i)
    torch.cuda.synchronize()
ev = []
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        ev.append((e.time_range.start, e.time_range.end, e.name))
ev.sort()
t0, t1 = ev[0][0], max(e[1] for e in ev)
busy, cur_s, cur_e = 0.0, ev[0][0], ev[0][1]
gaps = []
last_name = ev[0][2]
for s, e, n in ev[1:]:
    if s > cur_e:
        busy += cur_e - cur_s
        gaps.append((s - cur_e, last_name, n))
        cur_s, cur_e = s, e
        last_name = n
    elif e > cur_e:
        cur_e = e
        last_name = n
busy += cur_e - cur_s
span = t1 - t0
print("%d steps: span %.1f us/step, busy %.1f us/step, idle %.1f us/step, %d device events/step"
      % (N, span / N, busy / N, (span - busy) / N, len(ev) // N))
agg = {}
for g, a, b in gaps:
    k = (a[:60], b[:60])
    c = agg.setdefault(k, [0, 0.0])
    c[0] += 1
    c[1] += g
for k, (c, tot) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:25]:
    print("%7.1f us/step  x%.1f/step  after %-60s before %s" % (tot / N, c / N, k[0], k[1]))
