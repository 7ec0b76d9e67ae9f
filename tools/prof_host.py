"""cProfile of the e2e training step's host side (where do the Python milliseconds go)."""
import cProfile, pstats, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from nerf_shared_b200 import utils

dev = torch.device("cuda:0"); utils.device = dev
args = bench.Args()
K, poses, images = bench.synth_scene()
coarse, fine = utils.create_nerf_models(args)
opt = utils.get_optimizer(coarse, fine, args)
renderer = utils.get_renderer(args, {"near": 2.0, "far": 6.0}); renderer.train()
poses_d = torch.tensor(poses, device=dev); images_pinned = torch.from_numpy(images).pin_memory()
hwf = [bench.H, bench.W, float(K[0][0])]; i_train = np.arange(bench.N_IMAGES)
def step(i):
    batch_rays, target, _, _ = utils.sample_random_ray_batch(args, images_pinned, poses_d, None, 1024, False, None, i_train, hwf, K, 0, i + 1)
    rgb, _, _, extras = renderer.render_from_rays(bench.H, bench.W, K, chunk=args.chunk, rays=batch_rays, coarse_model=coarse, fine_model=fine, retraw=True)
    opt.zero_grad()
    loss = utils.img2mse(rgb, target) + utils.img2mse(extras["rgb0"], target)
    loss.backward(); opt.step()
    return float(loss.item())
for i in range(5): step(i)
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(20): step(i)
torch.cuda.synchronize(); print("ms/step", (time.perf_counter() - t0) / 20 * 1e3)
pr = cProfile.Profile(); pr.enable()
for i in range(20): step(i)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
