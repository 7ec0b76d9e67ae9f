"""Loss curve of main.py's training loop on an analytic scene (unit sphere at the origin shaded by its normal, white
background, 6 training views of 64x64, one held-out view).  `run("ours", ...)` drives train.GraphedTrainStep on cuda:0;
`run("reference", ...)` drives the UNMODIFIED reference (/root/reference, CPU, build container only) on the same ray
selections -- that is how tests/golden/reference_fit_curve.txt was made (tests/fit_curve.py --impl reference).
Random draws inside the renderer (perturb = 1) differ between the two, so the curves agree statistically, not step by
step.  Test infrastructure."""
import contextlib
import io
import os
import sys

import torch

from oracle import nerf_oracle as O  # seeded parameters / poses only

H = W = 64


def _scene(utils, dev, thetas):
    K = O.blender_intrinsics(H, W)
    ro, rd = [], []
    for th in thetas:
        o, d = utils.get_rays(H, W, K, torch.Tensor(O.pose_spherical(th, -30.0, 4.0)[:3, :4]).to(dev))
        ro.append(o.reshape(-1, 3).contiguous()), rd.append(d.reshape(-1, 3))
    ro, rd = torch.cat(ro), torch.cat(rd)
    a, b, c = (rd * rd).sum(-1), 2 * (ro * rd).sum(-1), (ro * ro).sum(-1) - 1.0
    disc = b * b - 4 * a * c
    t = (-b - torch.sqrt(disc.clamp_min(0))) / (2 * a)
    normal = torch.nn.functional.normalize(ro + t[:, None] * rd, dim=-1)
    return ro, rd, torch.where((disc > 0)[:, None], 0.5 + 0.5 * normal, torch.ones_like(normal))


def run(impl, steps, every=25, log=None, return_models=False, raw_noise_std=0.0):
    """-> ([(step, mean loss of the last `every` steps)], PSNR of the held-out view in dB)
    (return_models=True: also the trained coarse / fine modules)"""
    if impl == "reference":
        here = os.path.dirname(os.path.abspath(__file__))
        sys.path.insert(0, os.path.join(here, "golden"))
        import refimport
        nerf, render_utils, utils = refimport.load()
        dev = "cpu"
        torch.set_num_threads(os.cpu_count())
    else:
        from nerf_shared_b200 import nerf, render_utils, train, utils
        dev = "cuda"
    K = O.blender_intrinsics(H, W)
    ro, rd, colour = _scene(utils, dev, (-150.0, -90.0, -30.0, 30.0, 90.0, 150.0))
    to, td, truth = _scene(utils, dev, (0.0,))

    def model(seed):
        m = nerf.NeRF(D=8, W=256, output_ch=5, skips=[4], use_viewdirs=True, multires=10, multires_views=4)
        m.load_state_dict(O.make_params(seed), strict=True)
        return m.to(dev)

    torch.manual_seed(0)
    mc, mf = model(11), model(12)
    with contextlib.redirect_stdout(io.StringIO()):
        r = render_utils.Renderer(perturb=1.0, N_importance=128, N_samples=64, use_viewdirs=True, white_bkgd=True,
                                  raw_noise_std=raw_noise_std, near=2.0, far=6.0)
    r.train()
    if impl == "reference":
        opt = torch.optim.Adam(list(mc.parameters()) + list(mf.parameters()), lr=5e-4, betas=(0.9, 0.999))
    else:
        class A:
            lrate = 5e-4
        opt = utils.get_optimizer(mc, mf, A())
        step = train.GraphedTrainStep(r, mc, mf, opt, H, W, K, 1024)

    g = torch.Generator().manual_seed(5)
    acc, curve = [], []
    for i in range(1, steps + 1):
        sel = torch.randint(0, ro.shape[0], (1024,), generator=g).to(dev)
        rays, tgt = torch.stack([ro[sel], rd[sel]], 0).contiguous(), colour[sel].contiguous()
        if impl == "reference":     # main.py:77-104
            rgb, _, _, extras = r.render_from_rays(H, W, K, chunk=32768, rays=rays, coarse_model=mc, fine_model=mf,
                                                   retraw=True)
            opt.zero_grad()
            loss = utils.img2mse(rgb, tgt) + utils.img2mse(extras["rgb0"], tgt)
            loss.backward()
            opt.step()
        else:
            loss, _ = step(rays, tgt)
        acc.append(loss.detach().clone())
        if i % every == 0:
            curve.append((i, float(torch.stack(acc).mean())))
            acc = []
            if log is not None:
                log("%5d  %.5f" % curve[-1])
    r.perturb = 0.0
    r.raw_noise_std = 0.0
    with torch.no_grad():
        rgb = r.render_from_rays(H, W, K, chunk=32768, rays=torch.stack([to, td], 0), coarse_model=mc, fine_model=mf,
                                 retraw=False)[0]
    psnr = float(utils.mse2psnr(utils.img2mse(rgb, truth)))
    return (curve, psnr, mc, mf) if return_models else (curve, psnr)
