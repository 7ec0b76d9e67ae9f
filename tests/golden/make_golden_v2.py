"""Goldens with NON-DEGENERATE density, from the UNMODIFIED reference (/root/reference) -> reference_golden_v2.npz.

At nn.Linear's default initialisation every fine-network sigma is negative, so the round-1 goldens compare constant
white / black images and all-zero fine gradients.  This file adds three model sets whose density is positive on a
large fraction of samples:

  stress   matrices x1.59, seeds 13 (coarse) / 14 (fine)
  bias     default init (seeds 11 / 12) with alpha_linear.bias += 0.1
  trained  the reference's own weights after 300 steps of main.py's loop on the analytic sphere scene
           (tests/golden/trained_sphere.npz, made by make_trained_weights.py)

and, per set: render_rays on 256 rays, a training step with non-zero fine gradients, a pose step with a non-zero
grad_pose; for the trained set also the held-out 64x64 frame and the 400x400 (BASELINE.json configs[0]) frame.

Build container only:   python tests/golden/make_golden_v2.py [section ...]     (default: all sections)
Sections are merged into the existing file, so the slow ones (frame400: ~5 min of CPU) can be rerun alone."""
import hashlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import refimport  # noqa: E402

nerf, render_utils, utils = refimport.load()
from oracle import nerf_oracle as O  # noqa: E402
from tests import golden_sets as GS  # noqa: E402

torch.set_num_threads(os.cpu_count() or 1)
OUT = os.path.join(HERE, "reference_golden_v2.npz")
sections = sys.argv[1:] or ["rr", "train", "pose", "llff", "frame64", "frame400"]
out = {}
if os.path.exists(OUT):
    with np.load(OUT) as z:
        out = {k: z[k] for k in z.files}


def npy(t):
    return t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t)


def ref_model(params):
    m = nerf.NeRF(D=8, W=256, output_ch=5, skips=[4], use_viewdirs=True, multires=10, multires_views=4)
    m.load_state_dict(params, strict=True)
    return m


def renderer(**kw):
    import contextlib
    import io
    base = dict(perturb=0.0, N_importance=128, N_samples=64, use_viewdirs=True, white_bkgd=True, raw_noise_std=0.0,
                near=2.0, far=6.0)
    base.update(kw)
    with contextlib.redirect_stdout(io.StringIO()):
        return render_utils.Renderer(**base)


def depth_of(ret):
    return torch.sum(ret["weights"] * ret["z_vals"], -1)


sets = GS.available_sets()
print("model sets:", sets, "sections:", sections)

for name in sets:
    pc, pf = GS.params(name)
    mc, mf = ref_model(pc), ref_model(pf)
    rb = GS.ray_batch(name)                                   # [256, 11]

    if "rr" in sections:
        with torch.no_grad():
            ret = renderer().render_rays(rb, mc, mf, retraw=True, retweights=True)
        for k in ("rgb_map", "disp_map", "acc_map", "rgb0", "acc0", "z_std"):
            out["rr/%s/%s" % (name, k)] = npy(ret[k])
        out["rr/%s/depth" % name] = npy(depth_of(ret))
        out["rr/%s/sigma_pos_frac" % name] = np.float64((ret["raw"][..., 3] > 0).float().mean().item())
        out["rr/%s/acc_mean" % name] = np.float64(ret["acc_map"].mean().item())
        # random-sampling mode with the reference's own draws (perturb = 1, raw_noise_std = 1)
        torch.manual_seed(4321)
        with torch.no_grad():
            ret = renderer(perturb=1.0, raw_noise_std=1.0).render_rays(rb[:64], mc, mf, retraw=True, retweights=True)
        for k in ("rgb_map", "acc_map", "rgb0"):
            out["rr_rand/%s/%s" % (name, k)] = npy(ret[k])
        out["rr_rand/%s/depth" % name] = npy(depth_of(ret))
        print(name, "rr: sigma>0 on %.1f%% of fine samples, mean acc %.3f" %
              (100 * out["rr/%s/sigma_pos_frac" % name], out["rr/%s/acc_mean" % name]))

    if "train" in sections:
        r = renderer()
        r.train()
        tc, tf = ref_model(pc), ref_model(pf)
        sub = rb[GS.TRAIN_ROWS]
        ro, rd = sub[:, 0:3].contiguous(), sub[:, 3:6].contiguous()
        tgt = GS.train_target(name)
        opt = torch.optim.Adam(list(tc.parameters()) + list(tf.parameters()), lr=5e-4, betas=(0.9, 0.999))
        rgb, _, _, extras = r.render_from_rays(16, 16, GS.K16, chunk=32768, rays=torch.stack([ro, rd], 0), coarse_model=tc,
                                               fine_model=tf, retraw=True)
        opt.zero_grad()
        loss = utils.img2mse(rgb, tgt) + utils.img2mse(extras["rgb0"], tgt)
        loss.backward()
        out["train/%s/loss" % name] = npy(loss)
        out["train/%s/psnr" % name] = npy(utils.mse2psnr(utils.img2mse(rgb, tgt)))
        for tag, m in (("coarse", tc), ("fine", tf)):
            for k, p in m.named_parameters():
                g = p.grad
                out["train/%s/%s/gnorm/%s" % (name, tag, k)] = np.float64(g.double().norm().item())
                if g.numel() <= 384 or k.endswith(".bias"):
                    out["train/%s/%s/g/%s" % (name, tag, k)] = npy(g)
                else:
                    out["train/%s/%s/ghead/%s" % (name, tag, k)] = npy(g.reshape(-1)[:2048])
        opt.step()
        for tag, m in (("coarse", tc), ("fine", tf)):
            out["train/%s/%s/after_adam/alpha_w" % (name, tag)] = npy(m.alpha_linear.weight)
            out["train/%s/%s/after_adam/l7_b" % (name, tag)] = npy(m.pts_linears[7].bias)
        print(name, "train: loss %.5f, |g fine alpha_w| %.3e, |g fine l0_w| %.3e" %
              (float(loss), out["train/%s/fine/gnorm/alpha_linear.weight" % name],
               out["train/%s/fine/gnorm/pts_linears.0.weight" % name]))

    if "pose" in sections:
        fc, ff = ref_model(pc), ref_model(pf)
        for p in list(fc.parameters()) + list(ff.parameters()):
            p.requires_grad_(False)
        r = renderer()
        pose = torch.Tensor(GS.pose_of(name)).clone().requires_grad_(True)
        ro_p, rd_p = utils.get_rays(16, 16, GS.K16, pose)
        ys, xs = GS.POSE_SEL // 16, GS.POSE_SEL % 16
        rgb, _, _, _ = r.render_from_rays(16, 16, GS.K16, chunk=32768, rays=torch.stack([ro_p[ys, xs], rd_p[ys, xs]], 0),
                                          coarse_model=fc, fine_model=ff, retraw=True)
        lp = utils.img2mse(rgb, GS.pose_target(name))
        lp.backward()
        out["pose/%s/loss" % name], out["pose/%s/grad_pose" % name], out["pose/%s/rgb" % name] = npy(lp), npy(pose.grad), npy(rgb)
        print(name, "pose: loss %.5f, |grad_pose| %.3e" % (float(lp), float(pose.grad.norm())))

if "llff" in sections:
    pc, pf = GS.params("bias")
    mc, mf = ref_model(pc), ref_model(pf)
    K, c2w = O.llff_intrinsics(12, 16, 13.0), O.llff_pose(3)
    r = renderer(N_importance=64, white_bkgd=False, ndc=True, near=0.0, far=1.0)
    with torch.no_grad():
        rgb, disp, acc, extras = r.render(12, 16, K, mc, mf, chunk=64, c2w=torch.Tensor(c2w[:3, :4]), retraw=True)
    out["llff/bias/K"], out["llff/bias/c2w"] = K, c2w
    out["llff/bias/rgb"], out["llff/bias/disp"], out["llff/bias/acc"] = npy(rgb), npy(disp), npy(acc)
    out["llff/bias/rgb0"] = npy(extras["rgb0"])
    print("llff bias: rgb range %.3f..%.3f, mean acc %.3f" % (float(rgb.min()), float(rgb.max()), float(acc.mean())))

if "trained" in sets:
    pc, pf = GS.params("trained")
    mc, mf = ref_model(pc), ref_model(pf)
    for sec, HW in (("frame64", 64), ("frame400", 400)):
        if sec not in sections:
            continue
        import time
        K = O.blender_intrinsics(HW, HW)
        c2w = torch.Tensor(O.pose_spherical(0.0, -30.0, 4.0)[:3, :4])
        r = renderer()
        t0 = time.time()
        with torch.no_grad():
            rgb, disp, acc, extras = r.render(HW, HW, K, mc, mf, chunk=32768, c2w=c2w, retraw=True)
        dt = time.time() - t0
        w, z = None, None
        truth = GS.sphere_truth(HW)
        mse = torch.mean((rgb - truth) ** 2)
        out["%s/rgb8" % sec] = (255 * np.clip(npy(rgb), 0, 1)).astype(np.uint8)      # utils.to8b (utils.py:30)
        out["%s/psnr_vs_truth" % sec] = np.float64((-10.0 * torch.log(mse) / np.log(10.0)).item())
        out["%s/sha256_rgb" % sec] = np.frombuffer(hashlib.sha256(np.ascontiguousarray(npy(rgb)).tobytes()).digest(), dtype=np.uint8)
        pix = GS.frame_subsample(HW)
        out["%s/sub/rgb" % sec] = npy(rgb.reshape(-1, 3)[pix])
        out["%s/sub/acc" % sec] = npy(acc.reshape(-1)[pix])
        out["%s/sub/disp" % sec] = npy(disp.reshape(-1)[pix])
        out["%s/sub/rgb0" % sec] = npy(extras["rgb0"].reshape(-1, 3)[pix])
        if HW == 64:
            out["%s/rgb" % sec], out["%s/acc" % sec] = npy(rgb), npy(acc)
        out["%s/reference_cpu_seconds" % sec] = np.float64(dt)
        print("%s: PSNR vs truth %.3f dB, mean acc %.3f, %.1f s on %d threads" %
              (sec, out["%s/psnr_vs_truth" % sec], float(acc.mean()), dt, torch.get_num_threads()))

np.savez_compressed(OUT, **out)
print("wrote %d arrays, %.1f KB" % (len(out), os.path.getsize(OUT) / 1024))
