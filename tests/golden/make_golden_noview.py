"""Generate tests/golden/reference_noview_golden.npz by running the UNMODIFIED reference (/root/reference)
with use_viewdirs=False (nerf.py:93-94,131-132; render_utils.py:205,225: ray rows of 8 columns), plus the model of
configs/stonehenge.txt (multires 15 / 6).

Run in the build container only:  python tests/golden/make_golden_noview.py
Every array written here comes out of reference code; inputs come from oracle.nerf_oracle's seeded generators.
"""
import contextlib
import io
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

torch.set_num_threads(8)


def npy(t):
    return t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t)


def ref_model(params, output_ch):
    m = nerf.NeRF(D=8, W=256, output_ch=output_ch, skips=[4], use_viewdirs=False, multires=10, multires_views=4)
    m.load_state_dict(params, strict=True)
    return m


out = {}
rng = np.random.default_rng(21)
# output_ch = 5 is what create_nerf_models builds when N_importance > 0 (utils.py:122); 4 is the constructor default
p5c, p5f, p4 = O.make_params(31, output_ch=5), O.make_params(32, output_ch=5), O.make_params(33, output_ch=4)
pts = torch.from_numpy(rng.uniform(-3.0, 3.0, size=(5, 24, 3)).astype(np.float32))
with torch.no_grad():
    out["mlp/pts"] = npy(pts)
    out["mlp/raw5_seed31"] = npy(ref_model(p5c, 5)(pts, None))
    out["mlp/raw4_seed33"] = npy(ref_model(p4, 4)(pts, None))

# configs/stonehenge.txt: multires = 15, multires_views = 6 (93- and 39-wide encodings), use_viewdirs=True
p156 = O.make_params(41, input_ch=93, input_ch_views=39)
m156 = nerf.NeRF(D=8, W=256, output_ch=5, skips=[4], use_viewdirs=True, multires=15, multires_views=6)
m156.load_state_dict(p156, strict=True)
pts156 = torch.from_numpy(rng.uniform(-1.5, 1.5, size=(5, 24, 3)).astype(np.float32))
vd156 = torch.nn.functional.normalize(torch.from_numpy(rng.normal(size=(5, 3)).astype(np.float32)), dim=-1)
with torch.no_grad():
    out["mlp156/pts"], out["mlp156/viewdirs"] = npy(pts156), npy(vd156)
    out["mlp156/raw_seed41"] = npy(m156(pts156, vd156))

for tag, (H, W, K, c2w, kw) in {
    "blender": (6, 7, O.blender_intrinsics(6, 7), O.pose_spherical(10.0, -30.0, 4.0),
                dict(perturb=0.0, N_importance=128, white_bkgd=True, ndc=False, near=2.0, far=6.0)),
    "llff_ndc": (6, 8, O.llff_intrinsics(6, 8, 6.5), O.llff_pose(3),
                 dict(perturb=0.0, N_importance=64, white_bkgd=False, ndc=True, near=0.0, far=1.0)),
}.items():
    with contextlib.redirect_stdout(io.StringIO()):
        r = render_utils.Renderer(N_samples=64, use_viewdirs=False, raw_noise_std=0.0, lindisp=False, **kw)
    with torch.no_grad():
        rgb, disp, acc, extras = r.render(H, W, K, ref_model(p5c, 5), ref_model(p5f, 5), chunk=32,
                                          c2w=torch.Tensor(c2w[:3, :4]), retraw=True)
    out["render/%s/K" % tag], out["render/%s/c2w" % tag] = K, c2w
    out["render/%s/rgb" % tag], out["render/%s/disp" % tag], out["render/%s/acc" % tag] = npy(rgb), npy(disp), npy(acc)
    for k in ("rgb0", "disp0", "acc0", "z_std", "raw"):
        out["render/%s/%s" % (tag, k)] = npy(extras[k])

# ---- parameter gradients of both variants (the reference's autograd): d/dtheta sum(raw * G) with a fixed cotangent G,
#      and a full training step of the use_viewdirs=False pair through the Renderer (main.py:77-103)
def grad_section(tag, model, pts_, vd_, gen_seed):
    G = torch.from_numpy(np.random.default_rng(gen_seed).normal(size=tuple(pts_.shape[:-1]) + (model(pts_, vd_).shape[-1],)).astype(np.float32))
    model.zero_grad()
    (model(pts_, vd_) * G).sum().backward()
    out["grad/%s/G" % tag] = npy(G)
    for k, p in model.named_parameters():
        if p.grad is None:
            continue
        out["grad/%s/gnorm/%s" % (tag, k)] = np.float64(p.grad.double().norm().item())
        out["grad/%s/ghead/%s" % (tag, k)] = npy(p.grad.reshape(-1)[:512])


grad_section("noview5", ref_model(p5c, 5), pts, None, 51)
grad_section("wide156", m156, pts156, vd156, 52)

with contextlib.redirect_stdout(io.StringIO()):
    r = render_utils.Renderer(perturb=0.0, N_importance=128, N_samples=64, use_viewdirs=False, white_bkgd=True,
                              raw_noise_std=0.0, near=2.0, far=6.0)
r.train()
# (x 1.59 matrices: positive density on a fraction of the samples, non-zero fine gradients)
p5c_s, p5f_s = O.make_params(31, output_ch=5, scale=1.59), O.make_params(32, output_ch=5, scale=1.59)
mc, mf = ref_model(p5c_s, 5), ref_model(p5f_s, 5)
H = W = 8
K = O.blender_intrinsics(H, W)
ro, rd = utils.get_rays(H, W, K, torch.Tensor(O.pose_spherical(75.0, -30.0, 4.0)[:3, :4]))
ro, rd = ro.reshape(-1, 3)[:32].contiguous(), rd.reshape(-1, 3)[:32].contiguous()
tgt = torch.from_numpy(np.random.default_rng(53).uniform(0, 1, size=(32, 3)).astype(np.float32))
rgb, _, _, extras = r.render_from_rays(H, W, K, chunk=32768, rays=torch.stack([ro, rd], 0), coarse_model=mc, fine_model=mf, retraw=True)
loss = utils.img2mse(rgb, tgt) + utils.img2mse(extras["rgb0"], tgt)
loss.backward()
out["train_noview/rays_o"], out["train_noview/rays_d"], out["train_noview/target"] = npy(ro), npy(rd), npy(tgt)
out["train_noview/loss"] = npy(loss)
for tag, m in (("coarse", mc), ("fine", mf)):
    for k, p in m.named_parameters():
        if p.grad is not None:
            out["train_noview/%s/gnorm/%s" % (tag, k)] = np.float64(p.grad.double().norm().item())
print("noview train: loss %.5f, |g fine output_linear.weight| %.3e" % (float(loss), out["train_noview/fine/gnorm/output_linear.weight"]))

np.savez_compressed(os.path.join(HERE, "reference_noview_golden.npz"), **out)
print("wrote", len(out), "arrays,", os.path.getsize(os.path.join(HERE, "reference_noview_golden.npz")), "bytes")
