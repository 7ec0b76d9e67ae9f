"""Generate tests/golden/*.npz by running the UNMODIFIED reference (/root/reference).

Run in the build container only:  python tests/golden/make_golden.py
The outputs are small and committed; the GPU box never needs /root/reference.
Every array written here comes out of reference code (nerf_shared.{nerf,render_utils,utils});
inputs are produced by oracle.nerf_oracle's seeded generators (numpy PCG64, stable bits).
"""
import hashlib
import json
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


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(npy(a)).tobytes()).hexdigest()


def ref_model(params):
    m = nerf.NeRF(D=8, W=256, output_ch=5, skips=[4], use_viewdirs=True, multires=10, multires_views=4)
    missing = m.load_state_dict(params, strict=True)
    return m


def quiet_renderer(**kw):
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        return render_utils.Renderer(**kw)


out = {}
meta = {"torch": torch.__version__, "numpy": np.__version__, "sha256": {}}

# ------------------------------------------------------------------ rays
cases = {
    "blender16": (16, 16, O.blender_intrinsics(16, 16), O.pose_spherical(40.0, -30.0, 4.0)),
    "blender_13x21": (13, 21, O.blender_intrinsics(13, 21), O.pose_spherical(-117.0, -30.0, 4.0)),
    "llff_12x16": (12, 16, O.llff_intrinsics(12, 16, 407.5625 / 31.5), O.llff_pose(0)),
}
for name, (H, W, K, c2w) in cases.items():
    ro, rd = utils.get_rays(H, W, K, torch.Tensor(c2w[:3, :4]))
    out["rays/%s/K" % name] = K
    out["rays/%s/c2w" % name] = c2w
    out["rays/%s/rays_o" % name] = npy(ro.contiguous())
    out["rays/%s/rays_d" % name] = npy(rd)
    if name.startswith("llff"):
        no, nd = utils.ndc_rays(H, W, K[0][0], 1.0, ro, rd)
        out["rays/%s/ndc_o" % name] = npy(no)
        out["rays/%s/ndc_d" % name] = npy(nd)

# full-size checksums (BASELINE.json configs 1, 4, 5)
for name, (H, W, K, c2w) in {
    "C1_400": (400, 400, O.blender_intrinsics(400, 400), O.pose_spherical(40.0, -30.0, 4.0)),
    "C4_800": (800, 800, O.blender_intrinsics(800, 800), O.pose_spherical(40.0, -30.0, 4.0)),
    "C5_llff": (378, 504, O.llff_intrinsics(), O.llff_pose(0)),
}.items():
    ro, rd = utils.get_rays(H, W, K, torch.Tensor(c2w[:3, :4]))
    meta["sha256"]["rays_d/" + name] = sha(rd)
    meta["sha256"]["rays_o/" + name] = sha(ro.contiguous())
    if name == "C5_llff":
        no, nd = utils.ndc_rays(H, W, K[0][0], 1.0, ro, rd)
        meta["sha256"]["ndc_o/" + name] = sha(no)
        meta["sha256"]["ndc_d/" + name] = sha(nd)

# ------------------------------------------------------------------ encoder
rng = np.random.default_rng(5)
x = torch.from_numpy(rng.uniform(-4.5, 4.5, size=(64, 3)).astype(np.float32))
for L in (10, 4):
    fn, dim = nerf.get_embedder(L, 0)
    out["embed/L%d/x" % L] = npy(x)
    out["embed/L%d/y" % L] = npy(fn(x))
    assert dim == 3 + 6 * L

# ------------------------------------------------------------------ MLP forward
params_c = O.make_params(11)
params_f = O.make_params(12)
params_stress = O.make_params(13, scale=1.59)
model_c, model_f = ref_model(params_c), ref_model(params_f)
pts = torch.from_numpy(rng.uniform(-3.0, 3.0, size=(6, 16, 3)).astype(np.float32))
vd = torch.from_numpy(rng.normal(size=(6, 3)).astype(np.float32))
vd = vd / vd.norm(dim=-1, keepdim=True)
with torch.no_grad():
    out["mlp/pts"] = npy(pts)
    out["mlp/viewdirs"] = npy(vd)
    out["mlp/raw_seed11"] = npy(model_c(pts, vd))
    out["mlp/raw_seed13x1.59"] = npy(ref_model(params_stress)(pts, vd))
    out["mlp/density_seed11"] = npy(model_c.get_density(pts))
meta["sha256"]["params_seed11"] = sha(np.concatenate([npy(v).ravel() for v in params_c.values()]))
meta["sha256"]["params_seed12"] = sha(np.concatenate([npy(v).ravel() for v in params_f.values()]))

# ------------------------------------------------------------------ stratified z + compositing + sample_pdf, via render_rays
H = W = 8
K = O.blender_intrinsics(H, W)
c2w = torch.Tensor(O.pose_spherical(75.0, -30.0, 4.0)[:3, :4])
ro, rd = utils.get_rays(H, W, K, c2w)
ro, rd = ro.reshape(-1, 3)[:24].contiguous(), rd.reshape(-1, 3)[:24].contiguous()
vdirs = rd / torch.norm(rd, dim=-1, keepdim=True)
N = ro.shape[0]
ray_batch = torch.cat([ro, rd, 2.0 * torch.ones(N, 1), 6.0 * torch.ones(N, 1), vdirs], -1)
out["rr/ray_batch"] = npy(ray_batch)

for tag, kw in {
    "det": dict(perturb=0.0, raw_noise_std=0.0, white_bkgd=True, lindisp=False),
    "rand": dict(perturb=1.0, raw_noise_std=1.0, white_bkgd=True, lindisp=False),
    "rand_lindisp_black": dict(perturb=1.0, raw_noise_std=0.0, white_bkgd=False, lindisp=True),
}.items():
    r = quiet_renderer(N_importance=128, N_samples=64, use_viewdirs=True, near=2.0, far=6.0, **kw)
    torch.manual_seed(1234)
    with torch.no_grad():
        ret = r.render_rays(ray_batch, model_c, model_f, retraw=True, retweights=True)
    for k, v in ret.items():
        out["rr/%s/%s" % (tag, k)] = npy(v)

# stress weights (saturating alpha), deterministic sampling
r = quiet_renderer(perturb=0.0, N_importance=128, N_samples=64, use_viewdirs=True, white_bkgd=True,
                   raw_noise_std=0.0, near=2.0, far=6.0)
ms = ref_model(params_stress)
with torch.no_grad():
    ret = r.render_rays(ray_batch, ms, ms, retraw=True, retweights=True)
for k, v in ret.items():
    out["rr/stress/%s" % k] = npy(v)

# raw2outputs in isolation on a synthetic raw with large |sigma| (covers saturation, acc==0 -> nan disp)
raw = torch.from_numpy((rng.normal(size=(12, 64, 4)) * np.array([1, 1, 1, 3.0])).astype(np.float32))
raw[0, :, 3] = -1.0  # a ray with zero opacity everywhere: acc = 0, disp = nan (render_utils.py:284)
zz = torch.sort(torch.from_numpy(rng.uniform(2, 6, size=(12, 64)).astype(np.float32)), -1)[0]
dd = torch.from_numpy(rng.normal(size=(12, 3)).astype(np.float32))
for wb in (True, False):
    r = quiet_renderer(white_bkgd=wb)
    res = r.raw2outputs(raw, zz, dd)
    for nm, v in zip(["rgb", "disp", "acc", "weights", "depth"], res):
        out["r2o/wb%d/%s" % (int(wb), nm)] = npy(v)
out["r2o/raw"], out["r2o/z"], out["r2o/rays_d"] = npy(raw), npy(zz), npy(dd)

# sample_pdf in isolation: det and random u, with the cdf the reference built
bins = torch.sort(torch.from_numpy(rng.uniform(2, 6, size=(10, 63)).astype(np.float32)), -1)[0]
wts = torch.from_numpy(rng.uniform(0, 1, size=(10, 62)).astype(np.float32) ** 4)
wts[3] = 0.0  # all-zero weights: pdf is uniform through the +1e-5 (utils.py:76)
out["pdf/bins"], out["pdf/weights"] = npy(bins), npy(wts)
out["pdf/det128"] = npy(utils.sample_pdf(bins, wts, 128, det=True))
out["pdf/det64"] = npy(utils.sample_pdf(bins, wts, 64, det=True))
torch.manual_seed(77)
out["pdf/rand128"] = npy(utils.sample_pdf(bins, wts, 128, det=False))
torch.manual_seed(77)
out["pdf/rand128_u"] = npy(torch.rand(10, 128))
# the reference's own cdf / inds (recomputed with the reference's expressions, utils.py:76-79,101)
w_ = wts + 1e-5
pdf_ = w_ / torch.sum(w_, -1, keepdim=True)
cdf_ = torch.cat([torch.zeros_like(pdf_[..., :1]), torch.cumsum(pdf_, -1)], -1)
out["pdf/cdf"] = npy(cdf_)
out["pdf/inds_det128"] = npy(torch.searchsorted(cdf_, torch.linspace(0., 1., 128).expand(10, 128).contiguous(), right=True))
out["pdf/inds_rand128"] = npy(torch.searchsorted(cdf_, torch.from_numpy(out["pdf/rand128_u"]), right=True))

# ------------------------------------------------------------------ Renderer.render end to end (c2w given)
for tag, (H, W, K, c2w, kw) in {
    "blender": (6, 7, O.blender_intrinsics(6, 7), O.pose_spherical(10.0, -30.0, 4.0),
                dict(perturb=0.0, N_importance=128, white_bkgd=True, ndc=False, near=2.0, far=6.0)),
    "llff_ndc": (6, 8, O.llff_intrinsics(6, 8, 6.5), O.llff_pose(3),
                 dict(perturb=0.0, N_importance=64, white_bkgd=False, ndc=True, near=0.0, far=1.0)),
}.items():
    r = quiet_renderer(N_samples=64, use_viewdirs=True, raw_noise_std=0.0, lindisp=False, **kw)
    with torch.no_grad():
        rgb, disp, acc, extras = r.render(H, W, K, model_c, model_f, chunk=32, c2w=torch.Tensor(c2w[:3, :4]), retraw=True)
    out["render/%s/K" % tag], out["render/%s/c2w" % tag] = K, c2w
    out["render/%s/rgb" % tag], out["render/%s/disp" % tag], out["render/%s/acc" % tag] = npy(rgb), npy(disp), npy(acc)
    for k in ("rgb0", "disp0", "acc0", "z_std"):
        out["render/%s/%s" % (tag, k)] = npy(extras[k])

# ------------------------------------------------------------------ training step (main.py:77-104)
r = quiet_renderer(perturb=0.0, N_importance=128, N_samples=64, use_viewdirs=True, white_bkgd=True,
                   raw_noise_std=0.0, near=2.0, far=6.0)
r.train()
mc, mf = ref_model(params_c), ref_model(params_f)
tgt = torch.from_numpy(np.random.default_rng(1).uniform(0, 1, size=(N, 3)).astype(np.float32))
opt = torch.optim.Adam(list(mc.parameters()) + list(mf.parameters()), lr=5e-4, betas=(0.9, 0.999))
Kb = O.blender_intrinsics(8, 8)
rgb, _, _, extras = r.render_from_rays(8, 8, Kb, chunk=32768, rays=torch.stack([ro, rd], 0), coarse_model=mc,
                                       fine_model=mf, retraw=True)
opt.zero_grad()
loss = utils.img2mse(rgb, tgt) + utils.img2mse(extras["rgb0"], tgt)
loss.backward()
out["train/target"] = npy(tgt)
out["train/loss"] = npy(loss)
out["train/psnr"] = npy(utils.mse2psnr(utils.img2mse(rgb, tgt)))
for tag, m in (("coarse", mc), ("fine", mf)):
    for k, p in m.named_parameters():
        g = p.grad
        out["train/%s/gnorm/%s" % (tag, k)] = np.float64(g.double().norm().item())
        out["train/%s/ghead/%s" % (tag, k)] = npy(g.reshape(-1)[:64])
    # two full small/medium gradients for a dense check
    out["train/%s/g_alpha_w" % tag] = npy(m.alpha_linear.weight.grad)
    out["train/%s/g_rgb_w" % tag] = npy(m.rgb_linear.weight.grad)
    out["train/%s/g_l0_w" % tag] = npy(m.pts_linears[0].weight.grad)
opt.step()
for tag, m in (("coarse", mc), ("fine", mf)):
    out["train/%s/after_adam/alpha_w" % tag] = npy(m.alpha_linear.weight)
    out["train/%s/after_adam/l0_b" % tag] = npy(m.pts_linears[0].bias)

# pose-gradient step (demo_est_rel_pose.py:80-98): gradient wrt c2w through get_rays + render, frozen weights
for p in list(mc.parameters()) + list(mf.parameters()):
    p.requires_grad_(False)
Hp, Wp = 8, 8
pose = torch.Tensor(O.pose_spherical(75.0, -30.0, 4.0)).clone().requires_grad_(True)
ro_p, rd_p = utils.get_rays(Hp, Wp, Kb, pose)
sel = np.random.default_rng(2).choice(Hp * Wp, size=16, replace=False)
ys, xs = sel // Wp, sel % Wp
batch_rays = torch.stack([ro_p[ys, xs], rd_p[ys, xs]], 0)
rgb, _, _, _ = r.render_from_rays(Hp, Wp, Kb, chunk=32768, rays=batch_rays, coarse_model=mc, fine_model=mf, retraw=True)
tgt_p = torch.from_numpy(np.random.default_rng(3).uniform(0, 1, size=(16, 3)).astype(np.float32))
lp = utils.img2mse(rgb, tgt_p)
lp.backward()
out["pose/pose"], out["pose/sel"], out["pose/target"] = npy(pose), sel.astype(np.int64), npy(tgt_p)
out["pose/loss"], out["pose/grad_pose"], out["pose/rgb"] = npy(lp), npy(pose.grad), npy(rgb)

np.savez_compressed(os.path.join(HERE, "reference_golden.npz"), **out)
with open(os.path.join(HERE, "reference_golden_meta.json"), "w") as f:
    json.dump(meta, f, indent=1, sort_keys=True)
sz = os.path.getsize(os.path.join(HERE, "reference_golden.npz"))
print("wrote %d arrays, %.1f KB" % (len(out), sz / 1024))
