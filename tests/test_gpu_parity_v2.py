"""-m gpu: tolerance parity on NON-DEGENERATE density (tests/golden/reference_golden_v2.npz, outputs of the unmodified
reference): per-pixel rgb / depth / acc of the tensor-core (bf16 operand) path against the fp32 reference with the
outliers counted and bounded, against the bf16-emulating oracle (tight), PSNR within 0.1 dB on full frames, a training
step with non-zero fine gradients and a pose step with a non-zero pose gradient.  Every statistic is printed with a
`PARITY` prefix (kept under profiles/)."""
import os

import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O
from tests import golden_sets as GS
from tests.gpu_util import T, cuda, maxabs

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BAR = 2e-3            # BASELINE.json north_star: rgb / depth maps within 2e-3 max-abs per pixel


@pytest.fixture(scope="module")
def g2():
    with np.load(os.path.join(ROOT, "tests", "golden", "reference_golden_v2.npz")) as z:
        return {k: z[k] for k in z.files}


def _models(name, precision="bf16"):
    from nerf_shared_b200 import nerf
    out = []
    for p in GS.params(name):
        m = nerf.NeRF(D=8, W=256, output_ch=5, skips=[4], use_viewdirs=True, multires=10, multires_views=4)
        m.load_state_dict(p, strict=True)
        m.precision = precision
        out.append(m.cuda())
    return out


def _renderer(**kw):
    from nerf_shared_b200 import render_utils
    base = dict(perturb=0.0, N_importance=128, N_samples=64, use_viewdirs=True, white_bkgd=True, raw_noise_std=0.0,
                near=2.0, far=6.0)
    base.update(kw)
    return render_utils.Renderer(**base)


def _stats(tag, ours, ref, bar=BAR):
    d = (ours.detach().float().cpu() - torch.as_tensor(ref).float()).abs()
    d = d.reshape(d.shape[0], -1).max(-1)[0] if d.dim() > 1 else d          # per pixel
    n_bad = int((d > bar).sum())
    print("PARITY %-44s max %.3e  median %.3e  pixels over %.0e: %d / %d" % (tag, float(d.max()), float(d.median()), bar, n_bad, d.numel()))
    return float(d.max()), n_bad, d.numel()


# outlier budgets of the bf16 path against the fp32 reference, as fractions of the pixels (measured values are in
# profiles/r2_parity.txt).  "stress" saturates alpha: the 1e10 last interval makes a pixel a step function of
# sign(sigma_last), so a few rays flip there whatever the operand precision.
OUTLIER_FRAC = {"stress": 0.10, "bias": 0.0, "trained": 0.06}
TIGHT = {"stress": 5e-3, "bias": 1e-4, "trained": 5e-3}        # vs the bf16-emulating oracle, max-abs rgb


@pytest.mark.parametrize("name", GS.available_sets())
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_render_rays_nondegenerate(g2, name, precision):
    """Renderer.render_rays (render_utils.py:67-174) on 256 rays, deterministic sampling."""
    mc, mf = _models(name, precision)
    rb = GS.ray_batch(name).cuda()
    with torch.no_grad():
        ret = _renderer().render_rays(rb, mc, mf, retraw=True, retweights=True)
    depth = (ret["weights"] * ret["z_vals"]).sum(-1)
    g = lambda k: g2["rr/%s/%s" % (name, k)]
    res = {}
    for k, ours in (("rgb_map", ret["rgb_map"]), ("rgb0", ret["rgb0"]), ("acc_map", ret["acc_map"]), ("depth", depth)):
        res[k] = _stats("render_rays[%s,%s] %s vs reference" % (name, precision, k), ours, g(k))
    n = rb.shape[0]
    if precision == "fp32":
        for k in ("rgb_map", "rgb0", "acc_map"):
            assert res[k][0] <= 5e-4, (name, k, res[k])
        assert res["depth"][0] <= 2e-3, (name, res["depth"])
        return
    # tight target: the oracle with the kernel's rounding points (bf16 operands, fp32 accumulation, fp32 heads)
    pc, pf = GS.params(name)
    rbc = rb.cpu()
    with torch.no_grad():
        emu = O.render_rays(pc, pf, rbc[:, 0:3], rbc[:, 3:6], rbc[:, 8:11], 2.0, 6.0, bf16=True)
    emu["depth"] = emu["depth_map"]
    for k, ours in (("rgb_map", ret["rgb_map"]), ("rgb0", ret["rgb0"]), ("acc_map", ret["acc_map"]), ("depth", depth)):
        t = _stats("render_rays[%s,bf16] %s vs bf16-emulating oracle" % (name, k), ours, emu[k], 1e-4)
        e = _stats("  (bf16-emulating oracle vs reference, %s)" % k, emu[k], g(k))
        assert t[0] <= TIGHT[name] * (20.0 if k == "depth" else 1.0), (name, k, t)
        # against the fp32 reference the kernel is in the same outlier class as the emulation of its arithmetic:
        # what exceeds the 2e-3 bar does so because of bf16 operands (north_star precision), not because of the kernel
        assert res[k][1] <= e[1] + max(2, int(0.02 * n)), (name, k, res[k], e)
        if name != "stress":
            assert res[k][1] <= OUTLIER_FRAC[name] * n, (name, k, res[k])
    # image-level: PSNR of ours against the reference render of the same rays
    mse = float(((ret["rgb_map"].cpu() - T(g("rgb_map"))) ** 2).mean())
    print("PARITY render_rays[%s,bf16] PSNR(ours, reference) = %.2f dB" % (name, -10 * np.log10(max(mse, 1e-20))))


@pytest.mark.parametrize("name", GS.available_sets())
def test_render_rays_random_mode_nondegenerate(g2, name):
    """perturb = 1, raw_noise_std = 1 with the reference's own draws (torch.manual_seed(4321), CPU generator)."""
    mc, mf = _models(name, "bf16")
    rb = GS.ray_batch(name)[:64].cuda()
    r = _renderer(perturb=1.0, raw_noise_std=1.0)
    torch.manual_seed(4321)
    draws = {"t_rand": torch.rand(64, 64), "noise_c": torch.randn(64, 64), "u": torch.rand(64, 128), "noise_f": torch.randn(64, 192)}
    r._uniform = lambda shape, device, what: draws[what].to(device)
    r._normal = lambda shape, device, what: draws[what].to(device)
    with torch.no_grad():
        ret = r.render_rays(rb, mc, mf, retraw=True, retweights=True)
    depth = (ret["weights"] * ret["z_vals"]).sum(-1)
    for k, ours in (("rgb_map", ret["rgb_map"]), ("rgb0", ret["rgb0"]), ("acc_map", ret["acc_map"]), ("depth", depth)):
        mx, bad, n = _stats("render_rays_rand[%s,bf16] %s vs reference" % (name, k), ours, g2["rr_rand/%s/%s" % (name, k)])
        # raw_noise_std = 1 puts every sample's sigma + noise near the relu / 1e10-interval knife edge
        assert bad <= max(2, (0.15 if name == "stress" else 0.05) * n), (name, k, mx, bad)


def _rel_cos(g, e):
    g, e = g.detach().float().cpu().reshape(-1), torch.as_tensor(e).float().reshape(-1)
    return float((g - e).norm() / (e.norm() + 1e-30)), float((g * e).sum() / (g.norm() * e.norm() + 1e-30))


@pytest.mark.parametrize("name", GS.available_sets())
def test_training_step_nondegenerate(g2, name):
    """main.py:77-104 on 64 rays with a model whose FINE gradients are non-zero: loss / PSNR against the reference,
    every parameter gradient against the reference's norm and against the oracle's fp32 autograd (pinned to the same
    goldens by tests/test_oracle_golden_v2.py), then Adam."""
    from nerf_shared_b200 import utils
    mc, mf = _models(name)
    sub = GS.ray_batch(name)[GS.TRAIN_ROWS].cuda()
    ro, rd = sub[:, 0:3].contiguous(), sub[:, 3:6].contiguous()
    tgt = GS.train_target(name).cuda()
    r = _renderer()
    r.train()
    opt = torch.optim.Adam(list(mc.parameters()) + list(mf.parameters()), lr=5e-4, betas=(0.9, 0.999))
    rgb, _, _, extras = r.render_from_rays(16, 16, GS.K16, chunk=32768, rays=torch.stack([ro, rd], 0), coarse_model=mc,
                                           fine_model=mf, retraw=True)
    opt.zero_grad()
    loss = utils.img2mse(rgb, tgt) + utils.img2mse(extras["rgb0"], tgt)
    loss.backward()
    ref_loss = float(g2["train/%s/loss" % name])
    psnr = float(utils.mse2psnr(utils.img2mse(rgb, tgt)))
    print("PARITY train[%s] loss %.6f (reference %.6f, rel %.2e)  PSNR %.3f dB (reference %.3f)" %
          (name, float(loss), ref_loss, abs(float(loss) - ref_loss) / ref_loss, psnr, float(g2["train/%s/psnr" % name])))
    assert abs(float(loss) - ref_loss) <= 2e-3 * ref_loss
    assert abs(psnr - float(g2["train/%s/psnr" % name])) <= 0.1          # BASELINE.json: PSNR within 0.1 dB
    pc, pf = GS.params(name)
    _, gc, gf, _ = O.train_step_grads(pc, pf, ro.cpu(), rd.cpu(), tgt.cpu(), 2.0, 6.0)
    # the same autograd with the kernel's rounding points (bf16 GEMM operands and bf16 activation gradients): what
    # bf16 operands alone do to each layer's gradient -- the class the kernels have to stay in
    _, ec, ef, _ = O.train_step_grads(pc, pf, ro.cpu(), rd.cpu(), tgt.cpu(), 2.0, 6.0, bf16=True)
    for tag, m, og, eg in (("coarse", mc, gc, ec), ("fine", mf, gf, ef)):
        named = dict(m.named_parameters())
        worst = {}
        for k, p in named.items():
            ref_n = float(g2["train/%s/%s/gnorm/%s" % (name, tag, k)])
            assert ref_n > 0 and float(p.grad.norm()) > 0, (name, tag, k)   # the fixture exercises every parameter
            rel, cos = _rel_cos(p.grad, og[k])
            rel_e, _ = _rel_cos(eg[k], og[k])
            rel_t, _ = _rel_cos(p.grad, eg[k])
            w = worst.setdefault(k.split(".")[0], [0.0, 1.0, 0.0, 0.0])
            w[0], w[1], w[2], w[3] = max(w[0], rel), min(w[1], cos), max(w[2], rel_e), max(w[3], rel_t)
        for grp, (rel, cos, rel_e, rel_t) in sorted(worst.items()):
            print("PARITY train[%s] %-6s %-15s grad rel L2 vs fp32 autograd %.4f (cos %.5f) | bf16-emulated autograd vs fp32 "
                  "%.4f | ours vs bf16-emulated %.4f" % (name, tag, grp, rel, cos, rel_e, rel_t))
            assert rel <= 1.5 * rel_e + 0.03 and cos > 0.9, (name, tag, grp, rel, cos, rel_e)
        flat = lambda d: torch.cat([d[k].detach().float().cpu().reshape(-1) for k in named])
        g_all, o_all, e_all = flat({k: p.grad for k, p in named.items()}), flat(og), flat(eg)
        rel, cos = _rel_cos(g_all, o_all)
        rel_e, cos_e = _rel_cos(e_all, o_all)
        dn = abs(float(g_all.norm()) - float(o_all.norm())) / float(o_all.norm())
        print("PARITY train[%s] %-6s whole gradient vs fp32 autograd: rel L2 %.4f cos %.5f |norm| off by %.4f "
              "(bf16-emulated autograd: rel L2 %.4f cos %.5f)" % (name, tag, rel, cos, dn, rel_e, cos_e))
        # (bf16 operands alone move the whole gradient by 1-14 % on these fixtures -- the emulation column; the kernels
        #  have to stay with the emulation, not beat it)
        assert rel <= 1.2 * rel_e + 0.01 and cos >= cos_e - 0.005 and dn < 0.05, (name, tag, rel, cos, dn, rel_e, cos_e)
    opt.step()
    for tag, m in (("coarse", mc), ("fine", mf)):
        # Adam's first step is lr * sign(g) where |g| >> eps: compare the update direction on the sigma head
        e = T(g2["train/%s/%s/after_adam/alpha_w" % (name, tag)])
        before = GS.params(name)[0 if tag == "coarse" else 1]["alpha_linear.weight"]
        du, de = m.alpha_linear.weight.detach().cpu() - before, e - before
        big = de.abs() > 4e-4
        agree = float((torch.sign(du[big]) == torch.sign(de[big])).float().mean()) if bool(big.any()) else 1.0
        print("PARITY train[%s] %s Adam step on alpha_linear.weight: sign agreement %.4f on %d entries, max|d| %.2e" %
              (name, tag, agree, int(big.sum()), maxabs(m.alpha_linear.weight, e)))
        assert agree >= 0.95 and maxabs(m.alpha_linear.weight, e) <= 1.05e-3


@pytest.mark.parametrize("name", GS.available_sets())
def test_pose_step_nondegenerate(g2, name):
    """demo_est_rel_pose.py:80-98: gradient of the photometric loss with respect to the 4x4 pose, frozen weights."""
    from nerf_shared_b200 import utils
    mc, mf = _models(name)
    for p in list(mc.parameters()) + list(mf.parameters()):
        p.requires_grad_(False)
    pose = cuda(GS.pose_of(name)).clone().requires_grad_(True)
    ro, rd = utils.get_rays(16, 16, GS.K16, pose)
    ys, xs = GS.POSE_SEL // 16, GS.POSE_SEL % 16
    rgb = _renderer().render_from_rays(16, 16, GS.K16, chunk=32768, rays=torch.stack([ro[ys, xs], rd[ys, xs]], 0),
                                       coarse_model=mc, fine_model=mf, retraw=True)[0]
    loss = utils.img2mse(rgb, GS.pose_target(name).cuda())
    loss.backward()
    e = T(g2["pose/%s/grad_pose" % name])
    rel, cos = _rel_cos(pose.grad, e)
    ref_loss = float(g2["pose/%s/loss" % name])
    # bf16-emulated autograd of the same step on the CPU: the deviation bf16 operands alone produce
    pc, pf = GS.params(name)
    pe = torch.from_numpy(GS.pose_of(name)).clone().requires_grad_(True)
    eo, ed = O.get_rays(16, 16, GS.K16, pe)
    eo, ed = eo[ys, xs], ed[ys, xs]
    out = O.render_rays(pc, pf, eo, ed, O.unit_viewdirs(ed), 2.0, 6.0, bf16=True)
    O.mse(out["rgb_map"], GS.pose_target(name)).backward()
    rel_e, cos_e = _rel_cos(pe.grad, e)
    rel_t, cos_t = _rel_cos(pose.grad, pe.grad)
    print("PARITY pose[%s] loss %.6f (reference %.6f)  grad_pose vs reference: rel L2 %.4f cos %.5f | bf16-emulated autograd vs "
          "reference: %.4f / %.5f | ours vs bf16-emulated: %.4f / %.5f  |reference grad| %.3e" %
          (name, float(loss), ref_loss, rel, cos, rel_e, cos_e, rel_t, cos_t, float(e.norm())))
    _stats("pose[%s] rgb vs reference" % name, rgb, g2["pose/%s/rgb" % name])
    assert float(e.norm()) > 1e-3
    assert abs(float(loss) - ref_loss) <= 2e-3 * ref_loss
    assert rel <= 1.5 * rel_e + 0.05 and cos > 0.97, (name, rel, cos, rel_e)


def test_llff_ndc_render_nondegenerate(g2):
    mc, mf = _models("bias")
    r = _renderer(N_importance=64, white_bkgd=False, ndc=True, near=0.0, far=1.0)
    with torch.no_grad():
        rgb, disp, acc, extras = r.render(12, 16, g2["llff/bias/K"], mc, mf, chunk=64, c2w=cuda(g2["llff/bias/c2w"])[:3, :4])
    for k, ours in (("rgb", rgb), ("rgb0", extras["rgb0"]), ("acc", acc)):
        mx, bad, n = _stats("llff_ndc[bias,bf16] %s vs reference" % k, ours.reshape(-1, *ours.shape[2:]), g2["llff/bias/" + k].reshape(-1, *ours.shape[2:]))
        assert bad == 0, (k, mx, bad)


@pytest.mark.parametrize("HW", [64, 400])
def test_trained_model_full_frame(g2, HW):
    """A trained model (real surface) rendered as a full frame -- 64 x 64 held-out view and the 400 x 400 frame of
    BASELINE.json configs[0] -- against what the unmodified reference rendered with the same weights: 8-bit image,
    2,048 sampled pixels in fp32, and PSNR against the analytic ground truth within 0.1 dB of the reference's."""
    if "trained" not in GS.available_sets() or "frame%d/rgb8" % HW not in g2:
        pytest.skip("trained-weights goldens not generated")
    sec = "frame%d" % HW
    mc, mf = _models("trained")
    K = O.blender_intrinsics(HW, HW)
    c2w = cuda(O.pose_spherical(0.0, -30.0, 4.0))[:3, :4]
    with torch.no_grad():
        rgb, disp, acc, extras = _renderer().render(HW, HW, K, mc, mf, chunk=32768, c2w=c2w, retraw=False)
    truth = GS.sphere_truth(HW)
    psnr = float(-10.0 * torch.log10(torch.mean((rgb.cpu() - truth) ** 2)))
    ref_psnr = float(g2[sec + "/psnr_vs_truth"])
    print("PARITY %s[trained,bf16] PSNR vs ground truth %.4f dB (reference %.4f dB, delta %.4f)" % (sec, psnr, ref_psnr, psnr - ref_psnr))
    assert abs(psnr - ref_psnr) <= 0.1
    pix = GS.frame_subsample(HW).cuda()
    n = pix.numel()
    for k, ours in (("rgb", rgb.reshape(-1, 3)[pix]), ("rgb0", extras["rgb0"].reshape(-1, 3)[pix]), ("acc", acc.reshape(-1)[pix])):
        mx, bad, _ = _stats("%s[trained,bf16] %s (sampled pixels) vs reference" % (sec, k), ours, g2["%s/sub/%s" % (sec, k)])
        assert bad <= OUTLIER_FRAC["trained"] * n, (k, mx, bad)
    from nerf_shared_b200 import utils
    d8 = (utils.to8b_device(rgb).cpu().numpy().astype(np.int16) - g2[sec + "/rgb8"].astype(np.int16))
    off = np.abs(d8).max(-1)
    print("PARITY %s[trained,bf16] 8-bit frame: %.3f%% of pixels differ by more than one count (max %d)" %
          (sec, 100.0 * float((off > 1).mean()), int(off.max())))
    assert float((off > 1).mean()) <= 0.01
    assert float(np.ptp(g2[sec + "/rgb8"])) > 100          # the fixture is a real image
