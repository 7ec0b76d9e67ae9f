"""Pins oracle/nerf_oracle.py to the outputs of the unmodified reference (tests/golden)."""
import hashlib

import numpy as np
import torch

from oracle import nerf_oracle as O


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def sha(t):
    return hashlib.sha256(np.ascontiguousarray(t.detach().numpy()).tobytes()).hexdigest()


def bits_equal(a, b):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b.detach().numpy() if torch.is_tensor(b) else b)
    return a.shape == b.shape and a.tobytes() == b.tobytes()


def test_get_rays_bit_exact(golden):
    for name in ("blender16", "blender_13x21", "llff_12x16"):
        K, c2w = golden["rays/%s/K" % name], golden["rays/%s/c2w" % name]
        H, W = golden["rays/%s/rays_d" % name].shape[:2]
        ro, rd = O.get_rays(H, W, K, T(c2w)[:3, :4])
        assert bits_equal(golden["rays/%s/rays_d" % name], rd)
        assert bits_equal(golden["rays/%s/rays_o" % name], ro.contiguous())


def test_ndc_bit_exact(golden):
    name = "llff_12x16"
    K = golden["rays/%s/K" % name]
    ro, rd = T(golden["rays/%s/rays_o" % name]), T(golden["rays/%s/rays_d" % name])
    H, W = rd.shape[:2]
    no, nd = O.ndc_rays(H, W, K[0][0], 1.0, ro, rd)
    assert bits_equal(golden["rays/%s/ndc_o" % name], no)
    assert bits_equal(golden["rays/%s/ndc_d" % name], nd)


def test_full_size_ray_checksums(golden_meta):
    s = golden_meta["sha256"]
    c2w = T(O.pose_spherical(40.0, -30.0, 4.0))[:3, :4]
    for name, H in (("C1_400", 400), ("C4_800", 800)):
        ro, rd = O.get_rays(H, H, O.blender_intrinsics(H, H), c2w)
        assert sha(rd) == s["rays_d/" + name]
        assert sha(ro.contiguous()) == s["rays_o/" + name]
    K = O.llff_intrinsics()
    ro, rd = O.get_rays(378, 504, K, T(O.llff_pose(0))[:3, :4])
    assert sha(rd) == s["rays_d/C5_llff"]
    no, nd = O.ndc_rays(378, 504, K[0][0], 1.0, ro, rd)
    assert sha(no) == s["ndc_o/C5_llff"] and sha(nd) == s["ndc_d/C5_llff"]


def test_params_stable(golden_meta):
    for seed in (11, 12):
        p = O.make_params(seed)
        assert sha(torch.cat([v.reshape(-1) for v in p.values()])) == golden_meta["sha256"]["params_seed%d" % seed]
    n = sum(v.numel() for v in O.make_params(11).values())
    assert n == 595844


def test_positional_encoding_bit_exact(golden):
    for L in (10, 4):
        y = O.positional_encoding(T(golden["embed/L%d/x" % L]), L)
        assert bits_equal(golden["embed/L%d/y" % L], y)


def test_mlp_forward(golden):
    pts, vd = T(golden["mlp/pts"]), T(golden["mlp/viewdirs"])
    raw = O.nerf_query(O.make_params(11), pts, vd)
    np.testing.assert_allclose(raw.numpy(), golden["mlp/raw_seed11"], rtol=0, atol=2e-6)
    raw = O.nerf_query(O.make_params(13, scale=1.59), pts, vd)
    np.testing.assert_allclose(raw.numpy(), golden["mlp/raw_seed13x1.59"], rtol=0, atol=2e-5)
    dens = O.nerf_query(O.make_params(11), pts, torch.ones(pts.shape[0], 3))[..., -1]
    np.testing.assert_allclose(dens.numpy(), golden["mlp/density_seed11"], rtol=0, atol=2e-6)


def test_mlp_bf16_emulation_is_close(golden):
    pts, vd = T(golden["mlp/pts"]), T(golden["mlp/viewdirs"])
    raw = O.nerf_query(O.make_params(11), pts, vd, bf16=True)
    assert np.abs(raw.numpy() - golden["mlp/raw_seed11"]).max() < 2e-2


def _draws(tag, N, S, I, noise_std):
    """Replay the reference's torch RNG consumption order (render_utils.py:121,264; utils.py:86)."""
    torch.manual_seed(1234)
    t_rand = torch.rand(N, S)
    noise_c = torch.randn(N, S) * noise_std if noise_std > 0 else None
    u = torch.rand(N, I)
    noise_f = torch.randn(N, S + I) * noise_std if noise_std > 0 else None
    return t_rand, noise_c, u, noise_f


def test_render_rays_stage_by_stage(golden):
    rb = T(golden["rr/ray_batch"])
    ro, rd, near, far, vd = rb[:, 0:3], rb[:, 3:6], rb[:, 6:7], rb[:, 7:8], rb[:, 8:11]
    pc, pf = O.make_params(11), O.make_params(12)
    N = rb.shape[0]
    for tag, kw in {
        "det": dict(white_bkgd=True, lindisp=False),
        "rand": dict(white_bkgd=True, lindisp=False),
        "rand_lindisp_black": dict(white_bkgd=False, lindisp=True),
    }.items():
        if tag == "det":
            t_rand = noise_c = u = noise_f = None
        else:
            t_rand, noise_c, u, noise_f = _draws(tag, N, 64, 128, 1.0 if tag == "rand" else 0.0)
        out = O.render_rays(pc, pf, ro, rd, vd, near, far, t_rand=t_rand, u=u, noise_c=noise_c, noise_f=noise_f, **kw)
        g = lambda k: golden["rr/%s/%s" % (tag, k)]
        # sample depths follow the reference op by op: the merged z only differs through the
        # coarse weights (an ATen GEMM result), so compare with a tight tolerance
        np.testing.assert_allclose(out["z_vals"].numpy(), g("z_vals"), rtol=0, atol=1e-5)
        np.testing.assert_allclose(out["raw"].numpy(), g("raw"), rtol=0, atol=2e-5)
        np.testing.assert_allclose(out["weights"].numpy(), g("weights"), rtol=0, atol=1e-5)
        for k in ("rgb_map", "acc_map", "rgb0", "acc0", "z_std"):
            np.testing.assert_allclose(out[k].numpy(), g(k), rtol=0, atol=1e-5, err_msg=tag + k)
        for k in ("disp_map", "disp0"):
            np.testing.assert_allclose(out[k].numpy(), g(k), rtol=1e-4, atol=1e-5, err_msg=tag + k)


def test_stratified_z_bit_exact(golden):
    """z of the coarse pass does not depend on the MLP: recover it from the reference run by
    feeding the reference's own draws (bit-exact target of BASELINE.json)."""
    rb = T(golden["rr/ray_batch"])
    near, far = rb[:, 6:7], rb[:, 7:8]
    t_rand, _, _, _ = _draws("rand", rb.shape[0], 64, 128, 1.0)
    z = O.stratified_z(near, far, 64, t_rand, False)
    # the reference returns only the merged z; its coarse samples are a subset of it
    merged = golden["rr/rand/z_vals"]
    for r in range(z.shape[0]):
        assert np.isin(z[r].numpy(), merged[r]).all()
    z = O.stratified_z(near, far, 64, None, False)
    merged = golden["rr/det/z_vals"]
    for r in range(z.shape[0]):
        assert np.isin(z[r].numpy(), merged[r]).all()


def test_stress_weights(golden):
    rb = T(golden["rr/ray_batch"])
    ro, rd, near, far, vd = rb[:, 0:3], rb[:, 3:6], rb[:, 6:7], rb[:, 7:8], rb[:, 8:11]
    ps = O.make_params(13, scale=1.59)
    out = O.render_rays(ps, ps, ro, rd, vd, near, far)
    np.testing.assert_allclose(out["rgb_map"].numpy(), golden["rr/stress/rgb_map"], rtol=0, atol=2e-5)
    np.testing.assert_allclose(out["acc_map"].numpy(), golden["rr/stress/acc_map"], rtol=0, atol=2e-5)


def test_composite_isolated(golden):
    raw, z, d = T(golden["r2o/raw"]), T(golden["r2o/z"]), T(golden["r2o/rays_d"])
    for wb in (1, 0):
        res = O.composite(raw, z, d, bool(wb))
        for nm, v in zip(["rgb", "disp", "acc", "weights", "depth"], res):
            ref = golden["r2o/wb%d/%s" % (wb, nm)]
            assert bits_equal(ref, v), nm  # same ATen ops, same order
    assert np.isnan(golden["r2o/wb1/disp"][0])  # acc == 0 ray (render_utils.py:284)


def test_sample_pdf_isolated(golden):
    bins, w = T(golden["pdf/bins"]), T(golden["pdf/weights"])
    assert bits_equal(golden["pdf/cdf"], O.pdf_cdf(w))
    assert bits_equal(golden["pdf/det128"], O.sample_pdf(bins, w, 128))
    assert bits_equal(golden["pdf/det64"], O.sample_pdf(bins, w, 64))
    u = T(golden["pdf/rand128_u"])
    assert bits_equal(golden["pdf/rand128"], O.sample_pdf(bins, w, 128, u))
    cdf = T(golden["pdf/cdf"])
    _, inds = O.invert_cdf(bins, cdf, u)
    assert bits_equal(golden["pdf/inds_rand128"], inds)
    _, inds = O.invert_cdf(bins, cdf, torch.linspace(0., 1., 128).expand(10, 128))
    assert bits_equal(golden["pdf/inds_det128"], inds)
    assert inds.min() >= 1 and inds.max() <= 63


def test_render_image(golden):
    pc, pf = O.make_params(11), O.make_params(12)
    g = golden
    rgb, disp, acc = O.render_image(pc, pf, 6, 7, g["render/blender/K"], T(g["render/blender/c2w"])[:3, :4], 2.0, 6.0,
                                    n_importance=128, white_bkgd=True)
    np.testing.assert_allclose(rgb.numpy(), g["render/blender/rgb"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(acc.numpy(), g["render/blender/acc"], rtol=0, atol=1e-5)
    rgb, disp, acc = O.render_image(pc, pf, 6, 8, g["render/llff_ndc/K"], T(g["render/llff_ndc/c2w"])[:3, :4], 0.0, 1.0,
                                    ndc=True, n_importance=64, white_bkgd=False)
    np.testing.assert_allclose(rgb.numpy(), g["render/llff_ndc/rgb"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(acc.numpy(), g["render/llff_ndc/acc"], rtol=0, atol=1e-5)


def test_no_viewdirs_model_and_render(golden_noview):
    """use_viewdirs=False: trunk -> output_linear (nerf.py:131-132), 8-column ray rows (render_utils.py:225)"""
    g = golden_noview
    pts = T(g["mlp/pts"])
    np.testing.assert_allclose(O.nerf_query(O.make_params(31, output_ch=5), pts, None).numpy(), g["mlp/raw5_seed31"],
                               rtol=0, atol=2e-6)
    np.testing.assert_allclose(O.nerf_query(O.make_params(33, output_ch=4), pts, None).numpy(), g["mlp/raw4_seed33"],
                               rtol=0, atol=2e-6)
    pc, pf = O.make_params(31, output_ch=5), O.make_params(32, output_ch=5)
    rgb, disp, acc = O.render_image(pc, pf, 6, 7, g["render/blender/K"], T(g["render/blender/c2w"])[:3, :4], 2.0, 6.0,
                                    n_importance=128, white_bkgd=True)
    np.testing.assert_allclose(rgb.numpy(), g["render/blender/rgb"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(acc.numpy(), g["render/blender/acc"], rtol=0, atol=1e-5)
    rgb, disp, acc = O.render_image(pc, pf, 6, 8, g["render/llff_ndc/K"], T(g["render/llff_ndc/c2w"])[:3, :4], 0.0, 1.0,
                                    ndc=True, n_importance=64, white_bkgd=False)
    np.testing.assert_allclose(rgb.numpy(), g["render/llff_ndc/rgb"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(acc.numpy(), g["render/llff_ndc/acc"], rtol=0, atol=1e-5)


def test_wide_encoding_model(golden_noview):
    """configs/stonehenge.txt: multires 15 / 6"""
    g = golden_noview
    raw = O.nerf_query(O.make_params(41, input_ch=93, input_ch_views=39), T(g["mlp156/pts"]), T(g["mlp156/viewdirs"]),
                       multires=15, multires_views=6)
    np.testing.assert_allclose(raw.numpy(), g["mlp156/raw_seed41"], rtol=0, atol=2e-6)


def test_train_step_grads(golden):
    rb = T(golden["rr/ray_batch"])
    ro, rd = rb[:, 0:3].contiguous(), rb[:, 3:6].contiguous()
    tgt = T(golden["train/target"])
    pc, pf = O.make_params(11), O.make_params(12)
    loss, gc, gf, _ = O.train_step_grads(pc, pf, ro, rd, tgt, 2.0, 6.0)
    np.testing.assert_allclose(loss.numpy(), golden["train/loss"], rtol=1e-5)
    for tag, gr in (("coarse", gc), ("fine", gf)):
        for k, g in gr.items():
            ref = float(golden["train/%s/gnorm/%s" % (tag, k)])
            assert abs(g.double().norm().item() - ref) <= 1e-3 * ref + 1e-9, (tag, k)
            np.testing.assert_allclose(g.reshape(-1)[:64].numpy(), golden["train/%s/ghead/%s" % (tag, k)],
                                       rtol=1e-3, atol=1e-6)
        np.testing.assert_allclose(gr["alpha_linear.weight"].numpy(), golden["train/%s/g_alpha_w" % tag], rtol=1e-3, atol=1e-6)
    # Adam (utils.py:170; step 1, lr 5e-4)
    for tag, p, gr in (("coarse", pc, gc), ("fine", pf, gf)):
        for name, key in (("alpha_linear.weight", "alpha_w"), ("pts_linears.0.bias", "l0_b")):
            z = torch.zeros_like(p[name])
            new, _, _ = O.adam_step(p[name], gr[name], z, z, 1, 5e-4)
            np.testing.assert_allclose(new.numpy(), golden["train/%s/after_adam/%s" % (tag, key)], rtol=0, atol=2e-6)


def test_pose_gradient(golden):
    pc, pf = O.make_params(11), O.make_params(12)
    pose = T(golden["pose/pose"]).clone().requires_grad_(True)
    K = O.blender_intrinsics(8, 8)
    ro, rd = O.get_rays(8, 8, K, pose)
    sel = golden["pose/sel"]
    ys, xs = sel // 8, sel % 8
    ro, rd = ro[ys, xs], rd[ys, xs]
    out = O.render_rays(pc, pf, ro, rd, O.unit_viewdirs(rd), 2.0, 6.0)
    loss = O.mse(out["rgb_map"], T(golden["pose/target"]))
    loss.backward()
    np.testing.assert_allclose(loss.detach().numpy(), golden["pose/loss"], rtol=1e-5)
    np.testing.assert_allclose(pose.grad.numpy(), golden["pose/grad_pose"], rtol=2e-3, atol=1e-7)


def test_oracle_gradients_of_the_noview_and_wide_models(golden_noview):
    """autograd of the oracle's MLP for use_viewdirs=False (output_ch 5) and multires 15 / 6 against the reference's
    parameter gradients (tests/golden/make_golden_noview.py: d/dtheta sum(raw * G))."""
    import torch
    g = golden_noview
    cases = (("noview5", O.make_params(31, output_ch=5), torch.from_numpy(g["mlp/pts"]), None, {}),
             ("wide156", O.make_params(41, input_ch=93, input_ch_views=39), torch.from_numpy(g["mlp156/pts"]),
              torch.from_numpy(g["mlp156/viewdirs"]), dict(multires=15, multires_views=6)))
    for tag, params, pts, vd, kw in cases:
        p = {k: v.clone().requires_grad_(True) for k, v in params.items()}
        raw = O.nerf_query(p, pts, vd, **kw)
        (raw * torch.from_numpy(g["grad/%s/G" % tag])).sum().backward()
        checked = 0
        for k, v in p.items():
            key = "grad/%s/gnorm/%s" % (tag, k)
            if key not in g:
                assert v.grad is None or float(v.grad.abs().max()) == 0.0, (tag, k)   # (views_linears of the noview model is unused)
                continue
            ref_n = float(g[key])
            assert abs(float(v.grad.double().norm()) - ref_n) <= 1e-4 * ref_n + 1e-12, (tag, k)
            e = torch.from_numpy(g["grad/%s/ghead/%s" % (tag, k)])
            assert float((v.grad.reshape(-1)[:512] - e).norm()) <= 1e-4 * float(e.norm()) + 1e-9, (tag, k)
            checked += 1
        assert checked >= 18
