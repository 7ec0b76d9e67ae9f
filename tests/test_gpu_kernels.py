"""-m gpu: the streaming kernels (rays, sampling, compositing, resampling, Adam) through the C ABI
against the CPU oracle and the reference's golden vectors."""
import hashlib

import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O
from tests.gpu_util import T, bits_equal, cuda, maxabs, n_diff

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    from nerf_shared_b200 import _ops, utils
    return _ops, utils


def sha(t):
    return hashlib.sha256(np.ascontiguousarray(t.detach().cpu().numpy()).tobytes()).hexdigest()


def test_get_rays_bit_exact_golden(mods, golden):
    _, utils = mods
    for name in ("blender16", "blender_13x21", "llff_12x16"):
        K, c2w = golden["rays/%s/K" % name], golden["rays/%s/c2w" % name]
        H, W = golden["rays/%s/rays_d" % name].shape[:2]
        ro, rd = utils.get_rays(H, W, K, cuda(c2w)[:3, :4])
        assert bits_equal(golden["rays/%s/rays_d" % name], rd), n_diff(golden["rays/%s/rays_d" % name], rd)
        assert bits_equal(golden["rays/%s/rays_o" % name], ro)


def test_get_rays_full_size_checksums(mods, golden_meta):
    """BASELINE.json configs 1, 4, 5 at full size: sha256 of the bytes == the reference's."""
    _, utils = mods
    s = golden_meta["sha256"]
    c2w = cuda(O.pose_spherical(40.0, -30.0, 4.0))[:3, :4]
    for name, H in (("C1_400", 400), ("C4_800", 800)):
        ro, rd = utils.get_rays(H, H, O.blender_intrinsics(H, H), c2w)
        assert sha(rd) == s["rays_d/" + name]
        assert sha(ro) == s["rays_o/" + name]
    K = O.llff_intrinsics()
    ro, rd = utils.get_rays(378, 504, K, cuda(O.llff_pose(0))[:3, :4])
    assert sha(rd) == s["rays_d/C5_llff"]
    no, nd = utils.ndc_rays(378, 504, K[0][0], 1.0, ro, rd)
    assert sha(no) == s["ndc_o/C5_llff"] and sha(nd) == s["ndc_d/C5_llff"]


def test_get_rays_selected_pixels(mods):
    _, utils = mods
    H, W = 37, 53
    K = O.blender_intrinsics(H, W)
    c2w = cuda(O.pose_spherical(12.0, -30.0, 4.0))[:3, :4]
    ro, rd = utils.get_rays(H, W, K, c2w)
    pix = torch.from_numpy(np.random.default_rng(0).choice(H * W, 200, replace=False))
    so, sd = utils.get_rays_at(H, W, K, c2w, pix)
    assert bits_equal(rd.reshape(-1, 3)[pix.cuda()], sd) and bits_equal(ro.reshape(-1, 3)[pix.cuda()], so)


def test_ndc_bit_exact_golden(mods, golden):
    _, utils = mods
    name = "llff_12x16"
    K = golden["rays/%s/K" % name]
    ro, rd = cuda(golden["rays/%s/rays_o" % name]), cuda(golden["rays/%s/rays_d" % name])
    H, W = rd.shape[:2]
    no, nd = utils.ndc_rays(H, W, K[0][0], 1.0, ro, rd)
    assert bits_equal(golden["rays/%s/ndc_o" % name], no), n_diff(golden["rays/%s/ndc_o" % name], no)
    assert bits_equal(golden["rays/%s/ndc_d" % name], nd), n_diff(golden["rays/%s/ndc_d" % name], nd)


def test_rays_backward_vs_autograd(mods):
    """pose gradient through get_rays -> normalize -> ndc (demo_est_rel_pose.py:80-97 path)."""
    _, utils = mods
    _ops = mods[0]
    H, W = 9, 11
    K = O.llff_intrinsics(H, W, 9.5)
    pose = T(O.llff_pose(1)).clone()
    g = torch.Generator().manual_seed(0)
    w_o, w_d, w_v = (torch.randn(H, W, 3, generator=g) for _ in range(3))

    pc = pose.clone().requires_grad_(True)
    ro, rd = O.get_rays(H, W, K, pc)
    vd = rd / torch.norm(rd, dim=-1, keepdim=True)
    no, nd = O.ndc_rays(H, W, K[0][0], 1.0, ro, rd)
    ((no * w_o).sum() + (nd * w_d).sum() + (vd * w_v).sum()).backward()

    pg = pose.cuda().requires_grad_(True)
    ro, rd = utils.get_rays(H, W, K, pg)
    vd = _ops.Normalize.apply(rd)
    no, nd = utils.ndc_rays(H, W, K[0][0], 1.0, ro, rd)
    ((no * w_o.cuda()).sum() + (nd * w_d.cuda()).sum() + (vd * w_v.cuda()).sum()).backward()
    np.testing.assert_allclose(pg.grad.cpu().numpy(), pc.grad.numpy(), rtol=2e-4, atol=1e-4)


@pytest.mark.parametrize("lindisp", [False, True])
@pytest.mark.parametrize("perturb", [False, True])
def test_stratified_z_bit_exact(mods, perturb, lindisp):
    _ops, _ = mods
    N, S = 1000, 64
    rng = np.random.default_rng(3)
    rb = torch.zeros(N, 11)
    rb[:, 0:6] = T(rng.normal(size=(N, 6)).astype(np.float32))
    rb[:, 6] = T(rng.uniform(0.5, 2.5, N).astype(np.float32))
    rb[:, 7] = rb[:, 6] + T(rng.uniform(1.0, 5.0, N).astype(np.float32))
    t_rand = torch.rand(N, S, generator=torch.Generator().manual_seed(5)) if perturb else None
    z_ref = O.stratified_z(rb[:, 6:7], rb[:, 7:8], S, t_rand, lindisp)
    pts_ref = O.points_on_rays(rb[:, 0:3], rb[:, 3:6], z_ref)
    z, pts = _ops.sample_coarse(rb.cuda(), S, None if t_rand is None else t_rand.cuda(), lindisp, want_pts=True)
    assert bits_equal(z_ref.contiguous(), z), n_diff(z_ref.contiguous(), z)
    assert bits_equal(pts_ref.contiguous(), pts), n_diff(pts_ref.contiguous(), pts)
    pts2 = _ops.RayPoints.apply(rb[:, 0:3].cuda(), rb[:, 3:6].cuda(), z)
    assert bits_equal(pts_ref.contiguous(), pts2)


def test_stratified_z_golden_subset(mods, golden):
    """the reference's merged fine z contains its coarse z: ours must be a subset, bit for bit."""
    _ops, _ = mods
    rb = cuda(golden["rr/ray_batch"])
    z = _ops.sample_coarse(rb, 64, None, False).cpu().numpy()
    merged = golden["rr/det/z_vals"]
    for r in range(z.shape[0]):
        assert np.isin(z[r], merged[r]).all()


def test_composite_golden(mods, golden):
    _ops, _ = mods
    raw, z, d = cuda(golden["r2o/raw"]), cuda(golden["r2o/z"]), cuda(golden["r2o/rays_d"])
    for wb in (1, 0):
        res = _ops.Composite.apply(raw, z, d, None, bool(wb))
        for nm, v in zip(["rgb", "disp", "acc", "weights", "depth"], res):
            ref = golden["r2o/wb%d/%s" % (wb, nm)]
            np.testing.assert_allclose(v.cpu().numpy(), ref, rtol=2e-5, atol=2e-6, equal_nan=True, err_msg=nm)
    assert torch.isnan(res[1][0])  # acc == 0 ray keeps the reference's nan disparity


@pytest.mark.parametrize("S", [64, 192, 100, 33])
@pytest.mark.parametrize("white", [True, False])
def test_composite_fwd_bwd_vs_oracle(mods, S, white):
    _ops, _ = mods
    N = 257
    rng = np.random.default_rng(S)
    raw = T((rng.normal(size=(N, S, 4)) * np.array([1, 1, 1, 2.0])).astype(np.float32))
    z = torch.sort(T(rng.uniform(2, 6, size=(N, S)).astype(np.float32)), -1)[0]
    d = T(rng.normal(size=(N, 3)).astype(np.float32))
    noise = T(rng.normal(size=(N, S)).astype(np.float32))
    g_rgb, g_disp, g_acc, g_depth = (T(rng.normal(size=s).astype(np.float32)) for s in ((N, 3), (N,), (N,), (N,)))

    rc, dc = raw.clone().requires_grad_(True), d.clone().requires_grad_(True)
    out = O.composite(rc, z, dc, white, noise)
    loss = (out[0] * g_rgb).sum() + (out[2] * g_acc).sum() + (out[4] * g_depth).sum() + \
        (torch.nan_to_num(out[1], nan=0.0, posinf=0.0) * g_disp).clamp(-1e3, 1e3).sum() * 0
    loss.backward()

    rg, dg = raw.cuda().requires_grad_(True), d.cuda().requires_grad_(True)
    res = _ops.Composite.apply(rg, z.cuda(), dg, noise.cuda(), white)
    for a, b, nm in zip(res, out, ["rgb", "disp", "acc", "weights", "depth"]):
        np.testing.assert_allclose(a.detach().cpu().numpy(), b.detach().numpy(), rtol=1e-4, atol=1e-5, equal_nan=True,
                                   err_msg=nm)
    l2 = (res[0] * g_rgb.cuda()).sum() + (res[2] * g_acc.cuda()).sum() + (res[4] * g_depth.cuda()).sum()
    l2.backward()
    scale = float(rc.grad.abs().max())
    assert maxabs(rg.grad, rc.grad) <= 2e-4 * scale + 1e-6
    assert maxabs(dg.grad, dc.grad) <= 2e-4 * float(dc.grad.abs().max()) + 1e-6


def test_composite_disp_gradient(mods):
    _ops, _ = mods
    N, S = 64, 64
    rng = np.random.default_rng(9)
    raw = T((rng.normal(size=(N, S, 4)) + np.array([0, 0, 0, 1.0])).astype(np.float32))
    z = torch.sort(T(rng.uniform(2, 6, size=(N, S)).astype(np.float32)), -1)[0]
    d = T(rng.normal(size=(N, 3)).astype(np.float32))
    g = T(rng.normal(size=(N,)).astype(np.float32))
    rc = raw.clone().requires_grad_(True)
    (O.composite(rc, z, d, True)[1] * g).sum().backward()
    rg = raw.cuda().requires_grad_(True)
    (_ops.Composite.apply(rg, z.cuda(), d.cuda(), None, True)[1] * g.cuda()).sum().backward()
    assert maxabs(rg.grad, rc.grad) <= 5e-4 * float(rc.grad.abs().max()) + 1e-7


def test_invert_cdf_bit_exact_golden(mods, golden):
    """searchsorted indices and samples, given the reference's own cdf: bit-exact."""
    _ops, _ = mods
    bins, cdf = cuda(golden["pdf/bins"]), cuda(golden["pdf/cdf"])
    u = cuda(golden["pdf/rand128_u"])
    s, inds = _ops.invert_cdf(bins, cdf, u)
    assert bits_equal(golden["pdf/inds_rand128"], inds)
    assert bits_equal(golden["pdf/rand128"], s), n_diff(golden["pdf/rand128"], s)
    ud = torch.linspace(0., 1., 128).expand(10, 128).contiguous().cuda()
    s, inds = _ops.invert_cdf(bins, cdf, ud)
    assert bits_equal(golden["pdf/inds_det128"], inds)
    assert bits_equal(golden["pdf/det128"], s), n_diff(golden["pdf/det128"], s)


def test_invert_cdf_large_random(mods):
    _ops, _ = mods
    N, B, I = 4096, 63, 128
    g = torch.Generator().manual_seed(0)
    w = torch.rand(N, B - 1, generator=g) ** 3
    w[::7] = 0
    cdf = O.pdf_cdf(w)
    bins = torch.sort(torch.rand(N, B, generator=g) * 4 + 2, -1)[0]
    u = torch.rand(N, I, generator=g)
    u[:, 0], u[:, 1] = 0.0, 1.0   # edges: inds 1 .. B
    s_ref, i_ref = O.invert_cdf(bins, cdf, u)
    s, i = _ops.invert_cdf(bins.cuda(), cdf.cuda(), u.cuda())
    assert bits_equal(i_ref, i)
    assert bits_equal(s_ref, s), n_diff(s_ref, s)


def _close_except_knife_edges(got, ref, bin_width, atol=3e-5, max_frac=2e-3):
    """The cdf is a sum whose last ulp depends on the summation order (ATen's CPU order is itself
    ISA dependent), so a u that lands exactly on a cdf entry (u = 1.0 of the deterministic
    linspace above all) may fall in the neighbouring bin, and bins narrower than 1e-5 of
    probability amplify cdf rounding (utils.py:113).  Everything else must agree tightly; the
    exceptions must be rare and stay inside the neighbouring bin."""
    err = np.abs(got - ref)
    bad = err > atol
    assert bad.mean() <= max_frac, bad.mean()
    assert err.max() <= bin_width, err.max()


def test_sample_pdf_api_golden(mods, golden):
    _, utils = mods
    bins, w = cuda(golden["pdf/bins"]), cuda(golden["pdf/weights"])
    for n, key in ((128, "pdf/det128"), (64, "pdf/det64")):
        out = utils.sample_pdf(bins, w, n, det=True)
        _close_except_knife_edges(out.cpu().numpy(), golden[key], 0.08)


@pytest.mark.parametrize("det", [True, False])
@pytest.mark.parametrize("I", [128, 64])
def test_sample_pdf_merge_vs_oracle(mods, det, I):
    _ops, _ = mods
    N, S = 777, 64
    g = torch.Generator().manual_seed(I)
    near, far = torch.full((N, 1), 2.0), torch.full((N, 1), 6.0)
    z = O.stratified_z(near, far, S, torch.rand(N, S, generator=g))
    w = torch.rand(N, S, generator=g) ** 4
    w[5] = 0
    u = None if det else torch.rand(N, I, generator=g)
    z_ref, std_ref, smp_ref = O.hierarchical_z(z, w, I, u)
    z_out, z_std, smp = _ops.sample_pdf_merge(z.cuda(), w.cuda(), I, None if u is None else u.cuda(), want_samples=True)
    _close_except_knife_edges(smp.cpu().numpy(), smp_ref.numpy(), 0.08)
    _close_except_knife_edges(z_out.cpu().numpy(), z_ref.numpy(), 0.08)
    # z_std inherits the knife-edge sample (one of I values may sit one bin over)
    np.testing.assert_allclose(z_std.cpu().numpy(), std_ref.numpy(), rtol=0, atol=0.08 / np.sqrt(I) + 1e-5)
    assert np.median(np.abs(z_std.cpu().numpy() - std_ref.numpy())) < 1e-5
    zo = z_out.cpu()
    assert (zo[:, 1:] >= zo[:, :-1]).all()                       # sortedness
    # idempotence / permutation property: merged = sort(cat(z, samples)) exactly
    again = torch.sort(torch.cat([z.cuda(), smp], -1), -1)[0]
    assert bits_equal(again, z_out)


@pytest.mark.gpu
@pytest.mark.parametrize("det", [True, False])
@pytest.mark.parametrize("I", [128, 64])
@pytest.mark.parametrize("S", [64, 40])
def test_sample_pdf_merge_sorted_path_same_bits(mods, det, I, S):
    """The production path (no per-sample output: u sorted first, rank merge through a histogram) must give the
    same merged depths, bit for bit, as the path that inverts in u order and sorts afterwards."""
    _ops, _ = mods
    N = 1031
    g = torch.Generator().manual_seed(7 * I + S)
    near, far = torch.full((N, 1), 2.0), torch.full((N, 1), 6.0)
    z = O.stratified_z(near, far, S, torch.rand(N, S, generator=g))
    w = torch.rand(N, S, generator=g) ** 6
    w[3] = 0
    w[4, 10:] = 0                      # all mass in a few bins: many samples share one bin
    z[6] = 3.0                         # degenerate ray: every coarse depth equal
    u = None if det else torch.rand(N, I, generator=g)
    uc = None if u is None else u.cuda()
    z_a, std_a, smp = _ops.sample_pdf_merge(z.cuda(), w.cuda(), I, uc, want_samples=True)
    z_b, std_b = _ops.sample_pdf_merge(z.cuda(), w.cuda(), I, uc)
    assert bits_equal(z_a, z_b), n_diff(z_a, z_b)
    np.testing.assert_allclose(std_b.cpu().numpy(), std_a.cpu().numpy(), rtol=1e-5, atol=1e-6)
    assert bits_equal(torch.sort(torch.cat([z.cuda(), smp], -1), -1)[0], z_b)


def test_ray_points_backward(mods):
    _ops, _ = mods
    N, S = 300, 192
    g = torch.Generator().manual_seed(1)
    o, d, z, gp = torch.randn(N, 3, generator=g), torch.randn(N, 3, generator=g), torch.rand(N, S, generator=g), \
        torch.randn(N, S, 3, generator=g)
    oc, dc = o.clone().requires_grad_(True), d.clone().requires_grad_(True)
    (O.points_on_rays(oc, dc, z) * gp).sum().backward()
    og, dg = o.cuda().requires_grad_(True), d.cuda().requires_grad_(True)
    (_ops.RayPoints.apply(og, dg, z.cuda()) * gp.cuda()).sum().backward()
    np.testing.assert_allclose(og.grad.cpu().numpy(), oc.grad.numpy(), rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(dg.grad.cpu().numpy(), dc.grad.numpy(), rtol=1e-4, atol=1e-4)


def test_adam_step(mods):
    from nerf_shared_b200 import _lib
    lib = _lib.load()
    _lib.ensure_device("cuda:0")
    n = 100003
    g = torch.Generator().manual_seed(2)
    p, gr = torch.randn(n, generator=g), torch.randn(n, generator=g) * 1e-3
    m, v = torch.zeros(n), torch.zeros(n)
    pg, gg, mg, vg = p.cuda(), gr.cuda(), m.cuda(), v.cuda()
    for step in (1, 2, 3):
        p, m, v = O.adam_step(p, gr, m, v, step, 5e-4)
        _lib.check(lib.nb_adam_step(_lib.ptr(pg), _lib.ptr(gg), _lib.ptr(mg), _lib.ptr(vg), n, 5e-4, 0.9, 0.999, 1e-8,
                                    step, 1.0, _lib.stream_ptr()))
    np.testing.assert_allclose(pg.cpu().numpy(), p.numpy(), rtol=0, atol=1e-6)
    # against torch.optim.Adam itself
    q = torch.nn.Parameter(torch.randn(1000, generator=torch.Generator().manual_seed(4)))
    opt = torch.optim.Adam([q], lr=5e-4, betas=(0.9, 0.999))
    q2, m2, v2 = q.detach().clone().cuda(), torch.zeros(1000).cuda(), torch.zeros(1000).cuda()
    for step in (1, 2):
        q.grad = torch.full_like(q, 0.01 * step)
        opt.step()
        g2 = torch.full((1000,), 0.01 * step).cuda()
        _lib.check(lib.nb_adam_step(_lib.ptr(q2), _lib.ptr(g2), _lib.ptr(m2), _lib.ptr(v2), 1000, 5e-4, 0.9, 0.999, 1e-8,
                                    step, 1.0, _lib.stream_ptr()))
    np.testing.assert_allclose(q2.cpu().numpy(), q.detach().numpy(), rtol=0, atol=1e-6)


def test_empty_inputs(mods):
    _ops, utils = mods
    z = _ops.sample_coarse(torch.zeros(0, 11).cuda(), 64, None, False)
    assert z.shape == (0, 64)
    res = _ops.Composite.apply(torch.zeros(0, 64, 4).cuda(), z, torch.zeros(0, 3).cuda(), None, True)
    assert res[0].shape == (0, 3)


def test_mse_pair_matches_torch(mods):
    """fused img2mse(fine) + img2mse(coarse) + mse2psnr (main.py:87-99) against the reference lambdas and autograd."""
    _ops, utils = mods
    g = torch.Generator().manual_seed(3)
    rgb, rgb0, tgt = (torch.rand(1024, 3, generator=g) for _ in range(3))
    a, b = rgb.clone().requires_grad_(True), rgb0.clone().requires_grad_(True)
    ref = torch.mean((a - tgt) ** 2) + torch.mean((b - tgt) ** 2)
    ref.backward()
    ac, bc = rgb.cuda().requires_grad_(True), rgb0.cuda().requires_grad_(True)
    loss, m, m0, psnr = _ops.MsePair.apply(ac, bc, tgt.cuda())
    loss.backward()
    assert abs(float(loss) - float(ref)) <= 1e-6 * float(ref)
    assert abs(float(m) - float(torch.mean((rgb - tgt) ** 2))) <= 1e-7
    assert abs(float(psnr) - float(-10.0 * torch.log(torch.mean((rgb - tgt) ** 2)) / np.log(10.0))) <= 1e-4
    np.testing.assert_allclose(ac.grad.cpu().numpy(), a.grad.numpy(), rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(bc.grad.cpu().numpy(), b.grad.numpy(), rtol=1e-5, atol=1e-9)
    # fine map only (N_importance = 0), gradient through the fine term alone
    ac2 = rgb.cuda().requires_grad_(True)
    _, m_only, _, _ = _ops.MsePair.apply(ac2, None, tgt.cuda())
    (3.0 * m_only).backward()
    np.testing.assert_allclose(ac2.grad.cpu().numpy(), 3.0 * a.grad.numpy(), rtol=1e-5, atol=1e-9)


def test_ray_rows_same_bits_as_op_by_op(mods):
    """one-launch ray rows == normalise + fills + cat (render_utils.py:211,223-226), bit for bit."""
    _ops, _ = mods
    g = torch.Generator().manual_seed(5)
    o, d = torch.randn(1000, 3, generator=g).cuda(), torch.randn(1000, 3, generator=g).cuda()
    rows, vd = _ops.ray_rows(o, d, d, 2.0, 6.0)
    vd_ref = _ops.Normalize.apply(d)
    ref = torch.cat([o, d, torch.full((1000, 1), 2.0).cuda(), torch.full((1000, 1), 6.0).cuda(), vd_ref], -1)
    assert bits_equal(rows, ref) and bits_equal(vd, vd_ref)
    np.testing.assert_allclose(vd_ref.cpu().numpy(), O.unit_viewdirs(d.cpu()).numpy(), rtol=0, atol=2e-7)


def test_to8b_device_matches_numpy(mods):
    _, utils = mods
    g = torch.Generator().manual_seed(9)
    x = torch.rand(37, 41, 3, generator=g) * 1.4 - 0.2          # out-of-range values on both sides
    x[0, 0, 0], x[0, 0, 1], x[0, 0, 2] = 1.0, 0.0, 0.999999
    ref = (255 * np.clip(x.numpy(), 0, 1)).astype(np.uint8)
    got = utils.to8b_device(x.cuda()).cpu().numpy()
    assert got.dtype == np.uint8 and np.array_equal(got, ref)
    assert np.array_equal(utils.to8b_device(x.cuda().reshape(-1)[:1001]).cpu().numpy(), ref.reshape(-1)[:1001])


@pytest.mark.parametrize("S,white,with_noise", [(64, True, False), (192, True, True), (128, False, False), (100, True, False)])
def test_composite_train_equals_separate_kernels(S, white, with_noise):
    """nb_composite_train (compositing + MSE term + compositing backward in one pass) against nb_composite_fwd ->
    img2mse -> autograd -> nb_composite_bwd on the same inputs."""
    from nerf_shared_b200 import _ops
    n = 777
    g = torch.Generator().manual_seed(S)
    raw = (torch.randn(n, S, 4, generator=g) * torch.tensor([1.0, 1.0, 1.0, 2.0])).cuda()
    z = torch.sort(torch.rand(n, S, generator=g) * 4 + 2, -1)[0].cuda()
    rd = torch.randn(n, 3, generator=g).cuda()
    noise = torch.randn(n, S, generator=g).cuda() if with_noise else None
    tgt = torch.rand(n, 3, generator=g).cuda()
    raw_a = raw.clone().requires_grad_(True)
    rgb, disp, acc, w, depth = _ops.Composite.apply(raw_a, z, rd, noise, white)
    loss = torch.mean((rgb - tgt) ** 2)
    loss.backward()
    acc_sum = torch.zeros(1, device="cuda")
    rgb2, disp2, acc2, w2, d_raw = _ops.composite_train(raw, z, rd, noise, tgt, white, 1.0 / (3 * n), acc_sum, want_weights=True)
    # (S = 64 composites forward with the half-warp kernel: another reduction order, last-bit differences)
    assert maxabs(rgb2, rgb) <= 2e-6 and maxabs(acc2, acc) <= 2e-6 and maxabs(w2, w) <= 1e-6
    assert float(((disp2 - disp) / disp).nan_to_num().abs().max()) <= 1e-5
    assert abs(float(acc_sum) - float(loss)) <= 2e-6 * float(loss)
    scale = float(raw_a.grad.abs().max())
    assert float((d_raw - raw_a.grad).abs().max()) <= 2e-6 * scale, float((d_raw - raw_a.grad).abs().max()) / scale
