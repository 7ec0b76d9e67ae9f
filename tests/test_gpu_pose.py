"""-m gpu: gradients wrt the MLP's inputs and the camera pose (SURVEY.md section 3.3 / BASELINE.json
configs[4]: examples/relative_pose_estimation_demo/demo_est_rel_pose.py:74-98, weights frozen)."""
import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O
from tests.gpu_util import T, cuda, make_model, maxabs

pytestmark = pytest.mark.gpu


def _bf(x):
    return x.bfloat16().float()


def _freeze(m):
    for p in m.parameters():
        p.requires_grad_(False)
    return m


def _rel_cos(g, e):
    g, e = g.detach().double().cpu().reshape(-1), e.detach().double().cpu().reshape(-1)
    return float((g - e).norm() / (e.norm() + 1e-30)), float((g * e).sum() / (g.norm() * e.norm() + 1e-30))


@pytest.mark.parametrize("n_rays,S,frozen", [(5, 64, True), (3, 192, False), (7, 128, True), (1, 32, True)])
def test_input_grads_same_rounding(n_rays, S, frozen):
    """d_pts / d_viewdirs == encoder-Jacobian^T (dA_0 W_0 + dA_5 W_5[:, :63]) resp. dA_hv W_v[:, 256:],
    evaluated in float64 from the very bf16 dA tiles and bf16 weights the kernel reads"""
    from nerf_shared_b200 import _ops
    m = make_model(11)
    if frozen:
        _freeze(m)
    m._nb_keep_dA = True
    rng = np.random.default_rng(n_rays * 1000 + S)
    pts = cuda(rng.uniform(-3, 3, size=(n_rays, S, 3)).astype(np.float32)).requires_grad_(True)
    vd = torch.nn.functional.normalize(cuda(rng.normal(size=(n_rays, 3)).astype(np.float32)), dim=-1).requires_grad_(True)
    d_raw = cuda((rng.normal(size=(n_rays, S, 4)) * 1e-2).astype(np.float32))
    raw = m(pts, vd)
    raw.backward(d_raw)
    torch.cuda.synchronize()
    assert pts.grad is not None and vd.grad is not None
    if frozen:
        assert all(p.grad is None for p in m.parameters())
    M = n_rays * S
    dA = _ops.untile(m._nb_last_dA, _ops.DA_SLABS, M).double().cpu()
    sd = {k: _bf(v.detach()).double().cpu() for k, v in m.state_dict().items()}
    d_enc_p = dA[:, 0:256] @ sd["pts_linears.0.weight"] + dA[:, 1280:1536] @ sd["pts_linears.5.weight"][:, :63]
    d_enc_d = dA[:, 2304:2432] @ sd["views_linears.0.weight"][:, 256:]
    p64 = pts.detach().double().cpu().reshape(-1, 3).requires_grad_(True)
    v64 = vd.detach().double().cpu().requires_grad_(True)
    ep = O.positional_encoding(p64, 10)
    ed = O.positional_encoding(v64[:, None].expand(n_rays, S, 3).reshape(-1, 3), 4)
    ((ep * d_enc_p).sum() + (ed * d_enc_d).sum()).backward()
    ep_rel, _ = _rel_cos(pts.grad.reshape(-1, 3), p64.grad)
    ed_rel, _ = _rel_cos(vd.grad, v64.grad)
    print("input grads, same rounding: rel err pts %.3g dirs %.3g" % (ep_rel, ed_rel))
    # octave 9 multiplies sin/cos errors (2e-6 abs) by 512: 1e-3 relative on that term
    assert ep_rel < 3e-3 and ed_rel < 1e-4, (ep_rel, ed_rel)
    assert maxabs(pts.grad.reshape(-1, 3), p64.grad.float()) <= 3e-3 * float(p64.grad.abs().max())


def test_input_grads_vs_fp32_autograd():
    """the whole bf16 input-gradient path (dgrad chain + input GEMM + encoder derivative) against fp32
    autograd of the oracle MLP"""
    n_rays, S = 12, 128
    m = _freeze(make_model(12))
    rng = np.random.default_rng(5)
    pts_h = T(rng.uniform(-1.5, 1.5, size=(n_rays, S, 3)).astype(np.float32))
    vd_h = torch.nn.functional.normalize(T(rng.normal(size=(n_rays, 3)).astype(np.float32)), dim=-1)
    d_raw = T((rng.normal(size=(n_rays, S, 4)) * 1e-2).astype(np.float32))
    pts, vd = pts_h.cuda().requires_grad_(True), vd_h.cuda().requires_grad_(True)
    m(pts, vd).backward(d_raw.cuda())
    pr, vr = pts_h.clone().requires_grad_(True), vd_h.clone().requires_grad_(True)
    O.nerf_query(O.make_params(12), pr, vr).backward(d_raw)
    rel_p, cos_p = _rel_cos(pts.grad, pr.grad)
    rel_v, cos_v = _rel_cos(vd.grad, vr.grad)
    print("input grads vs fp32 autograd: pts rel %.3g cos %.5f | dirs rel %.3g cos %.5f" % (rel_p, cos_p, rel_v, cos_v))
    # same error profile as the trunk weight gradients (relu-mask flips of bf16 pre-activations near zero)
    assert rel_p < 0.2 and cos_p > 0.98, (rel_p, cos_p)
    assert rel_v < 0.08 and cos_v > 0.995, (rel_v, cos_v)


def _pose_loss_gpu(renderer, mc, mf, H, W, K, pose, pix, target):
    """demo_est_rel_pose.py:80-97: rays of the whole grid from the pose, a pixel batch, render, MSE"""
    from nerf_shared_b200 import utils
    rays_o, rays_d = utils.get_rays(H, W, K, pose)
    ro, rd = rays_o.reshape(-1, 3)[pix], rays_d.reshape(-1, 3)[pix]
    rgb, disp, acc, extras = renderer.render_from_rays(H, W, K, chunk=4096, rays=torch.stack([ro, rd], 0),
                                                       coarse_model=mc, fine_model=mf, retraw=True)
    return utils.img2mse(rgb, target), extras


@pytest.mark.parametrize("scene", ["blender", "llff_ndc"])
def test_pose_gradient_end_to_end(scene):
    """d loss / d c2w through Renderer (coarse -> resample (detached) -> fine) against fp32 autograd of the
    oracle given the same fine depths; frozen weights"""
    from nerf_shared_b200 import render_utils
    n = 96
    g = torch.Generator().manual_seed(3)
    if scene == "blender":
        H = W = 40
        K = O.blender_intrinsics(H, W)
        pose_h = T(O.pose_spherical(25.0, -30.0, 4.0))[:3, :4].clone()
        kw = dict(near=2.0, far=6.0, N_importance=128, white_bkgd=True, ndc=False)
    else:
        H, W = 36, 48
        K = O.llff_intrinsics(H, W, 38.8)
        pose_h = T(O.llff_pose(2))[:3, :4].clone()
        kw = dict(near=0.0, far=1.0, N_importance=64, white_bkgd=False, ndc=True)
    r = render_utils.Renderer(perturb=0.0, N_samples=64, use_viewdirs=True, raw_noise_std=0.0, lindisp=False, **kw)
    mc, mf = _freeze(make_model(11, scale=1.59)), _freeze(make_model(12, scale=1.59))
    pix = torch.randperm(H * W, generator=g)[:n]
    target = torch.rand(n, 3, generator=g)

    pose = pose_h.cuda().requires_grad_(True)
    loss, extras = _pose_loss_gpu(r, mc, mf, H, W, K, pose, pix.cuda(), target.cuda())
    loss.backward()
    assert pose.grad is not None and pose.grad.shape == (3, 4)

    # oracle: same pixels, the fine depths the GPU path used (resampling is detached in the reference,
    # render_utils.py:145, so z is a constant of the differentiated graph)
    r2 = render_utils.Renderer(perturb=0.0, N_samples=64, use_viewdirs=True, raw_noise_std=0.0, lindisp=False, **kw)
    with torch.no_grad():
        from nerf_shared_b200 import utils
        ro_g, rd_g = utils.get_rays(H, W, K, pose.detach())
        vd_g = torch.nn.functional.normalize(rd_g.reshape(-1, 3)[pix.cuda()], dim=-1)
        ro_g, rd_g = ro_g.reshape(-1, 3)[pix.cuda()], rd_g.reshape(-1, 3)[pix.cuda()]
        if kw["ndc"]:
            ro_g, rd_g = utils.ndc_rays(H, W, K[0][0], 1.0, ro_g, rd_g)
        rows = torch.cat([ro_g, rd_g, torch.full((n, 1), kw["near"], device="cuda"),
                          torch.full((n, 1), kw["far"], device="cuda"), vd_g], -1)
        z_fine = r2.render_rays(rows, mc, mf, retweights=True)["z_vals"].cpu()
    pr = pose_h.clone().requires_grad_(True)
    ro, rd = O.get_rays(H, W, K, pr)
    ro, rd = ro.reshape(-1, 3)[pix], rd.reshape(-1, 3)[pix]
    vd = rd / torch.norm(rd, dim=-1, keepdim=True)
    if kw["ndc"]:
        ro, rd = O.ndc_rays(H, W, float(K[0][0]), 1.0, ro, rd)
    raw = O.nerf_query(O.make_params(12, scale=1.59), O.points_on_rays(ro, rd, z_fine), vd)
    rgb = O.composite(raw, z_fine, rd, kw["white_bkgd"])[0]
    loss_ref = O.mse(rgb, target)
    loss_ref.backward()
    rel, cos = _rel_cos(pose.grad, pr.grad)
    print("%s pose gradient: loss %.6f vs %.6f, rel err %.3g, cosine %.5f\n gpu %s\n ref %s" % (
        scene, float(loss.detach()), float(loss_ref.detach()), rel, cos, pose.grad.cpu().numpy().round(5), pr.grad.numpy().round(5)))
    assert abs(float(loss) - float(loss_ref)) <= 2e-3 * max(float(loss_ref), 1e-3)
    assert rel < 0.15 and cos > 0.985, (rel, cos)


def test_pose_adam_steps_reduce_loss():
    """a few iterations of the demo's optimisation loop (Adam on a pose offset, frozen weights) run and
    lower the photometric loss against a render from the true pose"""
    from nerf_shared_b200 import render_utils
    H = W = 32
    K = O.blender_intrinsics(H, W)
    r = render_utils.Renderer(perturb=0.0, N_samples=64, N_importance=64, use_viewdirs=True, raw_noise_std=0.0,
                              white_bkgd=True, near=2.0, far=6.0)
    mc, mf = _freeze(make_model(11, scale=1.59)), _freeze(make_model(12, scale=1.59))
    true_pose = cuda(O.pose_spherical(25.0, -30.0, 4.0))[:3, :4]
    pix = torch.arange(H * W, device="cuda")
    with torch.no_grad():
        target = r.render(H, W, K, mc, mf, c2w=true_pose, retraw=False)[0].reshape(-1, 3)
    offset = torch.zeros(3, device="cuda", requires_grad=True)
    start = torch.tensor([0.012, -0.008, 0.01], device="cuda")
    opt = torch.optim.Adam([offset], lr=2e-3)
    losses = []
    for _ in range(12):
        pose = torch.cat([true_pose[:, :3], (true_pose[:, 3] + start + offset)[:, None]], -1)
        loss, _ = _pose_loss_gpu(r, mc, mf, H, W, K, pose, pix, target)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    print("pose optimisation losses:", ["%.3e" % l for l in losses])
    # a random-init field is bumpy at the scale of the offset: require progress, not convergence
    assert min(losses[3:]) < 0.8 * losses[0], losses


def test_graphed_pose_step_equals_the_eager_step():
    """train.GraphedPoseStep (selected-pixel rays -> render with frozen networks -> MSE -> backward to the pose, one CUDA
    graph) against the demo's eager sequence (demo_est_rel_pose.py:80-97: full-grid get_rays, index, render, img2mse,
    backward) on a model with a non-zero pose gradient."""
    from nerf_shared_b200 import render_utils, train, utils
    from tests.gpu_util import make_model
    H = W = 24
    K = O.blender_intrinsics(H, W)
    mc, mf = make_model(13, scale=1.59), make_model(14, scale=1.59)
    for p in list(mc.parameters()) + list(mf.parameters()):
        p.requires_grad_(False)
    r = render_utils.Renderer(perturb=0.0, N_importance=64, N_samples=64, use_viewdirs=True, white_bkgd=True,
                              raw_noise_std=0.0, near=2.0, far=6.0)
    step = train.GraphedPoseStep(r, mc, mf, H, W, K, 128)
    g = torch.Generator().manual_seed(0)
    for it in range(3):
        pose = cuda(O.pose_spherical(30.0 + 5 * it, -30.0, 4.0)).clone().requires_grad_(True)
        pix = torch.randperm(H * W, generator=g)[:128].cuda()
        tgt = torch.rand(128, 3, generator=g).cuda()
        loss_g, grad_g = step(pose.detach(), pix, tgt)
        loss_g, grad_g = float(loss_g), grad_g.clone()
        rays_o, rays_d = utils.get_rays(H, W, K, pose)
        ys, xs = (pix // W).cpu().numpy(), (pix % W).cpu().numpy()
        rgb = r.render_from_rays(H, W, K, chunk=32768, rays=torch.stack([rays_o[ys, xs], rays_d[ys, xs]], 0), coarse_model=mc,
                                 fine_model=mf, retraw=True)[0]
        loss = utils.img2mse(rgb, tgt)
        loss.backward()
        assert abs(float(loss) - loss_g) <= 1e-6 * abs(float(loss))
        assert float(pose.grad.abs().max()) > 1e-3
        assert float((pose.grad - grad_g).abs().max()) <= 1e-4 * float(pose.grad.abs().max()), (pose.grad, grad_g)
    assert step.launches_per_replay > 10


def test_lazy_ray_grid_equals_the_full_grid():
    """utils.get_rays with a pose that requires grad returns lazy [H, W, 3] grids: `rays[ys, xs]` (the demo's indexing,
    demo_est_rel_pose.py:83-84) touches only the selected pixels; values, gradients and every other use equal the
    materialised grid."""
    from nerf_shared_b200 import utils
    H, W = 37, 53
    K = O.blender_intrinsics(H, W)
    rng = np.random.default_rng(0)
    ys, xs = rng.integers(0, H, 200), rng.integers(0, W, 200)
    G = torch.randn(200, 3, generator=torch.Generator().manual_seed(1)).cuda()
    res = {}
    for lazy in ("1", "0"):
        import os
        os.environ["NB_LAZY_RAYS"] = lazy
        try:
            pose = cuda(O.pose_spherical(31.0, -30.0, 4.0)).clone().requires_grad_(True)
            ro, rd = utils.get_rays(H, W, K, pose)
            assert isinstance(ro, utils.LazyRayGrid) == (lazy == "1")
            assert tuple(ro.shape) == (H, W, 3) and ro.requires_grad
            so, sd = ro[ys, xs], rd[ys, xs]
            assert type(so) is torch.Tensor and tuple(so.shape) == (200, 3)
            ((so * G).sum() + (sd * G * 2).sum()).backward()
            full = rd.reshape(-1, 3)                       # any other use materialises the grid
            res[lazy] = (so.detach().clone(), sd.detach().clone(), pose.grad.clone(), full.detach().clone(), float(rd.sum()))
        finally:
            os.environ.pop("NB_LAZY_RAYS", None)
    a, b = res["1"], res["0"]
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and torch.equal(a[3], b[3]) and a[4] == b[4]
    assert float((a[2] - b[2]).abs().max()) <= 1e-5 * float(b[2].abs().max())
