"""-m gpu: the NeRF MLP kernels (UMMA self-test, fp32 CUDA-core path, fused tcgen05 path)."""
import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O
from tests.gpu_util import T, cuda, make_model, maxabs

pytestmark = pytest.mark.gpu


def _inputs(n_rays, S, seed=0, spread=3.0):
    rng = np.random.default_rng(seed)
    pts = T(rng.uniform(-spread, spread, size=(n_rays, S, 3)).astype(np.float32))
    vd = T(rng.normal(size=(n_rays, 3)).astype(np.float32))
    return pts, vd / vd.norm(dim=-1, keepdim=True)


@pytest.mark.parametrize("K", [64, 128, 256])
def test_umma_selftest(K):
    """one tcgen05 GEMM through the library's own swizzle + descriptors == A @ B^T"""
    from nerf_shared_b200 import _ops
    g = torch.Generator().manual_seed(K)
    A, B = torch.randn(128, K, generator=g), torch.randn(128, K, generator=g)
    D = _ops.debug_umma(A.cuda(), B.cuda()).cpu()
    ref = A.bfloat16().float() @ B.bfloat16().float().t()
    err = (D - ref).abs().max().item()
    if err > 1e-3:
        # diagnostics for a first bring-up: which rows / columns are right?
        ok = ((D - ref).abs() < 1e-3)
        print("rows ok:", ok.all(1).nonzero().flatten().tolist()[:32], "cols ok:", ok.all(0).nonzero().flatten().tolist()[:32])
        print("D[0,:8]", D[0, :8].tolist(), "ref", ref[0, :8].tolist())
    assert err <= 1e-3, err


@pytest.mark.parametrize("K", [64, 128])
def test_umma_mn_major_selftest(K):
    """the wgrad operand mode: D = A^T B with both operands MN-major (reduction over rows)"""
    from nerf_shared_b200 import _ops
    g = torch.Generator().manual_seed(K + 1)
    A, B = torch.randn(K, 128, generator=g), torch.randn(K, 128, generator=g)
    D = _ops.debug_umma_mn(A.cuda(), B.cuda()).cpu()
    ref = A.bfloat16().float().t() @ B.bfloat16().float()
    err = (D - ref).abs().max().item()
    if err > 1e-3:
        ok = ((D - ref).abs() < 1e-3)
        print("rows ok:", ok.all(1).nonzero().flatten().tolist()[:32], "cols ok:", ok.all(0).nonzero().flatten().tolist()[:32])
        print("D[0,:8]", D[0, :8].tolist(), "ref", ref[0, :8].tolist(), "D^T?", float((D.t() - ref).abs().max()))
    assert err <= 1e-3, err


def test_fp32_kernel_vs_golden(golden):
    m = make_model(11, precision="fp32")
    raw = m(cuda(golden["mlp/pts"]), cuda(golden["mlp/viewdirs"]))
    np.testing.assert_allclose(raw.detach().cpu().numpy(), golden["mlp/raw_seed11"], rtol=0, atol=2e-5)
    m = make_model(13, scale=1.59, precision="fp32")
    raw = m(cuda(golden["mlp/pts"]), cuda(golden["mlp/viewdirs"]))
    np.testing.assert_allclose(raw.detach().cpu().numpy(), golden["mlp/raw_seed13x1.59"], rtol=0, atol=2e-4)
    dens = make_model(11, precision="fp32").get_density(cuda(golden["mlp/pts"]))
    np.testing.assert_allclose(dens.detach().cpu().numpy(), golden["mlp/density_seed11"], rtol=0, atol=2e-5)


def test_fp32_kernel_ragged_sizes():
    m = make_model(11, precision="fp32")
    p = O.make_params(11)
    for n_rays, S in ((1, 1), (3, 7), (5, 64), (2, 192)):
        pts, vd = _inputs(n_rays, S, seed=n_rays)
        ref = O.nerf_query(p, pts, vd)
        out = m(pts.cuda(), vd.cuda())
        assert out.shape == ref.shape
        np.testing.assert_allclose(out.detach().cpu().numpy(), ref.numpy(), rtol=0, atol=2e-5)


def test_tc_layers_vs_bf16_oracle():
    """every stored layer activation of the fused kernel against the bf16-emulating oracle:
    localises a wrong layer / slab / swizzle at first bring-up."""
    from nerf_shared_b200 import _ops
    m = make_model(11)
    p = O.make_params(11)
    n_rays, S = 5, 64   # 320 samples: 2.5 tiles (ragged tail)
    pts, vd = _inputs(n_rays, S, seed=3)
    raw, acts = _ops.mlp_forward_tc(m, pts.cuda(), vd.cuda(), S, save_acts=True)
    torch.cuda.synchronize()
    ref_acts = []
    flat = pts.reshape(-1, 3)
    dirs = vd[:, None].expand(pts.shape).reshape(-1, 3)
    ref = O.mlp(p, O.positional_encoding(flat, 10), O.positional_encoding(dirs, 4), bf16=True, acts=ref_acts)
    acts = _ops.untile(acts, _ops.ACT_SLABS, flat.shape[0]).float().cpu()
    names = ["h%d" % i for i in range(8)] + ["feature", "hv"]
    offs = [i * 256 for i in range(8)] + [2048, 2304]
    enc = torch.cat([O.positional_encoding(flat, 10), torch.zeros(flat.shape[0], 1)], -1)
    assert (acts[:, 2432:2496] - enc.bfloat16().float()).abs().max() < 1.6e-2     # enc(pts) slab, 1 bf16 ulp at |x|<4
    encd = O.positional_encoding(dirs, 4).bfloat16().float()
    assert (acts[:, 2496:2523] - encd).abs().max() < 8e-3 and acts[:, 2523:2560].abs().max() == 0
    report = []
    for nm, off, r in zip(names, offs, ref_acts):
        got = acts[:, off:off + r.shape[1]]
        report.append((nm, float((got - r).abs().max()), float(r.abs().max())))
    print("layer max-abs errors:", report)
    for nm, err, mag in report:
        assert err <= 0.02 * max(mag, 1.0) + 1e-2, (nm, err, mag, report)
    assert maxabs(raw, ref) < 5e-3, (maxabs(raw, ref), report)


@pytest.mark.parametrize("n_rays,S", [(1, 64), (7, 64), (4, 192), (33, 100), (1024, 64), (512, 192)])
def test_tc_vs_bf16_oracle_and_fp32(n_rays, S):
    m = make_model(11)
    p = O.make_params(11)
    pts, vd = _inputs(n_rays, S, seed=S + n_rays)
    with torch.no_grad():
        out = m(pts.cuda(), vd.cuda())
    assert out.shape == (n_rays, S, 4)
    # fp32 device kernel as the reference at every size; CPU oracle on a bounded subset
    m32 = make_model(11, precision="fp32")
    ref32 = m32(pts.cuda(), vd.cuda())
    assert maxabs(out, ref32) < 2e-2
    sub = slice(0, min(n_rays, 8))
    ref_bf = O.nerf_query(p, pts[sub], vd[sub], bf16=True)
    assert maxabs(out[sub], ref_bf) < 5e-3
    ref = O.nerf_query(p, pts[sub], vd[sub])
    assert maxabs(ref32[sub], ref) < 5e-5


def test_tc_golden(golden):
    m = make_model(11)
    raw = m(cuda(golden["mlp/pts"]), cuda(golden["mlp/viewdirs"]))
    assert maxabs(raw, T(golden["mlp/raw_seed11"])) < 2e-2


def test_tc_repacks_after_weight_update():
    m = make_model(11)
    pts, vd = _inputs(4, 64, seed=1)
    with torch.no_grad():
        a = m(pts.cuda(), vd.cuda()).clone()
        m.rgb_linear.bias.add_(1.0)
        b = m(pts.cuda(), vd.cuda())
    np.testing.assert_allclose((b - a)[..., :3].cpu().numpy(), 1.0, atol=1e-5)
    np.testing.assert_allclose((b - a)[..., 3].cpu().numpy(), 0.0, atol=1e-6)


def test_tc_linearity_in_last_layer_full_size():
    """BASELINE-size property check (1024 rays x 192 samples): doubling rgb_linear doubles rgb-raw
    minus bias exactly in fp32 head arithmetic (size-independent property, no oracle needed)."""
    m = make_model(12)
    pts, vd = _inputs(1024, 192, seed=8)
    with torch.no_grad():
        m.rgb_linear.bias.zero_()
        a = m(pts.cuda(), vd.cuda())[..., :3].clone()
        m.rgb_linear.weight.mul_(2.0)
        b = m(pts.cuda(), vd.cuda())[..., :3]
    assert torch.equal(b, 2.0 * a)


def _grads_vs_oracle(model, params, pts, vd, G, tol, **kw):
    raw = model(pts.cuda(), None if vd is None else vd.cuda())
    (raw * G.cuda()).sum().backward()
    p = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    ref = O.nerf_query(p, pts, vd, **kw)
    assert maxabs(raw, ref) <= 2e-4
    (ref * G).sum().backward()
    worst = 0.0
    for k, t in model.named_parameters():
        e = p[k].grad
        if e is None or float(e.abs().max()) == 0.0:
            assert t.grad is None or float(t.grad.abs().max()) == 0.0, k
            continue
        rel = float((t.grad.cpu() - e).norm() / e.norm())
        worst = max(worst, rel)
        assert rel <= tol, (k, rel)
    return worst


def test_fp32_training_path_noview_and_wide_models(golden_noview):
    """csrc/mlp_f32_train.cu: forward + parameter gradients of the use_viewdirs=False model (output_ch 5 and 4) and of
    the multires 15 / 6 model (configs/stonehenge.txt) against the reference's gradients (golden) and the oracle's fp32
    autograd; ragged sample counts."""
    from nerf_shared_b200 import nerf
    g = golden_noview
    rng = np.random.default_rng(9)
    # reference goldens
    for tag, mk, params, pts, vd in (
            ("noview5", lambda: nerf.NeRF(D=8, W=256, output_ch=5, skips=[4], use_viewdirs=False), O.make_params(31, output_ch=5),
             T(g["mlp/pts"]), None),
            ("wide156", lambda: nerf.NeRF(D=8, W=256, output_ch=5, skips=[4], use_viewdirs=True, multires=15, multires_views=6),
             O.make_params(41, input_ch=93, input_ch_views=39), T(g["mlp156/pts"]), T(g["mlp156/viewdirs"]))):
        m = mk()
        m.load_state_dict(params, strict=True)
        m = m.cuda()
        assert m.precision == "fp32"
        raw = m(pts.cuda(), None if vd is None else vd.cuda())
        (raw * cuda(g["grad/%s/G" % tag])).sum().backward()
        n = 0
        for k, t in m.named_parameters():
            key = "grad/%s/ghead/%s" % (tag, k)
            if key not in g:
                continue
            e = T(g[key])
            rel = float((t.grad.reshape(-1)[:512].cpu() - e).norm() / (e.norm() + 1e-30))
            assert rel <= 2e-4, (tag, k, rel)
            n += 1
        assert n >= 18
    # oracle autograd at ragged sizes (not multiples of the 64-row GEMM tile)
    for out_ch in (5, 4):
        m = nerf.NeRF(D=8, W=256, output_ch=out_ch, skips=[4], use_viewdirs=False)
        params = O.make_params(60 + out_ch, output_ch=out_ch, scale=1.3)
        m.load_state_dict(params, strict=True)
        pts = torch.from_numpy(rng.uniform(-3, 3, size=(7, 19, 3)).astype(np.float32))
        G = torch.from_numpy(rng.normal(size=(7, 19, out_ch)).astype(np.float32))
        print("noview out_ch=%d worst rel grad err %.2e" % (out_ch, _grads_vs_oracle(m.cuda(), params, pts, None, G, 2e-4)))
    # the standard view-dependent model on the fp32 path (exact-operand reference of the bf16 kernels' gradients)
    m = make_model(13, scale=1.59, precision="fp32")
    pts = torch.from_numpy(rng.uniform(-3, 3, size=(5, 67, 3)).astype(np.float32))
    vd = torch.nn.functional.normalize(torch.from_numpy(rng.normal(size=(5, 3)).astype(np.float32)), dim=-1)
    G = torch.from_numpy(rng.normal(size=(5, 67, 4)).astype(np.float32))
    print("viewdirs fp32 worst rel grad err %.2e" % _grads_vs_oracle(m, O.make_params(13, scale=1.59), pts, vd, G, 2e-4))


def test_noview_training_step_through_the_renderer(golden_noview):
    """main.py:77-103 with a use_viewdirs=False coarse / fine pair (8-column ray rows, output_ch 5): loss and every
    parameter gradient norm against the reference's."""
    from nerf_shared_b200 import nerf, render_utils, utils
    g = golden_noview
    models = []
    for seed in (31, 32):
        m = nerf.NeRF(D=8, W=256, output_ch=5, skips=[4], use_viewdirs=False)
        m.load_state_dict(O.make_params(seed, output_ch=5, scale=1.59), strict=True)
        models.append(m.cuda())
    mc, mf = models
    r = render_utils.Renderer(perturb=0.0, N_importance=128, N_samples=64, use_viewdirs=False, white_bkgd=True,
                              raw_noise_std=0.0, near=2.0, far=6.0)
    r.train()
    rays = torch.stack([cuda(g["train_noview/rays_o"]), cuda(g["train_noview/rays_d"])], 0)
    tgt = cuda(g["train_noview/target"])
    rgb, _, _, extras = r.render_from_rays(8, 8, O.blender_intrinsics(8, 8), chunk=32768, rays=rays, coarse_model=mc, fine_model=mf, retraw=True)
    loss = utils.img2mse(rgb, tgt) + utils.img2mse(extras["rgb0"], tgt)
    loss.backward()
    assert abs(float(loss) - float(g["train_noview/loss"])) <= 1e-4 * float(g["train_noview/loss"])
    for tag, m in (("coarse", mc), ("fine", mf)):
        for k, p in m.named_parameters():
            key = "train_noview/%s/gnorm/%s" % (tag, k)
            if key in g:
                ref_n = float(g[key])
                assert ref_n > 0 and abs(float(p.grad.double().norm()) - ref_n) <= 2e-3 * ref_n, (tag, k)
