"""CPU oracle for the nerf_shared hot path  --  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

This file restates, on the CPU, the arithmetic of the reference's ray -> sample ->
encode -> MLP -> composite -> resample path (stanford-iprl-lab/nerf_shared, citations are
`file:line` into that tree).  It exists so that tests/, __graft_entry__.smoke() and
bench.py's `cpu_baseline` / `--impl reference` legs have something to check the CUDA
kernels against and to time on the host cores.  Nothing under nerf_shared_b200/ imports
it, and the product path fails loudly without its CUDA library.

Pinning: the reference ships no golden vectors or tests, so parity is
pinned by running the UNMODIFIED reference in the build container:
tests/golden/make_golden.py imports /root/reference/nerf_shared, runs it on seeded inputs
and commits its outputs under tests/golden/*.npz; tests/test_oracle_golden.py checks every
function below against those files (bit-exact where the reference result is an fp32
elementwise chain, 1e-6-ish where the reference uses an ATen reduction).

The reference does its arithmetic with PyTorch (ATen) fp32 CPU ops, so the oracle does
too: torch CPU tensors, one rounding per op, no fusion.  The structure is functional
(every random draw is an explicit argument) instead of the reference's classes.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor

# --------------------------------------------------------------------------------------
# rays
# --------------------------------------------------------------------------------------


def pixel_grid(H: int, W: int) -> Tuple[Tensor, Tensor]:
    """Pixel-centre free integer grid, i along x (width), j along y (height); utils.py:34-36."""
    xs = torch.linspace(0, W - 1, W)
    ys = torch.linspace(0, H - 1, H)
    i = xs[None, :].expand(H, W)
    j = ys[:, None].expand(H, W)
    return i, j


def get_rays(H: int, W: int, K, c2w: Tensor) -> Tuple[Tensor, Tensor]:
    """Camera rays of a pinhole camera; utils.py:33-42.

    dirs = [(i-cx)/fx, -(j-cy)/fy, -1]; rays_d[k] = (dirs0*R[k,0] + dirs1*R[k,1]) + dirs2*R[k,2]
    (ATen's 3-element inner sum is left-to-right, probed); rays_o = c2w[:3,3] broadcast.
    K entries enter as Python scalars, so everything is fp32 with true division (CPU ATen).
    """
    c2w = torch.as_tensor(c2w, dtype=torch.float32)
    i, j = pixel_grid(H, W)
    fx, fy, cx, cy = float(K[0][0]), float(K[1][1]), float(K[0][2]), float(K[1][2])
    d0 = (i - cx) / fx
    d1 = -((j - cy) / fy)
    d2 = -torch.ones_like(i)
    R = c2w[:3, :3]
    comps = []
    for k in range(3):
        comps.append((d0 * R[k, 0] + d1 * R[k, 1]) + d2 * R[k, 2])
    rays_d = torch.stack(comps, -1)
    rays_o = c2w[:3, 3].expand(rays_d.shape)
    return rays_o, rays_d


def ndc_rays(H: int, W: int, focal: float, near: float, rays_o: Tensor, rays_d: Tensor):
    """Forward-facing NDC reparametrisation; utils.py:54-71 (op order kept)."""
    t = -(near + rays_o[..., 2]) / rays_d[..., 2]
    o = rays_o + t[..., None] * rays_d
    sx = -1.0 / (W / (2.0 * focal))
    sy = -1.0 / (H / (2.0 * focal))
    o0 = sx * o[..., 0] / o[..., 2]
    o1 = sy * o[..., 1] / o[..., 2]
    o2 = 1.0 + 2.0 * near / o[..., 2]
    d0 = sx * (rays_d[..., 0] / rays_d[..., 2] - o[..., 0] / o[..., 2])
    d1 = sy * (rays_d[..., 1] / rays_d[..., 2] - o[..., 1] / o[..., 2])
    d2 = -2.0 * near / o[..., 2]
    return torch.stack([o0, o1, o2], -1), torch.stack([d0, d1, d2], -1)


def unit_viewdirs(rays_d: Tensor) -> Tensor:
    """render_utils.py:211-212."""
    v = rays_d / torch.norm(rays_d, dim=-1, keepdim=True)
    return v.reshape(-1, 3).float()


# --------------------------------------------------------------------------------------
# sampling along the ray
# --------------------------------------------------------------------------------------


def stratified_z(near: Tensor, far: Tensor, n_samples: int, t_rand: Optional[Tensor] = None,
                 lindisp: bool = False) -> Tensor:
    """Coarse sample depths; render_utils.py:105-129.

    near/far: [N,1].  t_rand: [N,S] uniform draws or None (perturb == 0).
    """
    t = torch.linspace(0.0, 1.0, steps=n_samples)
    if not lindisp:
        z = near * (1.0 - t) + far * t
    else:
        z = 1.0 / (1.0 / near * (1.0 - t) + 1.0 / far * t)
    z = z.expand(near.shape[0], n_samples)
    if t_rand is not None:
        mids = 0.5 * (z[..., 1:] + z[..., :-1])
        upper = torch.cat([mids, z[..., -1:]], -1)
        lower = torch.cat([z[..., :1], mids], -1)
        z = lower + (upper - lower) * t_rand
    return z


def points_on_rays(rays_o: Tensor, rays_d: Tensor, z: Tensor) -> Tensor:
    """render_utils.py:131,148."""
    return rays_o[..., None, :] + rays_d[..., None, :] * z[..., :, None]


def pdf_cdf(weights: Tensor) -> Tensor:
    """utils.py:76-79: cdf with a leading zero, len = len(weights)+1."""
    w = weights + 1e-5
    pdf = w / torch.sum(w, -1, keepdim=True)
    cdf = torch.cumsum(pdf, -1)
    return torch.cat([torch.zeros_like(cdf[..., :1]), cdf], -1)


def invert_cdf(bins: Tensor, cdf: Tensor, u: Tensor) -> Tuple[Tensor, Tensor]:
    """utils.py:100-115.  Returns (samples, inds); inds is the searchsorted(right=True) result."""
    u = u.contiguous()
    inds = torch.searchsorted(cdf.contiguous(), u, right=True)
    below = torch.clamp(inds - 1, min=0)
    above = torch.clamp(inds, max=cdf.shape[-1] - 1)
    cdf_b = torch.gather(cdf, -1, below)
    cdf_a = torch.gather(cdf, -1, above)
    bin_b = torch.gather(bins, -1, below)
    bin_a = torch.gather(bins, -1, above)
    denom = cdf_a - cdf_b
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    t = (u - cdf_b) / denom
    return bin_b + t * (bin_a - bin_b), inds


def sample_pdf(bins: Tensor, weights: Tensor, n_importance: int, u: Optional[Tensor] = None):
    """utils.py:74-117.  u=None means det=True (u = linspace(0,1,n))."""
    cdf = pdf_cdf(weights)
    if u is None:
        u = torch.linspace(0.0, 1.0, steps=n_importance).expand(list(cdf.shape[:-1]) + [n_importance])
    samples, _ = invert_cdf(bins, cdf, u)
    return samples


def hierarchical_z(z: Tensor, weights: Tensor, n_importance: int, u: Optional[Tensor] = None):
    """render_utils.py:141-147,168: returns (merged sorted z [N,S+I], z_std [N])."""
    z_mid = 0.5 * (z[..., 1:] + z[..., :-1])
    z_samples = sample_pdf(z_mid, weights[..., 1:-1], n_importance, u).detach()
    z_all, _ = torch.sort(torch.cat([z, z_samples], -1), -1)
    return z_all, torch.std(z_samples, dim=-1, unbiased=False), z_samples


# --------------------------------------------------------------------------------------
# positional encoding + MLP
# --------------------------------------------------------------------------------------


def positional_encoding(x: Tensor, n_freqs: int) -> Tensor:
    """nerf.py:16-41: [x, sin(2^0 x), cos(2^0 x), ..., sin(2^(L-1) x), cos(2^(L-1) x)]."""
    if n_freqs < 0:
        return x
    out = [x]
    if n_freqs > 0:
        freqs = 2.0 ** torch.linspace(0.0, n_freqs - 1, steps=n_freqs)
        for f in freqs:
            out.append(torch.sin(x * f))
            out.append(torch.cos(x * f))
    return torch.cat(out, -1)


PARAM_ORDER_VIEWDIRS = (
    ["pts_linears.%d" % i for i in range(8)] + ["feature_linear", "alpha_linear", "views_linears.0", "rgb_linear"])


def make_params(seed: int, D: int = 8, W: int = 256, input_ch: int = 63, input_ch_views: int = 27,
                skips: Sequence[int] = (4,), scale: float = 1.0, output_ch: int = 0) -> Dict[str, Tensor]:
    """Random-init weights with nn.Linear's distribution (U(+-1/sqrt(fan_in))), drawn from a
    numpy PCG64 stream so that the same bits come out on every machine / torch version.
    Keys are the reference state_dict keys (nerf.py:79-94).  output_ch > 0 builds the
    use_viewdirs=False model instead (views_linears.0 [W/2, W] and output_linear [output_ch, W],
    nerf.py:84,93-94)."""
    rng = np.random.default_rng(seed)
    shapes = []
    for i in range(D):
        fan_in = input_ch if i == 0 else (W + input_ch if (i - 1) in skips else W)
        shapes.append(("pts_linears.%d" % i, W, fan_in))
    if output_ch > 0:
        shapes += [("views_linears.0", W // 2, W), ("output_linear", output_ch, W)]
    else:
        shapes += [("views_linears.0", W // 2, input_ch_views + W), ("feature_linear", W, W),
                   ("alpha_linear", 1, W), ("rgb_linear", 3, W // 2)]
    p = {}
    for name, fo, fi in shapes:
        b = 1.0 / math.sqrt(fi)
        p[name + ".weight"] = torch.from_numpy((rng.uniform(-b, b, size=(fo, fi)) * scale).astype(np.float32))
        p[name + ".bias"] = torch.from_numpy(rng.uniform(-b, b, size=(fo,)).astype(np.float32))
    return p


def _bf16(x: Tensor) -> Tensor:
    return x.to(torch.bfloat16).to(torch.float32)


def mlp(params: Dict[str, Tensor], enc_pts: Tensor, enc_dirs: Optional[Tensor], D: int = 8,
        skips: Sequence[int] = (4,), bf16: bool = False, acts: Optional[list] = None) -> Tensor:
    """nerf.py:110-134 (use_viewdirs=True branch) -> [M,4] = (r,g,b,sigma) pre-activation.

    bf16=True emulates the CUDA kernel's numerics: every GEMM operand (weights and the
    activations fed to the next layer) rounded to bf16, products/accumulation and bias in
    fp32; the sigma head is taken in fp32 from the un-rounded layer-7 activation and the
    3-wide rgb head from the un-rounded view-branch activation (both run on CUDA cores in
    the kernel's epilogue)."""
    q = _bf16 if bf16 else (lambda t: t)

    def lin(name, x):
        return F.linear(x, q(params[name + ".weight"]), params[name + ".bias"])

    x_in = q(enc_pts)
    h = x_in
    for i in range(D):
        h = F.relu(lin("pts_linears.%d" % i, h))
        h_full = h
        h = q(h)
        if acts is not None:
            acts.append(h)
        if i in skips:
            h = torch.cat([x_in, h], -1)
            h_full = torch.cat([x_in, h_full], -1)
    if "output_linear.weight" in params:        # use_viewdirs=False (nerf.py:131-132); fp32 only
        assert not bf16
        return lin("output_linear", h)
    if bf16:
        alpha = F.linear(h_full, params["alpha_linear.weight"], params["alpha_linear.bias"])
    else:
        alpha = lin("alpha_linear", h)
    feat = q(lin("feature_linear", h))
    hv = F.relu(lin("views_linears.0", torch.cat([feat, q(enc_dirs)], -1)))
    if acts is not None:
        acts.extend([feat, q(hv)])
    if bf16:
        rgb = F.linear(hv, params["rgb_linear.weight"], params["rgb_linear.bias"])
    else:
        rgb = lin("rgb_linear", hv)
    return torch.cat([rgb, alpha], -1)


def nerf_query(params: Dict[str, Tensor], pts: Tensor, viewdirs: Tensor, multires: int = 10,
               multires_views: int = 4, bf16: bool = False, chunk: int = 65536) -> Tensor:
    """nerf.py:96-108: pts [N,S,3], viewdirs [N,3] -> raw [N,S,4]
    (use_viewdirs=False parameters: viewdirs is ignored, raw is [N,S,output_ch])."""
    flat = pts.reshape(-1, 3)
    noview = "output_linear.weight" in params
    dirs = None if noview else viewdirs[:, None].expand(pts.shape).reshape(-1, 3)
    outs = []
    for s in range(0, flat.shape[0], chunk):
        e = positional_encoding(flat[s:s + chunk], multires)
        ed = None if noview else positional_encoding(dirs[s:s + chunk], multires_views)
        outs.append(mlp(params, e, ed, bf16=bf16))
    out = torch.cat(outs, 0)
    return out.reshape(list(pts.shape[:-1]) + [out.shape[-1]])


# --------------------------------------------------------------------------------------
# volume compositing
# --------------------------------------------------------------------------------------


def composite(raw: Tensor, z: Tensor, rays_d: Tensor, white_bkgd: bool, noise: Optional[Tensor] = None):
    """render_utils.py:241-290 -> (rgb_map, disp_map, acc_map, weights, depth_map)."""
    dists = z[..., 1:] - z[..., :-1]
    dists = torch.cat([dists, torch.full_like(dists[..., :1], 1e10)], -1)
    dists = dists * torch.norm(rays_d[..., None, :], dim=-1)
    rgb = torch.sigmoid(raw[..., :3])
    sigma = raw[..., 3] if noise is None else raw[..., 3] + noise
    alpha = 1.0 - torch.exp(-F.relu(sigma) * dists)
    trans = torch.cumprod(torch.cat([torch.ones_like(alpha[:, :1]), 1.0 - alpha + 1e-10], -1), -1)[:, :-1]
    weights = alpha * trans
    rgb_map = torch.sum(weights[..., None] * rgb, -2)
    depth_map = torch.sum(weights * z, -1)
    acc_map = torch.sum(weights, -1)
    disp_map = 1.0 / torch.max(1e-10 * torch.ones_like(depth_map), depth_map / acc_map)
    if white_bkgd:
        rgb_map = rgb_map + (1.0 - acc_map[..., None])
    return rgb_map, disp_map, acc_map, weights, depth_map


# --------------------------------------------------------------------------------------
# the whole path
# --------------------------------------------------------------------------------------


def render_rays(params_c, params_f, rays_o, rays_d, viewdirs, near, far, *, n_samples=64,
                n_importance=128, white_bkgd=True, lindisp=False, t_rand=None, u=None,
                noise_c=None, noise_f=None, multires=10, multires_views=4, bf16=False):
    """render_utils.py:67-174 for a flat batch of rays.  near/far scalars or [N,1]."""
    N = rays_o.shape[0]
    near = near if torch.is_tensor(near) else torch.full((N, 1), float(near))
    far = far if torch.is_tensor(far) else torch.full((N, 1), float(far))
    z = stratified_z(near, far, n_samples, t_rand, lindisp)
    raw = nerf_query(params_c, points_on_rays(rays_o, rays_d, z), viewdirs, multires, multires_views, bf16)
    rgb, disp, acc, w, depth = composite(raw, z, rays_d, white_bkgd, noise_c)
    out = {"rgb_map": rgb, "disp_map": disp, "acc_map": acc, "raw": raw, "z_vals": z, "weights": w,
           "depth_map": depth}
    if n_importance > 0:
        z_all, z_std, z_samples = hierarchical_z(z, w, n_importance, u)
        pf = params_f if params_f is not None else params_c
        raw_f = nerf_query(pf, points_on_rays(rays_o, rays_d, z_all), viewdirs, multires, multires_views, bf16)
        rgb_f, disp_f, acc_f, w_f, depth_f = composite(raw_f, z_all, rays_d, white_bkgd, noise_f)
        out.update({"rgb0": rgb, "disp0": disp, "acc0": acc, "raw0": raw, "z_vals0": z, "weights0": w,
                    "rgb_map": rgb_f, "disp_map": disp_f, "acc_map": acc_f, "raw": raw_f, "z_vals": z_all,
                    "weights": w_f, "depth_map": depth_f, "z_std": z_std, "z_samples": z_samples})
    return out


def render_image(params_c, params_f, H, W, K, c2w, near, far, *, ndc=False, chunk=32768, **kw):
    """render_utils.py:176-238 with c2w given (full image)."""
    rays_o, rays_d = get_rays(H, W, K, c2w)
    viewdirs = unit_viewdirs(rays_d)
    if ndc:
        rays_o, rays_d = ndc_rays(H, W, float(K[0][0]), 1.0, rays_o, rays_d)
    ro = rays_o.reshape(-1, 3).float()
    rd = rays_d.reshape(-1, 3).float()
    parts = []
    for s in range(0, ro.shape[0], chunk):
        parts.append(render_rays(params_c, params_f, ro[s:s + chunk], rd[s:s + chunk], viewdirs[s:s + chunk],
                                 near, far, **kw))
    keys = ["rgb_map", "disp_map", "acc_map"]
    cat = {k: torch.cat([p[k] for p in parts], 0) for k in keys}
    return (cat["rgb_map"].reshape(H, W, 3), cat["disp_map"].reshape(H, W), cat["acc_map"].reshape(H, W))


def mse(x: Tensor, y: Tensor) -> Tensor:
    """utils.py:24."""
    return torch.mean((x - y) ** 2)


def psnr(m: Tensor) -> Tensor:
    """utils.py:27."""
    return -10.0 * torch.log(m) / math.log(10.0)


def train_step_grads(params_c, params_f, rays_o, rays_d, target, near, far, **kw):
    """main.py:77-103: loss = mse(rgb_fine) + mse(rgb_coarse); returns (loss, grads_c, grads_f, outputs)."""
    pc = {k: v.clone().requires_grad_(True) for k, v in params_c.items()}
    pf = {k: v.clone().requires_grad_(True) for k, v in params_f.items()}
    viewdirs = unit_viewdirs(rays_d)
    out = render_rays(pc, pf, rays_o, rays_d, viewdirs, near, far, **kw)
    loss = mse(out["rgb_map"], target) + mse(out["rgb0"], target)
    loss.backward()
    return loss.detach(), {k: v.grad for k, v in pc.items()}, {k: v.grad for k, v in pf.items()}, out


def adam_step(p: Tensor, g: Tensor, m: Tensor, v: Tensor, step: int, lr: float, b1=0.9, b2=0.999, eps=1e-8):
    """torch.optim.Adam (utils.py:170) single-tensor update, step is 1-based."""
    m = b1 * m + (1 - b1) * g
    v = b2 * v + (1 - b2) * g * g
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)) + eps
    return p - (lr / bc1) * (m / denom), m, v


# --------------------------------------------------------------------------------------
# synthetic workloads
# --------------------------------------------------------------------------------------


def pose_spherical(theta_deg: float, phi_deg: float, radius: float) -> np.ndarray:
    """Standard blender orbit pose (load_LINEMOD.py:29-34), float32 [4,4]."""
    def trans_t(t):
        return np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, t], [0, 0, 0, 1]], dtype=np.float32)

    def rot_phi(phi):
        return np.array([[1, 0, 0, 0], [0, np.cos(phi), -np.sin(phi), 0], [0, np.sin(phi), np.cos(phi), 0],
                         [0, 0, 0, 1]], dtype=np.float32)

    def rot_theta(th):
        return np.array([[np.cos(th), 0, -np.sin(th), 0], [0, 1, 0, 0], [np.sin(th), 0, np.cos(th), 0],
                         [0, 0, 0, 1]], dtype=np.float32)

    c2w = trans_t(radius)
    c2w = rot_phi(phi_deg / 180.0 * np.pi) @ c2w
    c2w = rot_theta(theta_deg / 180.0 * np.pi) @ c2w
    c2w = np.array([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]], dtype=np.float32) @ c2w
    return c2w.astype(np.float32)


def blender_intrinsics(H: int, W: int) -> np.ndarray:
    focal = 0.5 * W / np.tan(0.5 * 0.6911112070083618)
    return np.array([[focal, 0, 0.5 * W], [0, focal, 0.5 * H], [0, 0, 1]])


def llff_intrinsics(H: int = 378, W: int = 504, focal: float = 407.5625) -> np.ndarray:
    return np.array([[focal, 0, 0.5 * W], [0, focal, 0.5 * H], [0, 0, 1]])


def llff_pose(seed: int = 0) -> np.ndarray:
    rng = np.random.default_rng(seed)
    w = rng.normal(0, 0.02, 3)
    th = np.linalg.norm(w)
    k = w / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    R = np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * Kx @ Kx
    c2w = np.eye(4, dtype=np.float32)
    c2w[:3, :3] = R
    c2w[:3, 3] = rng.normal(0, 0.05, 3)
    return c2w.astype(np.float32)
