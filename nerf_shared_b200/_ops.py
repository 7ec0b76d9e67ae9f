"""torch.autograd.Function wrappers around the C ABI (include/nerf_b200.h).

Tensors stay ordinary torch tensors owned by the caching allocator; every kernel is launched on
torch's current stream.  Nothing here falls back to eager PyTorch math: a missing library or a
non-sm_100 device raises (see _lib.py).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import check, ensure_device, f32c, ptr, stream_ptr

_const_cache = {}


def linspace01(n: int, device) -> torch.Tensor:
    """torch.linspace(0, 1, n) evaluated by the CPU ATen kernel (the reference's oracle device;
    its bits are not reproducible by a closed form) and cached on `device`."""
    key = (int(n), str(device))
    t = _const_cache.get(key)
    if t is None:
        t = torch.linspace(0.0, 1.0, steps=int(n), device="cpu", dtype=torch.float32).to(device)
        _const_cache[key] = t
    return t


def _lib_for(t: torch.Tensor):
    ensure_device(t.device)
    return _lib.load()


class _Fn(torch.autograd.Function):
    """autograd.Function whose apply() materialises lazy ray grids first: Function.apply does not go through
    __torch_function__, so a LazyRayGrid would enter forward() as an opaque leaf and its gradient would never reach
    the pose."""

    @classmethod
    def apply(cls, *args):
        return super().apply(*[a._full() if isinstance(a, LazyRayGrid) else a for a in args])


# ----------------------------------------------------------------------------------------- rays
class GetRays(torch.autograd.Function):
    """utils.get_rays (utils.py:33-42); differentiable wrt c2w (demo_est_rel_pose.py:80-82)."""

    @staticmethod
    def forward(ctx, c2w, H, W, fx, fy, cx, cy, pix):
        lib = _lib_for(c2w)
        c = f32c(c2w)
        n = H * W if pix is None else pix.numel()
        rays_o = torch.empty((n, 3), device=c.device, dtype=torch.float32)
        rays_d = torch.empty((n, 3), device=c.device, dtype=torch.float32)
        with torch.cuda.device(c.device):
            check(lib.nb_get_rays_sel_fwd(H, W, fx, fy, cx, cy, ptr(c), c.stride(0), ptr(pix), n, ptr(rays_o),
                                          ptr(rays_d), None, stream_ptr()), "nb_get_rays")
        ctx.geom = (H, W, fx, fy, cx, cy)
        ctx.pix = pix
        ctx.c2w_shape = tuple(c2w.shape)
        if pix is None:
            rays_o, rays_d = rays_o.view(H, W, 3), rays_d.view(H, W, 3)
        return rays_o, rays_d

    @staticmethod
    def backward(ctx, g_o, g_d):
        H, W, fx, fy, cx, cy = ctx.geom
        ref = g_d if g_d is not None else g_o
        lib = _lib_for(ref)
        n = H * W if ctx.pix is None else ctx.pix.numel()
        go = f32c(g_o).view(-1, 3) if g_o is not None else None
        gd = f32c(g_d).view(-1, 3) if g_d is not None else None
        d34 = torch.zeros((3, 4), device=ref.device, dtype=torch.float32)
        with torch.cuda.device(ref.device):
            check(lib.nb_get_rays_bwd(H, W, fx, fy, cx, cy, ptr(ctx.pix), n, ptr(go), ptr(gd), ptr(d34), stream_ptr()),
                  "nb_get_rays_bwd")
        g = torch.zeros(ctx.c2w_shape, device=ref.device, dtype=torch.float32)
        g[:3, :4] = d34
        return g, None, None, None, None, None, None, None


import numpy as np  # noqa: E402


class LazyRayGrid(torch.Tensor):
    """rays_o or rays_d of a pinhole camera as a [H, W, 3] tensor that is computed on demand (see get_rays)."""

    @staticmethod
    def make(geom, c2w, which, shared):
        t = torch.Tensor._make_wrapper_subclass(LazyRayGrid, (geom[0], geom[1], 3), dtype=torch.float32, device=c2w.device,
                                                requires_grad=True)
        t._nb = (geom, c2w, which, shared)
        return t

    @classmethod
    def __torch_dispatch__(cls, func, types, args=(), kwargs=None):
        # only reached by calls that bypass the Python API (the autograd layer is already behind us here): values only
        unwrap = lambda a: a._full().detach() if isinstance(a, LazyRayGrid) else a
        return func(*torch.utils._pytree.tree_map(unwrap, args), **torch.utils._pytree.tree_map(unwrap, kwargs or {}))

    def _full(self):
        geom, c2w, which, shared = self._nb
        if "full" not in shared:
            shared["full"] = GetRays.apply(c2w, *geom, None)
        return shared["full"][which]

    def _select(self, ys, xs):
        geom, c2w, which, shared = self._nb
        key = (id(ys), id(xs))
        if shared.get("sel_key") != key:          # rays_o[ys, xs] and rays_d[ys, xs] share one launch
            if torch.is_tensor(ys):
                pix = (ys.to(device=c2w.device, dtype=torch.int64) * geom[1] + xs.to(device=c2w.device, dtype=torch.int64))
            else:                              # NumPy pixel coordinates (the demo): one small host array, one copy
                pix = torch.from_numpy(ys.astype(np.int64) * geom[1] + xs.astype(np.int64)).to(c2w.device, non_blocking=True)
            shared["sel"] = GetRays.apply(c2w, *geom, pix.reshape(-1).contiguous())
            shared["sel_key"], shared["sel_keep"] = key, (ys, xs)
        return shared["sel"][which]

    def __getitem__(self, idx):
        # a Python-level slot: the [ys, xs] selection never enters torch's generic indexing path (which costs ~0.4 ms per
        # call on a wrapper subclass with NumPy index arrays)
        if isinstance(idx, tuple) and len(idx) == 2 and _is_int_index(idx[0]) and _is_int_index(idx[1]) \
                and len(idx[0]) == len(idx[1]):
            return self._select(idx[0], idx[1])
        return self._full()[idx]

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        if getattr(func, "__name__", "") == "__get__" or func in (torch.Tensor.size, torch.Tensor.dim, torch.Tensor.numel):
            with torch._C.DisableTorchFunctionSubclass():       # metadata (shape, dtype, device, ...) lives on the wrapper
                return func(*args, **kwargs)
        if func is torch.Tensor.__getitem__ and isinstance(args[0], LazyRayGrid) and isinstance(args[1], tuple) \
                and len(args[1]) == 2 and all(_is_int_index(a) for a in args[1]):
            ys, xs = args[1]
            if np.shape(ys) == np.shape(xs) and np.ndim(ys) == 1:
                return args[0]._select(ys, xs)
        with torch._C.DisableTorchFunctionSubclass():
            unwrap = lambda a: a._full() if isinstance(a, LazyRayGrid) else a
            args = torch.utils._pytree.tree_map(unwrap, args)
            kwargs = torch.utils._pytree.tree_map(unwrap, kwargs)
            return func(*args, **kwargs)


def _is_int_index(a):
    if torch.is_tensor(a):
        return a.dtype in (torch.int64, torch.int32) and a.dim() == 1
    return isinstance(a, np.ndarray) and a.dtype.kind in "iu" and a.ndim == 1


class Normalize(_Fn):
    """v / |v| over the last dim (render_utils.py:211)."""

    @staticmethod
    def forward(ctx, v):
        lib = _lib_for(v)
        vc = f32c(v).view(-1, 3)
        out = torch.empty_like(vc)
        with torch.cuda.device(v.device):
            check(lib.nb_normalize_fwd(ptr(vc), vc.shape[0], ptr(out), stream_ptr()), "nb_normalize_fwd")
        ctx.save_for_backward(vc)
        ctx.shape = v.shape
        return out.view(v.shape)

    @staticmethod
    def backward(ctx, g):
        (vc,) = ctx.saved_tensors
        lib = _lib_for(vc)
        gc = f32c(g).view(-1, 3)
        dv = torch.empty_like(vc)
        with torch.cuda.device(vc.device):
            check(lib.nb_normalize_bwd(ptr(vc), ptr(gc), vc.shape[0], ptr(dv), stream_ptr()), "nb_normalize_bwd")
        return dv.view(ctx.shape)


class NdcRays(_Fn):
    """utils.ndc_rays (utils.py:54-71)."""

    @staticmethod
    def forward(ctx, rays_o, rays_d, H, W, focal, near):
        lib = _lib_for(rays_d)
        o, d = f32c(rays_o).view(-1, 3), f32c(rays_d).view(-1, 3)
        if o.data_ptr() == rays_o.data_ptr() and not rays_o.is_contiguous():
            o = o.contiguous()
        oo, od = torch.empty_like(o), torch.empty_like(d)
        with torch.cuda.device(d.device):
            check(lib.nb_ndc_rays_fwd(H, W, float(focal), float(near), ptr(o), ptr(d), o.shape[0], ptr(oo), ptr(od),
                                      stream_ptr()), "nb_ndc_rays_fwd")
        ctx.save_for_backward(o, d)
        ctx.args = (H, W, float(focal), float(near))
        ctx.shape = rays_d.shape
        return oo.view(rays_d.shape), od.view(rays_d.shape)

    @staticmethod
    def backward(ctx, g_o, g_d):
        o, d = ctx.saved_tensors
        lib = _lib_for(d)
        H, W, focal, near = ctx.args
        go = f32c(g_o).view(-1, 3) if g_o is not None else None
        gd = f32c(g_d).view(-1, 3) if g_d is not None else None
        do, dd = torch.empty_like(o), torch.empty_like(d)
        with torch.cuda.device(d.device):
            check(lib.nb_ndc_rays_bwd(H, W, focal, near, ptr(o), ptr(d), ptr(go), ptr(gd), o.shape[0], ptr(do), ptr(dd),
                                      stream_ptr()), "nb_ndc_rays_bwd")
        return do.view(ctx.shape), dd.view(ctx.shape), None, None, None, None


# ------------------------------------------------------------------------------------- sampling
def sample_coarse(ray_batch: torch.Tensor, n_samples: int, t_rand, lindisp: bool, want_pts: bool = False):
    """render_utils.py:105-131 -> z [N,S] (and pts [N,S,3]); no gradient (near/far are constants)."""
    lib = _lib_for(ray_batch)
    rb = f32c(ray_batch.detach())
    n, ld = rb.shape
    z = torch.empty((n, n_samples), device=rb.device, dtype=torch.float32)
    pts = torch.empty((n, n_samples, 3), device=rb.device, dtype=torch.float32) if want_pts else None
    tr = f32c(t_rand) if t_rand is not None else None
    if tr is not None:
        assert tuple(tr.shape) == (n, n_samples)
    with torch.cuda.device(rb.device):
        check(lib.nb_sample_coarse(ptr(rb), n, ld, ptr(linspace01(n_samples, rb.device)), ptr(tr), n_samples,
                                   int(bool(lindisp)), ptr(z), ptr(pts), stream_ptr()), "nb_sample_coarse")
    return (z, pts) if want_pts else z


class RayPoints(_Fn):
    """pts = o + d * z (render_utils.py:131,148)."""

    @staticmethod
    def forward(ctx, rays_o, rays_d, z):
        lib = _lib_for(z)
        o, d, zc = f32c(rays_o), f32c(rays_d), f32c(z)
        n, S = zc.shape
        pts = torch.empty((n, S, 3), device=zc.device, dtype=torch.float32)
        with torch.cuda.device(zc.device):
            check(lib.nb_ray_points_fwd(ptr(o), ptr(d), ptr(zc), n, S, ptr(pts), stream_ptr()), "nb_ray_points_fwd")
        ctx.save_for_backward(zc)
        return pts

    @staticmethod
    def backward(ctx, g):
        (zc,) = ctx.saved_tensors
        lib = _lib_for(zc)
        n, S = zc.shape
        gc = f32c(g)
        need_o, need_d = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        do = torch.empty((n, 3), device=zc.device, dtype=torch.float32) if need_o else None
        dd = torch.empty((n, 3), device=zc.device, dtype=torch.float32) if need_d else None
        if need_o or need_d:
            with torch.cuda.device(zc.device):
                check(lib.nb_ray_points_bwd(ptr(zc), ptr(gc), n, S, ptr(do), ptr(dd), stream_ptr()), "nb_ray_points_bwd")
        return do, dd, None


def sample_pdf_merge(z, weights, n_importance: int, u=None, want_samples: bool = False):
    """utils.sample_pdf + sort/cat/std (utils.py:74-117, render_utils.py:141-147,168); detached."""
    lib = _lib_for(z)
    zc, wc = f32c(z.detach()), f32c(weights.detach())
    n, S = zc.shape
    z_out = torch.empty((n, S + n_importance), device=zc.device, dtype=torch.float32)
    z_std = torch.empty((n,), device=zc.device, dtype=torch.float32)
    z_samples = torch.empty((n, n_importance), device=zc.device, dtype=torch.float32) if want_samples else None
    uc = f32c(u) if u is not None else None
    with torch.cuda.device(zc.device):
        check(lib.nb_sample_pdf_merge(ptr(zc), ptr(wc), ptr(uc), ptr(linspace01(n_importance, zc.device)), n, S,
                                      n_importance, ptr(z_out), ptr(z_std), ptr(z_samples), stream_ptr()),
              "nb_sample_pdf_merge")
    return (z_out, z_std, z_samples) if want_samples else (z_out, z_std)


def invert_cdf(bins, cdf, u):
    """utils.py:100-115 from an explicit cdf -> (samples, searchsorted indices)."""
    lib = _lib_for(cdf)
    b, c, uc = f32c(bins), f32c(cdf), f32c(u)
    n, B = c.shape
    I = uc.shape[-1]
    samples = torch.empty((n, I), device=c.device, dtype=torch.float32)
    inds = torch.empty((n, I), device=c.device, dtype=torch.int64)
    with torch.cuda.device(c.device):
        check(lib.nb_invert_cdf(ptr(b), ptr(c), ptr(uc), n, B, I, ptr(samples), ptr(inds), stream_ptr()), "nb_invert_cdf")
    return samples, inds


def ray_rows(rays_o, rays_d, view_src, near, far):
    """-> ([n,11] ray rows, [n,3] unit view directions) for rays that carry no gradient (render_utils.py:211,223-226)."""
    lib = _lib_for(rays_o)
    o, d, v = f32c(rays_o.detach()), f32c(rays_d.detach()), f32c(view_src.detach())
    n = o.shape[0]
    rows = torch.empty((n, 11), device=o.device, dtype=torch.float32)
    vd = torch.empty((n, 3), device=o.device, dtype=torch.float32)
    with torch.cuda.device(o.device):
        check(lib.nb_ray_rows(ptr(o), ptr(d), ptr(v), float(near), float(far), n, ptr(rows), ptr(vd), stream_ptr()), "nb_ray_rows")
    return rows, vd


class MsePair(_Fn):
    """(loss, img_loss, img_loss0, psnr) of the training step: img2mse on the fine and coarse maps, their sum and
    mse2psnr of the fine term (main.py:87-99, utils.py:24,27) in one launch; one launch backward."""

    @staticmethod
    def forward(ctx, rgb, rgb0, target):
        lib = _lib_for(rgb)
        a, t = f32c(rgb), f32c(target)
        b = f32c(rgb0) if rgb0 is not None else None
        if a.shape != t.shape or (b is not None and b.shape != t.shape):
            raise ValueError("MsePair: shapes differ: %s %s %s" % (tuple(a.shape), None if b is None else tuple(b.shape), tuple(t.shape)))
        out = torch.empty(4, device=a.device, dtype=torch.float32)
        with torch.cuda.device(a.device):
            check(lib.nb_mse_pair_fwd(ptr(a), ptr(b), ptr(t), a.numel(), ptr(out), stream_ptr()), "nb_mse_pair_fwd")
        ctx.save_for_backward(a, b, t)
        ctx.set_materialize_grads(False)
        loss, m, m0, psnr = out[2], out[0], out[1], out[3]
        ctx.mark_non_differentiable(psnr)
        return loss, m, m0, psnr

    @staticmethod
    def backward(ctx, g_loss, g_m, g_m0, _g_psnr):
        a, b, t = ctx.saved_tensors
        lib = _lib_for(a)
        gs = [f32c(x) if x is not None else None for x in (g_m, g_m0, g_loss)]
        d_a = torch.empty_like(a)
        d_b = torch.empty_like(b) if (b is not None and ctx.needs_input_grad[1]) else None
        with torch.cuda.device(a.device):
            check(lib.nb_mse_pair_bwd(ptr(a), ptr(b), ptr(t), a.numel(), ptr(gs[0]), ptr(gs[1]), ptr(gs[2]), ptr(d_a), ptr(d_b),
                                      stream_ptr()), "nb_mse_pair_bwd")
        return d_a, d_b, None


# ----------------------------------------------------------------------------------- compositing
class Composite(_Fn):
    """Renderer.raw2outputs (render_utils.py:241-290)."""

    @staticmethod
    def forward(ctx, raw, z, rays_d, noise, white_bkgd):
        lib = _lib_for(raw)
        r, zc, d = f32c(raw), f32c(z), f32c(rays_d)
        nz = f32c(noise) if noise is not None else None
        n, S = zc.shape
        dev = r.device
        rgb = torch.empty((n, 3), device=dev, dtype=torch.float32)
        disp = torch.empty((n,), device=dev, dtype=torch.float32)
        acc = torch.empty((n,), device=dev, dtype=torch.float32)
        depth = torch.empty((n,), device=dev, dtype=torch.float32)
        weights = torch.empty((n, S), device=dev, dtype=torch.float32)
        with torch.cuda.device(dev):
            check(lib.nb_composite_fwd(ptr(r), ptr(zc), ptr(d), ptr(nz), n, S, int(bool(white_bkgd)), ptr(rgb), ptr(disp),
                                       ptr(acc), ptr(weights), ptr(depth), stream_ptr()), "nb_composite_fwd")
        ctx.save_for_backward(r, zc, d, nz)
        ctx.white = int(bool(white_bkgd))
        ctx.set_materialize_grads(False)     # unused outputs arrive as None, not as zero tensors the kernel would read
        ctx.mark_non_differentiable(weights)
        return rgb, disp, acc, weights, depth

    @staticmethod
    def backward(ctx, g_rgb, g_disp, g_acc, g_w, g_depth):
        r, zc, d, nz = ctx.saved_tensors
        lib = _lib_for(r)
        n, S = zc.shape
        c = lambda g: f32c(g) if g is not None else None
        d_raw = torch.empty_like(r)
        d_rd = torch.empty_like(d) if ctx.needs_input_grad[2] else None
        with torch.cuda.device(r.device):
            check(lib.nb_composite_bwd(ptr(r), ptr(zc), ptr(d), ptr(nz), n, S, ctx.white, ptr(c(g_rgb)), ptr(c(g_disp)),
                                       ptr(c(g_acc)), ptr(c(g_depth)), ptr(d_raw), ptr(d_rd), stream_ptr()),
                  "nb_composite_bwd")
        return d_raw, None, d_rd, None, None


# ------------------------------------------------------------------------------------------ MLP
def mlp_params(model):
    """The 24 parameters in the C ABI's order (cached on the module: ModuleList indexing is slow)."""
    cached = model.__dict__.get("_nb_params")
    if cached is not None and cached[0] is model.pts_linears[0].weight and cached[-1] is model.rgb_linear.bias:
        return cached
    pl = list(model.pts_linears)
    ps = ([l.weight for l in pl] + [l.bias for l in pl] +
          [model.feature_linear.weight, model.feature_linear.bias, model.alpha_linear.weight,
           model.alpha_linear.bias, model.views_linears[0].weight, model.views_linears[0].bias,
           model.rgb_linear.weight, model.rgb_linear.bias])
    model.__dict__["_nb_params"] = ps
    return ps


def mlp_weight_struct(model) -> _lib.NbMlpWeights:
    """Fill NbMlpWeights from a nerf.NeRF module's fp32 parameters (no copies); cached per set of
    parameter storages."""
    ps = mlp_params(model)
    key = tuple(p.data_ptr() for p in ps)
    cached = model.__dict__.get("_nb_wstruct")
    if cached is not None and cached[0] == key:
        return cached[1]
    w = _lib.NbMlpWeights()
    for i in range(8):
        w.pts_w[i] = key[i]
        w.pts_b[i] = key[8 + i]
    w.feature_w, w.feature_b, w.alpha_w, w.alpha_b = key[16], key[17], key[18], key[19]
    w.views_w, w.views_b, w.rgb_w, w.rgb_b = key[20], key[21], key[22], key[23]
    w.L_pts, w.L_dirs = int(model.multires), int(model.multires_views)
    model.__dict__["_nb_wstruct"] = (key, w)
    return w


def packed_weights(model, backward: bool = False) -> torch.Tensor:
    """bf16 operand blob + fp32 side table for the tensor-core kernels (backward=True: the
    transposed slabs of the dgrad chain), re-packed whenever a parameter changed (its autograd
    version counter or storage moved)."""
    params = mlp_params(model)
    dev = params[0].device
    # Staleness key, evaluated on every forward / backward: version counters only grow, so their sum changes whenever
    # any parameter is written in place (optimizer step, load_state_dict, `p += ...` under no_grad); the storage of the
    # first and the last parameter catches a re-homing of the parameters (FlatAdam moves them into its flat buffer).
    v = 0
    for p in params:
        v += p._version
    key = (v, params[0].data_ptr(), params[23].data_ptr())
    slot = "_nb_packed_bwd" if backward else "_nb_packed"
    cache = model.__dict__.get(slot)
    if cache is not None and cache[0] == key and cache[1].device == dev:
        return cache[1]
    for p in params:
        if p.dtype != torch.float32 or not p.is_contiguous():
            raise _lib.NerfB200Error("NeRF parameters must be contiguous fp32 tensors")
    lib = _lib_for(params[0])
    size_fn = lib.nb_mlp_bwd_packed_bytes if backward else lib.nb_mlp_packed_bytes
    nbytes = int(size_fn(int(model.multires), int(model.multires_views)))
    buf = cache[1] if (cache is not None and cache[1].device == dev and cache[1].numel() == nbytes) else \
        torch.empty((nbytes,), device=dev, dtype=torch.uint8)
    w = mlp_weight_struct(model)
    with torch.cuda.device(dev):
        if backward:
            check(lib.nb_mlp_pack_bwd(C.byref(w), ptr(buf), stream_ptr()), "nb_mlp_pack_bwd")
        else:
            check(lib.nb_mlp_pack(C.byref(w), ptr(buf), stream_ptr()), "nb_mlp_pack")
    model.__dict__[slot] = (key, buf)
    return buf


def mlp_forward_f32(model, pts, viewdirs, S):
    lib = _lib_for(pts)
    p, v = f32c(pts).view(-1, 3), f32c(viewdirs).view(-1, 3)
    M = p.shape[0]
    raw = torch.empty((M, 4), device=p.device, dtype=torch.float32)
    w = mlp_weight_struct(model)
    with torch.cuda.device(p.device):
        check(lib.nb_mlp_fwd_f32(C.byref(w), ptr(p), ptr(v), M, S, ptr(raw), stream_ptr()), "nb_mlp_fwd_f32")
    return raw


def mlp_forward_f32_noview(model, pts):
    """use_viewdirs=False model (nerf.py:93-94,131-132): pts [M,3] -> raw [M,output_ch]; fp32, forward only."""
    lib = _lib_for(pts)
    p = f32c(pts.detach()).view(-1, 3)
    M = p.shape[0]
    pl = list(model.pts_linears)
    ps = [l.weight for l in pl] + [l.bias for l in pl] + [model.output_linear.weight, model.output_linear.bias]
    for t in ps:
        if t.dtype != torch.float32 or not t.is_contiguous() or t.device != p.device:
            raise _lib.NerfB200Error("NeRF parameters must be contiguous fp32 tensors on the device of the points")
    w = _lib.NbMlpWeights()
    for i in range(8):
        w.pts_w[i], w.pts_b[i] = ps[i].data_ptr(), ps[8 + i].data_ptr()
    w.L_pts, w.L_dirs = int(model.multires), 0
    out_ch = int(ps[16].shape[0])
    raw = torch.empty((M, out_ch), device=p.device, dtype=torch.float32)
    with torch.cuda.device(p.device):
        check(lib.nb_mlp_fwd_f32_noview(C.byref(w), ptr(ps[16]), ptr(ps[17]), out_ch, ptr(p), M, ptr(raw), stream_ptr()),
              "nb_mlp_fwd_f32_noview")
    return raw


def mlp_forward_tc(model, pts, viewdirs, S, save_acts=False):
    lib = _lib_for(pts)
    p, v = f32c(pts).view(-1, 3), f32c(viewdirs).view(-1, 3)
    M = p.shape[0]
    packed = packed_weights(model)
    raw = torch.empty((M, 4), device=p.device, dtype=torch.float32)
    acts = None
    if save_acts:
        acts = torch.empty((int(lib.nb_mlp_act_bytes(M)),), device=p.device, dtype=torch.uint8)
    with torch.cuda.device(p.device):
        check(lib.nb_mlp_fwd_tc(ptr(packed), int(model.multires), int(model.multires_views), ptr(p), ptr(v), M, S,
                                ptr(raw), ptr(acts), stream_ptr()), "nb_mlp_fwd_tc")
    return raw, acts


class NerfMLP(_Fn):
    """NeRF.forward (nerf.py:96-134): pts [M,3], viewdirs [M/S,3] -> raw [M,4]."""

    @staticmethod
    def forward(ctx, pts, viewdirs, S, model, track, *params):
        # `track` = torch.is_grad_enabled() at the call site (grad mode is always off in here)
        need_grad = bool(track) and any(ctx.needs_input_grad)
        if model.precision == "fp32":
            # inference / parity path: exact fp32 operands, no backward kernels
            raw = mlp_forward_f32(model, pts, viewdirs, S)
            ctx.mark_non_differentiable(raw)
            return raw
        raw, acts = mlp_forward_tc(model, pts, viewdirs, S, save_acts=need_grad)
        if need_grad:
            ctx.model, ctx.S = model, S
            ctx.save_for_backward(f32c(pts).view(-1, 3), f32c(viewdirs).view(-1, 3), acts)
        return raw

    @staticmethod
    def backward(ctx, g_raw):
        from . import _bwd
        return _bwd.nerf_mlp_backward(ctx, g_raw)


ACT_SLABS, DA_SLABS = 40, 38


def untile(buf: torch.Tensor, n_slabs: int, M: int) -> torch.Tensor:
    """Decode a tiled/swizzled slab buffer (acts or dA) to a logical [M, n_slabs*64] bf16 matrix
    (tests / debugging): chunk g of row r sits at physical chunk g ^ (r % 8)."""
    tiles = (M + 127) // 128
    # (the activation buffer carries the relu' bit masks after its slab tiles: decode the slab part only)
    x = buf[:tiles * n_slabs * 16384].view(torch.bfloat16).view(tiles, n_slabs, 128, 8, 8)
    r = torch.arange(128, device=buf.device)
    g = torch.arange(8, device=buf.device)
    phys = (g[None, :] ^ (r[:, None] % 8))                       # [128, 8]
    idx = phys[None, None, :, :, None].expand(tiles, n_slabs, 128, 8, 8)
    logical = torch.gather(x, 3, idx)                            # [tiles, slabs, 128, 8, 8]
    return logical.permute(0, 2, 1, 3, 4).reshape(tiles * 128, n_slabs * 64)[:M]


def debug_umma_mn(A: torch.Tensor, B: torch.Tensor) -> torch.Tensor:
    lib = _lib_for(A)
    a, b = f32c(A), f32c(B)
    D = torch.empty((128, 128), device=a.device, dtype=torch.float32)
    with torch.cuda.device(a.device):
        check(lib.nb_debug_umma_mn(ptr(a), ptr(b), a.shape[0], ptr(D), stream_ptr()), "nb_debug_umma_mn")
    return D


def debug_umma(A: torch.Tensor, B: torch.Tensor) -> torch.Tensor:
    lib = _lib_for(A)
    a, b = f32c(A), f32c(B)
    D = torch.empty((128, 128), device=a.device, dtype=torch.float32)
    with torch.cuda.device(a.device):
        check(lib.nb_debug_umma(ptr(a), ptr(b), a.shape[1], ptr(D), stream_ptr()), "nb_debug_umma")
    return D


def composite_train(raw, z, rays_d, noise, target, white_bkgd, inv_n3, loss_sum, want_weights=False):
    """Compositing forward + this map's MSE term + its compositing backward in one launch (nb_composite_train):
    -> (rgb_map, disp, acc, weights or None, d_raw); loss_sum (1 float, zeroed by the caller) accumulates the term."""
    lib = _lib_for(raw)
    r, zc, d, t = f32c(raw), f32c(z), f32c(rays_d), f32c(target)
    nz = f32c(noise) if noise is not None else None
    n, S = zc.shape
    dev = r.device
    rgb = torch.empty((n, 3), device=dev, dtype=torch.float32)
    disp = torch.empty((n,), device=dev, dtype=torch.float32)
    acc = torch.empty((n,), device=dev, dtype=torch.float32)
    weights = torch.empty((n, S), device=dev, dtype=torch.float32) if want_weights else None
    d_raw = torch.empty_like(r)
    with torch.cuda.device(dev):
        check(lib.nb_composite_train(ptr(r), ptr(zc), ptr(d), ptr(nz), ptr(t), n, S, int(bool(white_bkgd)), float(inv_n3),
                                     ptr(rgb), ptr(disp), ptr(acc), ptr(weights), ptr(d_raw), ptr(loss_sum), stream_ptr()),
              "nb_composite_train")
    return rgb, disp, acc, weights, d_raw


# ------------------------------------------------------------------------ fp32 training path (csrc/mlp_f32_train.cu)
def _f32_params(model):
    """Parameters in the order nb_mlp_f32_bwd's offsets use; (list, output_linear or None)."""
    if model.use_viewdirs:
        return mlp_params(model), None
    pl = list(model.pts_linears)
    return [l.weight for l in pl] + [l.bias for l in pl] + [model.output_linear.weight, model.output_linear.bias], model.output_linear


class NerfMLPf32(_Fn):
    """NeRF.forward with exact fp32 operands and a backward (layer-by-layer CUDA-core GEMMs, saved fp32 activations):
    encodings up to multires 15 / 6 and use_viewdirs=False models; parameter gradients only."""

    @staticmethod
    def forward(ctx, pts, viewdirs, S, model, *params):
        lib = _lib_for(pts)
        p = f32c(pts).view(-1, 3)
        M = p.shape[0]
        ps, out_lin = _f32_params(model)
        for t in ps:
            if t.dtype != torch.float32 or not t.is_contiguous() or t.device != p.device:
                raise _lib.NerfB200Error("NeRF parameters must be contiguous fp32 tensors on the device of the points")
        w = _lib.NbMlpWeights()
        for i in range(8):
            w.pts_w[i], w.pts_b[i] = ps[i].data_ptr(), ps[8 + i].data_ptr()
        w.L_pts = int(model.multires)
        noview = out_lin is not None
        if noview:
            w.L_dirs, out_ch = 0, int(out_lin.weight.shape[0])
            v = None
        else:
            w.feature_w, w.feature_b, w.alpha_w, w.alpha_b = (ps[16].data_ptr(), ps[17].data_ptr(), ps[18].data_ptr(), ps[19].data_ptr())
            w.views_w, w.views_b, w.rgb_w, w.rgb_b = ps[20].data_ptr(), ps[21].data_ptr(), ps[22].data_ptr(), ps[23].data_ptr()
            w.L_dirs, out_ch = int(model.multires_views), 4
            v = f32c(viewdirs).view(-1, 3)
        acts = torch.empty((int(lib.nb_mlp_f32_act_floats(w.L_pts, w.L_dirs, int(noview), M)),), device=p.device, dtype=torch.float32)
        raw = torch.empty((M, out_ch), device=p.device, dtype=torch.float32)
        with torch.cuda.device(p.device):
            check(lib.nb_mlp_f32_fwd_train(C.byref(w), ptr(out_lin.weight) if noview else None, ptr(out_lin.bias) if noview else None,
                                           out_ch if noview else 0, ptr(p), ptr(v), M, int(S), ptr(raw), ptr(acts), stream_ptr()),
                  "nb_mlp_f32_fwd_train")
        ctx.w, ctx.model, ctx.M, ctx.noview, ctx.out_ch = w, model, M, noview, out_ch
        ctx.save_for_backward(acts)
        return raw

    @staticmethod
    def backward(ctx, g_raw):
        (acts,) = ctx.saved_tensors
        lib = _lib_for(acts)
        model, M = ctx.model, ctx.M
        ps, out_lin = _f32_params(model)
        offs, total = [], 0
        for t in ps:
            offs.append(total)
            total += t.numel()
        offs = offs + [0] * (24 - len(offs))
        grads = torch.zeros((total,), device=acts.device, dtype=torch.float32)
        scratch = torch.empty((2 * M * 256,), device=acts.device, dtype=torch.float32)
        g = f32c(g_raw).view(M, ctx.out_ch)
        with torch.cuda.device(acts.device):
            check(lib.nb_mlp_f32_bwd(C.byref(ctx.w), ptr(out_lin.weight) if ctx.noview else None, ctx.out_ch if ctx.noview else 0,
                                     ptr(acts), ptr(g), M, (C.c_int64 * 24)(*offs), ptr(grads), ptr(scratch), stream_ptr()),
                  "nb_mlp_f32_bwd")
        out = []
        for i, t in enumerate(ps):
            out.append(grads[offs[i]:offs[i] + t.numel()].view_as(t) if ctx.needs_input_grad[4 + i] else None)
        return (None, None, None, None) + tuple(out)


def mlp_f32_trainable(model, pts, viewdirs, S):
    """raw [M, 4 or output_ch] through the fp32 training path; gradients flow to the parameters only."""
    if pts.requires_grad or (viewdirs is not None and viewdirs.requires_grad):
        raise NotImplementedError("the fp32 training path differentiates with respect to the network parameters only "
                                  "(pose / input gradients: use the tensor-core path, precision='bf16')")
    ps, _ = _f32_params(model)
    return NerfMLPf32.apply(pts, viewdirs, S, model, *ps)


# ------------------------------------------------------------------------ inference: Renderer.render_rays as one C-ABI call
def tc_model_ok(m):
    """The tensor-core kernels' configuration (every shipped config but stonehenge's 15 / 6 encodings)."""
    return (getattr(m, "use_viewdirs", False) and m.precision == "bf16" and m.D == 8 and m.W == 256 and list(m.skips) == [4]
            and m.i_embed == 0 and m.multires <= 10 and m.multires_views <= 4)


def render_rays_inference(r, coarse, fine, rays_o, rays_d, view_src, step, retraw):
    """All rays of a render() call through nb_render_rays_fwd, `step` rays per call, outputs written straight into the
    full-size result tensors (no per-chunk concatenation).  Random draws come from the renderer's hooks in the reference's
    order per chunk: t_rand, coarse noise, u, fine noise."""
    lib = _lib_for(rays_o)
    o, d, v = f32c(rays_o), f32c(rays_d), f32c(view_src)
    n, dev = o.shape[0], o.device
    S, I = int(r.N_samples), int(r.N_importance)
    F = S + I
    f = lambda *shape: torch.empty(shape, device=dev, dtype=torch.float32)
    rgb, disp, acc = f(n, 3), f(n), f(n)
    rgb0, disp0, acc0, z_std = (f(n, 3), f(n), f(n), f(n)) if I > 0 else (None, None, None, None)
    raw = f(n, F, 4) if retraw else None
    pc = packed_weights(coarse)
    pf = packed_weights(fine) if (fine is not None and I > 0) else None
    L, Lv = int(coarse.multires), int(coarse.multires_views)
    step = max(1, min(int(step), n))
    ws_bytes = int(lib.nb_render_rays_workspace_bytes(step, S, I))
    ws = torch.empty((ws_bytes,), device=dev, dtype=torch.uint8)
    sl = lambda t, b, e: None if t is None else t[b:e]
    with torch.cuda.device(dev):
        for b in range(0, n, step):
            e = min(b + step, n)
            m = e - b
            t_rand = f32c(r._uniform((m, S), dev, "t_rand")) if r.perturb > 0.0 else None
            noise_c = f32c(r._normal((m, S), dev, "noise_c") * r.raw_noise_std) if r.raw_noise_std > 0.0 else None
            u = f32c(r._uniform((m, I), dev, "u")) if (I > 0 and r.perturb != 0.0) else None
            noise_f = f32c(r._normal((m, F), dev, "noise_f") * r.raw_noise_std) if (I > 0 and r.raw_noise_std > 0.0) else None
            check(lib.nb_render_rays_fwd(ptr(pc), ptr(pf), L, Lv, ptr(o[b:e]), ptr(d[b:e]), ptr(v[b:e]), float(r.near), float(r.far),
                                         m, S, I, ptr(linspace01(S, dev)), ptr(t_rand), ptr(linspace01(I, dev)) if I > 0 else None,
                                         ptr(u), ptr(noise_c), ptr(noise_f), int(bool(r.lindisp)), int(bool(r.white_bkgd)), ptr(ws),
                                         ws_bytes, ptr(sl(rgb, b, e)), ptr(sl(disp, b, e)), ptr(sl(acc, b, e)), ptr(sl(rgb0, b, e)),
                                         ptr(sl(disp0, b, e)), ptr(sl(acc0, b, e)), ptr(sl(z_std, b, e)), ptr(sl(raw, b, e)),
                                         stream_ptr()), "nb_render_rays_fwd")
    return rgb, disp, acc, rgb0, disp0, acc0, z_std, raw
