"""Renderer.render for a training batch as ONE autograd node, lazily captured into CUDA graphs.

main.py:77-104 calls render_from_rays / img2mse / backward / optimizer.step call by call; driven through the per-stage
autograd wrappers of _ops.py that is ~30 small Python-level operations per step and the loop is host-bound (1.9 ms for
a step whose kernels take 1.3 ms).  When the rays carry no gradient (training: they come from the dataset) the whole
coarse + fine render is issued here as a straight sequence of C-ABI launches in forward() and another in backward():
same kernels, same order, same random draws (Renderer._uniform / _normal hooks).  The coarse network's backward
depends only on the coarse map's gradient, so backward() issues it on a side stream beside the fine network's.

With the flat optimizer of utils.get_optimizer (gradients accumulate in place) the two sequences are captured once per
(renderer configuration, models, batch size) into a forward and a backward CUDA graph and replayed afterwards: the
unchanged main.py loop then costs a few host operations per step and runs at the speed of its kernels
(NB_EAGER_GRAPH=0 disables the capture).  Pose estimation (rays with gradients) keeps the per-stage wrappers.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _bwd, _lib, _ops
from ._lib import check, f32c, ptr, stream_ptr

_side_streams = {}
GRAPHS_ENABLED = True      # train.GraphedTrainStep switches the lazy capture off while it warms up / captures its own graph


def _side(dev):
    s = _side_streams.get(dev)
    if s is None:
        s = _side_streams[dev] = torch.cuda.Stream(device=dev)
    return s


def eligible(renderer, coarse, fine, rays_o, rays_d, n_rays, chunk_cap):
    """The fused node covers the training configuration of every shipped config (two D8/W256 view-dependent models on the
    tensor-core path, hierarchical sampling, one launch-sized batch, rays without gradient)."""
    if not torch.is_grad_enabled() or fine is None or renderer.N_importance <= 0 or not renderer.use_viewdirs:
        return False
    if n_rays == 0 or n_rays > chunk_cap or coarse is fine:
        return False
    for m in (coarse, fine):
        if not (getattr(m, "use_viewdirs", False) and m.precision == "bf16" and m.D == 8 and m.W == 256
                and list(m.skips) == [4] and m.i_embed == 0 and m.multires <= 10 and m.multires_views <= 4):
            return False
    return (rays_o.requires_grad or rays_d.requires_grad or any(p.requires_grad for p in coarse.parameters())
            or any(p.requires_grad for p in fine.parameters()))


def _forward_impl(r, coarse, fine, o, d, v, packed=None):
    """render_utils.py:204-290 for one batch -> (outputs, tensors the backward needs).  packed: (coarse, fine) operand
    blobs to (re)pack into unconditionally (graph capture); None = the cached blobs of _ops.packed_weights."""
    lib = _ops._lib_for(o)
    n, dev = o.shape[0], o.device
    S, I = int(r.N_samples), int(r.N_importance)
    F = S + I
    f = lambda *shape: torch.empty(shape, device=dev, dtype=torch.float32)
    rows, vd = f(n, 11), f(n, 3)
    z_c, pts_c, raw_c = f(n, S), f(n, S, 3), f(n, S, 4)
    maps_c = (f(n, 3), f(n), f(n))
    w_c = f(n, S)
    z_f, z_std, pts_f, raw_f = f(n, F), f(n), f(n, F, 3), f(n, F, 4)
    maps_f = (f(n, 3), f(n), f(n))
    acts_c = torch.empty((int(lib.nb_mlp_act_bytes(n * S)),), device=dev, dtype=torch.uint8)
    acts_f = torch.empty((int(lib.nb_mlp_act_bytes(n * F)),), device=dev, dtype=torch.uint8)
    white = int(bool(r.white_bkgd))
    L, Lv = int(coarse.multires), int(coarse.multires_views)

    def weights(model, k):
        if packed is None:
            return _ops.packed_weights(model)
        check(lib.nb_mlp_pack(C.byref(_ops.mlp_weight_struct(model)), ptr(packed[k]), stream_ptr()), "nb_mlp_pack")
        return packed[k]

    # random draws in the reference's order: t_rand, coarse noise, u, fine noise (render_utils.py:121,264; utils.py:86)
    t_rand = f32c(r._uniform((n, S), dev, "t_rand")) if r.perturb > 0.0 else None
    with torch.cuda.device(dev):
        check(lib.nb_ray_rows(ptr(o), ptr(d), ptr(v), float(r.near), float(r.far), n, ptr(rows), ptr(vd), stream_ptr()), "nb_ray_rows")
        check(lib.nb_sample_coarse(ptr(rows), n, 11, ptr(_ops.linspace01(S, dev)), ptr(t_rand), S, int(bool(r.lindisp)),
                                   ptr(z_c), ptr(pts_c), stream_ptr()), "nb_sample_coarse")
        check(lib.nb_mlp_fwd_tc(ptr(weights(coarse, 0)), L, Lv, ptr(pts_c), ptr(vd), n * S, S, ptr(raw_c), ptr(acts_c),
                                stream_ptr()), "nb_mlp_fwd_tc")
        noise_c = f32c(r._normal((n, S), dev, "noise_c") * r.raw_noise_std) if r.raw_noise_std > 0.0 else None
        check(lib.nb_composite_fwd(ptr(raw_c), ptr(z_c), ptr(d), ptr(noise_c), n, S, white, ptr(maps_c[0]), ptr(maps_c[1]),
                                   ptr(maps_c[2]), ptr(w_c), None, stream_ptr()), "nb_composite_fwd")
        u = f32c(r._uniform((n, I), dev, "u")) if r.perturb != 0.0 else None
        check(lib.nb_sample_pdf_merge(ptr(z_c), ptr(w_c), ptr(u), ptr(_ops.linspace01(I, dev)), n, S, I, ptr(z_f), ptr(z_std),
                                      None, stream_ptr()), "nb_sample_pdf_merge")
        check(lib.nb_ray_points_fwd(ptr(o), ptr(d), ptr(z_f), n, F, ptr(pts_f), stream_ptr()), "nb_ray_points_fwd")
        check(lib.nb_mlp_fwd_tc(ptr(weights(fine, 1)), L, Lv, ptr(pts_f), ptr(vd), n * F, F, ptr(raw_f), ptr(acts_f),
                                stream_ptr()), "nb_mlp_fwd_tc")
        noise_f = f32c(r._normal((n, F), dev, "noise_f") * r.raw_noise_std) if r.raw_noise_std > 0.0 else None
        check(lib.nb_composite_fwd(ptr(raw_f), ptr(z_f), ptr(d), ptr(noise_f), n, F, white, ptr(maps_f[0]), ptr(maps_f[1]),
                                   ptr(maps_f[2]), None, None, stream_ptr()), "nb_composite_fwd")
    outs = (maps_f[0], maps_f[1], maps_f[2], maps_c[0], maps_c[1], maps_c[2], z_std, raw_f)
    saved = (d, vd, z_c, raw_c, acts_c, pts_c, z_f, raw_f, acts_f, pts_f, noise_c, noise_f, v)
    return outs, saved


def _backward_impl(models, white, dims, saved, grads, need_c, need_f, packed_bwd=None, ray_grads=False):
    """grads = (g_rgb, g_disp, g_acc, g_rgb0, g_disp0, g_acc0, g_raw) -> (24 coarse, 24 fine parameter gradients,
    d_rays_o, d_rays_d, d_view_src).  ray_grads: also differentiate with respect to the rays (pose estimation,
    demo_est_rel_pose.py:80-97): through compositing (|rays_d| in the interval lengths), the MLP inputs (points and view
    directions, nb_mlp_bwd_inputs), pts = o + d z and the view-direction normalisation."""
    d, vd, z_c, raw_c, acts_c, pts_c, z_f, raw_f, acts_f, pts_f, noise_c, noise_f, v = saved
    g_rgb, g_disp, g_acc, g_rgb0, g_disp0, g_acc0, g_raw = grads
    coarse, fine = models
    n, S, F = dims
    lib = _ops._lib_for(d)
    dev = d.device
    c = lambda g: f32c(g) if g is not None else None
    main, side = torch.cuda.current_stream(dev), _side(dev)

    def branch(k, model, raw, z, noise, acts, pts, n_s, g3, need_p):
        """-> (24 parameter gradients, d_o, d_d, d_viewdirs) of one pass"""
        if (not any(need_p) and not ray_grads) or (all(g is None for g in g3[:3]) and g3[3] is None):
            return [None] * 24, None, None, None
        d_raw = torch.empty_like(raw)
        d_rd = torch.empty_like(d) if ray_grads else None
        pk = None
        with torch.cuda.device(dev):
            check(lib.nb_composite_bwd(ptr(raw), ptr(z), ptr(d), ptr(noise), n, n_s, white, ptr(c(g3[0])), ptr(c(g3[1])),
                                       ptr(c(g3[2])), None, ptr(d_raw), ptr(d_rd), stream_ptr()), "nb_composite_bwd")
            if packed_bwd is not None:
                check(lib.nb_mlp_pack_bwd(C.byref(_ops.mlp_weight_struct(model)), ptr(packed_bwd[k]), stream_ptr()), "nb_mlp_pack_bwd")
                pk = packed_bwd[k]
        if g3[3] is not None:                  # a gradient that reached `raw` directly (extras['raw'])
            d_raw = d_raw + g3[3].reshape(d_raw.shape)
        d_pts, d_vd, gp = _bwd.mlp_backward(model, pts.view(-1, 3), vd, acts, n_s, d_raw, ray_grads, need_p, packed_bwd=pk)
        if not ray_grads:
            return gp, None, None, None
        d_o = torch.empty_like(d)
        d_dd = torch.empty_like(d)
        with torch.cuda.device(dev):
            check(lib.nb_ray_points_bwd(ptr(z), ptr(d_pts), n, n_s, ptr(d_o), ptr(d_dd), stream_ptr()), "nb_ray_points_bwd")
        return gp, d_o, d_dd + d_rd, d_vd

    # the coarse map's gradient does not depend on the fine pass (its resampled depths are detached,
    # render_utils.py:145): the coarse backward runs as a second branch beside the fine one
    side.wait_stream(main)
    with torch.cuda.stream(side):
        out_c = branch(0, coarse, raw_c, z_c, noise_c, acts_c, pts_c, S, (g_rgb0, g_disp0, g_acc0, None), need_c)
        for g in list(out_c[0]) + list(out_c[1:]):
            if g is not None:
                g.record_stream(main)          # allocated on the side stream, consumed on this one
    out_f = branch(1, fine, raw_f, z_f, noise_f, acts_f, pts_f, F, (g_rgb, g_disp, g_acc, g_raw), need_f)
    main.wait_stream(side)
    d_o = d_d = d_v = None
    if ray_grads:
        parts = [o for o in (out_f, out_c) if o[1] is not None]
        if parts:
            d_o = parts[0][1] if len(parts) == 1 else parts[0][1] + parts[1][1]
            d_d = parts[0][2] if len(parts) == 1 else parts[0][2] + parts[1][2]
            d_vd = parts[0][3] if len(parts) == 1 else parts[0][3] + parts[1][3]
            d_v = torch.empty_like(v)              # viewdirs = view_src / |view_src| (render_utils.py:211)
            with torch.cuda.device(dev):
                check(lib.nb_normalize_bwd(ptr(v), ptr(f32c(d_vd)), n, ptr(d_v), stream_ptr()), "nb_normalize_bwd")
    return out_c[0], out_f[0], d_o, d_d, d_v


# ---------------------------------------------------------------------------------------------- lazily captured graphs
class _Graphs:
    """Forward / backward CUDA graphs of the fused node for one (renderer configuration, models, batch size, mode).
    mode "train": rays without gradient, weight gradients accumulate in place into the flat optimizer's buffer;
    mode "pose": rays with gradient, frozen weights (demo_est_rel_pose.py: the networks are loaded not trainable)."""

    def __init__(self, r, coarse, fine, n, dev, mode):
        self.r, self.models, self.n, self.dev, self.mode = r, (coarse, fine), n, dev, mode
        lib = _ops._lib_for(next(coarse.parameters()))
        L, Lv = int(coarse.multires), int(coarse.multires_views)
        b = lambda nbytes: torch.empty((int(nbytes),), device=dev, dtype=torch.uint8)
        self.packed = (b(lib.nb_mlp_packed_bytes(L, Lv)), b(lib.nb_mlp_packed_bytes(L, Lv)))
        self.packed_bwd = (b(lib.nb_mlp_bwd_packed_bytes(L, Lv)), b(lib.nb_mlp_bwd_packed_bytes(L, Lv)))
        self.inp = torch.zeros((3, n, 3), device=dev, dtype=torch.float32)          # rays_o, rays_d, view_src
        self.g_in = torch.zeros((2, n, 3), device=dev, dtype=torch.float32)         # d loss / d rgb_map, d loss / d rgb0
        self.fwd = None
        self.bwd = {}                     # keyed by whether the coarse map carries a gradient
        self.bwd_out = {}
        self.outs = self.saved = None
        self.generation = 0
        self.pending = False
        self.pool = torch.cuda.graph_pool_handle()

    def _capture(self, fn):
        # one eager run first (lazy initialisation, allocator warm-up), on a side stream as graph capture requires
        s = torch.cuda.Stream(device=self.dev)
        s.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(s):
            fn()
        torch.cuda.current_stream(self.dev).wait_stream(s)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, pool=self.pool):
            out = fn()
        return g, out

    def forward(self, o, d, v):
        # main.py hands over the two halves of one [2, n, 3] batch and takes the view directions from rays_d: one copy
        # for both rays when they are adjacent in memory, one device-side copy for the directions
        if (d.data_ptr() == o.data_ptr() + 4 * o.numel() and o.is_contiguous() and d.is_contiguous()
                and o.untyped_storage().data_ptr() == d.untyped_storage().data_ptr()):
            self.inp[:2].copy_(torch.as_strided(o, (2, self.n, 3), (3 * self.n, 3, 1)))
        else:
            self.inp[0].copy_(o)
            self.inp[1].copy_(d)
        self.inp[2].copy_(self.inp[1] if v.data_ptr() == d.data_ptr() else v)
        if self.fwd is None:
            rng = torch.cuda.get_rng_state(self.dev)     # the warm-up run consumes draws: the first replay gets the ones
            with torch.no_grad():                        # an uncaptured first call would have made
                self.fwd, (self.outs, self.saved) = self._capture(
                    lambda: _forward_impl(self.r, self.models[0], self.models[1], self.inp[0], self.inp[1], self.inp[2],
                                          packed=self.packed))
            torch.cuda.set_rng_state(rng, self.dev)
        self.fwd.replay()
        self.generation += 1
        # the caller gets its own copies (the static outputs are overwritten by the next replay): one allocation and
        # one multi-tensor copy launch for all of them
        return self._copies(self.outs)

    @staticmethod
    def _copies(src):
        flat = torch.empty((sum(t.numel() for t in src),), device=src[0].device, dtype=src[0].dtype)
        dst, o = [], 0
        for t in src:
            dst.append(flat[o:o + t.numel()].view(t.shape))
            o += t.numel()
        torch._foreach_copy_(dst, list(src))
        return tuple(dst)

    def backward(self, g_rgb, g_rgb0, white, dims):
        """-> (d_rays_o, d_rays_d, d_view_src) in pose mode, None in train mode (gradients accumulated in place)."""
        if g_rgb0 is not None:
            torch._foreach_copy_([self.g_in[0], self.g_in[1]], [g_rgb, g_rgb0])
        else:
            self.g_in[0].copy_(g_rgb)
        key = g_rgb0 is not None
        pose = self.mode == "pose"
        need = [not pose] * 24
        if key not in self.bwd:
            with torch.no_grad():
                g, out = self._capture(
                    lambda: _backward_impl(self.models, white, dims, self.saved,
                                           (self.g_in[0], None, None, self.g_in[1] if key else None, None, None, None),
                                           need, need, packed_bwd=self.packed_bwd, ray_grads=pose))
            self.bwd[key], self.bwd_out[key] = g, out[2:]
            if not pose:
                return None       # the capture's eager warm-up run already accumulated this step's gradients once
        self.bwd[key].replay()
        return self._copies(self.bwd_out[key]) if pose else None


def _graphs_for(r, coarse, fine, n, dev, ray_grads):
    """The cached graphs if this call may use them, else None."""
    if not GRAPHS_ENABLED or os.environ.get("NB_EAGER_GRAPH", "1") == "0" or torch.cuda.is_current_stream_capturing():
        return None
    if "_uniform" in r.__dict__ or "_normal" in r.__dict__:          # tests feed fixed draws through the hooks
        return None
    params = [p for m in (coarse, fine) for p in _ops.mlp_params(m)]
    if ray_grads:
        if any(p.requires_grad for p in params):
            return None                                                # pose graphs: frozen networks only
        mode, extra = "pose", sum(p._version for p in params)          # (weights edited in place: new graphs)
    else:
        tc, tf = coarse.__dict__.get("_nb_grad_target"), fine.__dict__.get("_nb_grad_target")
        if tc is None or tf is None or tc[0] is not tf[0] or tc[0].device != dev:
            return None                                                # gradients must accumulate in place (FlatAdam)
        if not all(p.requires_grad for p in params):
            return None
        mode, extra = "train", tc[0].data_ptr()
    key = (mode, id(coarse), id(fine), n, int(r.N_samples), int(r.N_importance), float(r.perturb), float(r.raw_noise_std),
           bool(r.white_bkgd), bool(r.lindisp), float(r.near), float(r.far), extra)
    cache = r.__dict__.setdefault("_nb_graphs", {})
    e = cache.get(key)
    if e is None:
        if len(cache) >= 4:                 # batch sizes / models keep changing: do not hoard graphs
            return None
        e = cache[key] = _Graphs(r, coarse, fine, n, dev, mode)
    if e.pending:
        # a second forward before the first one's backward (gradient accumulation over several batches): that call
        # runs uncaptured, on its own buffers, and leaves the graph's saved activations alone
        e.pending = False
        return None
    return e


class FusedRender(torch.autograd.Function):
    """(rays_o, rays_d [n,3] already NDC-warped if the renderer does that, view_src [n,3]) ->
    rgb, disp, acc, rgb0, disp0, acc0, z_std, raw (render_utils.py:67-174 for one batch).  Differentiable with respect to
    the networks' parameters and, for pose estimation, the three ray inputs."""

    @staticmethod
    def forward(ctx, renderer, coarse, fine, rays_o, rays_d, view_src, *params):
        o, d, v = f32c(rays_o), f32c(rays_d), f32c(view_src)
        n, dev = o.shape[0], o.device
        S, F = int(renderer.N_samples), int(renderer.N_samples) + int(renderer.N_importance)
        ctx.models, ctx.white, ctx.dims = (coarse, fine), int(bool(renderer.white_bkgd)), (n, S, F)
        ctx.ray_grads = any(ctx.needs_input_grad[3:6])
        ctx.set_materialize_grads(False)
        e = _graphs_for(renderer, coarse, fine, n, dev, ctx.ray_grads)
        ctx.graphs = e
        if e is not None:
            outs = e.forward(o, d, v)
            e.pending = True
            ctx.generation = e.generation
        else:
            outs, saved = _forward_impl(renderer, coarse, fine, o, d, v)
            ctx.save_for_backward(*saved)
        ctx.mark_non_differentiable(outs[6])
        return outs

    @staticmethod
    def backward(ctx, g_rgb, g_disp, g_acc, g_rgb0, g_disp0, g_acc0, _g_zstd, g_raw):
        need = list(ctx.needs_input_grad[6:])
        need_c, need_f = need[:24], need[24:]
        e = ctx.graphs

        def ray_out(d_o, d_d, d_v):
            ni = ctx.needs_input_grad
            return (d_o if ni[3] else None, d_d if ni[4] else None, d_v if ni[5] else None)

        if e is not None:
            if e.generation != ctx.generation:
                raise _lib.NerfB200Error(
                    "the activations of this render were overwritten by a later render of the same shape before backward() "
                    "ran (the fused node keeps one set per batch size); set NB_EAGER_GRAPH=0 to keep one set per call")
            e.pending = False
            saved = e.saved
            simple = g_rgb is not None and g_disp is None and g_acc is None and g_disp0 is None and g_acc0 is None and g_raw is None
            if simple and e.mode == "train" and g_rgb0 is not None and all(need):
                # p.grad = None since the last step: fresh gradients wanted (decided here, outside the captured sequence)
                for m in ctx.models:
                    _bwd.prepare_inplace(m, [True] * 24)
                e.backward(g_rgb, g_rgb0, ctx.white, ctx.dims)
                return (None,) * (6 + 48)
            if simple and e.mode == "pose" and not any(need):
                return (None,) * 3 + ray_out(*e.backward(g_rgb, g_rgb0, ctx.white, ctx.dims)) + (None,) * 48
        else:
            saved = ctx.saved_tensors
        grads_c, grads_f, d_o, d_d, d_v = _backward_impl(ctx.models, ctx.white, ctx.dims, saved,
                                                         (g_rgb, g_disp, g_acc, g_rgb0, g_disp0, g_acc0, g_raw), need_c, need_f,
                                                         ray_grads=ctx.ray_grads)
        return (None,) * 3 + ray_out(d_o, d_d, d_v) + tuple(grads_c) + tuple(grads_f)
