"""Backward of NerfMLP (autograd of nerf.py:110-134): tensor-core dgrad chain + wgrad GEMMs."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib, _ops
from ._lib import check, f32c, ptr, stream_ptr


def grad_layout(model):
    """(offsets[24], total) of the flat fp32 gradient buffer, in _ops.mlp_params order."""
    offs, total = [], 0
    for p in _ops.mlp_params(model):
        offs.append(total)
        total += p.numel()
    return offs, total


def packed_ingrad(model) -> torch.Tensor:
    """Transposed bf16 slabs of the three weight blocks that touch the encoded inputs (first layer,
    skip connection, view branch); cached like _ops.packed_weights."""
    params = _ops.mlp_params(model)
    dev = params[0].device
    lib = _ops._lib_for(params[0])
    key = tuple((p.data_ptr(), p._version) for p in (params[0], params[5], params[20]))
    cache = model.__dict__.get("_nb_packed_in")
    if cache is not None and cache[0] == key and cache[1].device == dev:
        return cache[1]
    nbytes = int(lib.nb_mlp_ingrad_packed_bytes(int(model.multires), int(model.multires_views)))
    buf = torch.empty((nbytes,), device=dev, dtype=torch.uint8)
    w = _ops.mlp_weight_struct(model)
    with torch.cuda.device(dev):
        check(lib.nb_mlp_pack_ingrad(C.byref(w), ptr(buf), stream_ptr()), "nb_mlp_pack_ingrad")
    model.__dict__["_nb_packed_in"] = (key, buf)
    return buf


def prepare_inplace(model, need_param):
    """In-place gradient target (FlatAdam): parameters whose .grad was cleared since the last backward (model.zero_grad(),
    p.grad = None -- torch's set_to_none default) get their slice of the flat buffer zeroed and the view re-attached, so
    that the weight-gradient kernels start from fresh gradients and the optimizer sees the result.  -> True if the model
    has an in-place target."""
    target = model.__dict__.get("_nb_grad_target")
    if target is None:
        return False
    grads, offs = target
    for i, p in enumerate(_ops.mlp_params(model)):
        if p.grad is None and need_param[i]:
            view = grads[offs[i]:offs[i] + p.numel()].view_as(p)
            view.zero_()
            p.grad = view
    return True


def mlp_backward(model, pts, viewdirs, acts, S, g_raw, need_in, need_param, packed_bwd=None):
    """The MLP's backward kernels for one forward call: -> (d_pts, d_viewdirs, [24 parameter gradients]).
    need_param[i]: whether parameter i (in _ops.mlp_params order) wants a gradient.  With a FlatAdam optimizer
    registered on the model the weight-gradient kernels accumulate in place and the list holds None."""
    lib = _ops._lib_for(pts)
    M = pts.shape[0]
    params = _ops.mlp_params(model)
    need_w = any(need_param)
    if not (need_w or need_in):
        return None, None, [None] * len(params)
    dev = pts.device
    grads, offs, in_place, cleared = None, [0] * 24, False, []
    if need_w:
        # a FlatAdam optimizer registers its flat gradient buffer on the model: the wgrad kernels then
        # accumulate in place (every p.grad is a view of that buffer) and autograd gets no tensors
        target = model.__dict__.get("_nb_grad_target")
        in_place = target is not None and target[0].device == dev
        if in_place:
            grads, offs = target
            # model.zero_grad() / p.grad = None since the last backward (set_to_none is torch's default): the user
            # asked for fresh gradients, so clear this model's slices before accumulating into them, and hand the
            # views back afterwards -- otherwise p.grad stays None and the optimizer would discard the result
            cleared = [i for i, p in enumerate(params) if p.grad is None and need_param[i]]
            for i in cleared:
                grads[offs[i]:offs[i] + params[i].numel()].zero_()
        else:
            offs, total = grad_layout(model)
            grads = torch.zeros((total,), device=dev, dtype=torch.float32)
    dA = torch.empty((int(lib.nb_mlp_da_bytes(M)),), device=dev, dtype=torch.uint8)
    if packed_bwd is None:               # (a CUDA-graph caller packs into its own buffer inside the capture)
        packed_bwd = _ops.packed_weights(model, backward=True)
    off_arr = (C.c_int64 * 24)(*offs)
    g = f32c(g_raw).view(-1, 4)
    L, Lv = int(model.multires), int(model.multires_views)
    d_pts = d_vd = None
    with torch.cuda.device(dev):
        # frozen weights (grads == NULL): only the dgrad chain runs (pose estimation, demo_est_rel_pose.py:375)
        import os
        stream_min = int(os.environ.get("NB_STREAM_BWD", "0"))       # samples from which dgrad and wgrad run side by side
        if stream_min > 0 and grads is not None and M >= stream_min:
            ready = torch.empty(((M + 127) // 128,), device=dev, dtype=torch.int32)
            check(lib.nb_mlp_bwd_tc_streamed(ptr(packed_bwd), L, Lv, ptr(acts), ptr(g), M, ptr(dA), off_arr, ptr(grads), ptr(ready),
                                             int(os.environ.get("NB_STREAM_DGRAD", "0")), int(os.environ.get("NB_STREAM_WGRAD", "0")),
                                             stream_ptr()), "nb_mlp_bwd_tc_streamed")
        else:
            check(lib.nb_mlp_bwd_tc(ptr(packed_bwd), L, Lv, ptr(acts), ptr(g), M, ptr(dA), off_arr, ptr(grads), stream_ptr()),
                  "nb_mlp_bwd_tc")
        if need_in:
            d_pts = torch.empty((M, 3), device=dev, dtype=torch.float32)
            d_vd = torch.zeros((M // S, 3), device=dev, dtype=torch.float32)
            check(lib.nb_mlp_bwd_inputs(ptr(packed_ingrad(model)), L, Lv, ptr(dA), ptr(pts), ptr(viewdirs), M, S,
                                        ptr(d_pts), ptr(d_vd), stream_ptr()), "nb_mlp_bwd_inputs")
    for i in cleared:
        params[i].grad = grads[offs[i]:offs[i] + params[i].numel()].view_as(params[i])
    out = []
    for i, p in enumerate(params):
        use = need_w and need_param[i] and not in_place
        out.append(grads[offs[i]:offs[i] + p.numel()].view_as(p) if use else None)
    model._nb_last_dA = dA if getattr(model, "_nb_keep_dA", False) else None
    return d_pts, d_vd, out


def nerf_mlp_backward(ctx, g_raw):
    pts, viewdirs, acts = ctx.saved_tensors
    d_pts, d_vd, out = mlp_backward(ctx.model, pts, viewdirs, acts, ctx.S, g_raw,
                                    ctx.needs_input_grad[0] or ctx.needs_input_grad[1], list(ctx.needs_input_grad[5:]))
    return (d_pts if ctx.needs_input_grad[0] else None, d_vd if ctx.needs_input_grad[1] else None, None, None,
            None) + tuple(out)
