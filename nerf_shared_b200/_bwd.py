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


def nerf_mlp_backward(ctx, g_raw):
    model, S = ctx.model, ctx.S
    pts, viewdirs, acts = ctx.saved_tensors
    lib = _ops._lib_for(pts)
    M = pts.shape[0]
    if ctx.needs_input_grad[0] or ctx.needs_input_grad[1]:
        raise NotImplementedError("gradients wrt sample positions / view directions (pose estimation path) are not "
                                  "implemented yet in libnerf_b200")
    params = _ops.mlp_params(model)
    need_w = any(ctx.needs_input_grad[5:])
    if not need_w:
        return (None, None, None, None, None) + (None,) * len(params)
    dev = pts.device
    # a FlatAdam optimizer registers its flat gradient buffer on the model: the wgrad kernels then
    # accumulate in place (every p.grad is a view of that buffer) and autograd gets no tensors
    target = model.__dict__.get("_nb_grad_target")
    in_place = target is not None and target[0].device == dev
    if in_place:
        grads, offs = target
    else:
        offs, total = grad_layout(model)
        grads = torch.zeros((total,), device=dev, dtype=torch.float32)
    dA = torch.empty((int(lib.nb_mlp_da_bytes(M)),), device=dev, dtype=torch.uint8)
    packed_bwd = _ops.packed_weights(model, backward=True)
    off_arr = (C.c_int64 * 24)(*offs)
    g = f32c(g_raw).view(-1, 4)
    with torch.cuda.device(dev):
        check(lib.nb_mlp_bwd_tc(ptr(packed_bwd), int(model.multires), int(model.multires_views), ptr(acts), ptr(g), M,
                                ptr(dA), off_arr, ptr(grads), stream_ptr()), "nb_mlp_bwd_tc")
    out = []
    for i, p in enumerate(params):
        use = ctx.needs_input_grad[5 + i] and not in_place
        out.append(grads[offs[i]:offs[i] + p.numel()].view_as(p) if use else None)
    ctx.model._nb_last_dA = dA if getattr(ctx.model, "_nb_keep_dA", False) else None
    return (None, None, None, None, None) + tuple(out)
