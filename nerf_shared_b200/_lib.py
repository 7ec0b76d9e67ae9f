"""ctypes binding of libnerf_b200.so (the C ABI declared in include/nerf_b200.h).

There is no CPU or eager-PyTorch fallback: if the shared library is missing, or the device is
not an sm_100 GPU, the first kernel call raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnerf_b200.so")

_lib = None
_lock = threading.Lock()
_inited_devices = set()

vp, i32, i64, f32, f64 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double


class NbMlpWeights(C.Structure):
    """Mirror of `struct NbMlpWeights` (include/nerf_b200.h)."""
    _fields_ = [
        ("pts_w", vp * 8), ("pts_b", vp * 8),
        ("feature_w", vp), ("feature_b", vp),
        ("alpha_w", vp), ("alpha_b", vp),
        ("views_w", vp), ("views_b", vp),
        ("rgb_w", vp), ("rgb_b", vp),
        ("L_pts", i32), ("L_dirs", i32),
    ]


# name -> (restype, argtypes); must list every NB_API symbol of include/nerf_b200.h
SIGNATURES = {
    "nb_version": (i32, []),
    "nb_last_error": (C.c_char_p, []),
    "nb_init": (i32, [i32]),
    "nb_shutdown": (i32, []),
    "nb_set_launch_limits": (i32, [i32, i32]),
    "nb_launch_count": (i64, []),
    "nb_get_rays_fwd": (i32, [i32, i32, f32, f32, f32, f32, vp, i32, vp, vp, vp, vp]),
    "nb_get_rays_sel_fwd": (i32, [i32, i32, f32, f32, f32, f32, vp, i32, vp, i64, vp, vp, vp, vp]),
    "nb_get_rays_bwd": (i32, [i32, i32, f32, f32, f32, f32, vp, i64, vp, vp, vp, vp]),
    "nb_normalize_fwd": (i32, [vp, i64, vp, vp]),
    "nb_ray_rows": (i32, [vp, vp, vp, f32, f32, i64, vp, vp, vp]),
    "nb_to8b": (i32, [vp, i64, vp, vp]),
    "nb_render_rays_workspace_bytes": (i64, [i64, i32, i32]),
    "nb_render_rays_fwd": (i32, [vp, vp, i32, i32, vp, vp, vp, f32, f32, i64, i32, i32, vp, vp, vp, vp, vp, vp, i32, i32, vp, i64,
                                 vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "nb_render_rays_fwd_noalloc": (i32, [vp, vp, i32, i32, vp, vp, vp, f32, f32, i64, i32, i32, vp, vp, vp, vp, vp, vp, i32, i32,
                                         vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "nb_composite_train": (i32, [vp, vp, vp, vp, vp, i64, i32, i32, f32, vp, vp, vp, vp, vp, vp, vp]),
    "nb_loss_finalize": (i32, [vp, vp, vp, vp]),
    "nb_mse_pair_fwd": (i32, [vp, vp, vp, i64, vp, vp]),
    "nb_mse_pair_bwd": (i32, [vp, vp, vp, i64, vp, vp, vp, vp, vp, vp]),
    "nb_normalize_bwd": (i32, [vp, vp, i64, vp, vp]),
    "nb_ndc_rays_fwd": (i32, [i32, i32, f64, f64, vp, vp, i64, vp, vp, vp]),
    "nb_ndc_rays_bwd": (i32, [i32, i32, f64, f64, vp, vp, vp, vp, i64, vp, vp, vp]),
    "nb_sample_coarse": (i32, [vp, i64, i32, vp, vp, i32, i32, vp, vp, vp]),
    "nb_ray_points_fwd": (i32, [vp, vp, vp, i64, i32, vp, vp]),
    "nb_ray_points_bwd": (i32, [vp, vp, i64, i32, vp, vp, vp]),
    "nb_sample_pdf_merge": (i32, [vp, vp, vp, vp, i64, i32, i32, vp, vp, vp, vp]),
    "nb_sample_pdf": (i32, [vp, vp, vp, vp, i64, i32, i32, vp, vp]),
    "nb_invert_cdf": (i32, [vp, vp, vp, i64, i32, i32, vp, vp, vp]),
    "nb_composite_fwd": (i32, [vp, vp, vp, vp, i64, i32, i32, vp, vp, vp, vp, vp, vp]),
    "nb_composite_bwd": (i32, [vp, vp, vp, vp, i64, i32, i32, vp, vp, vp, vp, vp, vp, vp]),
    "nb_mlp_fwd_f32": (i32, [C.POINTER(NbMlpWeights), vp, vp, i64, i32, vp, vp]),
    "nb_mlp_fwd_f32_noview": (i32, [C.POINTER(NbMlpWeights), vp, vp, i32, vp, i64, vp, vp]),
    "nb_mlp_f32_act_floats": (i64, [i32, i32, i32, i64]),
    "nb_mlp_f32_fwd_train": (i32, [C.POINTER(NbMlpWeights), vp, vp, i32, vp, vp, i64, i32, vp, vp, vp]),
    "nb_mlp_f32_bwd": (i32, [C.POINTER(NbMlpWeights), vp, i32, vp, vp, i64, C.POINTER(i64), vp, vp, vp]),
    "nb_mlp_packed_bytes": (i64, [i32, i32]),
    "nb_mlp_pack": (i32, [C.POINTER(NbMlpWeights), vp, vp]),
    "nb_mlp_fwd_tc": (i32, [vp, i32, i32, vp, vp, i64, i32, vp, vp, vp]),
    "nb_mlp_act_bytes": (i64, [i64]),
    "nb_mlp_bwd_packed_bytes": (i64, [i32, i32]),
    "nb_mlp_pack_bwd": (i32, [C.POINTER(NbMlpWeights), vp, vp]),
    "nb_mlp_da_bytes": (i64, [i64]),
    "nb_mlp_bwd_tc": (i32, [vp, i32, i32, vp, vp, i64, vp, C.POINTER(i64), vp, vp]),
    "nb_mlp_bwd_tc_parts": (i32, [vp, i32, i32, vp, vp, i64, vp, C.POINTER(i64), vp, i32, vp]),
    "nb_mlp_bwd_tc_streamed": (i32, [vp, i32, i32, vp, vp, i64, vp, C.POINTER(i64), vp, vp, i32, i32, vp]),
    "nb_mlp_ingrad_packed_bytes": (i64, [i32, i32]),
    "nb_mlp_pack_ingrad": (i32, [C.POINTER(NbMlpWeights), vp, vp]),
    "nb_mlp_bwd_inputs": (i32, [vp, i32, i32, vp, vp, vp, i64, i32, vp, vp, vp]),
    "nb_debug_umma_mn": (i32, [vp, vp, i32, vp, vp]),
    "nb_adam_step": (i32, [vp, vp, vp, vp, i64, f32, f32, f32, f32, i32, f32, vp]),
    "nb_adam_hyper": (i32, [vp, f32, f32, f32, f32, i32, f32, i32, vp]),
    "nb_train_step_prologue": (i32, [vp, f32, f32, f32, f32, i32, f32, i32, vp, vp, i64, vp, vp, i64, vp]),
    "nb_adam_step_dev": (i32, [vp, vp, vp, vp, i64, vp, vp]),
    "nb_adam_allreduce_p2p": (i32, [vp, vp, vp, i64, vp, C.POINTER(vp), C.POINTER(vp), vp, i32, i32, vp, vp]),
    "nb_debug_wgrad_trace": (i32, [C.POINTER(C.c_uint64), i32]),
    "nb_debug_umma": (i32, [vp, vp, i32, vp, vp]),
}


class NerfB200Error(RuntimeError):
    pass


def load():
    """dlopen the library and declare every prototype (no GPU needed for this step)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise NerfB200Error(
                "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or make -C nerf_shared_b200/csrc). nerf_shared_b200 has no CPU / eager fallback." % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = _NvtxLib(lib) if os.environ.get("NB_NVTX", "0") not in ("0", "") else lib
    return _lib


class _NvtxLib:
    """NB_NVTX=1: every C-ABI call is bracketed by an NVTX range named after the entry point, so that nsys / ncu
    timelines show the library's calls (ranges are pushed on the calling thread, around the asynchronous launch)."""

    def __init__(self, lib):
        self._lib = lib
        self._cache = {}

    def __getattr__(self, name):
        fn = self._cache.get(name)
        if fn is None:
            raw = getattr(self._lib, name)
            if name in ("nb_last_error", "nb_launch_count", "nb_version"):
                fn = raw
            else:
                def fn(*a, _raw=raw, _name=name):
                    torch.cuda.nvtx.range_push(_name)
                    try:
                        return _raw(*a)
                    finally:
                        torch.cuda.nvtx.range_pop()
            self._cache[name] = fn
        return fn


def last_error() -> str:
    return load().nb_last_error().decode("utf-8", "replace")


def check(rc: int, what: str = ""):
    if rc != 0:
        raise NerfB200Error("%s failed (rc=%d): %s" % (what or "libnerf_b200 call", rc, last_error()))


def ensure_device(device) -> int:
    """nb_init() once per device; raises unless it is an sm_100 GPU."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise NerfB200Error("nerf_shared_b200 runs on sm_100a GPUs only (got a %s tensor); there is no CPU path" % dev.type)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    if idx not in _inited_devices:
        check(load().nb_init(idx), "nb_init(%d)" % idx)
        _inited_devices.add(idx)
    return idx


def stream_ptr():
    """cudaStream_t of torch's current stream on the current device (the capture stream inside
    torch.cuda.graph); the raw getter is ~20x cheaper than building a torch.cuda.Stream object."""
    return vp(torch._C._cuda_getCurrentRawStream(torch.cuda.current_device()))


def ptr(t):
    """Device pointer of a contiguous tensor (None -> NULL)."""
    if t is None:
        return vp(0)
    assert t.is_contiguous(), "libnerf_b200 takes contiguous tensors"
    return vp(t.data_ptr())


def f32c(t):
    """fp32 contiguous CUDA view/copy of t."""
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def launch_count() -> int:
    return int(load().nb_launch_count())
