"""-m gpu: nb_render_rays_fwd_noalloc (the inference path on the library's built-in scratch array; an entry point with
no counterpart in the reference) against nb_render_rays_fwd on a caller's
workspace, through the C-ABI.  Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O
from tests.gpu_util import cuda, make_model

pytestmark = pytest.mark.gpu

S, I = 64, 128
NOALLOC_MAX_BYTES = 16 << 20          # NB_RENDER_RAYS_NOALLOC_MAX_BYTES in include/nerf_b200.h


def _inputs(n, seed, random_draws):
    rng = np.random.default_rng(seed)
    K = O.blender_intrinsics(64, 64)
    c2w = torch.from_numpy(O.pose_spherical(20.0 + seed, -30.0, 4.0))[:3, :4]
    ro, rd = O.get_rays(64, 64, K, c2w)
    pix = rng.integers(0, 64 * 64, size=n)
    ro, rd = ro.reshape(-1, 3)[pix].contiguous(), rd.reshape(-1, 3)[pix].contiguous()
    ro, rd = ro.float(), rd.float()
    d = {"o": cuda(ro), "d": cuda(rd), "v": cuda(rd.clone())}
    F = S + I
    if random_draws:
        d["t_rand"] = cuda(rng.random((n, S), dtype=np.float32))
        d["u"] = cuda(rng.random((n, I), dtype=np.float32))
        d["noise_c"] = cuda(rng.standard_normal((n, S)).astype(np.float32))
        d["noise_f"] = cuda(rng.standard_normal((n, F)).astype(np.float32))
    return d


def _outputs(n):
    f = lambda *s: torch.full(s, float("nan"), device="cuda", dtype=torch.float32)
    return [f(n, 3), f(n), f(n), f(n, 3), f(n), f(n), f(n), f(n, S + I, 4)]


def _call(lib, which, pc, pf, x, n, white_bkgd, ws=None, out=None):
    from nerf_shared_b200 import _lib, _ops
    p = _lib.ptr
    out = _outputs(n) if out is None else out
    head = [p(pc), p(pf), 10, 4, p(x["o"]), p(x["d"]), p(x["v"]), 2.0, 6.0, n, S, I, p(_ops.linspace01(S, "cuda")), p(x.get("t_rand")),
            p(_ops.linspace01(I, "cuda")), p(x.get("u")), p(x.get("noise_c")), p(x.get("noise_f")), 0, int(white_bkgd)]
    tail = [p(t) for t in out] + [_lib.stream_ptr()]
    if which == "noalloc":
        rc = lib.nb_render_rays_fwd_noalloc(*head, *tail)
    else:
        rc = lib.nb_render_rays_fwd(*head, p(ws), int(ws.numel()), *tail)
    return rc, out


def test_noalloc_entry_equals_the_workspace_entry_on_two_calls_in_a_row():
    from nerf_shared_b200 import _lib, _ops
    lib = _lib.load()
    mc, mf = make_model(13, scale=1.59), make_model(14, scale=1.59)
    pc, pf = _ops.packed_weights(mc), _ops.packed_weights(mf)
    # call 1: a large chunk with random draws; call 2: a small ragged chunk of other rays, deterministic sampling -- it
    # runs on whatever call 1 left in the shared scratch array
    cases = [(1500, 3, True, True), (37, 4, False, False), (1500, 5, True, False)]
    free_before = free_after = None
    for k, (n, seed, random_draws, white) in enumerate(cases):
        x = _inputs(n, seed, random_draws)
        need = int(lib.nb_render_rays_workspace_bytes(n, S, I))
        assert need <= NOALLOC_MAX_BYTES
        ws = torch.empty((need,), device="cuda", dtype=torch.uint8)
        rc, ref = _call(lib, "workspace", pc, pf, x, n, white, ws)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        out = _outputs(n)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc, got = _call(lib, "noalloc", pc, pf, x, n, white, out=out)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        free_after = torch.cuda.mem_get_info()[0]
        for a, b in zip(got, ref):
            assert not torch.isnan(b).all()
            assert torch.equal(a.nan_to_num(), b.nan_to_num()), (k, n)       # same kernels, same order: the same bits
        assert float(ref[2].max()) > 0.5                                    # not a blank render
    # the third call (every kernel's module loaded, inputs and outputs allocated before) took no device memory: a workspace
    # allocated by the call would be 12.8 MB here (4 MiB of slack for other contexts on the device)
    assert free_after >= free_before - (4 << 20), (free_before, free_after)


def test_noalloc_entry_rejects_sizes_beyond_its_scratch():
    from nerf_shared_b200 import _lib, _ops
    lib = _lib.load()
    mc = make_model(13, scale=1.59)
    pc = _ops.packed_weights(mc)
    n = 4096
    assert int(lib.nb_render_rays_workspace_bytes(n, S, I)) > NOALLOC_MAX_BYTES
    x = _inputs(n, 1, False)
    rc, out = _call(lib, "noalloc", pc, None, x, n, True)
    assert rc == -1                                                         # NB_E_ARG
    msg = _lib.last_error()
    assert "nb_render_rays_fwd_noalloc" in msg and "NB_RENDER_RAYS_NOALLOC_MAX_BYTES" in msg
    torch.cuda.synchronize()
    assert bool(torch.isnan(out[0]).all())                                  # nothing was launched
