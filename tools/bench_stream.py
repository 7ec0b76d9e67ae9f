"""HBM-roofline check of the streaming kernels (everything on the path that is not the MLP).

    python tools/bench_stream.py [n_rays]        # default 262144 rays (one Renderer launch of an 800x800 frame)

For each kernel: CUDA-event time of the launch alone (inputs larger than L2 or L2 flushed between
launches), algorithmic bytes per ray (SURVEY.md section 8d / DESIGN.md section 4) and the fraction
of the measured copy bandwidth (MEASURED_PEAKS.json).  Prints a table and one JSON line.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_shared_b200 import _lib, _ops, utils  # noqa: E402


def hbm_peak():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p)).get("hbm_gbs", 6650.0)), "measured"
    return 6650.0, "fallback"


def timed(fn, flush, reps=10):
    reps = int(os.environ.get("NB_STREAM_REPS", reps))            # NB_STREAM_REPS=1 NB_STREAM_WARMUP=0: one launch of
    for _ in range(int(os.environ.get("NB_STREAM_WARMUP", 3))):  # each kernel, for an ncu --set full capture
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps


def run(n=262144, S=64, I=128, quiet=False):
    dev = torch.device("cuda:0")
    F = S + I
    peak, src = hbm_peak()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    g = torch.Generator(device=dev).manual_seed(0)
    rays_o = torch.randn(n, 3, device=dev, generator=g)
    rays_d = torch.randn(n, 3, device=dev, generator=g)
    vd = torch.nn.functional.normalize(rays_d, dim=-1)
    rows = torch.cat([rays_o, rays_d, torch.full((n, 1), 2.0, device=dev), torch.full((n, 1), 6.0, device=dev), vd], -1).contiguous()
    t_rand = torch.rand(n, S, device=dev, generator=g)
    z = _ops.sample_coarse(rows, S, t_rand, False)
    raw_c = torch.randn(n, S, 4, device=dev, generator=g)
    raw_f = torch.randn(n, F, 4, device=dev, generator=g)
    u = torch.rand(n, I, device=dev, generator=g)
    with torch.no_grad():
        _, _, _, w_c, _ = _ops.Composite.apply(raw_c, z, rays_d, None, True)
        z_f, _ = _ops.sample_pdf_merge(z, w_c, I, u)
    g_rgb = torch.randn(n, 3, device=dev, generator=g)
    H = W = 512
    K = [[500.0, 0, 256.0], [0, 500.0, 256.0], [0, 0, 1]]
    c2w = torch.eye(4, device=dev)[:3, :4].contiguous()
    lib = _lib.load()

    def comp_fwd(raw, zz, want_w):
        SS = zz.shape[1]
        rgb = torch.empty(n, 3, device=dev); disp = torch.empty(n, device=dev); acc = torch.empty(n, device=dev)
        depth = torch.empty(n, device=dev)
        wts = torch.empty(n, SS, device=dev) if want_w else None
        def f():
            _lib.check(lib.nb_composite_fwd(_lib.ptr(raw), _lib.ptr(zz), _lib.ptr(rays_d), None, n, SS, 1, _lib.ptr(rgb),
                                            _lib.ptr(disp), _lib.ptr(acc), _lib.ptr(wts), _lib.ptr(depth), _lib.stream_ptr()))
        return f

    def comp_bwd(raw, zz):
        SS = zz.shape[1]
        d_raw = torch.empty_like(raw)
        def f():
            _lib.check(lib.nb_composite_bwd(_lib.ptr(raw), _lib.ptr(zz), _lib.ptr(rays_d), None, n, SS, 1, _lib.ptr(g_rgb), None,
                                            None, None, _lib.ptr(d_raw), None, _lib.stream_ptr()))
        return f

    cases = [
        # name, fn, algorithmic bytes per ray
        ("get_rays (512x512 grid)", lambda: utils.get_rays(H, W, K, c2w), None),
        ("sample_coarse (z only, perturb)", lambda: _ops.sample_coarse(rows, S, t_rand, False), 44 + 4 * S + 4 * S),
        ("ray_points fwd S=64", lambda: _ops.RayPoints.apply(rays_o, rays_d, z), 24 + 4 * S + 12 * S),
        ("ray_points fwd S=192", lambda: _ops.RayPoints.apply(rays_o, rays_d, z_f), 24 + 4 * F + 12 * F),
        ("composite_fwd coarse (weights out)", comp_fwd(raw_c, z, True), 20 * S + 12 + 4 * S + 24),
        ("composite_fwd fine", comp_fwd(raw_f, z_f, False), 20 * F + 12 + 24),
        ("composite_bwd coarse", comp_bwd(raw_c, z), 20 * S + 12 + 12 + 16 * S),
        ("composite_bwd fine", comp_bwd(raw_f, z_f), 20 * F + 12 + 12 + 16 * F),
        ("sample_pdf_merge (random u)", lambda: _ops.sample_pdf_merge(z, w_c, I, u), 8 * S + 4 * I + 4 * F + 4),
        ("sample_pdf_merge (det)", lambda: _ops.sample_pdf_merge(z, w_c, I, None), 8 * S + 4 * F + 4),
    ]
    out = {}
    with torch.no_grad():
        for name, fn, bpr in cases:
            ms = timed(fn, flush)
            if bpr is None:  # get_rays: H*W pixels, 24 B written each
                nbytes = H * W * 24
            else:
                nbytes = n * bpr
            gbs = nbytes / ms / 1e6
            out[name] = {"ms": round(ms, 4), "bytes": int(nbytes), "GBps": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peak, 3)}
            if not quiet:
                print("%-38s %8.4f ms  %8.1f MB  %7.1f GB/s  %5.1f%% of %s copy peak %.0f GB/s" % (
                    name, ms, nbytes / 1e6, gbs, 100 * gbs / peak, src, peak))
    return {"n_rays": n, "hbm_peak_gbs": peak, "peak_source": src, "kernels": out}


if __name__ == "__main__":
    res = run(int(sys.argv[1]) if len(sys.argv) > 1 else 262144)
    print(json.dumps(res))
