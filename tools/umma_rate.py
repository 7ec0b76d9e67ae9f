"""Tensor-pipe ceiling of this B200 for the MLP kernels' MMA shapes (nb_debug_umma_rate)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_shared_b200 import _lib
lib = _lib.load()
import ctypes as _C
lib.nb_debug_umma_rate.restype = _C.c_int   # opt-in symbol: make -C nerf_shared_b200/csrc EXTRA_SRCS=umma_rate.cu
lib.nb_debug_umma_rate.argtypes = [_C.c_int, _C.c_int, _C.c_int, _C.c_void_p, _C.c_void_p]; _lib.ensure_device("cuda:0")
names = {5: "cg2 ring3 commit-wait", 6: "cg2 ring1 commit-wait", 7: "cg2 ring3 via producer", 0: "cg2 M256 N256", 1: "cg1 M128 N256", 2: "cg2 M256 N128", 3: "cg1 M128 N128", 4: "cg2 M256 N256 commit/4"}
for ctas in (2, 148):
    for mode in (0, 4, 5, 6, 7):
        for iters in (4096,):
            cyc = torch.zeros(2 * 148, dtype=torch.int64, device="cuda")
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            _lib.check(lib.nb_debug_umma_rate(mode, iters, ctas, _lib.ptr(cyc), _lib.stream_ptr()))
            torch.cuda.synchronize()
            e0.record()
            _lib.check(lib.nb_debug_umma_rate(mode, iters, ctas, _lib.ptr(cyc), _lib.stream_ptr()))
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            cg = 2 if mode in (0, 2, 4, 5, 6, 7) else 1
            units = ctas // cg
            c = cyc.view(-1, 2)[:units].cpu()
            M, N = (256 if cg == 2 else 128), (128 if mode in (2, 3) else 256)
            flop = 2.0 * M * N * 16 * iters * units
            tot = c[:, 0].float()
            print("%-24s ctas=%3d: %.1f cyc/MMA (issue %.1f) -> %.0f FLOP/clk/SM ; kernel %.3f ms = %.0f TFLOP/s (eff. clock %.0f MHz)" % (
                names[mode], ctas, float(tot.mean()) / iters, float(c[:, 1].float().mean()) / iters,
                2.0 * M * N * 16 / cg / (float(tot.mean()) / iters), ms, flop / ms / 1e9, float(tot.mean()) / ms / 1e3))
