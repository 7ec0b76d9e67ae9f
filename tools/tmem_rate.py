"""tcgen05.ld throughput (nb_debug_umma_rate modes 8 / 9)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_shared_b200 import _lib
lib = _lib.load()
import ctypes as _C
lib.nb_debug_umma_rate.restype = _C.c_int   # opt-in symbol: make -C nerf_shared_b200/csrc EXTRA_SRCS=umma_rate.cu
lib.nb_debug_umma_rate.argtypes = [_C.c_int, _C.c_int, _C.c_int, _C.c_void_p, _C.c_void_p]; _lib.ensure_device("cuda:0")
for mode, nw in ((8, 4), (9, 1)):
    iters = 256
    cyc = torch.zeros(2 * 148, dtype=torch.int64, device="cuda")
    for _ in range(2):
        _lib.check(lib.nb_debug_umma_rate(mode, iters, 2, _lib.ptr(cyc), _lib.stream_ptr()))
    torch.cuda.synchronize()
    c = float(cyc[0])
    print("mode %d (%d warps): %.0f cycles per 256-column read of 32 lanes; %.1f B/clk per warp, %.1f B/clk per SM" % (
        mode, nw, c / iters, 32768 * iters / c, nw * 32768 * iters / c))
