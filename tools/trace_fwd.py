"""Timeline of CTA 0's first quads in the fused forward kernel (library built with EXTRA_FLAGS=-DNB_TRACE)."""
import ctypes, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_shared_b200 import _lib, _ops
from tests.gpu_util import make_model
m = make_model(11)
n_rays, S = 8192, 192
pts = torch.rand(n_rays * S, 3, device="cuda") * 6 - 3
vd = torch.nn.functional.normalize(torch.randn(n_rays, 3, device="cuda"), dim=-1)
save = len(sys.argv) > 1 and sys.argv[1] == "save"
for _ in range(3):
    _ops.mlp_forward_tc(m, pts, vd, S, save_acts=save)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 640)()
lib = ctypes.CDLL(_lib.LIB_PATH)
assert lib.nb_trace_read(ctypes.cast(buf, ctypes.c_void_p)) == 0
t = np.array(buf[:], dtype=np.int64).reshape(4, 10, 2, 8)
t0 = t[1, 0, 0, 0]
print("quad iteration 1 and 2 of CTA 0 (cycles relative to quad 1 layer 0 tile 0 MMA start); save_acts=%s" % save)
print("q l x | mma:act_ok  mma:issued | epi:acc_ok  epi:done  epi:arrived | mma_span  wait_for_acc->epi  epi_time  arrive")
for q in (1, 2):
    for l in range(10):
        for x in range(2):
            e = t[q, l, x] - t0
            print("%d %d %d | %9d %9d | %9d %9d %9d | issue %5d  mma_done-issue %5d  drain_wait %5d  epi %5d  arr %4d" % (
                q, l, x, e[0], e[1], e[2], e[3], e[4], e[1] - e[0], e[2] - e[1], e[5] - e[2], e[3] - e[5], e[4] - e[3]))
