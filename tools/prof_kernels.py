"""One launch of each tensor-core kernel at the sizes the bench uses, for `ncu --set full -k regex:mlp_`.

    python tools/prof_kernels.py [n_rays]        # default 1024 rays x 192 fine samples (configs[1])

Launch order after the warm-up pass: mlp_fwd_tc (save_acts), mlp_dgrad, mlp_wgrad, mlp_small_grads,
then one inference-sized forward (32768 rays x 192, no saved activations).
"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_shared_b200 import _bwd, _lib, _ops
from tests.gpu_util import make_model

n_rays = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
S = 192
M = n_rays * S
dev = torch.device("cuda:0")
m = make_model(12)
lib = _lib.load()
pts = torch.rand(M, 3, device=dev) * 6 - 3
vd = torch.nn.functional.normalize(torch.randn(n_rays, 3, device=dev), dim=-1)
d_raw = torch.randn(M, 4, device=dev) * 1e-3
offs, total = _bwd.grad_layout(m)
grads = torch.zeros(total, device=dev)
dA = torch.empty(int(lib.nb_mlp_da_bytes(M)), dtype=torch.uint8, device=dev)
pk = _ops.packed_weights(m, backward=True)
off_arr = (C.c_int64 * 24)(*offs)


def one_pass():
    raw, acts = _ops.mlp_forward_tc(m, pts, vd, S, save_acts=True)
    _lib.check(lib.nb_mlp_bwd_tc(_lib.ptr(pk), 10, 4, _lib.ptr(acts), _lib.ptr(d_raw), M, _lib.ptr(dA), off_arr,
                                 _lib.ptr(grads), _lib.stream_ptr()))
    return raw


one_pass()
torch.cuda.synchronize()
one_pass()
torch.cuda.synchronize()
Mi = 32768 * S
pts_i = torch.rand(Mi, 3, device=dev) * 6 - 3
vd_i = torch.nn.functional.normalize(torch.randn(32768, 3, device=dev), dim=-1)
_ops.mlp_forward_tc(m, pts_i, vd_i, S, save_acts=False)
torch.cuda.synchronize()
print("prof_kernels: ok")
