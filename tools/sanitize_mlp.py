"""compute-sanitizer target: forward (saving activations), dgrad, wgrad, head gradients and the input-gradient kernel of
the tensor-core MLP at a ragged 2.5-tile size (320 samples) and a 9-tile size, plus the fused compositing kernels.
    compute-sanitizer --tool memcheck  python tools/sanitize_mlp.py
    compute-sanitizer --tool racecheck python tools/sanitize_mlp.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_shared_b200 import _ops
from tests.gpu_util import make_model

dev = torch.device("cuda:0")
m = make_model(12)
for n_rays, S in ((5, 64), (6, 192)):
    pts = (torch.rand(n_rays, S, 3, device=dev) * 6 - 3).requires_grad_(True)
    vd = torch.nn.functional.normalize(torch.randn(n_rays, 3, device=dev), dim=-1).requires_grad_(True)
    raw = m(pts, vd)
    raw.backward(torch.randn_like(raw) * 1e-2)
    z = torch.sort(torch.rand(n_rays, S, device=dev) * 4 + 2, -1)[0]
    acc = torch.zeros(1, device=dev)
    _ops.composite_train(raw.detach(), z, torch.randn(n_rays, 3, device=dev), None, torch.rand(n_rays, 3, device=dev), True,
                         1.0 / (3 * n_rays), acc, want_weights=True)
torch.cuda.synchronize()
print("sanitize_mlp: ok, grad norm %.3e" % float(m.pts_linears[0].weight.grad.norm()))
