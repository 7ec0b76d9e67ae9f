"""Per-CTA duration of the weight-gradient kernel at the fine-pass size (1024 rays x 192 samples).
usage: python tools/wgrad_skew.py [n_rays] [S]"""
import ctypes as C
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from nerf_shared_b200 import _lib          # noqa: E402
from tests.gpu_util import cuda, make_model  # noqa: E402

n_rays = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
S = int(sys.argv[2]) if len(sys.argv) > 2 else 192
lib = _lib.load()
m = make_model(11)
rng = np.random.default_rng(0)
pts = cuda(rng.uniform(-3, 3, size=(n_rays, S, 3)).astype(np.float32))
vd = torch.nn.functional.normalize(cuda(rng.normal(size=(n_rays, 3)).astype(np.float32)), dim=-1)
d_raw = cuda((rng.normal(size=(n_rays, S, 4)) * 1e-2).astype(np.float32))
for _ in range(3):
    m.zero_grad()
    m(pts, vd).backward(d_raw)
torch.cuda.synchronize()
buf = (C.c_uint64 * 296)()
assert lib.nb_debug_wgrad_trace(buf, 296) == 0
t = np.array(buf[:], dtype=np.int64).reshape(148, 2)
t0 = t[:, 0].min()
start, end = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3
print("kernel span %.1f us; CTA end: min %.1f  median %.1f  max %.1f us; mean busy fraction %.3f"
      % (end.max(), end.min(), np.median(end), end.max(), float(np.mean(end - start) / end.max())))
for b in range(0, 148, 4):
    print("  ".join("cta %3d: %6.1f" % (b + i, end[b + i]) for i in range(4) if b + i < 148))
