"""One launch of sample_pdf_merge per mode (random u, deterministic u) at 262,144 rays, for an ncu --set full capture.
    ncu --set full --clock-control none --import-source on -k regex:sample_pdf_merge -o gpurun_out/r2_prof_pdf python tools/prof_sample_pdf.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_shared_b200 import _ops  # noqa: E402

n, S, I = 262144, 64, 128
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
rays_o = torch.randn(n, 3, device=dev, generator=g)
rays_d = torch.randn(n, 3, device=dev, generator=g)
vd = torch.nn.functional.normalize(rays_d, dim=-1)
rows = torch.cat([rays_o, rays_d, torch.full((n, 1), 2.0, device=dev), torch.full((n, 1), 6.0, device=dev), vd], -1).contiguous()
z = _ops.sample_coarse(rows, S, torch.rand(n, S, device=dev, generator=g), False)
raw_c = torch.randn(n, S, 4, device=dev, generator=g)
u = torch.rand(n, I, device=dev, generator=g)
with torch.no_grad():
    _, _, _, w_c, _ = _ops.Composite.apply(raw_c, z, rays_d, None, True)
    for _ in range(2):
        _ops.sample_pdf_merge(z, w_c, I, u)
        _ops.sample_pdf_merge(z, w_c, I, None)
torch.cuda.synchronize()
print("done")
