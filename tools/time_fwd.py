"""Ad-hoc timing of the fused MLP forward kernel (CUDA events, kernel alone)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_shared_b200 import _lib, _ops
if os.environ.get("NB_LIB"):          # ablation builds (tools/exp_build.sh)
    _lib.LIB_PATH = os.path.abspath(os.environ["NB_LIB"])
from tests.gpu_util import make_model

FLOP_PER_SAMPLE = 1186816
peaks = json.load(open("MEASURED_PEAKS.json")) if os.path.exists("MEASURED_PEAKS.json") else {"bf16_tflops": 1590.0}
m = make_model(11)
for n_rays, S in ((1024, 64), (1024, 192)) if os.environ.get("NB_LIB") else ((1024, 64), (1024, 192), (32768, 64), (32768, 192), (160000, 192)):
    M = n_rays * S
    pts = (torch.rand(M, 3, device="cuda") * 6 - 3)
    vd = torch.nn.functional.normalize(torch.randn(n_rays, 3, device="cuda"), dim=-1)
    for save in (False, True):
        if save and M > 8_000_000:
            continue
        for _ in range(3):
            _ops.mlp_forward_tc(m, pts, vd, S, save_acts=save)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        reps = 10
        ev[0].record()
        for _ in range(reps):
            _ops.mlp_forward_tc(m, pts, vd, S, save_acts=save)
        ev[1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1]) / reps
        tf = FLOP_PER_SAMPLE * M / ms / 1e9
        print("rays=%d S=%d save_acts=%d: %.3f ms  %.1f TFLOP/s  (%.1f%% of measured burst peak %.0f)  %.2f Msamples/s" % (
            n_rays, S, save, ms, tf, 100 * tf / peaks["bf16_tflops"], peaks["bf16_tflops"], M / ms / 1e3))
m32 = make_model(11, precision="fp32")
M = 1024 * 64
pts = (torch.rand(M, 3, device="cuda") * 6 - 3)
vd = torch.nn.functional.normalize(torch.randn(1024, 3, device="cuda"), dim=-1)
for _ in range(2):
    _ops.mlp_forward_f32(m32, pts, vd, 64)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); _ops.mlp_forward_f32(m32, pts, vd, 64); e1.record(); torch.cuda.synchronize()
print("fp32 simt: %.3f ms %.2f TFLOP/s" % (e0.elapsed_time(e1), FLOP_PER_SAMPLE * M / e0.elapsed_time(e1) / 1e9))
