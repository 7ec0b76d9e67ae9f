"""GPU box: train both networks on the analytic sphere scene of tests/fit_scene.py with this package (main.py's loop,
graphed) until the density forms a surface, and save the state_dicts -> gpurun_out/trained_sphere.npz.

    python tests/golden/train_on_gpu.py [steps]        (then copy the file to tests/golden/trained_sphere.npz)

The weights are only an INPUT of the parity goldens: tests/golden/make_golden_v2.py feeds them to the unmodified
reference in the build container and records what the reference renders / differentiates with them.  (300 steps of the
reference on the CPU -- 50 minutes -- stop on the loss plateau with an all-transparent fine network, which is why the
weights are made here.)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import fit_scene  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
curve, psnr, mc, mf = fit_scene.run("ours", steps, every=250, log=lambda s: print(s, flush=True), return_models=True,
                                     raw_noise_std=1.0)   # the seeds-11/12 initialisation has sigma < 0 everywhere: noise lets density start
out = {"meta/steps": np.int64(steps), "meta/heldout_psnr": np.float64(psnr)}
for tag, m in (("coarse", mc), ("fine", mf)):
    for k, v in m.state_dict().items():
        out["%s/%s" % (tag, k)] = v.detach().float().cpu().numpy()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez(os.path.join(ROOT, "gpurun_out", "trained_sphere.npz"), **out)
print("held-out PSNR %.3f dB after %d steps; wrote %d arrays" % (psnr, steps, len(out)))
