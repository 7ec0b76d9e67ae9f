"""Train the UNMODIFIED reference (/root/reference, CPU) for 300 steps of main.py's loop on the analytic sphere scene of
tests/fit_scene.py and save both networks' state_dicts -> tests/golden/trained_sphere_300.npz.

Build container only (about 15 minutes on 8 cores):  python tests/golden/make_trained_weights.py
These are the "weights after >= 300 steps" of the parity plan: a model whose density is positive on a real surface, so
that compositing, resampling and the gradients of the fine network are exercised away from the sigma ~ 0 knife edge of
random initialisation.  Stored as fp32 (bit-exact state_dict values)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from tests import fit_scene  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
curve, psnr, mc, mf = fit_scene.run("reference", steps, every=25, log=lambda s: print(s, flush=True), return_models=True)
out = {"meta/steps": np.int64(steps), "meta/heldout_psnr": np.float64(psnr)}
for tag, m in (("coarse", mc), ("fine", mf)):
    for k, v in m.state_dict().items():
        out["%s/%s" % (tag, k)] = v.detach().cpu().numpy()
np.savez(os.path.join(HERE, "trained_sphere_300.npz"), **out)
print("held-out PSNR %.3f dB; wrote %d arrays" % (psnr, len(out)))
