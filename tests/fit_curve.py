"""CLI of tests/fit_scene.py: the loss curve of the training loop on the analytic sphere scene.
    python tests/fit_curve.py --impl ours --steps 300 > gpurun_out/fit_ours.txt                  (GPU box)
    python tests/fit_curve.py --impl reference --steps 300 > tests/golden/reference_fit_curve.txt  (build container, CPU)
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import fit_scene  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--impl", default="ours")
ap.add_argument("--steps", type=int, default=300)
ap.add_argument("--every", type=int, default=25)
args = ap.parse_args()
print("# impl=%s  step  mean loss of the last %d steps (img2mse(rgb) + img2mse(rgb0), main.py:99)" % (args.impl, args.every))
_, psnr = fit_scene.run(args.impl, args.steps, args.every, log=lambda s: print(s, flush=True))
print("# held-out view (theta = 0): PSNR %.2f dB" % psnr)
