"""Run the reference's main.py TEXT, unmodified, for a bounded number of iterations against this repo's drop-in
packages (`nerf_shared`, root `utils.py` / `render_utils.py`).  Test infrastructure (tests/test_gpu_dropin.py).

    python tests/run_reference_script.py <n_iters> <path to main.py> [main.py's own arguments ...]

main.py hard-codes 200,001 iterations (main.py:64): the only intervention is tqdm.trange clipped to n_iters.  Prints one
`DROPIN_MAIN {...}` JSON line with the wall time per iteration (host included) measured over the last half."""
import json
import os
import runpy
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

n_iters, script = int(sys.argv[1]), sys.argv[2]
sys.argv = [script] + sys.argv[3:]

import torch  # noqa: E402
import tqdm  # noqa: E402

_trange = tqdm.trange
_marks = []


def clipped(a, b, *k, **kw):
    def gen():
        for i in _trange(a, min(b, a + n_iters), *k, **kw):
            if i - a == n_iters // 2:
                torch.cuda.synchronize()
                _marks.append((i, time.perf_counter()))
            yield i
        torch.cuda.synchronize()
        _marks.append((min(b, a + n_iters), time.perf_counter()))
    return gen()


tqdm.trange = clipped
runpy.run_path(script, run_name="__main__")
(i0, t0), (i1, t1) = _marks[0], _marks[-1]
from nerf_shared_b200 import _lib  # noqa: E402
print("DROPIN_MAIN " + json.dumps({"iters": n_iters, "ms_per_iter": (t1 - t0) * 1e3 / max(i1 - i0, 1),
                                   "timed_iters": i1 - i0, "kernel_launches": _lib.launch_count(),
                                   "default_tensor_type": torch.tensor([0.0]).device.type}), flush=True)
