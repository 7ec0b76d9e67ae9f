import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (runs under `pytest -m gpu` on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    """Outputs of the unmodified reference, produced by tests/golden/make_golden.py."""
    path = os.path.join(ROOT, "tests", "golden", "reference_golden.npz")
    with np.load(path) as z:
        return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def golden_noview():
    """The same for use_viewdirs=False (tests/golden/make_golden_noview.py)."""
    with np.load(os.path.join(ROOT, "tests", "golden", "reference_noview_golden.npz")) as z:
        return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def golden_meta():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "reference_golden_meta.json")) as f:
        return json.load(f)
