"""Drop-in alias: `from nerf_shared import utils, nerf, render_utils, config_parser` (the
reference's import lines, main.py:7-8) resolve to the B200 implementation in nerf_shared_b200."""
import importlib
import sys

__all__ = ["config_parser", "nerf", "render_utils", "utils", "datasets"]

for _name in __all__:
    try:
        _mod = importlib.import_module("nerf_shared_b200." + _name)
    except ModuleNotFoundError as e:  # optional host-side modules (e.g. configargparse-free parser)
        if "nerf_shared_b200" in str(e):
            continue
        raise
    sys.modules[__name__ + "." + _name] = _mod
    globals()[_name] = _mod
