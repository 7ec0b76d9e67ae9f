"""Import the UNMODIFIED reference (/root/reference/nerf_shared) in this container.

Test infrastructure only (golden-vector generation).  The reference needs three
modules that are not installed here and do no hot-path arithmetic
(torchtyping: nerf.py:3, imageio: render_utils.py:1 / utils.py:1,
configargparse: config_parser.py:4); they are stubbed before the import.
`/root/reference` does not exist on the GPU box, so nothing under tests/ that
runs there imports this file.
"""
import os
import sys
import types

REF_ROOT = os.environ.get("NERF_REFERENCE_ROOT", "/root/reference")


def _stub_modules():
    if "torchtyping" not in sys.modules:
        m = types.ModuleType("torchtyping")

        class _TT:
            def __class_getitem__(cls, item):
                return cls

        m.TensorType = _TT
        m.TensorDetail = object
        sys.modules["torchtyping"] = m
    if "imageio" not in sys.modules:
        sys.modules["imageio"] = types.ModuleType("imageio")
    if "configargparse" not in sys.modules:
        import argparse
        m = types.ModuleType("configargparse")
        m.ArgumentParser = argparse.ArgumentParser
        sys.modules["configargparse"] = m
    # typeguard 4 chokes on the torchtyping annotation of get_density (nerf.py:136-139)
    import typeguard
    typeguard.typechecked = lambda f=None, **kw: f if f is not None else (lambda g: g)


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "nerf_shared"))


def load():
    """Returns (nerf, render_utils, utils) modules of the reference."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    _stub_modules()
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    # our own repo also ships a package called nerf_shared (the drop-in); make sure the
    # reference one wins inside this process
    for k in [k for k in sys.modules if k == "nerf_shared" or k.startswith("nerf_shared.")]:
        del sys.modules[k]
    import importlib
    nerf = importlib.import_module("nerf_shared.nerf")
    render_utils = importlib.import_module("nerf_shared.render_utils")
    utils = importlib.import_module("nerf_shared.utils")
    assert os.path.realpath(nerf.__file__).startswith(os.path.realpath(REF_ROOT))
    return nerf, render_utils, utils
