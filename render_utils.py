"""Root-level shim for `from render_utils import *` (demo_est_rel_pose.py:13)."""
from nerf_shared_b200.render_utils import *     # noqa: F401,F403
from nerf_shared_b200.render_utils import Renderer, DEBUG, imageio, np, torch, os, utils, nerf  # noqa: F401
