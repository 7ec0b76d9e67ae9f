"""Root-level shim: the reference's pose-estimation demo does `sys.path.append(<repo root>);
from utils import *` (examples/relative_pose_estimation_demo/demo_est_rel_pose.py:11-12) and relies
on the star import leaking `nn`, `F`, `np`, `torch`, `device`, ... -- forward everything."""
from nerf_shared_b200.utils import *            # noqa: F401,F403
from nerf_shared_b200.utils import (device, img2mse, mse2psnr, to8b, nn, F, np, torch, os)  # noqa: F401
from nerf_shared_b200 import utils              # noqa: F401  (demo calls utils.load_datasets(...))
from nerf_shared_b200 import nerf               # noqa: F401
