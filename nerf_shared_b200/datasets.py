"""Dataset dispatch (reference utils.load_datasets, utils.py:216-313).

Disk loaders (Blender JSON + PNG, LLFF poses_bounds.npy, LINEMOD, DeepVoxels) are host I/O outside the
hot path (out of scope); benchmarks and tests use synthetic data of the
same shapes.  `load_datasets` keeps the reference's return tuple for the one dataset type that needs no
files, `--dataset_type synthetic`, and says so for the others."""
from __future__ import annotations

import numpy as np


def synthetic_blender(H=400, W=400, n_images=8, seed=0):
    """Blender-shaped synthetic scene: camera_angle_x = 0.6911, poses on the
    pose_spherical(theta, -30, 4) orbit, uniform-random target images."""
    rng = np.random.default_rng(seed)
    focal = 0.5 * W / np.tan(0.5 * 0.6911112070083618)
    K = np.array([[focal, 0, 0.5 * W], [0, focal, 0.5 * H], [0, 0, 1]])
    poses = np.stack([_pose_spherical(float(t), -30.0, 4.0) for t in rng.uniform(-180, 180, n_images)], 0)
    images = np.random.default_rng(seed + 1).random((n_images, H, W, 3), dtype=np.float32)
    return images, poses.astype(np.float32), [H, W, focal], K


def synthetic_llff(H=378, W=504, n_images=4, seed=0):
    """LLFF-shaped synthetic scene (configs/fern.txt at factor 8): focal = 3260.5 / 8,
    forward-facing poses = identity rotation perturbed by small rotations, N(0, 0.05^2) translations."""
    rng = np.random.default_rng(seed)
    focal = 3260.5 / 8.0 * (W / 504.0)
    K = np.array([[focal, 0, 0.5 * W], [0, focal, 0.5 * H], [0, 0, 1]])
    poses = []
    for _ in range(n_images):
        w = rng.normal(0.0, 0.02, 3)                       # small rotation vector
        th = np.linalg.norm(w) + 1e-12
        k = w / th
        Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
        R = np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * (Kx @ Kx)
        c2w = np.eye(4)
        c2w[:3, :3] = R
        c2w[:3, 3] = rng.normal(0.0, 0.05, 3)
        poses.append(c2w)
    images = np.random.default_rng(seed + 1).random((n_images, H, W, 3), dtype=np.float32)
    return images, np.stack(poses, 0).astype(np.float32), [H, W, focal], K


def _pose_spherical(theta, phi, radius):
    # load_blender.py:10-32 / load_LINEMOD.py:29-34
    t = np.eye(4); t[2, 3] = radius
    p = np.deg2rad(phi); rp = np.eye(4); rp[1, 1] = np.cos(p); rp[1, 2] = -np.sin(p); rp[2, 1] = np.sin(p); rp[2, 2] = np.cos(p)
    h = np.deg2rad(theta); rt = np.eye(4); rt[0, 0] = np.cos(h); rt[0, 2] = -np.sin(h); rt[2, 0] = np.sin(h); rt[2, 2] = np.cos(h)
    c2w = rt @ rp @ t
    return np.array([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]]) @ c2w


def load_datasets(args):
    """-> images, poses, render_poses, hwf, i_split, K, bds_dict (utils.py:216-313)."""
    kind = getattr(args, "dataset_type", "synthetic")
    if kind != "synthetic":
        raise NotImplementedError(
            "dataset_type=%r reads files from disk; the disk loaders are outside the accelerated hot path of this "
            "build (SURVEY.md section 2, rows 10-13). Use --dataset_type synthetic or pass arrays directly." % kind)
    H = W = 400 if not getattr(args, "half_res", False) else 200
    images, poses, hwf, K = synthetic_blender(H, W)
    n = images.shape[0]
    i_split = [np.arange(0, n - 2), np.arange(n - 2, n - 1), np.arange(n - 1, n)]
    render_poses = np.stack([_pose_spherical(a, -30.0, 4.0) for a in np.linspace(-180, 180, 8 + 1)[:-1]], 0).astype(np.float32)
    return images, poses, render_poses, hwf, i_split, K, {"near": 2.0, "far": 6.0}
