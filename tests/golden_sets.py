"""Seeded inputs of the non-degenerate-density goldens (tests/golden/make_golden_v2.py) -- shared by the generator
(build container, against the unmodified reference) and the tests (oracle on the CPU, CUDA path on the GPU box).
Test infrastructure."""
import os

import numpy as np
import torch

from oracle import nerf_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
TRAINED = os.path.join(HERE, "golden", "trained_sphere.npz")
K16 = O.blender_intrinsics(16, 16)
TRAIN_ROWS = torch.arange(0, 256, 4)                       # 64 of the 256 rays, spread over the frame
POSE_SEL = np.random.default_rng(2).choice(256, size=32, replace=False)


def available_sets():
    return ["stress", "bias"] + (["trained"] if os.path.exists(TRAINED) else [])


def params(name):
    """(coarse, fine) state_dicts of a model set."""
    if name == "stress":
        return O.make_params(13, scale=1.59), O.make_params(14, scale=1.59)
    if name == "bias":
        pc, pf = O.make_params(11), O.make_params(12)
        for p in (pc, pf):
            p["alpha_linear.bias"] = p["alpha_linear.bias"] + 0.1
        return pc, pf
    if name == "trained":
        with np.load(TRAINED) as z:
            return tuple({k[len(tag) + 1:]: torch.from_numpy(z[k]) for k in z.files if k.startswith(tag + "/")}
                         for tag in ("coarse", "fine"))
    raise KeyError(name)


def pose_of(name):
    return O.pose_spherical(0.0 if name == "trained" else 75.0, -30.0, 4.0)


def ray_batch(name):
    """[256, 11] rows (o, d, near, far, viewdir) of a 16 x 16 Blender-shaped frame."""
    ro, rd = O.get_rays(16, 16, K16, torch.from_numpy(pose_of(name))[:3, :4])
    ro, rd = ro.reshape(-1, 3).contiguous(), rd.reshape(-1, 3).contiguous()
    n = ro.shape[0]
    return torch.cat([ro, rd, torch.full((n, 1), 2.0), torch.full((n, 1), 6.0), O.unit_viewdirs(rd)], -1)


def train_target(name):
    return torch.from_numpy(np.random.default_rng(21).uniform(0, 1, size=(64, 3)).astype(np.float32))


def pose_target(name):
    return torch.from_numpy(np.random.default_rng(23).uniform(0, 1, size=(32, 3)).astype(np.float32))


def frame_subsample(HW):
    return torch.from_numpy(np.random.default_rng(31).choice(HW * HW, size=min(2048, HW * HW), replace=False))


def sphere_truth(HW):
    """Colour of the analytic scene of tests/fit_scene.py (unit sphere shaded by its normal, white background) seen
    from the held-out pose, [HW, HW, 3]."""
    K = O.blender_intrinsics(HW, HW)
    ro, rd = O.get_rays(HW, HW, K, torch.from_numpy(O.pose_spherical(0.0, -30.0, 4.0))[:3, :4])
    ro, rd = ro.reshape(-1, 3), rd.reshape(-1, 3)
    a, b, c = (rd * rd).sum(-1), 2 * (ro * rd).sum(-1), (ro * ro).sum(-1) - 1.0
    disc = b * b - 4 * a * c
    t = (-b - torch.sqrt(disc.clamp_min(0))) / (2 * a)
    normal = torch.nn.functional.normalize(ro + t[:, None] * rd, dim=-1)
    return torch.where((disc > 0)[:, None], 0.5 + 0.5 * normal, torch.ones_like(normal)).reshape(HW, HW, 3)
