"""Tiny on-disk scenes in the reference's four dataset formats, written deterministically from a seed (test
infrastructure: tests/test_datasets.py and tests/golden/make_golden_datasets.py build the same files)."""
import json
import os

import cv2
import numpy as np


def _write_png(path, img):
    """img uint8 [H,W,3|4] in RGB(A) order."""
    os.makedirs(os.path.dirname(path), exist_ok=True)
    out = cv2.cvtColor(img, cv2.COLOR_RGBA2BGRA) if img.shape[-1] == 4 else cv2.cvtColor(img, cv2.COLOR_RGB2BGR)
    assert cv2.imwrite(path, out)


def _pose(rng):
    a = rng.normal(size=(3, 3))
    q, _ = np.linalg.qr(a)
    m = np.eye(4)
    m[:3, :3] = q
    m[:3, 3] = rng.normal(scale=2.0, size=3)
    return m


def blender(root, seed=0, H=8, W=8, counts=(3, 2, 4)):
    rng = np.random.default_rng(seed)
    for split, n in zip(("train", "val", "test"), counts):
        frames = []
        for i in range(n):
            rel = "./%s/r_%d" % (split, i)
            _write_png(os.path.join(root, rel + ".png"), rng.integers(0, 256, size=(H, W, 4), dtype=np.uint8))
            frames.append({"file_path": rel, "rotation": 0.1, "transform_matrix": _pose(rng).tolist()})
        meta = {"camera_angle_x": 0.6911112070083618, "near": 2.0 + 0.25 * len(split), "far": 6.0 + 0.5 * len(split),
                "frames": frames}
        with open(os.path.join(root, "transforms_%s.json" % split), "w") as f:
            json.dump(meta, f)
    return root


def linemod(root, seed=1, H=8, W=10, counts=(3, 2, 3)):
    rng = np.random.default_rng(seed)
    K = [[572.4, 0.0, 0.5 * W], [0.0, 573.5, 0.5 * H], [0.0, 0.0, 1.0]]
    for split, n in zip(("train", "val", "test"), counts):
        frames = []
        for i in range(n):
            path = os.path.join(root, split, "%04d.png" % i)
            _write_png(path, rng.integers(0, 256, size=(H, W, 4), dtype=np.uint8))
            frames.append({"file_path": path, "transform_matrix": _pose(rng).tolist(), "intrinsic_matrix": K})
        meta = {"near": 0.31 + 0.1 * len(split), "far": 1.7 + 0.2 * len(split), "frames": frames}
        with open(os.path.join(root, "transforms_%s.json" % split), "w") as f:
            json.dump(meta, f)
    return root


def llff(root, seed=2, n=6, H=24, W=32, factor=4):
    rng = np.random.default_rng(seed)
    rows = []
    for i in range(n):
        full = rng.integers(0, 256, size=(H, W, 3), dtype=np.uint8)
        _write_png(os.path.join(root, "images", "img_%02d.png" % i), full)
        small = cv2.resize(full, (W // factor, H // factor), interpolation=cv2.INTER_AREA)
        _write_png(os.path.join(root, "images_%d" % factor, "img_%02d.png" % i), small)
        p = _pose(rng)[:3, :4]
        p[:, 3] *= 0.2                                   # forward-facing: cameras close together
        hwf = np.array([[H], [W], [0.9 * W]], dtype=np.float64)
        rows.append(np.concatenate([np.concatenate([p, hwf], 1).reshape(-1), [1.2 + 0.1 * i, 9.0 + i]]))
    np.save(os.path.join(root, "poses_bounds.npy"), np.stack(rows, 0))
    return root


def deepvoxels(root, scene="cube", seed=3, counts=(4, 16, 9)):
    rng = np.random.default_rng(seed)
    for split, n in zip(("train", "validation", "test"), counts):
        d = os.path.join(root, split, scene)
        for i in range(n):
            _write_png(os.path.join(d, "rgb", "%06d.png" % i), rng.integers(0, 256, size=(4, 4, 3), dtype=np.uint8))
            os.makedirs(os.path.join(d, "pose"), exist_ok=True)
            with open(os.path.join(d, "pose", "%06d.txt" % i), "w") as f:
                f.write(" ".join("%.9g" % v for v in _pose(rng).reshape(-1)))
        if split == "train":
            with open(os.path.join(d, "intrinsics.txt"), "w") as f:
                f.write("525.0 319.5 239.5 0.\n0. 0. 0.\n1.\n1.\n480 640\n")
    return root
