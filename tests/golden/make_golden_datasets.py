"""Run the UNMODIFIED reference dataset loaders (/root/reference/nerf_shared/load_*.py, utils.load_datasets) on the tiny
fixture scenes of tests/dataset_fixtures.py -> tests/golden/datasets_golden.npz.

Build container only:  python tests/golden/make_golden_datasets.py
imageio is not installed here; the stub module gets a cv2-backed imread (PNG decoding, same bytes)."""
import os
import sys
import tempfile
import types

import cv2
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
import refimport  # noqa: E402

refimport._stub_modules()


def _imread(path, **kw):
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if img.ndim == 3 and img.shape[2] == 4:
        return cv2.cvtColor(img, cv2.COLOR_BGRA2RGBA)
    return cv2.cvtColor(img, cv2.COLOR_BGR2RGB)


sys.modules["imageio"].imread = _imread
nerf, render_utils, utils = refimport.load()
from tests import dataset_fixtures as fx  # noqa: E402

out = {}


def npy(x):
    return x.detach().cpu().numpy() if torch.is_tensor(x) else np.asarray(x)


def put(tag, images, poses, render_poses, hwf, i_split, K, bds):
    out[tag + "/images"] = npy(images)
    out[tag + "/poses"] = npy(poses)
    out[tag + "/render_poses"] = npy(render_poses)
    out[tag + "/hwf"] = np.array([float(v) for v in hwf], dtype=np.float64)
    for name, idx in zip(("i_train", "i_val", "i_test"), i_split):
        out[tag + "/" + name] = np.asarray(idx).astype(np.int64).reshape(-1)
    out[tag + "/K"] = np.asarray(K, dtype=np.float64)
    out[tag + "/near_far"] = np.array([float(bds["near"]), float(bds["far"])], dtype=np.float64)


def args(**kw):
    base = dict(dataset_type="blender", datadir="", factor=4, spherify=False, llffhold=2, no_ndc=False, half_res=False,
                testskip=1, white_bkgd=True, shape="cube", render_test=False)
    base.update(kw)
    return types.SimpleNamespace(**base)


with tempfile.TemporaryDirectory() as tmp:
    d = fx.blender(os.path.join(tmp, "blender"))
    put("blender", *utils.load_datasets(args(dataset_type="blender", datadir=d)))
    put("blender_half_skip2_black", *utils.load_datasets(args(dataset_type="blender", datadir=d, half_res=True, testskip=2,
                                                              white_bkgd=False, render_test=True)))
    d = fx.linemod(os.path.join(tmp, "linemod"))
    put("linemod", *utils.load_datasets(args(dataset_type="LINEMOD", datadir=d)))
    d = fx.llff(os.path.join(tmp, "llff"))
    put("llff", *utils.load_datasets(args(dataset_type="llff", datadir=d)))
    put("llff_no_ndc_spherify", *utils.load_datasets(args(dataset_type="llff", datadir=d, no_ndc=True, spherify=True, llffhold=0)))
    d = fx.deepvoxels(os.path.join(tmp, "dv"))
    put("deepvoxels", *utils.load_datasets(args(dataset_type="deepvoxels", datadir=d, testskip=4)))

np.savez_compressed(os.path.join(HERE, "datasets_golden.npz"), **out)
print("wrote %d arrays, %.1f KB" % (len(out), os.path.getsize(os.path.join(HERE, "datasets_golden.npz")) / 1024))
