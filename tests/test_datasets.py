"""CPU: the four disk loaders (nerf_shared_b200/datasets.py) against what the UNMODIFIED reference loaders returned on
the same fixture scenes (tests/golden/datasets_golden.npz, made by tests/golden/make_golden_datasets.py), and the
config-file front-end on every shipped reference config key."""
import os
import types

import numpy as np
import pytest

from tests import dataset_fixtures as fx

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def gold():
    with np.load(os.path.join(ROOT, "tests", "golden", "datasets_golden.npz")) as z:
        return {k: z[k] for k in z.files}


def _args(**kw):
    base = dict(dataset_type="blender", datadir="", factor=4, spherify=False, llffhold=2, no_ndc=False, half_res=False,
                testskip=1, white_bkgd=True, shape="cube", render_test=False)
    base.update(kw)
    return types.SimpleNamespace(**base)


def _check(gold, tag, res, pose_tol=0.0, path_tol=1e-6):
    images, poses, render_poses, hwf, i_split, K, bds = res
    g = lambda k: gold[tag + "/" + k]
    assert np.asarray(images).shape == g("images").shape, tag
    assert np.abs(np.asarray(images, dtype=np.float64) - g("images")).max() <= 1e-7, tag
    assert np.abs(np.asarray(poses, dtype=np.float64) - g("poses")).max() <= pose_tol, tag
    assert np.asarray(render_poses).shape == g("render_poses").shape, tag
    assert np.abs(np.asarray(render_poses, dtype=np.float64) - g("render_poses")).max() <= path_tol, tag
    assert np.allclose(np.array([float(v) for v in hwf]), g("hwf"), rtol=1e-12, atol=0), tag
    assert isinstance(hwf[0], int) and isinstance(hwf[1], int)
    for name, idx in zip(("i_train", "i_val", "i_test"), i_split):
        assert np.array_equal(np.asarray(idx).reshape(-1), g(name)), (tag, name)
    assert np.allclose(np.asarray(K, dtype=np.float64), g("K"), rtol=1e-12, atol=0), tag
    assert np.allclose([float(bds["near"]), float(bds["far"])], g("near_far"), rtol=1e-6, atol=0), tag


def test_blender_loader(tmp_path, gold):
    from nerf_shared_b200 import datasets
    d = fx.blender(str(tmp_path / "blender"))
    _check(gold, "blender", datasets.load_datasets(_args(dataset_type="blender", datadir=d)))
    # near / far are the JSON's (load_blender.py:57), not constants
    assert gold["blender/near_far"][0] == 3.0 and gold["blender/near_far"][1] == 8.0
    _check(gold, "blender_half_skip2_black",
           datasets.load_datasets(_args(dataset_type="blender", datadir=d, half_res=True, testskip=2, white_bkgd=False,
                                        render_test=True)))


def test_linemod_loader(tmp_path, gold):
    from nerf_shared_b200 import datasets
    d = fx.linemod(str(tmp_path / "linemod"))
    _check(gold, "linemod", datasets.load_datasets(_args(dataset_type="LINEMOD", datadir=d)))


def test_llff_loader(tmp_path, gold):
    from nerf_shared_b200 import datasets
    d = fx.llff(str(tmp_path / "llff"))
    _check(gold, "llff", datasets.load_datasets(_args(dataset_type="llff", datadir=d)), pose_tol=1e-6, path_tol=1e-5)
    _check(gold, "llff_no_ndc_spherify",
           datasets.load_datasets(_args(dataset_type="llff", datadir=d, no_ndc=True, spherify=True, llffhold=0)),
           pose_tol=1e-6, path_tol=1e-5)


def test_llff_downsampled_directory_is_created_when_missing(tmp_path):
    import shutil
    from nerf_shared_b200 import datasets
    d = fx.llff(str(tmp_path / "llff"))
    shutil.rmtree(os.path.join(d, "images_4"))
    images, poses, bds, render_poses, i_test = datasets.load_llff_data(d, factor=4)
    assert images.shape == (6, 6, 8, 3) and os.path.isdir(os.path.join(d, "images_4"))
    assert poses[0, 0, 4] == 6 and poses[0, 1, 4] == 8 and abs(poses[0, 2, 4] - 0.9 * 32 / 4) < 1e-5


def test_deepvoxels_loader(tmp_path, gold):
    from nerf_shared_b200 import datasets
    d = fx.deepvoxels(str(tmp_path / "dv"))
    _check(gold, "deepvoxels", datasets.load_datasets(_args(dataset_type="deepvoxels", datadir=d, testskip=4)))


def test_unknown_dataset_type_returns_none(capsys):
    from nerf_shared_b200 import datasets
    assert datasets.load_datasets(_args(dataset_type="nope")) is None       # utils.py:283-285
    assert "Unknown dataset type" in capsys.readouterr().out


def test_config_files_parse_like_the_reference(tmp_path):
    """`training` is store_true (config_parser.py:14): a config without `training = True` does not train;
    `tensorboard = False` stays False."""
    from nerf_shared_b200 import config_parser
    cfg = tmp_path / "a.txt"
    cfg.write_text("expname = x\ndatadir = ./data\ndataset_type = blender\nno_batching = True\nuse_viewdirs = True\n"
                   "white_bkgd = True\nN_samples = 64\nN_importance = 128\nhalf_res = True\ntensorboard = False\n")
    a = config_parser.config_parser().parse_args(["--config", str(cfg)])
    assert a.training is False and a.tensorboard is False and a.no_batching and a.half_res and a.N_importance == 128
    cfg.write_text(cfg.read_text() + "training = True\ntensorboard = True\n")
    a = config_parser.config_parser().parse_args(["--config", str(cfg)])
    assert a.training is True and a.tensorboard is True
    assert config_parser.config_parser().parse_args([]).training is False
