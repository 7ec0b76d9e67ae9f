"""-m gpu: the drop-in promise on hardware -- the reference's own caller scripts, text
unmodified, run against this package: main.py's training loop on a Blender-format scene on disk, and the pose-estimation
demo's optimisation loop through the root-level `utils` / `render_utils` shims under the CUDA default tensor type.

The scripts are read from oracle/_ref/ (staged by __graft_entry__.build() from /root/reference in the build container;
git-ignored, travels with the gpurun snapshot) or from $NERF_REFERENCE_ROOT."""
import json
import os
import subprocess
import sys
import types

import numpy as np
import pytest
import torch

from oracle import nerf_oracle as O
from tests import dataset_fixtures as fx

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _script(name, rel):
    for p in (os.path.join(ROOT, "oracle", "_ref", name),
              os.path.join(os.environ.get("NERF_REFERENCE_ROOT", "/root/reference"), rel)):
        if os.path.isfile(p):
            return p
    pytest.skip("reference script %s not staged (run __graft_entry__.build() where /root/reference exists)" % name)


def _run_main(tmp_path, extra_cfg="", n_iters=24, images_hw=40, batching=False):
    main_py = _script("main.py", "main.py")
    data = fx.blender(str(tmp_path / "scene"), H=images_hw, W=images_hw, counts=(4, 1, 2))
    cfg = tmp_path / "scene.txt"
    cfg.write_text("expname = dropin\nbasedir = %s\ndatadir = %s\ndataset_type = blender\ntraining = True\n"
                   "%suse_viewdirs = True\nwhite_bkgd = True\nlrate_decay = 500\nN_samples = 64\n"
                   "N_importance = 128\nN_rand = 1024\ni_print = 6\ni_weights = 12\ni_testset = 12\n%s"
                   % (tmp_path / "logs", data, "" if batching else "no_batching = True\n", extra_cfg))
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "run_reference_script.py"), str(n_iters), main_py,
                        "--config", str(cfg)], capture_output=True, text=True, cwd=str(tmp_path), env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + "\n" + r.stderr[-3000:]
    return r, tmp_path / "logs" / "dropin"


def test_reference_main_py_runs_unchanged(tmp_path):
    """main.py:16-141 as shipped: config file -> utils.load_datasets (Blender JSON + PNGs on disk) -> create_nerf_models
    -> get_optimizer -> load_checkpoint -> get_renderer -> batch_training_data -> 24 iterations of sample_random_ray_batch /
    render_from_rays / img2mse / backward / optimizer.step / lr decay, with a checkpoint and a test-set render at
    iteration 12, all under torch.set_default_tensor_type('torch.cuda.FloatTensor') (main.py:151)."""
    r, logdir = _run_main(tmp_path)
    line = [l for l in r.stdout.splitlines() if l.startswith("DROPIN_MAIN ")]
    assert line, r.stdout[-2000:]
    info = json.loads(line[-1][len("DROPIN_MAIN "):])
    print("main.py unchanged:", info)
    # (launches issued from Python; the steady-state steps replay the two lazily captured graphs of _fused.py)
    assert info["default_tensor_type"] == "cuda" and info["kernel_launches"] > 100
    losses = [float(l.split("Loss:")[1].split()[0]) for l in (r.stdout + r.stderr).splitlines() if "[TRAIN] Iter:" in l]
    assert len(losses) == 4 and all(np.isfinite(losses)), losses
    assert os.path.isfile(logdir / "000012.tar") and os.path.isfile(logdir / "000024.tar")
    pngs = sorted(os.listdir(logdir / "testset_000012"))
    assert pngs == ["000.png"], pngs            # testskip = 8 (default) keeps one of the two test frames
    ck = torch.load(logdir / "000024.tar", map_location="cpu")
    assert set(ck) == {"global_step", "coarse_model_state_dict", "fine_model_state_dict", "optimizer_state_dict"}
    assert ck["global_step"] == 23 and  ck["coarse_model_state_dict"]["pts_linears.5.weight"].shape == (256, 319)
    # the random targets are unlearnable, but Adam must have moved the weights and the moments must be populated
    st = ck["optimizer_state_dict"]["state"]
    assert len(st) == 48 and float(st[0]["exp_avg_sq"].abs().sum()) > 0


def test_reference_main_py_batching_branch(tmp_path):
    """use_batching (no `no_batching` in the config: utils.py:366-381,397-409): rays of all training images are
    precomputed with get_rays_np, shuffled, and consumed N_rand at a time, reshuffling after an epoch
    (4 images x 16 x 16 = 1,024 rays = one batch per epoch)."""
    r, logdir = _run_main(tmp_path, n_iters=12, images_hw=16, batching=True)
    assert "Shuffle data after an epoch!" in r.stdout
    assert os.path.isfile(logdir / "000012.tar")


def test_reference_pose_demo_loop_runs_unchanged():
    """examples/relative_pose_estimation_demo/demo_est_rel_pose.py: its estimate_relative_pose() -- SIFT interest regions,
    camera_transf (an nn.Module found through `from utils import *`), 300 Adam steps of get_rays(full grid, pose with
    grad) -> 512-pixel batch -> render_from_rays (frozen weights) -> img2mse -> backward -- executed from the script's
    own text through the root shims, under the CUDA default tensor type its __main__ sets (demo:361)."""
    demo = _script("demo_est_rel_pose.py", os.path.join("examples", "relative_pose_estimation_demo", "demo_est_rel_pose.py"))
    for name in ("torchtyping", "configargparse"):         # imported at the top of the demo, not used by the loop
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.TensorDetail = m.TensorType = m.ArgumentParser = object
            sys.modules[name] = m
    fake = os.path.join(ROOT, "examples", "relative_pose_estimation_demo", "demo_est_rel_pose.py")   # dirname^3 = repo root
    ns = {"__name__": "demo_est_rel_pose", "__file__": fake}
    saved_path = list(sys.path)
    torch.set_default_tensor_type("torch.cuda.FloatTensor")
    try:
        exec(compile(open(demo).read(), fake, "exec"), ns)
        assert ns["Renderer"].__module__ == "nerf_shared_b200.render_utils"       # the star imports hit the shims
        assert ns["get_rays"].__module__ == "nerf_shared_b200.utils"
        from tests.gpu_util import make_model
        H = W = 48
        K = O.blender_intrinsics(H, W)
        mc, mf = make_model(13, scale=1.59), make_model(14, scale=1.59)
        for m in (mc, mf):
            for p in m.parameters():
                p.requires_grad_(False)
        renderer = ns["Renderer"](perturb=0.0, N_importance=64, N_samples=64, use_viewdirs=True, white_bkgd=True,
                                  raw_noise_std=0.0, near=2.0, far=6.0)
        gt = O.pose_spherical(30.0, -30.0, 4.0)
        with torch.no_grad():
            rgb = renderer.render_from_pose(H, W, K, 32768, torch.Tensor(gt[:3, :4]).cuda(), mc, mf, retraw=False)[0]
        obs = (255 * np.clip(rgb.cpu().numpy(), 0, 1)).astype(np.uint8)
        # SIFT needs texture: overlay a deterministic checker so that interest regions exist
        yy, xx = np.mgrid[0:H, 0:W]
        obs = np.where((((yy // 6) + (xx // 6)) % 2 == 0)[..., None], obs, 255 - obs).astype(np.uint8)
        start = ns["rot_theta"](2.0 / 180.0 * np.pi) @ gt
        args = types.SimpleNamespace(chunk=32768)
        extra = {"model_name": "t", "output_dir": "/tmp/nerf_b200_dropin_demo/", "dil_iter": 3, "batch_size": 512,
                 "kernel_size": 5, "lrate": 0.01}
        from nerf_shared_b200 import _lib
        l0 = _lib.launch_count()
        ns["estimate_relative_pose"](mc, mf, renderer, obs, start, K, args, extra, obs_img_pose=gt, obs_img=None)
        torch.cuda.synchronize()
        assert _lib.launch_count() - l0 >= 300 * 2           # 300 iterations ran on the library (the renders replay captured graphs)
    finally:
        torch.set_default_tensor_type("torch.FloatTensor")
        sys.path[:] = saved_path


def test_reference_main_py_trains_the_wide_encoding_config(tmp_path):
    """configs/stonehenge.txt's network (multires = 15, multires_views = 6: 93- / 39-wide encodings) through the unchanged
    main.py: the models select the fp32 kernels (forward and backward) on their own."""
    r, logdir = _run_main(tmp_path, extra_cfg="multires = 15\nmultires_views = 6\ni_print = 2\n", n_iters=6, images_hw=40)
    text = r.stdout + r.stderr
    losses = [float(l.split("Loss:")[1].split()[0]) for l in text.splitlines() if "[TRAIN] Iter:" in l]
    assert len(losses) >= 3 and all(np.isfinite(losses)), (losses, text[-1500:])
