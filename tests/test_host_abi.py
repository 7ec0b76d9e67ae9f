"""CPU-only: the C-ABI library loads, exports every symbol include/nerf_b200.h declares, the Python
mirror keeps the reference's API surface, and the product path refuses to run without a GPU."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    g.build()
    from nerf_shared_b200 import _lib
    return _lib


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "nerf_b200.h")).read()
    return sorted(set(re.findall(r"NB_API\s+[\w\s\*]+?\b(nb_\w+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(built):
    lib = ctypes.CDLL(built.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), "libnerf_b200.so does not export %s" % s
    assert set(built.SIGNATURES) == set(syms), set(built.SIGNATURES) ^ set(syms)


def test_version_and_no_cpu_path(built):
    lib = built.load()
    assert lib.nb_version() == 100
    if not torch.cuda.is_available():
        assert lib.nb_init(0) != 0
        assert b"no CUDA device" in lib.nb_last_error() or b"CPU" in lib.nb_last_error()
    from nerf_shared_b200 import nerf
    m = nerf.NeRF(use_viewdirs=True)
    with pytest.raises(built.NerfB200Error):
        m(torch.zeros(2, 4, 3), torch.zeros(2, 3))   # CPU tensors: no fallback


def test_state_dict_keys_and_shapes_match_reference():
    from nerf_shared_b200 import nerf
    from oracle import nerf_oracle as O
    m = nerf.NeRF(D=8, W=256, output_ch=5, skips=[4], use_viewdirs=True)
    ref = O.make_params(0)
    sd = m.state_dict()
    assert set(sd.keys()) == set(ref.keys())
    for k in ref:
        assert tuple(sd[k].shape) == tuple(ref[k].shape), k
    assert m.input_ch == 63 and m.input_ch_views == 27
    assert sum(p.numel() for p in m.parameters()) == 595844


def test_api_surface():
    import inspect

    from nerf_shared import nerf, render_utils, utils   # the reference's import spelling (main.py:7-8)
    for name in ("device img2mse mse2psnr to8b get_rays get_rays_np ndc_rays sample_pdf create_nerf_models create_nerf "
                 "get_renderer get_optimizer load_checkpoint load_datasets copy_log_dir batch_training_data "
                 "sample_random_ray_batch save_checkpoints print_statistics render_path render_training_video "
                 "render_test_poses").split():
        assert hasattr(utils, name), name
    R = render_utils.Renderer
    sig = inspect.signature(R.__init__)
    assert list(sig.parameters)[1:] == ["perturb", "N_importance", "N_samples", "use_viewdirs", "white_bkgd",
                                        "raw_noise_std", "ndc", "lindisp", "near", "far"]
    assert list(inspect.signature(R.render).parameters)[1:] == ["H", "W", "K", "coarse_model", "fine_model", "chunk",
                                                                "rays", "retraw", "c2w", "c2w_staticcam"]
    assert list(inspect.signature(R.render_rays).parameters)[1:] == ["ray_batch", "coarse_model", "fine_model", "retraw",
                                                                     "retweights", "verbose", "pytest"]
    assert list(inspect.signature(nerf.NeRF.__init__).parameters)[1:] == ["D", "W", "output_ch", "skips", "use_viewdirs",
                                                                          "multires", "multires_views", "i_embed"]
    for meth in ("render_from_pose", "render_from_rays", "render_batch", "raw2outputs", "render_from_batch_poses", "render_path"):
        assert hasattr(R, meth)
    fn, dim = nerf.get_embedder(10, 0)
    assert dim == 63
    e = nerf.Embedder(include_input=True, input_dims=3, max_freq_log2=3, num_freqs=4, log_sampling=True,
                      periodic_fns=[torch.sin, torch.cos])
    assert e.out_dim == 27 and len(e.embed_fns) == 9


def test_embedder_matches_golden(golden):
    from nerf_shared_b200 import nerf
    for L in (10, 4):
        fn, _ = nerf.get_embedder(L, 0)
        y = fn(torch.from_numpy(golden["embed/L%d/x" % L]))
        assert torch.equal(y, torch.from_numpy(golden["embed/L%d/y" % L]))


def test_mlp_method_matches_oracle(golden):
    """NeRF.MLP (host-side API method) on embedded inputs == reference forward"""
    from nerf_shared_b200 import nerf
    from oracle import nerf_oracle as O
    m = nerf.NeRF(D=8, W=256, output_ch=5, skips=[4], use_viewdirs=True)
    m.load_state_dict(O.make_params(11))
    pts, vd = torch.from_numpy(golden["mlp/pts"]), torch.from_numpy(golden["mlp/viewdirs"])
    x = torch.cat([O.positional_encoding(pts.reshape(-1, 3), 10),
                   O.positional_encoding(vd[:, None].expand(pts.shape).reshape(-1, 3), 4)], -1)
    with torch.no_grad():
        out = m.MLP(x).reshape(6, 16, 4)
    assert (out - torch.from_numpy(golden["mlp/raw_seed11"])).abs().max() < 2e-6


def test_synthetic_scenes_have_the_survey_shapes():
    """Blender-shaped (400x400, focal from camera_angle_x) and LLFF-shaped (504x378, fern focal)
    synthetic scenes used by bench.py; poses are rigid transforms."""
    import numpy as np
    from nerf_shared_b200 import datasets
    images, poses, hwf, K = datasets.synthetic_blender(40, 40, 3)
    assert images.shape == (3, 40, 40, 3) and poses.shape == (3, 4, 4) and abs(hwf[2] - 0.5 * 40 / np.tan(0.5 * 0.6911112070083618)) < 1e-9
    images, poses, hwf, K = datasets.synthetic_llff(378, 504, 2)
    assert images.shape == (2, 378, 504, 3) and poses.shape == (2, 4, 4)
    assert abs(K[0][0] - 3260.5 / 8) < 1e-9 and K[0][2] == 252 and K[1][2] == 189
    for p in poses:
        R = p[:3, :3].astype(np.float64)
        np.testing.assert_allclose(R @ R.T, np.eye(3), atol=1e-6)
        assert abs(np.linalg.det(R) - 1.0) < 1e-6 and np.all(np.abs(p[:3, 3]) < 0.5)


def test_graphed_step_refuses_cpu():
    """train.GraphedTrainStep is part of the product path: no CPU fallback (it needs the FlatAdam buffers on a GPU)."""
    import pytest
    import torch
    from nerf_shared_b200 import train
    assert hasattr(train.GraphedTrainStep, "_step_overlap")
    if not torch.cuda.is_available():
        class _Opt:
            flat_p = torch.zeros(1)
        class _R:
            N_importance = 128
        with pytest.raises(Exception):
            train.GraphedTrainStep(_R(), None, object(), _Opt(), 8, 8, None, 16)   # torch.cuda.Stream on a CPU-only host
