"""Helpers shared by the -m gpu tests."""
import numpy as np
import torch

from oracle import nerf_oracle as O


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def cuda(a):
    return (a if torch.is_tensor(a) else T(a)).cuda()


def bits_equal(a, b):
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    b = b.detach().cpu().numpy() if torch.is_tensor(b) else np.asarray(b)
    return a.shape == b.shape and np.ascontiguousarray(a).tobytes() == np.ascontiguousarray(b).tobytes()


def n_diff(a, b):
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    b = b.detach().cpu().numpy() if torch.is_tensor(b) else np.asarray(b)
    return int((a.view(np.uint32) != b.view(np.uint32)).sum()), a.size


def make_model(seed, scale=1.0, precision="bf16"):
    """nerf_shared_b200 NeRF carrying the oracle's seeded parameters."""
    from nerf_shared_b200 import nerf
    m = nerf.NeRF(D=8, W=256, output_ch=5, skips=[4], use_viewdirs=True, multires=10, multires_views=4)
    m.load_state_dict(O.make_params(seed, scale=scale), strict=True)
    m.precision = precision
    return m.cuda()


def maxabs(a, b):
    return float((a.detach().float().cpu() - b.detach().float().cpu()).abs().max())
