"""Positional encoder and NeRF model behind the reference's API (nerf_shared/nerf.py).

Same class / function names, constructor arguments, attributes and state_dict keys as the
reference (nerf.py:11-143), so checkpoints load both ways; the arithmetic of `NeRF.forward`
runs in libnerf_b200's fused sm_100a kernels (encoding + 8x256 trunk + heads in one launch).
"""
from __future__ import annotations

import os

import torch
import torch.nn as nn

from . import _ops

_DEFAULT_PRECISION = os.environ.get("NERF_B200_PRECISION", "bf16")


def set_default_precision(p: str):
    """'bf16' = tcgen05 tensor-core kernels (bf16 operands, fp32 accumulate); 'fp32' = CUDA-core
    fp32 kernel (parity / debugging, forward only)."""
    global _DEFAULT_PRECISION
    assert p in ("bf16", "fp32")
    _DEFAULT_PRECISION = p


class Embedder:
    """[x, sin(f0 x), cos(f0 x), ..., sin(f_{L-1} x), cos(f_{L-1} x)] (reference nerf.py:11-41).

    Kept as a host-side object for API compatibility (`.embed`, `.out_dim`, `.embed_fns`,
    `.kwargs`); NeRF.forward never calls it -- the fused kernel builds the encoding in registers."""

    def __init__(self, **kwargs):
        self.kwargs = kwargs
        self.create_embedding_fn()

    def create_embedding_fn(self):
        kw = self.kwargs
        dims, n = kw["input_dims"], kw["num_freqs"]
        if kw["log_sampling"]:
            bands = 2.0 ** torch.linspace(0.0, kw["max_freq_log2"], steps=n, device="cpu")
        else:
            bands = torch.linspace(2.0 ** 0.0, 2.0 ** kw["max_freq_log2"], steps=n, device="cpu")
        fns = [(lambda x: x)] if kw["include_input"] else []
        for f in bands.tolist():
            for periodic in kw["periodic_fns"]:
                fns.append(lambda x, periodic=periodic, f=f: periodic(x * f))
        self.embed_fns = fns
        self.out_dim = dims * len(fns)
        self.freq_bands = bands

    def embed(self, inputs):
        return torch.cat([fn(inputs) for fn in self.embed_fns], -1)


def get_embedder(multires, i=0):
    if i == -1:
        return nn.Identity(), 3
    eo = Embedder(include_input=True, input_dims=3, max_freq_log2=multires - 1, num_freqs=multires,
                  log_sampling=True, periodic_fns=[torch.sin, torch.cos])
    return (lambda x, eo=eo: eo.embed(x)), eo.out_dim


class NeRF(nn.Module):
    def __init__(self, D=8, W=256, output_ch=4, skips=[4], use_viewdirs=False, multires=10, multires_views=4,
                 i_embed=0):
        super().__init__()
        self.D, self.W, self.skips, self.use_viewdirs = D, W, skips, use_viewdirs
        self.multires, self.multires_views, self.i_embed = multires, multires_views, i_embed
        self.precision = _DEFAULT_PRECISION
        if multires > 10 or (use_viewdirs and multires_views > 4) or not use_viewdirs:
            # encodings wider than the tensor-core tiling (63 / 27 columns; configs/stonehenge.txt uses 93 / 39) and the
            # trunk -> output_linear variant run on the fp32 CUDA-core kernels, forward and backward
            self.precision = "fp32"
        self.embed_fn, self.input_ch = get_embedder(multires, i_embed)
        self.input_ch_views, self.embeddirs_fn = 0, None
        if use_viewdirs:
            self.embeddirs_fn, self.input_ch_views = get_embedder(multires_views, i_embed)
        # identical parameter shapes / state_dict keys to the reference (nerf.py:79-94)
        layers = [nn.Linear(self.input_ch, W)]
        for i in range(D - 1):
            layers.append(nn.Linear(W + self.input_ch, W) if i in skips else nn.Linear(W, W))
        self.pts_linears = nn.ModuleList(layers)
        self.views_linears = nn.ModuleList([nn.Linear(self.input_ch_views + W, W // 2)])
        if use_viewdirs:
            self.feature_linear = nn.Linear(W, W)
            self.alpha_linear = nn.Linear(W, 1)
            self.rgb_linear = nn.Linear(W // 2, 3)
        else:
            self.output_linear = nn.Linear(W, output_ch)

    def _check_kernel_config(self):
        if not (self.D == 8 and self.W == 256 and list(self.skips) == [4] and self.i_embed == 0):
            raise NotImplementedError(
                "libnerf_b200 implements the network every reference config uses (D=8, W=256, skips=[4], "
                "i_embed=0); got D=%s W=%s skips=%s i_embed=%s" % (self.D, self.W, self.skips, self.i_embed))

    def _forward_noview(self, inputs):
        """use_viewdirs=False (nerf.py:93-94,131-132): trunk -> output_linear on the fp32 CUDA-core kernels: the fused
        forward kernel without gradients, the layer-by-layer training path (csrc/mlp_f32_train.cu) with them."""
        if torch.is_grad_enabled() and (inputs.requires_grad or any(p.requires_grad for p in self.parameters())):
            raw = _ops.mlp_f32_trainable(self, inputs.reshape(-1, 3), None, 1)
        else:
            raw = _ops.mlp_forward_f32_noview(self, inputs.reshape(-1, 3))
        return raw.reshape(list(inputs.shape[:-1]) + [raw.shape[-1]])

    def forward(self, inputs, viewdirs, netchunk=1024 * 64):
        """inputs [N, S, 3], viewdirs [N, 3] -> [N, S, 4] (rgb pre-sigmoid, sigma pre-relu);
        use_viewdirs=False models: viewdirs=None -> [N, S, output_ch].

        `netchunk` is accepted for signature compatibility (nerf.py:96,106) and ignored: the fused
        kernel keeps activations on-chip, so there is nothing to chunk."""
        self._check_kernel_config()
        if not self.use_viewdirs:
            if viewdirs is not None:   # the reference fails the same way: embeddirs_fn is None (nerf.py:100-103)
                raise TypeError("a use_viewdirs=False model has no view-direction encoder; pass viewdirs=None")
            return self._forward_noview(inputs)
        if viewdirs is None:
            raise NotImplementedError("use_viewdirs=True models need viewdirs")
        n_rays = viewdirs.reshape(-1, 3).shape[0]
        pts = inputs.reshape(-1, 3)
        if pts.shape[0] % max(n_rays, 1) != 0:
            raise ValueError("inputs %s do not broadcast against viewdirs %s" % (tuple(inputs.shape), tuple(viewdirs.shape)))
        S = pts.shape[0] // max(n_rays, 1)
        if self.precision == "fp32" and torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            # exact-operand training path (also the only one for multires > 10 / multires_views > 4)
            raw = _ops.mlp_f32_trainable(self, pts, viewdirs.reshape(-1, 3), S)
            return raw.reshape(list(inputs.shape[:-1]) + [4])
        raw = _ops.NerfMLP.apply(pts, viewdirs.reshape(-1, 3), S, self, torch.is_grad_enabled(), *_ops.mlp_params(self))
        return raw.reshape(list(inputs.shape[:-1]) + [4])

    def MLP(self, x):
        """Layer stack on an already-embedded input [.., input_ch + input_ch_views] (nerf.py:110-134).
        Plain torch modules; kept for API compatibility, not used by forward()."""
        import torch.nn.functional as F
        input_pts, input_views = torch.split(x, [self.input_ch, self.input_ch_views], dim=-1)
        h = input_pts
        for i, layer in enumerate(self.pts_linears):
            h = F.relu(layer(h))
            if i in self.skips:
                h = torch.cat([input_pts, h], -1)
        if not self.use_viewdirs:
            return self.output_linear(h)
        alpha = self.alpha_linear(h)
        h = torch.cat([self.feature_linear(h), input_views], -1)
        for layer in self.views_linears:
            h = F.relu(layer(h))
        return torch.cat([self.rgb_linear(h), alpha], -1)

    def get_density(self, points, chunk=1024 * 64):
        """points [..., P, 3] -> sigma [..., P] with an all-ones view direction (nerf.py:136-143)."""
        view_dir = torch.ones_like(points[..., 0, :])
        return self.forward(points, view_dir, chunk)[..., -1]
