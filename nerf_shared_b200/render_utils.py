"""Volume renderer behind the reference's API (nerf_shared/render_utils.py:13-319).

`Renderer` keeps the reference's constructor, method names, argument orders, defaults and return
structures; every stage runs in libnerf_b200 (sm_100a CUDA) through the autograd wrappers in
_ops.py:  rays -> stratified z -> fused encoding+MLP -> compositing -> inverse-CDF resampling ->
fine MLP -> compositing.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import _fused, _ops
from . import nerf   # noqa: F401  (the reference module exposes it; demos star-import this file)
from . import utils

DEBUG = False

try:  # only used to write PNG/MP4 files (render_from_batch_poses); not on the compute path
    import imageio
except Exception:  # pragma: no cover
    imageio = None

# rays per internal launch: `chunk` of the reference API is an upper bound on memory, results
# do not depend on it (render_utils.py:184-185); the fused kernels need far less scratch per
# ray, so small chunks are coalesced up to this many rays.
MIN_CHUNK = 1 << 18


def _write_image(path, rgb8):
    if imageio is not None and hasattr(imageio, "imwrite"):
        imageio.imwrite(path, rgb8)
    else:
        import cv2
        cv2.imwrite(path, rgb8[..., ::-1])


class Renderer(torch.nn.Module):
    def __init__(self, perturb=True, N_importance=128, N_samples=64, use_viewdirs=True, white_bkgd=True,
                 raw_noise_std=0.0, ndc=False, lindisp=False, near=0.0, far=1.0):
        super().__init__()
        self.perturb, self.N_importance, self.N_samples = perturb, N_importance, N_samples
        self.use_viewdirs, self.white_bkgd, self.raw_noise_std = use_viewdirs, white_bkgd, raw_noise_std
        self.ndc, self.lindisp, self.near, self.far = ndc, lindisp, near, far
        print({k: v for k, v in self.__dict__.items() if not k.startswith("_") and k != "training"})

    # ------------------------------------------------------------------ random draws
    # The reference draws from the global torch RNG in this order: t_rand (render_utils.py:121),
    # coarse noise (:264), u (utils.py:86), fine noise (:264).  Tests override these hooks to
    # feed the oracle's draws.
    def _uniform(self, shape, device, what):
        return torch.rand(shape, device=device, dtype=torch.float32)

    def _normal(self, shape, device, what):
        return torch.randn(shape, device=device, dtype=torch.float32)

    _row_parts = None
    _after_coarse = None   # train.GraphedTrainStep: called with the coarse map before the fine pass is issued

    def _columns_of(self, ray_batch):
        """contiguous (rays_o, rays_d, viewdirs) of `ray_batch` when it is a row range of the rows render() built."""
        if self._row_parts is None:
            return None
        rows, parts = self._row_parts
        if (ray_batch.dim() != 2 or ray_batch.shape[1] != 11 or not ray_batch.is_contiguous()
                or ray_batch.untyped_storage().data_ptr() != rows.untyped_storage().data_ptr()):
            return None
        i0 = (ray_batch.storage_offset() - rows.storage_offset()) // 11
        return tuple(t[i0:i0 + ray_batch.shape[0]] for t in parts)

    # ------------------------------------------------------------------ entry points
    def render_from_pose(self, H, W, K, chunk, c2w, coarse_model, fine_model, retraw=True):
        return tuple(self.render(H, W, K, coarse_model, fine_model, chunk=chunk, c2w=c2w, retraw=retraw))

    def render_from_rays(self, H, W, K, chunk, rays, coarse_model, fine_model, retraw=True):
        return tuple(self.render(H, W, K, coarse_model, fine_model, chunk=chunk, rays=rays, retraw=retraw))

    def render_path(self):
        pass

    def render_batch(self, coarse_model, fine_model, rays_flat, chunk=1024 * 32, retraw=False):
        step = max(int(chunk), MIN_CHUNK)
        parts = [self.render_rays(rays_flat[i:i + step], coarse_model, fine_model, retraw)
                 for i in range(0, rays_flat.shape[0], step)]
        if len(parts) == 1:
            return parts[0]
        return {k: torch.cat([p[k] for p in parts], 0) for k in parts[0]}

    def render_rays(self, ray_batch, coarse_model, fine_model, retraw=False, retweights=False, verbose=False,
                    pytest=False):
        """ray_batch rows: o(3) d(3) near far [viewdirs(3)] -> dict as the reference (render_utils.py:67-174)."""
        n = ray_batch.shape[0]
        dev = ray_batch.device
        parts = self._columns_of(ray_batch)
        if parts is not None:
            rays_o, rays_d, viewdirs = parts
        else:                                  # 8 columns: no view directions (render_utils.py:100)
            rays_o, rays_d = ray_batch[:, 0:3], ray_batch[:, 3:6]
            viewdirs = ray_batch[:, 8:11] if ray_batch.shape[-1] > 8 else None
        S, I = int(self.N_samples), int(self.N_importance)

        t_rand = self._uniform((n, S), dev, "t_rand") if self.perturb > 0.0 else None
        if torch.is_grad_enabled() and (rays_o.requires_grad or rays_d.requires_grad):
            z_vals = _ops.sample_coarse(ray_batch, S, t_rand, self.lindisp)
            pts = _ops.RayPoints.apply(rays_o, rays_d, z_vals)      # pose estimation: differentiable in o, d
        else:                                                        # depths and points from one launch
            z_vals, pts = _ops.sample_coarse(ray_batch, S, t_rand, self.lindisp, want_pts=True)
        raw = coarse_model(pts, viewdirs)
        rgb_map, disp_map, acc_map, weights, depth_map = self.raw2outputs(raw, z_vals, rays_d, _what="noise_c")

        ret = {}
        if I > 0:
            ret["rgb0"], ret["disp0"], ret["acc0"] = rgb_map, disp_map, acc_map
            if self._after_coarse is not None:
                self._after_coarse(rgb_map, (ray_batch, rays_o, rays_d, viewdirs, t_rand, z_vals, pts, raw, weights))
            u = self._uniform((n, I), dev, "u") if self.perturb != 0.0 else None   # det = (perturb == 0)
            z_vals, z_std = _ops.sample_pdf_merge(z_vals, weights, I, u)
            pts = _ops.RayPoints.apply(rays_o, rays_d, z_vals)
            raw = (coarse_model if fine_model is None else fine_model)(pts, viewdirs)
            rgb_map, disp_map, acc_map, weights, depth_map = self.raw2outputs(raw, z_vals, rays_d, _what="noise_f")
            ret["z_std"] = z_std
        ret["rgb_map"], ret["disp_map"], ret["acc_map"] = rgb_map, disp_map, acc_map
        if retraw:
            ret["raw"] = raw
        if retweights:
            ret["weights"], ret["z_vals"] = weights, z_vals
        # (the reference's per-key isnan/isinf host syncs, render_utils.py:170-172, only feed a
        #  DEBUG print and are dropped)
        return ret

    def render(self, H, W, K, coarse_model, fine_model, chunk=1024 * 32, rays=None, retraw=True, c2w=None,
               c2w_staticcam=None):
        """-> [rgb_map, disp_map, acc_map, extras] shaped like the rays (render_utils.py:176-238)."""
        if c2w is not None:
            rays_o, rays_d = utils.get_rays(H, W, K, c2w)
        else:
            rays_o, rays_d = rays
        viewdirs = rays_d if self.use_viewdirs else None
        if self.use_viewdirs and c2w_staticcam is not None:      # render_utils.py:207-210
            rays_o, rays_d = utils.get_rays(H, W, K, c2w_staticcam)
        out_shape = list(rays_d.shape[:-1])
        if viewdirs is None:
            return self._render_no_viewdirs(H, W, K, coarse_model, fine_model, chunk, rays_o, rays_d, retraw, out_shape)
        needs_grad = torch.is_grad_enabled() and any(torch.is_tensor(t) and t.requires_grad for t in (rays_o, rays_d, viewdirs))
        view_src = viewdirs.reshape(-1, 3)
        if self.ndc:             # (an autograd op when the rays carry a gradient)
            rays_o, rays_d = utils.ndc_rays(H, W, K[0][0], 1.0, rays_o, rays_d)
        rays_o = rays_o.reshape(-1, 3).float().contiguous()
        rays_d = rays_d.reshape(-1, 3).float().contiguous()
        no_grads = not (torch.is_grad_enabled() and (needs_grad or any(p.requires_grad for p in coarse_model.parameters())
                                                     or (fine_model is not None and any(p.requires_grad for p in fine_model.parameters()))))
        if (no_grads and self._after_coarse is None and _ops.tc_model_ok(coarse_model) and rays_d.shape[0] > 0
                and (fine_model is None or _ops.tc_model_ok(fine_model)) and (self.N_importance > 0 or not retraw) and os.environ.get("NB_FUSED_INFERENCE", "1") != "0"):
            # inference: every chunk of rays is ONE C-ABI call (nb_render_rays_fwd), results land in the full-size outputs
            rgb, disp, acc, rgb0, disp0, acc0, z_std, raw = _ops.render_rays_inference(
                self, coarse_model, fine_model, rays_o, rays_d, view_src.float().contiguous(), max(int(chunk), MIN_CHUNK), retraw)
            shp = lambda t: t.reshape(out_shape + list(t.shape[1:]))
            extras = {}
            if rgb0 is not None:
                extras = {"rgb0": shp(rgb0), "disp0": shp(disp0), "acc0": shp(acc0), "z_std": shp(z_std)}
            if retraw:
                extras["raw"] = shp(raw)
            return [shp(rgb), shp(disp), shp(acc), extras]
        if self._after_coarse is None and _fused.eligible(self, coarse_model, fine_model, rays_o, rays_d, rays_d.shape[0],
                                                          max(int(chunk), MIN_CHUNK)):
            # a training batch (rays without gradient) or a pose-estimation batch (rays with gradient): the whole coarse +
            # fine render as ONE autograd node, replaying lazily captured CUDA graphs where it can (_fused.py)
            out = _fused.FusedRender.apply(self, coarse_model, fine_model, rays_o, rays_d, view_src.float().contiguous(),
                                           *_ops.mlp_params(coarse_model), *_ops.mlp_params(fine_model))
            rgb, disp, acc, rgb0, disp0, acc0, z_std, raw = out
            shp = lambda t: t.reshape(out_shape + list(t.shape[1:]))
            extras = {"rgb0": shp(rgb0), "disp0": shp(disp0), "acc0": shp(acc0), "z_std": shp(z_std)}
            if retraw:
                extras["raw"] = shp(raw)
            return [shp(rgb), shp(disp), shp(acc), extras]
        if needs_grad:   # pose estimation outside the fused node's configurations: every step is a differentiable op
            viewdirs = _ops.Normalize.apply(viewdirs).reshape(-1, 3)
            bounds = rays_d.new_empty((rays_d.shape[0], 2))
            bounds[:, 0] = float(self.near)
            bounds[:, 1] = float(self.far)
            ray_rows = torch.cat([rays_o, rays_d, bounds, viewdirs], -1)
            parts = None
        else:            # one kernel: normalise + near/far + concatenate
            ray_rows, viewdirs = _ops.ray_rows(rays_o, rays_d, view_src, self.near, self.far)
            parts = (rays_o, rays_d, viewdirs)

        # render_rays would have to copy the o / d / viewdirs columns back out of the rows; render() still holds them
        self._row_parts = None if parts is None else (ray_rows, parts)
        try:
            res = self.render_batch(coarse_model, fine_model, ray_rows, chunk, retraw)
        finally:
            self._row_parts = None
        res = {k: v.reshape(out_shape + list(v.shape[1:])) for k, v in res.items()}
        head = ("rgb_map", "disp_map", "acc_map")
        return [res[k] for k in head] + [{k: v for k, v in res.items() if k not in head}]

    def _render_no_viewdirs(self, H, W, K, coarse_model, fine_model, chunk, rays_o, rays_d, retraw, out_shape):
        """render() for use_viewdirs=False: ray rows are [N,8] = o, d, near, far (render_utils.py:219-226).
        The models of this mode run forward-only (nerf.NeRF._forward_noview)."""
        if self.ndc:
            rays_o, rays_d = utils.ndc_rays(H, W, K[0][0], 1.0, rays_o, rays_d)
        rays_o = rays_o.reshape(-1, 3).float()
        rays_d = rays_d.reshape(-1, 3).float()
        bounds = rays_d.new_empty((rays_d.shape[0], 2))
        bounds[:, 0] = float(self.near)
        bounds[:, 1] = float(self.far)
        res = self.render_batch(coarse_model, fine_model, torch.cat([rays_o, rays_d, bounds], -1), chunk, retraw)
        res = {k: v.reshape(out_shape + list(v.shape[1:])) for k, v in res.items()}
        head = ("rgb_map", "disp_map", "acc_map")
        return [res[k] for k in head] + [{k: v for k, v in res.items() if k not in head}]

    def raw2outputs(self, raw, z_vals, rays_d, pytest=False, _what="noise"):
        """-> rgb_map, disp_map, acc_map, weights, depth_map (render_utils.py:241-290)."""
        if raw.shape[-1] != 4:      # output_ch=5 of a use_viewdirs=False model: channels 0-3 are read (:260,265)
            raw = raw[..., :4].contiguous()
        noise = None
        if self.raw_noise_std > 0.0:
            noise = self._normal(tuple(raw.shape[:-1]), raw.device, _what) * self.raw_noise_std
        return _ops.Composite.apply(raw, z_vals, rays_d, noise, bool(self.white_bkgd))

    def render_from_batch_poses(self, H, W, K, chunk, batch_c2w, coarse_model, fine_model, retraw, save_directory,
                                b_combine_as_video=False, tb_writer=None):
        """Render a list of poses to PNG files / a video / TensorBoard (render_utils.py:293-319).

        The frame tail is asynchronous: frame i is converted to uint8 on the device (3 bytes per pixel cross PCIe, not
        12) and copied into one of two pinned host buffers on a side stream; the host then waits for frame i-1 only
        and encodes its PNG while the GPU is already rendering frame i (FrameDownloader)."""
        os.makedirs(save_directory, exist_ok=True)
        frames = []

        def finish(i, arrs):
            frames.append(arrs[0])
            _write_image(os.path.join(save_directory, "{:03d}.png".format(i)), arrs[0])

        down = FrameDownloader(finish)
        with torch.no_grad():
            for i, c2w in enumerate(batch_c2w):
                c2w = c2w if torch.is_tensor(c2w) else torch.as_tensor(np.asarray(c2w), dtype=torch.float32)
                rgb, _, _, _ = self.render_from_pose(H, W, K, chunk=chunk, c2w=c2w, coarse_model=coarse_model,
                                                     fine_model=fine_model, retraw=False)
                down.submit(i, (utils.to8b_device(rgb),))
            down.drain()
        if b_combine_as_video and imageio is not None and hasattr(imageio, "mimwrite"):
            imageio.mimwrite(os.path.join(save_directory, "video.mp4"), frames, fps=30, quality=8)
        if tb_writer is not None:
            tb_writer.add_images("Test/Images", torch.from_numpy(np.stack(frames, 0)), dataformats="NHWC")


class FrameDownloader:
    """Double-buffered device -> pinned-host copies of per-frame results on a side stream.

    submit(i, tensors) queues the copy of frame i's tensors behind the kernels that produce them and then hands frame
    i-1 (whose copy has had a whole frame's render time to finish) to `on_frame(i-1, [numpy arrays])`; the host-side
    work of a frame (PNG encoding, stacking) therefore overlaps the next frame's kernels instead of a per-frame
    `.cpu()` stall.  drain() delivers the last frame."""

    def __init__(self, on_frame):
        self.on_frame = on_frame
        self.slots = [None, None]
        self.pending = None
        self.stream = None

    def _slot(self, k, tensors):
        s = self.slots[k]
        if s is None or any(h.shape != t.shape or h.dtype != t.dtype for h, t in zip(s["host"], tensors)) \
                or len(s["host"]) != len(tensors):
            s = {"host": [torch.empty(t.shape, dtype=t.dtype, device="cpu", pin_memory=True) for t in tensors],
                 "ev": torch.cuda.Event()}
            self.slots[k] = s
        return s

    def submit(self, i, tensors):
        dev = tensors[0].device
        if self.stream is None:
            self.stream = torch.cuda.Stream(device=dev)
        k = 0 if self.pending is None else 1 - self.pending[1]
        s = self._slot(k, tensors)
        self.stream.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(self.stream):
            for h, t in zip(s["host"], tensors):
                h.copy_(t, non_blocking=True)
                t.record_stream(self.stream)
            s["ev"].record(self.stream)
        prev, self.pending = self.pending, (i, k)
        if prev is not None:
            self._deliver(prev)

    def _deliver(self, item):
        i, k = item
        s = self.slots[k]
        s["ev"].synchronize()
        self.on_frame(i, [h.numpy().copy() for h in s["host"]])

    def drain(self):
        if self.pending is not None:
            self._deliver(self.pending)
            self.pending = None
