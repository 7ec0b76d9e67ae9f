"""Ray math, sampling and training glue behind the reference's API (nerf_shared/utils.py).

get_rays / ndc_rays / sample_pdf run in libnerf_b200 (sm_100a CUDA); the factory / checkpoint /
logging helpers are thin Python with the reference's names and argument orders so that the
reference's main.py and pose-estimation demo run unchanged.
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.nn as nn              # noqa: F401  (the demos rely on `from utils import *` leaking nn, F)
import torch.nn.functional as F    # noqa: F401

from . import _lib, _ops

device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
DEBUG = False

# ----------------------------------------------------------------------------- misc (utils.py:24-30)
def img2mse(x, y):
    """mean((x - y)^2) (utils.py:24).  Two fp32 CUDA tensors of one shape take one launch forward and one backward
    (nb_mse_pair_*) instead of three elementwise kernels and their autograd nodes each way."""
    if (torch.is_tensor(x) and torch.is_tensor(y) and x.is_cuda and y.is_cuda and x.dtype == torch.float32
            and y.dtype == torch.float32 and x.shape == y.shape and 0 < x.numel() <= (1 << 18) and not y.requires_grad):   # (one-block kernel: batches, not whole frames)
        return _ops.MsePair.apply(x, None, y)[1]
    return torch.mean((x - y) ** 2)


_NEG10_OVER_LN10 = -10.0 / float(np.log(10.0))


def mse2psnr(x):
    """-10 log10(x) (utils.py:27), shaped [1] for a scalar like the reference's division by a 1-element tensor; the
    constant is a Python float (the reference builds a tensor from a list on every call: a host-to-device copy)."""
    r = torch.log(x) * _NEG10_OVER_LN10
    return r.reshape(1) if r.dim() == 0 else r


to8b = lambda x: (255 * np.clip(x, 0, 1)).astype(np.uint8)


def to8b_device(x):
    """to8b of a CUDA float tensor on the device -> uint8 CUDA tensor (a frame leaves the GPU as bytes)."""
    _lib.ensure_device(x.device)        # raises for CPU tensors: there is no host path here (utils.to8b is the NumPy one)
    lib = _lib.load()
    xc = _lib.f32c(x.detach())
    out = torch.empty(xc.shape, device=xc.device, dtype=torch.uint8)
    with torch.cuda.device(xc.device):
        _lib.check(lib.nb_to8b(_lib.ptr(xc), xc.numel(), _lib.ptr(out), _lib.stream_ptr()), "nb_to8b")
    return out


def _on_gpu(t, like=None):
    t = torch.as_tensor(t, dtype=torch.float32) if not torch.is_tensor(t) else t
    if t.device.type != "cuda":
        t = t.to(like.device if (like is not None and like.device.type == "cuda") else device)
    return t


def _intrinsics(K):
    return float(K[0][0]), float(K[1][1]), float(K[0][2]), float(K[1][2])


# ----------------------------------------------------------------------------- rays (utils.py:33-71)
def get_rays(H, W, K, c2w):
    """-> rays_o, rays_d, each [H, W, 3]; differentiable wrt c2w.  Bit-exact with the reference's
    CPU result (one sm_100a kernel instead of ~10 ATen launches).

    When c2w requires grad (pose estimation, demo_est_rel_pose.py:80-82: the demo builds the whole H x W grid from the
    pose it optimises and keeps 512 of its rays) the two results are LAZY grids: `rays[ys, xs]` with integer index
    arrays generates just those pixels' rays (forward and backward touch 512 rays, not 190,512); any other use
    materialises the full grid, with identical values."""
    fx, fy, cx, cy = _intrinsics(K)
    c2w = _on_gpu(c2w)
    if torch.is_grad_enabled() and c2w.requires_grad and os.environ.get("NB_LAZY_RAYS", "1") != "0":
        geom = (int(H), int(W), fx, fy, cx, cy)
        shared = {}
        return LazyRayGrid.make(geom, c2w, 0, shared), LazyRayGrid.make(geom, c2w, 1, shared)
    return _ops.GetRays.apply(c2w, int(H), int(W), fx, fy, cx, cy, None)


LazyRayGrid = _ops.LazyRayGrid



def get_rays_at(H, W, K, c2w, pix):
    """Rays of selected pixels only: pix = y*W + x (int64 tensor) -> rays_o, rays_d [n, 3]."""
    fx, fy, cx, cy = _intrinsics(K)
    c2w = _on_gpu(c2w)
    pix = torch.as_tensor(pix, dtype=torch.int64).to(c2w.device).contiguous()
    return _ops.GetRays.apply(c2w, int(H), int(W), fx, fy, cx, cy, pix)


def get_rays_np(H, W, K, c2w):
    """Host NumPy rays (utils.py:45-52); start-up only (batch_training_data)."""
    i, j = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32), indexing="xy")
    dirs = np.stack([(i - K[0][2]) / K[0][0], -(j - K[1][2]) / K[1][1], -np.ones_like(i)], -1)
    rays_d = np.sum(dirs[..., None, :] * c2w[:3, :3], -1)
    return np.broadcast_to(c2w[:3, -1], rays_d.shape), rays_d


def ndc_rays(H, W, focal, near, rays_o, rays_d):
    rays_d = _on_gpu(rays_d)
    return _ops.NdcRays.apply(_on_gpu(rays_o, rays_d), rays_d, int(H), int(W), float(focal), float(near))


def sample_pdf(bins, weights, N_samples, det=False, pytest=False):
    """Inverse-CDF samples (utils.py:74-117): bins [..., B], weights [..., B-1] -> [..., N_samples]."""
    bins = _on_gpu(bins)
    lead = bins.shape[:-1]
    b = bins.reshape(-1, bins.shape[-1]).float().contiguous()
    w = _on_gpu(weights, bins).reshape(b.shape[0], -1).float().contiguous()
    n, B = b.shape
    if det:
        u = None
    elif pytest:
        np.random.seed(0)
        u = torch.tensor(np.random.rand(n, N_samples), dtype=torch.float32, device=b.device)
    else:
        u = torch.rand((n, N_samples), device=b.device, dtype=torch.float32)
    lib = _ops._lib_for(b)
    out = torch.empty((n, N_samples), device=b.device, dtype=torch.float32)
    with torch.cuda.device(b.device):
        _ops.check(lib.nb_sample_pdf(_ops.ptr(b), _ops.ptr(w), _ops.ptr(u), _ops.ptr(_ops.linspace01(N_samples, b.device)),
                                     n, B, int(N_samples), _ops.ptr(out), _ops.stream_ptr()), "nb_sample_pdf")
    return out.reshape(list(lead) + [N_samples])


# ----------------------------------------------------------------------------- factories (utils.py:119-172)
def create_nerf_models(args):
    from . import nerf
    output_ch = 5 if args.N_importance > 0 else 4
    kw = dict(output_ch=output_ch, skips=[4], use_viewdirs=args.use_viewdirs, multires=args.multires,
              multires_views=args.multires_views, i_embed=args.i_embed)
    coarse = nerf.NeRF(D=args.netdepth, W=args.netwidth, **kw).to(device)
    fine = nerf.NeRF(D=args.netdepth_fine, W=args.netwidth_fine, **kw).to(device) if args.N_importance > 0 else None
    return coarse, fine


create_nerf = create_nerf_models   # README.md:104 / BASELINE.json call it create_nerf


def get_renderer(args, bds_dict):
    from . import render_utils
    use_ndc = args.dataset_type == "llff" and not args.no_ndc
    if not use_ndc:
        print("Not ndc!")
    kw = dict(perturb=args.perturb, N_importance=args.N_importance, N_samples=args.N_samples,
              use_viewdirs=args.use_viewdirs, white_bkgd=args.white_bkgd, raw_noise_std=args.raw_noise_std,
              ndc=use_ndc, lindisp=args.lindisp)
    kw.update(bds_dict)
    return render_utils.Renderer(**kw)


def get_optimizer(coarse_model, fine_model, args):
    """Adam(lr=args.lrate, betas=(0.9, 0.999)) over both models (utils.py:163-172) as one flat fused
    sm_100a kernel; falls back to nothing: parameters must live on the GPU."""
    from . import optim
    params = list(coarse_model.parameters()) + (list(fine_model.parameters()) if fine_model is not None else [])
    opt = optim.FlatAdam(params, lr=args.lrate, betas=(0.9, 0.999))
    for m in (coarse_model, fine_model):
        if m is not None and getattr(m, "use_viewdirs", False):
            offs = [opt.grad_offset(p) for p in _ops.mlp_params(m)]
            if all(o is not None for o in offs):
                m.__dict__["_nb_grad_target"] = (opt.flat_g, offs)
    return opt


# ----------------------------------------------------------------------------- checkpoints (utils.py:174-214,444-456)
def load_checkpoint(coarse_model, fine_model, optimizer, args, b_load_ckpnt_as_trainable=False, checkpoint_index=None):
    """Same file format as the reference: {'global_step', '{coarse,fine}_model_state_dict',
    'optimizer_state_dict'} in '{basedir}/{expname}/{step:06d}.tar'.  Returns the start step."""
    if args.ft_path is not None and args.ft_path != "None":
        ckpts = [args.ft_path]
    else:
        d = os.path.join(args.basedir, args.expname)
        ckpts = [os.path.join(d, f) for f in sorted(os.listdir(d)) if "tar" in f] if os.path.isdir(d) else []
    print("Found ckpts", ckpts)
    if not ckpts or args.no_reload:
        return 0
    path = ckpts[-1 if checkpoint_index is None else checkpoint_index]
    print("Reloading from", path)
    ckpt = torch.load(path, map_location=device)
    optimizer.load_state_dict(ckpt["optimizer_state_dict"])
    coarse_model.load_state_dict(ckpt["coarse_model_state_dict"], strict=False)
    models = [coarse_model]
    if fine_model is not None:
        fine_model.load_state_dict(ckpt["fine_model_state_dict"])
        models.append(fine_model)
    for m in models:
        for p in m.parameters():
            p.requires_grad = b_load_ckpnt_as_trainable
    return ckpt["global_step"]


def save_checkpoints(args, coarse_model, fine_model, optimizer, global_step, i):
    path = os.path.join(args.basedir, args.expname, "{:06d}.tar".format(i))
    os.makedirs(os.path.dirname(path), exist_ok=True)
    torch.save({"global_step": global_step,
                "coarse_model_state_dict": coarse_model.state_dict(),
                "fine_model_state_dict": fine_model.state_dict() if fine_model is not None else None,
                "optimizer_state_dict": optimizer.state_dict()}, path)
    print("Saved checkpoints at", path)


# ----------------------------------------------------------------------------- datasets (utils.py:216-313)
def load_datasets(args):
    from . import datasets
    return datasets.load_datasets(args)


def copy_log_dir(args):
    d = os.path.join(args.basedir, args.expname)
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "args.txt"), "w") as f:
        for k in sorted(vars(args)):
            f.write("{} = {}\n".format(k, getattr(args, k)))
    if getattr(args, "config", None) is not None:
        with open(os.path.join(d, "config.txt"), "w") as f, open(args.config, "r") as src:
            f.write(src.read())



_sel_rng = None
_stages = {}


def _choice_without_replacement(n, k):
    """k distinct indices in [0, n), uniform -- what np.random.choice(n, k, replace=False) draws
    (utils.py:435), without permuting all n indices every iteration: a Generator seeded from the global
    NumPy stream (so np.random.seed() still controls it) samples in O(k)."""
    global _sel_rng
    if _sel_rng is None:
        _sel_rng = np.random.default_rng(np.random.randint(0, 2 ** 31 - 1))
    return _sel_rng.choice(n, size=k, replace=False)


def _pinned_stage(n):
    """Rotating pinned staging buffers per batch size: pixel ids (int64) and target colours (fp32) of one batch in ONE
    buffer, so that a batch is one H2D copy.  A stage is refilled only after the H2D copy that read it last time has
    completed (`ev`, recorded by the caller after issuing it): the training loop never synchronises per step, so
    without the event the host could run more than a ring ahead of the GPU and tear a batch.  Allocated with
    device="cpu" explicitly: main.py:151 makes CUDA the default tensor type."""
    st = _stages.get(n)
    if st is None:
        bufs = []
        for _ in range(4):
            raw = torch.empty(20 * n, dtype=torch.uint8, device="cpu", pin_memory=True)
            pix, tgt = raw[:8 * n].view(torch.int64), raw[8 * n:].view(torch.float32).view(n, 3)
            bufs.append({"raw": raw, "pix": pix, "tgt": tgt, "pix_np": pix.numpy(), "tgt_np": tgt.numpy(), "ev": None})
        st = {"i": 0, "bufs": bufs}
        _stages[n] = st
    st["i"] = (st["i"] + 1) % len(st["bufs"])
    b = st["bufs"][st["i"]]
    if b["ev"] is not None:
        b["ev"].synchronize()
    return b

# ----------------------------------------------------------------------------- batching (utils.py:360-442)
def batch_training_data(args, poses, hwf, K, images, i_train):
    """-> images, poses, rays_rgb, use_batching, N_rand, i_batch (same tuple as the reference)."""
    H, W, _ = hwf
    use_batching = not args.no_batching
    rays_rgb, i_batch = torch.empty(0, device=device), None
    if use_batching:
        rays = np.stack([np.stack(get_rays_np(H, W, K, p), 0) for p in poses[:, :3, :4]], 0)   # [N, 2, H, W, 3]
        rr = np.concatenate([rays, images[:, None]], 1).transpose(0, 2, 3, 1, 4)                # [N, H, W, 3, 3]
        rr = np.stack([rr[i] for i in i_train], 0).reshape(-1, 3, 3).astype(np.float32)
        np.random.shuffle(rr)
        rays_rgb, i_batch = torch.tensor(rr, device=device), 0
        images = torch.tensor(images, dtype=torch.float32, device=device)
    poses = torch.tensor(poses, dtype=torch.float32, device=device)
    return images, poses, rays_rgb, use_batching, args.N_rand, i_batch


def sample_random_ray_batch(args, images, poses, rays_rgb, N_rand, use_batching, i_batch, i_train, hwf, K, start, i):
    """-> batch_rays [2, N_rand, 3], target_s [N_rand, 3], rays_rgb, i_batch."""
    H, W, _ = hwf
    if use_batching:
        batch = rays_rgb[i_batch:i_batch + N_rand].transpose(0, 1)
        batch_rays, target_s = batch[:2], batch[2]
        i_batch += N_rand
        if i_batch >= rays_rgb.shape[0]:
            print("Shuffle data after an epoch!")
            rays_rgb = rays_rgb[torch.randperm(rays_rgb.shape[0], device=rays_rgb.device)]
            i_batch = 0
        return batch_rays, target_s, rays_rgb, i_batch
    # one random image; only the selected pixels' rays are generated (the reference builds the
    # whole H x W grid every iteration, utils.py:419, and keeps N_rand of it)
    img_i = np.random.choice(i_train)
    pose = poses[img_i, :3, :4]
    if i < args.precrop_iters:
        dH, dW = int(H // 2 * args.precrop_frac), int(W // 2 * args.precrop_frac)
        ys, xs = np.arange(H // 2 - dH, H // 2 + dH), np.arange(W // 2 - dW, W // 2 + dW)
        if i == start:
            print(f"[Config] Center cropping of size {2*dH} x {2*dW} is enabled until iter {args.precrop_iters}")
    else:
        ys, xs = np.arange(H), np.arange(W)
    sel = _choice_without_replacement(len(ys) * len(xs), N_rand)
    py, px = ys[sel // len(xs)], xs[sel % len(xs)]
    pix_h = torch.from_numpy(py.astype(np.int64) * W + px.astype(np.int64))
    tgt = images[img_i]
    if torch.is_tensor(tgt) and tgt.device.type == "cuda":
        pix = pix_h.to(device, non_blocking=True)
        target_s = tgt.reshape(-1, tgt.shape[-1])[pix]
    else:
        # host-resident images (no_batching keeps them in NumPy / pinned memory, utils.py:383-390):
        # gather the N_rand target pixels on the host and ship 12 B per ray, not the whole image
        tnp = tgt.numpy() if torch.is_tensor(tgt) else tgt
        stage = _pinned_stage(N_rand)
        np.add(py * W, px, out=stage["pix_np"], casting="unsafe")            # written straight into pinned memory
        if tnp.dtype == np.float32 and tnp.shape[-1] == 3 and tnp.flags.c_contiguous:
            np.take(tnp.reshape(-1, 3), stage["pix_np"], axis=0, out=stage["tgt_np"])
        else:
            stage["tgt_np"][...] = tnp[py, px]
        both = stage["raw"].to(device, non_blocking=True)                     # one H2D copy per batch
        if stage["ev"] is None:
            stage["ev"] = torch.cuda.Event()
        stage["ev"].record(torch.cuda.current_stream(both.device))
        pix = both[:8 * N_rand].view(torch.int64)
        target_s = both[8 * N_rand:].view(torch.float32).view(N_rand, 3)
        return get_ray_batch_at(H, W, K, pose, pix), target_s, rays_rgb, i_batch
    rays_o, rays_d = get_rays_at(H, W, K, pose, pix)
    return torch.stack([rays_o, rays_d], 0), target_s, rays_rgb, i_batch


def get_ray_batch_at(H, W, K, c2w, pix):
    """[2, n, 3] = stack(get_rays_at(...)) for a pose without gradient: the kernel writes both halves of the batch."""
    c2w = _on_gpu(c2w)
    if c2w.requires_grad:
        return torch.stack(get_rays_at(H, W, K, c2w, pix), 0)
    fx, fy, cx, cy = _intrinsics(K)
    n = pix.numel()
    c = _ops.f32c(c2w)
    batch = torch.empty((2, n, 3), device=c.device, dtype=torch.float32)
    with torch.cuda.device(c.device):
        _ops.check(_ops._lib_for(c).nb_get_rays_sel_fwd(int(H), int(W), fx, fy, cx, cy, _ops.ptr(c), c.stride(0), _ops.ptr(pix), n,
                                                        _ops.ptr(batch[0]), _ops.ptr(batch[1]), None, _ops.stream_ptr()), "nb_get_rays")
    return batch


# ----------------------------------------------------------------------------- logging / video (utils.py:330-358,458-494)
def print_statistics(args, loss, psnr, i, tb_writer=None):
    import tqdm
    tqdm.tqdm.write(f"[TRAIN] Iter: {i} Loss: {loss.item()}  PSNR: {psnr.item()}")
    if tb_writer is not None:
        tb_writer.add_scalar("Test/Loss", loss, i)
        tb_writer.add_scalar("Test/PSNR", psnr, i)


def render_path(render_poses, hwf, K, chunk, renderer_and_models, gt_imgs=None, savedir=None, render_factor=0):
    """Render a camera path -> (rgbs [N,H,W,3], disps [N,H,W]) as NumPy.  The reference's
    render_path / render_training_video / render_test_poses call each other with mismatched
    signatures (utils.py:354,463,485 vs 330) and cannot run; this is the working equivalent.
    renderer_and_models = (renderer, coarse_model, fine_model)."""
    renderer, coarse, fine = renderer_and_models
    H, W, focal = hwf
    if render_factor != 0:
        H, W = H // render_factor, W // render_factor
        K = np.array(K, dtype=np.float64) / render_factor
        K[2][2] = 1.0
    from .render_utils import FrameDownloader, _write_image
    rgbs, disps = [], []

    def finish(n, arrs):                       # frame n's host work runs while the GPU renders frame n + 1
        rgbs.append(arrs[0])
        disps.append(arrs[1])
        if savedir is not None:
            _write_image(os.path.join(savedir, "{:03d}.png".format(n)), arrs[2])

    down = FrameDownloader(finish)
    with torch.no_grad():
        for n, c2w in enumerate(render_poses):
            c2w = c2w if torch.is_tensor(c2w) else torch.as_tensor(np.asarray(c2w), dtype=torch.float32)
            rgb, disp, _, _ = renderer.render_from_pose(H, W, K, chunk, c2w[:3, :4], coarse, fine, retraw=False)
            down.submit(n, (rgb, disp) + ((to8b_device(rgb),) if savedir is not None else ()))   # the PNG's bytes are made on the device
        down.drain()
    return np.stack(rgbs, 0), np.stack(disps, 0)


def render_training_video(args, render_poses, hwf, K, render_kwargs_test, i):
    rgbs, disps = render_path(render_poses, hwf, K, args.chunk, render_kwargs_test)
    base = os.path.join(args.basedir, args.expname, "{}_spiral_{:06d}_".format(args.expname, i))
    import imageio
    imageio.mimwrite(base + "rgb.mp4", to8b(rgbs), fps=30, quality=8)
    imageio.mimwrite(base + "disp.mp4", to8b(disps / np.max(disps)), fps=30, quality=8)


def render_test_poses(args, images, poses, hwf, K, render_kwargs_test, i_split, i):
    i_test = i_split[2]
    d = os.path.join(args.basedir, args.expname, "testset_{:06d}".format(i))
    os.makedirs(d, exist_ok=True)
    render_path(poses[i_test], hwf, K, args.chunk, render_kwargs_test, gt_imgs=images[i_test], savedir=d)
    print("Saved test set")
