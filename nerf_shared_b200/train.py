"""The training step of the reference's main loop as ONE CUDA graph.

`GraphedTrainStep` captures main.py:77-112 -- Renderer.render_from_rays (coarse + fine), the two
MSE terms, loss.backward(), the optional gradient all-reduce and Adam -- through the same public
objects main.py uses (Renderer, NeRF, the optimizer from utils.get_optimizer).  A replay launches
every kernel of the step with no Python between them; the per-step host work is copying the ray
batch into the static input buffers and one 1-thread kernel that writes Adam's scalars.

    step = train.GraphedTrainStep(renderer, coarse, fine, optimizer, H, W, K, N_rand)
    for i in range(N_iters):
        batch_rays, target_s, *_ = utils.sample_random_ray_batch(...)
        loss, psnr = step(batch_rays, target_s)          # device scalars, no sync
        for g in optimizer.param_groups: g['lr'] = new_lrate   # main.py:108-112, read at the next call
"""
from __future__ import annotations

import os

import torch

from . import _ops, render_utils, utils


class GraphedTrainStep:
    def __init__(self, renderer, coarse_model, fine_model, optimizer, H, W, K, n_rays, chunk=1024 * 32, group=None,
                 warmup=2, overlap=None, mode=None):
        """mode: "autograd" (default: the step driven through the autograd wrappers, exactly the calls main.py makes),
        "static" (explicit C-ABI sequence on static buffers with the compositing + loss + compositing-backward fusion,
        _static.StaticStep), "pipeline" (the static sequence software-pipelined across steps: a call finishes the
        PREVIOUS batch and starts the coarse pass of the one it is given; flush() finishes the last one).  Measured on
        B200 (profiles/r2_step_modes.txt): the three are within 4 % of each other -- the step is bound by the SM time of
        its six tensor-core launches, not by their order -- so the default stays the autograd-driven one.
        NB_STEP overrides the default."""
        self.renderer, self.coarse, self.fine, self.opt = renderer, coarse_model, fine_model, optimizer
        self.H, self.W, self.K, self.chunk = H, W, K, chunk
        self.group = group
        self.world = torch.distributed.get_world_size(group) if (
            torch.distributed.is_available() and torch.distributed.is_initialized()) else 1
        dev = optimizer.flat_p.device
        if self.world > 1:
            # SUM all-reduce of equal ray shards, averaged inside the Adam kernel; every rank starts from rank 0's weights
            optimizer.grad_scale = 1.0 / self.world
            torch.distributed.broadcast(optimizer.flat_p, src=torch.distributed.get_global_rank(group, 0) if group is not None else 0,
                                        group=group)
            optimizer.bump_versions()
        self.rays = torch.zeros((2, n_rays, 3), device=dev, dtype=torch.float32)
        self.target = torch.zeros((n_rays, 3), device=dev, dtype=torch.float32)
        self.graph = None
        self._warmup = warmup
        # overlap: the coarse backward depends only on the coarse map (main.py:99 adds the two MSE terms), so it is
        # captured as a second branch of the graph beside the fine forward / backward.  The MLP kernels are
        # persistent (one CTA per SM, round-robin tiles): a branch's CTAs start on the SMs the other branch's last
        # wave leaves idle (1,024 rays: 1.73 and 5.19 waves instead of 2 and 6).
        if overlap is None:
            overlap = os.environ.get("NB_OVERLAP", "1") != "0"
        self.overlap = bool(overlap) and fine_model is not None and renderer.N_importance > 0 \
            and n_rays <= render_utils.MIN_CHUNK
        self._coarse_stream = torch.cuda.Stream(device=dev) if self.overlap else None
        # multi-GPU: the coarse model's gradients are complete when the side branch ends, long before the fine wgrad;
        # their half of the flat buffer is all-reduced on the side stream under the fine backward (NB_SPLIT_AR=0: one
        # all-reduce after the join).  Needs the [coarse | fine] layout utils.get_optimizer builds.
        self._n_coarse = 0
        if self.overlap and self.world > 1 and os.environ.get("NB_SPLIT_AR", "1") != "0":
            n_c = sum(p.numel() for p in coarse_model.parameters())
            first_fine = next(iter(fine_model.parameters()))
            if optimizer.grad_offset(first_fine) == n_c and 0 < n_c < optimizer.flat_g.numel() and n_c % 4 == 0:
                self._n_coarse = n_c
        self.loss = self.psnr = None
        self.launches_per_replay = 0
        # all-reduce fused with the optimizer kernel over peer memory (optim.FlatAdam._setup_p2p), default group only
        self._fused_ar = self.world > 1 and group is None and getattr(optimizer, "p2p", None) is not None \
            and os.environ.get("NB_FUSED_AR", "1") != "0"
        if mode is None and overlap is not None and not overlap:
            mode = "autograd"                 # the single-stream form exists for the autograd-driven step only
        mode = mode or os.environ.get("NB_STEP", "autograd")
        if mode not in ("static", "pipeline", "autograd"):
            raise ValueError("GraphedTrainStep mode must be static, pipeline or autograd (got %r)" % (mode,))
        self.static = None
        if mode != "autograd" and n_rays <= render_utils.MIN_CHUNK:
            from . import _lib, _static
            try:
                self.static = _static.StaticStep(renderer, coarse_model, fine_model, optimizer, H, W, K, n_rays, group=group,
                                                 pipeline=(mode == "pipeline"))
            except _lib.NerfB200Error:
                if mode == "pipeline":
                    raise
                self.static = None            # configurations outside the explicit sequence run through autograd
        self.mode = "autograd" if self.static is None else mode
        if self.static is not None:
            self.rays, self.target = self.static.inbox.rays, self.static.inbox.target
            self.overlap = True               # the explicit sequence always runs the coarse backward as a second branch

    def _reduce_and_step(self, already_reduced=0):
        """Gradient exchange + Adam at the end of a step.  With the optimizer's symmetric gradient buffer (NCCL, world > 1)
        both are ONE kernel over NVLink peer memory (FlatAdam.step_allreduce); otherwise an NCCL all-reduce of the part
        of the flat buffer that is not reduced yet, then the Adam launch."""
        opt = self.opt
        if self._fused_ar:
            opt.step_allreduce(already_reduced, opt.flat_g.numel() - already_reduced, instance=0)
            return
        if self.world > 1:
            torch.distributed.all_reduce(opt.flat_g[already_reduced:], group=self.group)
        opt.step()

    # main.py:77-104 with the optimizer in graph mode
    def _step(self):
        if self.static is not None:
            return self.static.body_pipelined() if self.mode == "pipeline" else self.static.body_sequential()
        if self.overlap:
            return self._step_overlap()
        opt = self.opt
        rgb, _, _, extras = self.renderer.render_from_rays(self.H, self.W, self.K, chunk=self.chunk, rays=self.rays,
                                                           coarse_model=self.coarse, fine_model=self.fine, retraw=True)
        opt.zero_grad()
        # img2mse(rgb) + img2mse(rgb0) and mse2psnr (main.py:87-99) as one launch, one launch backward
        loss, _, _, psnr = _ops.MsePair.apply(rgb, extras.get("rgb0"), self.target)
        loss.backward()
        self._reduce_and_step()
        return loss.detach(), psnr

    def _step_overlap(self):
        """Same arithmetic as _step, two streams: `side` carries the coarse pass and (later) its backward, the
        current stream the fine pass.  Tensors that cross streams stay referenced until the join, so the caching
        allocator cannot hand their memory to the other branch."""
        opt, r = self.opt, self.renderer
        dev = opt.flat_p.device
        main, side = torch.cuda.current_stream(dev), self._coarse_stream
        alive, box = [], {}
        opt.zero_grad()                       # before the fork: both wgrad launches accumulate into flat_g
        side.wait_stream(main)

        def after_coarse(rgb0, tensors):      # called by render_rays once the coarse map exists (still on `side`)
            alive.append(tensors)
            box["m0"] = _ops.MsePair.apply(rgb0, None, self.target)[1]
            main.wait_stream(side)
            torch.cuda.set_stream(main)       # the fine pass continues on the main stream

        r._after_coarse = after_coarse
        torch.cuda.set_stream(side)
        try:
            rgb, _, _, extras = r.render_from_rays(self.H, self.W, self.K, chunk=self.chunk, rays=self.rays,
                                                   coarse_model=self.coarse, fine_model=self.fine, retraw=True)
        finally:
            r._after_coarse = None
            torch.cuda.set_stream(main)
        _, m, _, psnr = _ops.MsePair.apply(rgb, None, self.target)
        # autograd runs a node on the stream its forward ran on; the caller's stream is switched as well so that
        # the engine has nothing to synchronise between the two branches
        torch.cuda.set_stream(side)
        try:
            box["m0"].backward()
            if self._n_coarse:                # coarse gradients are final here: reduce them under the fine backward
                if self._fused_ar:            # (fused with the coarse network's Adam update: nothing later reads its weights)
                    opt.step_allreduce(0, self._n_coarse, instance=1)
                else:
                    torch.distributed.all_reduce(opt.flat_g[:self._n_coarse], group=self.group)
        finally:
            torch.cuda.set_stream(main)
        m.backward()
        main.wait_stream(side)                # join
        alive.clear()                         # only now may the coarse pass's tensors go back to the allocator
        self._reduce_and_step(self._n_coarse)
        loss = m.detach() + box["m0"].detach()     # main.py:99 (after the update: nothing waits for the sum)
        return loss, psnr

    def _capture(self):
        opt = self.opt
        dev = opt.flat_p.device
        keep = [t.clone() for t in (opt.flat_p, opt.exp_avg, opt.exp_avg_sq)]
        step0 = opt._step
        opt.in_graph = True
        from . import _fused
        lazy, _fused.GRAPHS_ENABLED = _fused.GRAPHS_ENABLED, False      # no nested captures inside this one's warm-up
        try:
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                if self.mode == "pipeline":
                    self.static.prime()
                for _ in range(self._warmup):       # allocator / packing caches / lazy init, then undone
                    opt.push_hyper()
                    self._step()
                    opt.bump_versions()
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            with torch.no_grad():
                for t, k in zip((opt.flat_p, opt.exp_avg, opt.exp_avg_sq), keep):
                    t.copy_(k)
            opt._step = step0
            opt.bump_versions()                     # the capture must contain the weight re-packing kernels
            from . import _lib
            g = torch.cuda.CUDAGraph()
            l0 = _lib.launch_count()
            with torch.cuda.graph(g):
                self.loss, self.psnr = self._step()
            self.launches_per_replay = _lib.launch_count() - l0     # libnerf_b200 kernels in the graph
            self.graph = g
            if self.static is not None:
                self.static.primed = False          # the warm-up's pending coarse pass used weights that were rolled back
        finally:
            opt.in_graph = False
            _fused.GRAPHS_ENABLED = lazy

    def __call__(self, batch_rays, target):
        shp = tuple(torch.as_tensor(batch_rays).shape) if not torch.is_tensor(batch_rays) else tuple(batch_rays.shape)
        if shp != tuple(self.rays.shape) or tuple(target.shape) != tuple(self.target.shape):
            raise ValueError("GraphedTrainStep was built for rays %s / targets %s (n_rays is fixed by the captured graph); "
                             "got %s / %s" % (tuple(self.rays.shape), tuple(self.target.shape), shp, tuple(target.shape)))
        if self.graph is None:
            self.rays.copy_(torch.as_tensor(batch_rays), non_blocking=True)
            self.target.copy_(target, non_blocking=True)
            self._capture()
        dev = self.rays.device
        if self.mode == "pipeline" and not self.static.primed:
            # first batch (or first after flush()): only its coarse pass runs; its loss comes out of the next call
            self.rays.copy_(torch.as_tensor(batch_rays) if not torch.is_tensor(batch_rays) else batch_rays, non_blocking=True)
            self.target.copy_(target, non_blocking=True)
            self.static.prime()
            self.loss.fill_(float("nan"))
            self.psnr.fill_(float("nan"))
            return self.loss, self.psnr
        if (torch.is_tensor(batch_rays) and batch_rays.device == dev and batch_rays.dtype == torch.float32
                and batch_rays.is_contiguous() and target.device == dev and target.dtype == torch.float32
                and target.is_contiguous()):
            self.opt.push_hyper_and_inputs(batch_rays, self.rays, target, self.target)   # one launch
        else:
            self.rays.copy_(torch.as_tensor(batch_rays) if not torch.is_tensor(batch_rays) else batch_rays, non_blocking=True)
            self.target.copy_(target, non_blocking=True)
            self.opt.push_hyper()
        self.graph.replay()
        self.opt.bump_versions()
        return self.loss, self.psnr

    def flush(self):
        """pipeline mode: finish the step whose coarse pass is pending (the replay's coarse pass for a "next" batch runs
        on the previous inbox and is discarded).  -> (loss, psnr) of the last batch, or None if nothing is pending."""
        if self.mode != "pipeline" or self.graph is None or not self.static.primed:
            return None
        self.opt.push_hyper()
        self.graph.replay()
        self.opt.bump_versions()
        self.static.primed = False
        return self.loss, self.psnr


class ChunkedBatchStep:
    """One optimizer step on a ray batch too large to keep its saved activations resident (BASELINE.json configs[2]:
    65,536 rays per step are 16.8 M samples x 10 KB of bf16 activations + activation gradients = 168 GB): the batch is
    cut into chunks of `chunk_rays`, every chunk runs main.py:77-103 (render_from_rays -> the two MSE terms -> backward)
    with its loss weighted by its share of the batch, the weight gradients accumulate in the flat buffer, and ONE
    all-reduce + Adam step follows -- the gradient of the mean loss over the whole (global) batch.  The chunks replay
    the lazily captured graphs of _fused.py, so a chunk costs two graph launches.

        step = ChunkedBatchStep(renderer, coarse, fine, optimizer, H, W, K, chunk_rays=8192)
        loss = step(batch_rays, target)           # batch_rays [2, n, 3], target [n, 3]: this rank's shard
    """

    def __init__(self, renderer, coarse_model, fine_model, optimizer, H, W, K, chunk_rays=8192, group=None):
        self.renderer, self.coarse, self.fine, self.opt = renderer, coarse_model, fine_model, optimizer
        self.H, self.W, self.K, self.chunk_rays, self.group = H, W, K, int(chunk_rays), group
        self.world = torch.distributed.get_world_size(group) if (
            torch.distributed.is_available() and torch.distributed.is_initialized()) else 1
        if self.world > 1:
            optimizer.grad_scale = 1.0 / self.world

    def __call__(self, batch_rays, target):
        from . import utils
        n = batch_rays.shape[1]
        opt = self.opt
        opt.zero_grad()
        total = None
        for b in range(0, n, self.chunk_rays):
            e = min(b + self.chunk_rays, n)
            rays = batch_rays[:, b:e].contiguous()
            tgt = target[b:e].contiguous()
            rgb, _, _, extras = self.renderer.render_from_rays(self.H, self.W, self.K, chunk=self.chunk_rays, rays=rays,
                                                               coarse_model=self.coarse, fine_model=self.fine, retraw=False)
            loss = (utils.img2mse(rgb, tgt) + utils.img2mse(extras["rgb0"], tgt)) * ((e - b) / n)
            loss.backward()
            total = loss.detach() if total is None else total + loss.detach()
        if self.world > 1 and self.group is None and getattr(opt, "p2p", None) is not None:
            opt.push_hyper()                  # (advances the step counter: the fused kernel reads its scalars from the device)
            opt.step_allreduce()
            opt.bump_versions()
        else:
            if self.world > 1:
                torch.distributed.all_reduce(opt.flat_g, group=self.group)
            opt.step()
        return total


class GraphedPoseStep:
    """The pose-estimation step of examples/relative_pose_estimation_demo/demo_est_rel_pose.py:80-97 as ONE CUDA graph:
    rays of the selected pixels from a pose that requires grad (utils.get_rays_at: only those pixels, not the H x W grid
    the demo builds and differentiates) -> Renderer.render_from_rays with frozen networks -> img2mse -> backward to the
    pose.  Captured through the same public objects the demo uses; a call copies (pose, pixel ids, target colours) into
    the static inputs and replays.

        step = train.GraphedPoseStep(renderer, coarse, fine, H, W, K, n_rays=512)
        pose = cam_transf(start_pose)                       # any differentiable parametrisation of the 4x4 pose
        loss, grad_pose = step(pose.detach(), pix, target)  # device tensors, overwritten by the next call
        pose.backward(grad_pose)                            # chain rule into the parametrisation; optimizer.step()
    """

    def __init__(self, renderer, coarse_model, fine_model, H, W, K, n_rays, chunk=1024 * 32, warmup=2):
        self.renderer, self.coarse, self.fine = renderer, coarse_model, fine_model
        self.H, self.W, self.K, self.chunk = H, W, K, chunk
        dev = next(coarse_model.parameters()).device
        self.pose = torch.eye(4, device=dev, dtype=torch.float32).requires_grad_(True)
        self.pix = torch.zeros(n_rays, device=dev, dtype=torch.int64)
        self.target = torch.zeros((n_rays, 3), device=dev, dtype=torch.float32)
        self.graph, self.loss, self.grad = None, None, None
        self._warmup = warmup
        self.launches_per_replay = 0

    def _step(self):
        self.pose.grad = None
        rays_o, rays_d = utils.get_rays_at(self.H, self.W, self.K, self.pose, self.pix)
        rgb = self.renderer.render_from_rays(self.H, self.W, self.K, chunk=self.chunk, rays=torch.stack([rays_o, rays_d], 0),
                                             coarse_model=self.coarse, fine_model=self.fine, retraw=True)[0]
        loss = utils.img2mse(rgb, self.target)
        loss.backward()
        return loss.detach(), self.pose.grad

    def _capture(self):
        from . import _fused, _lib
        dev = self.pose.device
        lazy, _fused.GRAPHS_ENABLED = _fused.GRAPHS_ENABLED, False      # no nested captures inside this one's warm-up
        try:
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(self._warmup):
                    self._step()
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            g = torch.cuda.CUDAGraph()
            l0 = _lib.launch_count()
            with torch.cuda.graph(g):
                self.loss, self.grad = self._step()
            self.launches_per_replay = _lib.launch_count() - l0
            self.graph = g
        finally:
            _fused.GRAPHS_ENABLED = lazy

    def __call__(self, pose, pix, target):
        with torch.no_grad():
            p = torch.as_tensor(pose, dtype=torch.float32)
            if tuple(p.shape) == (3, 4):
                self.pose[:3].copy_(p, non_blocking=True)
            else:
                self.pose.copy_(p, non_blocking=True)
            self.pix.copy_(torch.as_tensor(pix), non_blocking=True)
            self.target.copy_(target, non_blocking=True)
        if self.graph is None:
            self._capture()
        self.graph.replay()
        return self.loss, self.grad
