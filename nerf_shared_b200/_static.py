"""The training step of main.py:77-112 as an explicit sequence of C-ABI launches on static buffers (no autograd).

`train.GraphedTrainStep` captures either this sequence or the autograd-driven one.  What the explicit form adds:

* `nb_composite_train`: compositing, the map's MSE term and the compositing backward of that term are one launch per
  pass (the loss of main.py:87-99 is a sum of per-map MSE terms, so each map's gradient is known inside the ray);
* the coarse network's whole backward depends only on the coarse map, so it runs on a side stream beside the fine
  pass (as in the autograd-driven two-branch graph);
* `pipeline=True`: SOFTWARE PIPELINING ACROSS STEPS.  The coarse network's update needs only the coarse gradients,
  and the next step's coarse pass needs only the updated coarse network and the next ray batch.  One replay therefore
  runs   main: fine pass of step k (resampling, forward, loss, backward, all-reduce, Adam)
         side: coarse backward of step k, coarse all-reduce + Adam, then the coarse pass of step k+1
  so that the head of a step (ray rows, stratified depths, coarse forward, coarse compositing: alone on the GPU in
  the sequential form) runs beside the bandwidth-bound fine weight-gradient kernel of the previous step, on disjoint
  SMs (`nb_set_launch_limits`).  The parameter trajectory is the sequential one: every kernel sees exactly the
  weights it would see in main.py's order.  The caller hands batch k+1 to the call that finishes step k; `flush()`
  finishes the last step.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _bwd, _lib, _ops
from ._lib import check, ptr, stream_ptr


def _env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class _Stage:
    """One ray batch and what the coarse pass leaves for the fine pass, in ONE flat fp32 allocation (a stage is
    handed from the side stream to the main stream with a single device copy):
    rays [2,N,3] | target [N,3] | o [N,3] | d [N,3] | viewdirs [N,3] | z_coarse [N,S] | weights_coarse [N,S]."""

    def __init__(self, n, S, dev, ndc):
        sizes = [6 * n, 3 * n, 3 * n, 3 * n, 3 * n, n * S, n * S]
        self.flat = torch.zeros(sum(sizes), device=dev, dtype=torch.float32)
        parts, o = [], 0
        for sz in sizes:
            parts.append(self.flat[o:o + sz])
            o += sz
        self.rays = parts[0].view(2, n, 3)
        self.target = parts[1].view(n, 3)
        self.vd = parts[4].view(n, 3)
        self.z_c, self.w_c = parts[5].view(n, S), parts[6].view(n, S)
        if ndc:
            self.o, self.d = parts[2].view(n, 3), parts[3].view(n, 3)
        else:                      # no NDC warp: the rays are used as given
            self.o, self.d = self.rays[0], self.rays[1]


class StaticStep:
    def __init__(self, renderer, coarse, fine, opt, H, W, K, n_rays, group=None, pipeline=False):
        r = renderer
        if not (r.use_viewdirs and r.N_importance > 0 and fine is not None):
            raise _lib.NerfB200Error("the static training step covers use_viewdirs=True, N_importance > 0, two models")
        for m in (coarse, fine):
            m._check_kernel_config()
            if m.precision != "bf16" or not m.use_viewdirs:
                raise _lib.NerfB200Error("the static training step runs the tensor-core (bf16 operand) kernels")
        if any(not p.requires_grad for p in opt._params):
            raise _lib.NerfB200Error("the static training step updates every parameter (frozen parameters: use the autograd path)")
        self.r, self.coarse, self.fine, self.opt, self.group = r, coarse, fine, opt, group
        self.H, self.W, self.K = int(H), int(W), K
        self.N, self.S, self.I = int(n_rays), int(r.N_samples), int(r.N_importance)
        self.F = self.S + self.I
        self.world = torch.distributed.get_world_size(group) if (
            torch.distributed.is_available() and torch.distributed.is_initialized()) else 1
        self.pipeline = bool(pipeline)
        dev = opt.flat_p.device
        self.dev = dev
        self.lib = _ops._lib_for(opt.flat_p)
        # gradient layout: utils.get_optimizer lays the flat buffer out as [coarse | fine]
        self.offs, self.n_model = {}, {}
        for tag, m in (("c", coarse), ("f", fine)):
            offs = [opt.grad_offset(p) for p in _ops.mlp_params(m)]
            if any(o is None for o in offs):
                raise _lib.NerfB200Error("the optimizer does not hold this model's parameters (use utils.get_optimizer)")
            self.offs[tag] = (C.c_int64 * 24)(*offs)
            self.n_model[tag] = sum(p.numel() for p in m.parameters())
        first_fine = min(opt.grad_offset(p) for p in fine.parameters())
        if first_fine != self.n_model["c"] or self.n_model["c"] + self.n_model["f"] != opt.flat_g.numel():
            raise _lib.NerfB200Error("expected the [coarse | fine] flat layout utils.get_optimizer builds")
        self.n_c = self.n_model["c"]
        self._alloc()
        self.side = torch.cuda.Stream(device=dev)
        self.primed = False
        # SM budgets while the next step's coarse pass runs beside the fine weight-gradient kernel (pipeline mode)
        self.fused_ar = self.world > 1 and group is None and getattr(opt, "p2p", None) is not None \
            and os.environ.get("NB_FUSED_AR", "1") != "0" and not self.pipeline
        self.lim_wgrad = _env_int("NB_PIPE_WGRAD_CTAS", 104)
        self.lim_fwd = _env_int("NB_PIPE_FWD_CLUSTERS", 22)
        self.wait_for_wgrad = os.environ.get("NB_PIPE_WAIT", "1") != "0"

    # ------------------------------------------------------------------------------------------ buffers
    def _alloc(self):
        N, S, F, dev, lib = self.N, self.S, self.F, self.dev, self.lib
        f = lambda *shape: torch.empty(shape, device=dev, dtype=torch.float32)
        b = lambda nbytes: torch.empty((int(nbytes),), device=dev, dtype=torch.uint8)
        ndc = bool(self.r.ndc)
        self.cur = _Stage(N, S, dev, ndc)
        self.nxt = _Stage(N, S, dev, ndc) if self.pipeline else self.cur
        # rays / targets of the call (pipeline mode: a separate inbox, only those two parts are used)
        self.inbox = _Stage(N, S, dev, False) if self.pipeline else self.cur
        self.rows = f(N, 11)
        self.t_rand = f(N, S) if self.r.perturb > 0.0 else None
        self.u = f(N, self.I) if self.r.perturb != 0.0 else None
        noisy = self.r.raw_noise_std > 0.0
        self.pts = {"c": f(N, S, 3), "f": f(N, F, 3)}
        self.raw = {"c": f(N, S, 4), "f": f(N, F, 4)}
        self.d_raw = {"c": f(N, S, 4), "f": f(N, F, 4)}
        self.noise = {"c": f(N, S) if noisy else None, "f": f(N, F) if noisy else None}
        self.acts = {"c": b(lib.nb_mlp_act_bytes(N * S)), "f": b(lib.nb_mlp_act_bytes(N * F))}
        self.dA = {"c": b(lib.nb_mlp_da_bytes(N * S)), "f": b(lib.nb_mlp_da_bytes(N * F))}
        self.maps = {t: (f(N, 3), f(N), f(N)) for t in ("c", "f")}          # rgb, disp, acc
        self.z_f, self.z_std = f(N, F), f(N)
        L, Lv = int(self.coarse.multires), int(self.coarse.multires_views)
        self.packed = {t: b(lib.nb_mlp_packed_bytes(L, Lv)) for t in ("c", "f")}
        self.packed_bwd = {t: b(lib.nb_mlp_bwd_packed_bytes(L, Lv)) for t in ("c", "f")}
        self.acc = torch.zeros(2, device=dev, dtype=torch.float32)          # loss accumulators: [fine, coarse]
        self.out_f, self.out_c = torch.zeros(4, device=dev), torch.zeros(4, device=dev)
        self.loss = torch.zeros((), device=dev)
        self.models = {"c": self.coarse, "f": self.fine}

    # ------------------------------------------------------------------------------------------ phases
    def _pack(self, tag, backward):
        w = _ops.mlp_weight_struct(self.models[tag])
        if backward:
            check(self.lib.nb_mlp_pack_bwd(C.byref(w), ptr(self.packed_bwd[tag]), stream_ptr()), "nb_mlp_pack_bwd")
        else:
            check(self.lib.nb_mlp_pack(C.byref(w), ptr(self.packed[tag]), stream_ptr()), "nb_mlp_pack")

    def _mlp_fwd(self, tag, vd, n_samples):
        m = self.models[tag]
        check(self.lib.nb_mlp_fwd_tc(ptr(self.packed[tag]), int(m.multires), int(m.multires_views), ptr(self.pts[tag]), ptr(vd),
                                     self.N * n_samples, n_samples, ptr(self.raw[tag]), ptr(self.acts[tag]), stream_ptr()),
              "nb_mlp_fwd_tc")

    def _mlp_bwd(self, tag, n_samples, parts=3):
        m = self.models[tag]
        check(self.lib.nb_mlp_bwd_tc_parts(ptr(self.packed_bwd[tag]), int(m.multires), int(m.multires_views), ptr(self.acts[tag]),
                                           ptr(self.d_raw[tag]), self.N * n_samples, ptr(self.dA[tag]), self.offs[tag],
                                           ptr(self.opt.flat_g), parts, stream_ptr()), "nb_mlp_bwd_tc")

    def _composite_train(self, tag, st, z, n_samples, weights):
        r = self.r
        noise = self.noise[tag]
        if noise is not None:                 # render_utils.py:263-264: randn * raw_noise_std
            torch.randn(noise.shape, device=self.dev, dtype=torch.float32, out=noise)
            noise.mul_(float(r.raw_noise_std))
        rgb, disp, acc = self.maps[tag]
        inv_n3 = 1.0 / (3.0 * self.N)
        check(self.lib.nb_composite_train(ptr(self.raw[tag]), ptr(z), ptr(st.d), ptr(noise), ptr(st.target), self.N, n_samples,
                                          int(bool(r.white_bkgd)), inv_n3, ptr(rgb), ptr(disp), ptr(acc), ptr(weights),
                                          ptr(self.d_raw[tag]), ptr(self.acc[0:1] if tag == "f" else self.acc[1:2]),
                                          stream_ptr()), "nb_composite_train")

    def _coarse_forward(self, st):
        """render_utils.py:204-131,241-290 for the coarse network + its MSE term and d(term)/d raw."""
        r, lib, N, S = self.r, self.lib, self.N, self.S
        if r.ndc:
            check(lib.nb_ndc_rays_fwd(self.H, self.W, float(self.K[0][0]), 1.0, ptr(st.rays[0]), ptr(st.rays[1]), N,
                                      ptr(st.o), ptr(st.d), stream_ptr()), "nb_ndc_rays_fwd")
        check(lib.nb_ray_rows(ptr(st.o), ptr(st.d), ptr(st.rays[1]), float(r.near), float(r.far), N, ptr(self.rows),
                              ptr(st.vd), stream_ptr()), "nb_ray_rows")
        if self.t_rand is not None:
            torch.rand(self.t_rand.shape, device=self.dev, dtype=torch.float32, out=self.t_rand)
        check(lib.nb_sample_coarse(ptr(self.rows), N, 11, ptr(_ops.linspace01(S, self.dev)), ptr(self.t_rand), S,
                                   int(bool(r.lindisp)), ptr(st.z_c), ptr(self.pts["c"]), stream_ptr()), "nb_sample_coarse")
        self._pack("c", False)
        self._mlp_fwd("c", st.vd, S)
        self._composite_train("c", st, st.z_c, S, st.w_c)

    def _fine_forward(self, st, hook=None):
        lib, N, S, I, F = self.lib, self.N, self.S, self.I, self.F
        if self.u is not None:
            torch.rand(self.u.shape, device=self.dev, dtype=torch.float32, out=self.u)
        check(lib.nb_sample_pdf_merge(ptr(st.z_c), ptr(st.w_c), ptr(self.u), ptr(_ops.linspace01(I, self.dev)), N, S, I,
                                      ptr(self.z_f), ptr(self.z_std), None, stream_ptr()), "nb_sample_pdf_merge")
        check(lib.nb_ray_points_fwd(ptr(st.o), ptr(st.d), ptr(self.z_f), N, F, ptr(self.pts["f"]), stream_ptr()), "nb_ray_points_fwd")
        self._pack("f", False)
        if hook is not None and hook[0] == "pack":
            hook[1]()
        self._mlp_fwd("f", st.vd, F)
        if hook is not None and hook[0] == "fwd":
            hook[1]()
        self._composite_train("f", st, self.z_f, F, None)

    def _adam(self, lo, n):
        o = self.opt
        check(self.lib.nb_adam_step_dev(ptr(o.flat_p[lo:]), ptr(o.flat_g[lo:]), ptr(o.exp_avg[lo:]), ptr(o.exp_avg_sq[lo:]), n,
                                        ptr(o._hyper), stream_ptr()), "nb_adam_step_dev")

    def _allreduce(self, t):
        if self.world > 1:
            torch.distributed.all_reduce(t, group=self.group)

    # ------------------------------------------------------------------------------------------ one step
    def body_sequential(self):
        """main.py:77-112 in its own order; the coarse backward is a second branch beside the fine pass."""
        opt, dev = self.opt, self.dev
        main, side = torch.cuda.current_stream(dev), self.side
        st = self.cur                             # (the call wrote rays / targets straight into it)
        opt.flat_g.zero_()
        self._coarse_forward(st)
        # The coarse backward has slack until the join, the fine pass is the critical path: it is queued behind the
        # fine forward's launch so that its persistent CTAs cannot occupy the SMs the resampling kernels and the fine
        # forward are about to need (NB_COARSE_BWD_AFTER = none | pack | fwd)
        when = os.environ.get("NB_COARSE_BWD_AFTER", "fwd")
        hook = None
        if when != "none":
            ev = torch.cuda.Event()
            hook = (when, lambda: ev.record(main))
        else:
            side.wait_stream(main)
        self._fine_forward(st, hook)
        if when != "none":
            side.wait_event(ev)
        with torch.cuda.stream(side):
            self._pack("c", True)
            self._mlp_bwd("c", self.S)
            if not self.fused_ar:
                self._allreduce(opt.flat_g[:self.n_c])
        self._pack("f", True)
        self._mlp_bwd("f", self.F)
        main.wait_stream(side)
        if self.fused_ar:
            opt.step_allreduce()              # all-reduce + Adam: one kernel over NVLink peer memory
        else:
            self._allreduce(opt.flat_g[self.n_c:])
            self._adam(0, opt.flat_g.numel())
        check(self.lib.nb_loss_finalize(ptr(self.acc[0:1]), ptr(self.acc[1:2]), ptr(self.out_f), stream_ptr()), "nb_loss_finalize")
        return self.out_f[2], self.out_f[3]

    def prime(self):
        """pipeline mode, first batch: its coarse pass (what a replay does for the NEXT batch)."""
        st = self.nxt
        st.rays.copy_(self.inbox.rays)
        st.target.copy_(self.inbox.target)
        self.acc[1:2].zero_()                 # (a rolled-back warm-up may have left its coarse term here)
        self._coarse_forward(st)
        self.primed = True

    def body_pipelined(self):
        """One replay: main = fine pass of the pending step; side = its coarse backward + coarse update, then the coarse
        pass of the batch in the inbox."""
        opt, dev, lib = self.opt, self.dev, self.lib
        main, side = torch.cuda.current_stream(dev), self.side
        n_c, n = self.n_c, opt.flat_g.numel()
        self.cur.flat.copy_(self.nxt.flat)           # rays, targets, view directions, coarse depths / weights of the pending step
        side.wait_stream(main)
        ev = torch.cuda.Event()
        with torch.cuda.stream(side):
            check(lib.nb_loss_finalize(ptr(self.acc[1:2]), None, ptr(self.out_c), stream_ptr()), "nb_loss_finalize")   # coarse MSE of the pending step
            opt.flat_g[:n_c].zero_()
            self._pack("c", True)
            self._mlp_bwd("c", self.S)
            self._allreduce(opt.flat_g[:n_c])
            ev_ar = torch.cuda.Event()
            ev_ar.record(side)                       # one communicator: the two all-reduces are ordered explicitly
            self._adam(0, n_c)
        opt.flat_g[n_c:].zero_()
        self._fine_forward(self.cur)
        self._pack("f", True)
        self._mlp_bwd("f", self.F, parts=1)
        ev.record(main)
        check(lib.nb_set_launch_limits(0, self.lim_wgrad), "nb_set_launch_limits")
        try:
            self._mlp_bwd("f", self.F, parts=2)
        finally:
            lib.nb_set_launch_limits(0, 0)
        with torch.cuda.stream(side):
            if self.wait_for_wgrad:
                side.wait_event(ev)                  # the next coarse pass runs beside the fine weight-gradient kernel
            check(lib.nb_set_launch_limits(self.lim_fwd, 0), "nb_set_launch_limits")
            try:
                st = self.nxt
                st.rays.copy_(self.inbox.rays)
                st.target.copy_(self.inbox.target)
                self._coarse_forward(st)
            finally:
                lib.nb_set_launch_limits(0, 0)
        if self.world > 1:
            main.wait_event(ev_ar)
        self._allreduce(opt.flat_g[n_c:])
        self._adam(n_c, n - n_c)
        check(lib.nb_loss_finalize(ptr(self.acc[0:1]), None, ptr(self.out_f), stream_ptr()), "nb_loss_finalize")
        main.wait_stream(side)
        torch.add(self.out_f[0], self.out_c[0], out=self.loss)       # main.py:99
        return self.loss, self.out_f[3]
