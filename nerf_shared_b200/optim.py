"""Flat fused Adam (torch.optim.Adam semantics, reference utils.py:163-172 / main.py:104-112).

All parameters, gradients and both moment buffers live in four flat fp32 tensors; one sm_100a
kernel (nb_adam_step) updates everything.  The models' weight-gradient kernels accumulate
straight into the flat gradient buffer (every `p.grad` is a view of it), which is also the single
buffer the multi-GPU path all-reduces.  state_dict()/load_state_dict() keep torch.optim.Adam's
format so reference checkpoints load both ways.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr


class FlatAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=5e-4, betas=(0.9, 0.999), eps=1e-8):
        params = list(params)
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False, maximize=False, foreach=None,
                        capturable=False, differentiable=False, fused=None)
        super().__init__(params, defaults)
        ps = [p for g in self.param_groups for p in g["params"]]
        if not ps:
            raise ValueError("FlatAdam got no parameters")
        dev = ps[0].device
        for p in ps:
            if p.dtype != torch.float32 or p.device != dev:
                raise ValueError("FlatAdam needs fp32 parameters on one device")
        self._params = ps
        self._offsets, total = [], 0
        for p in ps:
            self._offsets.append(total)
            total += p.numel()
        self._n = total
        n_pad = (total + 3) // 4 * 4              # the fused all-reduce + Adam kernel moves float4s
        self._n_pad = n_pad
        self._p_store = torch.zeros(n_pad, device=dev, dtype=torch.float32)
        self._m_store = torch.zeros(n_pad, device=dev, dtype=torch.float32)
        self._v_store = torch.zeros(n_pad, device=dev, dtype=torch.float32)
        self.p2p = None                           # multi-GPU: symmetric gradient buffer + peer pointers (_setup_p2p)
        self._g_store = self._setup_p2p(n_pad, dev)
        if self._g_store is None:
            self._g_store = torch.zeros(n_pad, device=dev, dtype=torch.float32)
        self.flat_p, self.flat_g = self._p_store[:total], self._g_store[:total]
        self.exp_avg, self.exp_avg_sq = self._m_store[:total], self._v_store[:total]
        self._step = 0
        self._epoch = 0            # flag epoch of the fused all-reduce kernel: one per pushed step, never rolled back
        self.grad_scale = 1.0      # e.g. 1/world_size after a SUM all-reduce of flat_g
        # CUDA-graph mode (train.GraphedTrainStep): step() launches the kernel that reads its scalars
        # from this device buffer and does no host bookkeeping; push_hyper() advances the step
        self._hyper = torch.zeros(8, device=dev, dtype=torch.float32)
        self.in_graph = False
        with torch.no_grad():
            for p, o in zip(ps, self._offsets):
                n = p.numel()
                self.flat_p[o:o + n].copy_(p.detach().reshape(-1))
                p.data = self.flat_p[o:o + n].view(p.shape)
                p.grad = self.flat_g[o:o + n].view(p.shape)
        self._views = [p.grad for p in ps]          # identity check in _grads_are_flat (48 `is` tests, no data_ptr calls)
        self._by_id = {id(p): o for p, o in zip(ps, self._offsets)}

    # ---- multi-GPU: gradient all-reduce fused with the update over NVLink peer memory (csrc/adam.cu) ------------------
    def _setup_p2p(self, n_pad, dev):
        """With torch.distributed initialised on NCCL and more than one rank, the gradient buffer is allocated as
        symmetric memory (every rank maps every other rank's buffer), so that nb_adam_allreduce_p2p can reduce the
        gradients with peer loads / stores inside the optimizer kernel.  Collective: every rank builds its optimizer at
        the same point of the program (SPMD).  Returns the [n_pad] gradient tensor, or None (single process, gloo,
        NB_P2P_ADAM=0, or no symmetric memory on this system: NCCL all-reduce + nb_adam_step_dev are used then)."""
        import os
        import torch.distributed as dist
        if os.environ.get("NB_P2P_ADAM", "1") == "0" or dev.type != "cuda":
            return None
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() < 2 or dist.get_backend() != "nccl":
            return None
        try:
            import ctypes as C
            import torch.distributed._symmetric_memory as symm
            world, rank = dist.get_world_size(), dist.get_rank()
            if world > 16:
                return None
            buf = symm.empty(n_pad + 64, dtype=torch.float32, device=dev)       # gradients | 32 flag ints (+ slack)
            buf.zero_()
            hdl = symm.rendezvous(buf, dist.group.WORLD.group_name)
            ptrs = [int(x) for x in hdl.buffer_ptrs]
            if len(ptrs) != world or any(x == 0 for x in ptrs):
                return None
            grads = (C.c_void_p * world)(*ptrs)
            flags = (C.c_void_p * world)(*[x + 4 * n_pad for x in ptrs])
            counter = torch.zeros(2, device=dev, dtype=torch.int32)
            self._p2p_ptrs = ptrs
            mc = 0
            if os.environ.get("NB_NVLS", "1") != "0":
                try:
                    mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
                except Exception:
                    mc = 0
            self._p2p_mc = mc
            torch.cuda.synchronize(dev)
            dist.barrier()                       # every rank's flags are zero before anybody can signal
            self.p2p = {"buf": buf, "hdl": hdl, "grads": grads, "flags": flags, "counter": counter, "rank": rank, "world": world,
                        "nvls": bool(self._p2p_mc)}
            return buf[:n_pad]
        except Exception as e:                   # pragma: no cover  (no NVLink peer mapping: NCCL path)
            if os.environ.get("NB_P2P_ADAM") == "1":
                raise
            print("nerf_shared_b200: symmetric-memory gradient buffer unavailable (%s: %s); using NCCL all-reduce" %
                  (type(e).__name__, e))
            self.p2p = None
            return None

    def step_allreduce(self, lo=0, n=None, instance=0):
        """world > 1, inside a step whose scalars were pushed (push_hyper*): ONE launch that all-reduces the flat
        gradient over NVLink peer memory and applies Adam with 1/world scaling (grad_scale).  Every rank must call it once
        per step, in the same order relative to push_hyper.
        (lo, n): a sub-range of the flat buffers (multiples of 4 elements) -- a step may update the coarse network's range
        as soon as its gradients are final and the fine network's at the end; two launches that can overlap in time
        must use different `instance`s (0 / 1: separate flags and counters)."""
        if self.p2p is None:
            raise _lib.NerfB200Error("step_allreduce() needs the symmetric gradient buffer (NCCL, world > 1)")
        import ctypes as C
        q = self.p2p
        if n is None:
            n = self._n_pad - lo
        elif lo + n == self._n:
            n = self._n_pad - lo                  # the tail range takes the padding along
        if lo % 4 or n % 4 or instance not in (0, 1):
            raise ValueError("step_allreduce: range must be a multiple of 4 elements, instance 0 or 1")
        key = (lo, instance)
        arrs = q.setdefault("ranges", {}).get(key)
        if arrs is None:
            w = q["world"]
            arrs = ((C.c_void_p * w)(*[x + 4 * lo for x in self._p2p_ptrs]),
                    (C.c_void_p * w)(*[x + 4 * (self._n_pad + 32 * instance) for x in self._p2p_ptrs]))
            q["ranges"][key] = arrs
        lib = _lib.load()
        with torch.cuda.device(self._p_store.device):
            check(lib.nb_adam_allreduce_p2p(ptr(self._p_store[lo:]), ptr(self._m_store[lo:]), ptr(self._v_store[lo:]), n,
                                            ptr(self._hyper), arrs[0], arrs[1],
                                            C.c_void_p(self._p2p_mc + 4 * lo if self._p2p_mc else 0), q["rank"], q["world"],
                                            ptr(q["counter"][instance:]), stream_ptr()), "nb_adam_allreduce_p2p")

    # the weight-gradient kernels look this up to accumulate in place (see _bwd.py)
    def grad_offset(self, p):
        return self._by_id.get(id(p))

    def _grads_are_flat(self):
        for p, v in zip(self._params, self._views):
            if p.grad is not v:
                g = p.grad                         # (a re-attached view of the same slice is as good)
                if g is None or g.data_ptr() != v.data_ptr() or g.shape != v.shape:
                    return False
        return True

    def zero_grad(self, set_to_none: bool = False):
        # gradients stay views of the flat buffer; one memset
        if not self._grads_are_flat():
            self._regather()
        self.flat_g.zero_()

    def _regather(self):
        """Someone replaced a p.grad (set_to_none, external assignment): bring the flat buffer and the views back in
        line.  A parameter whose .grad is None has had no backward since it was cleared -- the in-place
        weight-gradient path (_bwd.py) re-attaches the views of every parameter it writes -- so its slice is zeroed,
        which is what "no gradient" means to the flat kernel."""
        with torch.no_grad():
            for p, o in zip(self._params, self._offsets):
                view = self.flat_g[o:o + p.numel()].view(p.shape)
                if p.grad is None:
                    view.zero_()
                elif p.grad.data_ptr() != view.data_ptr():
                    view.copy_(p.grad)
                p.grad = view
            self._views = [p.grad for p in self._params]

    def grad_view(self, p):
        """The slice of the flat gradient buffer that is `p.grad` (None for foreign parameters)."""
        o = self._by_id.get(id(p))
        return None if o is None else self.flat_g[o:o + p.numel()].view(p.shape)

    def _segments(self):
        """[(offset, count)] of the trainable parameters, neighbours merged.  torch.optim.Adam skips parameters
        without a gradient (requires_grad=False); one launch per segment keeps their weights and moments untouched."""
        key = tuple(p.requires_grad for p in self._params)
        if getattr(self, "_seg_key", None) != key:
            segs = []
            for p, o in zip(self._params, self._offsets):
                if not p.requires_grad:
                    continue
                if segs and segs[-1][0] + segs[-1][1] == o:
                    segs[-1] = (segs[-1][0], segs[-1][1] + p.numel())
                else:
                    segs.append((o, p.numel()))
            self._seg_key, self._segs = key, segs
        return self._segs

    def _lr(self):
        g0 = self.param_groups[0]
        if any(g["lr"] != g0["lr"] for g in self.param_groups):
            raise NotImplementedError("FlatAdam uses one learning rate for all parameter groups")
        return g0

    def push_hyper(self):
        """Advance the step counter and write this step's scalars (lr from param_groups, bias corrections,
        grad_scale) to the device buffer the graph-captured update reads; one 1-thread kernel whose
        arguments are taken at launch, so there is no host buffer to race with."""
        g0 = self._lr()
        self._step += 1
        self._epoch += 1
        b1, b2 = g0["betas"]
        lib = _lib.load()
        with torch.cuda.device(self.flat_p.device):
            check(lib.nb_adam_hyper(ptr(self._hyper), float(g0["lr"]), float(b1), float(b2), float(g0["eps"]),
                                    int(self._step), float(self.grad_scale), int(self._epoch), stream_ptr()), "nb_adam_hyper")

    def push_hyper_and_inputs(self, src1, dst1, src2, dst2):
        """push_hyper() and the copies src1 -> dst1, src2 -> dst2 (contiguous fp32 CUDA tensors of equal sizes) as one
        launch: what train.GraphedTrainStep does in front of every replay."""
        g0 = self._lr()
        self._step += 1
        self._epoch += 1
        b1, b2 = g0["betas"]
        lib = _lib.load()
        with torch.cuda.device(self.flat_p.device):
            check(lib.nb_train_step_prologue(ptr(self._hyper), float(g0["lr"]), float(b1), float(b2), float(g0["eps"]),
                                             int(self._step), float(self.grad_scale), int(self._epoch), ptr(src1), ptr(dst1), src1.numel(),
                                             ptr(src2), ptr(dst2), src2.numel(), stream_ptr()), "nb_train_step_prologue")

    def bump_versions(self):
        """The kernels write parameters through raw pointers: bump the autograd version counters (no
        launch) so caches keyed on p._version (_ops.packed_weights) see the update."""
        try:
            torch.autograd.graph.increment_version(self._params)
        except TypeError:
            for p in self._params:
                torch.autograd.graph.increment_version(p)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        lib = _lib.load()
        if self.in_graph:
            with torch.cuda.device(self.flat_p.device):
                for o, n in self._segments():
                    check(lib.nb_adam_step_dev(ptr(self.flat_p[o:]), ptr(self.flat_g[o:]), ptr(self.exp_avg[o:]),
                                               ptr(self.exp_avg_sq[o:]), n, ptr(self._hyper), stream_ptr()), "nb_adam_step_dev")
            return loss
        if not self._grads_are_flat():
            self._regather()
        g0 = self._lr()
        self._step += 1
        b1, b2 = g0["betas"]
        _lib.ensure_device(self.flat_p.device)
        with torch.cuda.device(self.flat_p.device):
            for o, n in self._segments():
                check(lib.nb_adam_step(ptr(self.flat_p[o:]), ptr(self.flat_g[o:]), ptr(self.exp_avg[o:]),
                                       ptr(self.exp_avg_sq[o:]), n, float(g0["lr"]), float(b1), float(b2), float(g0["eps"]),
                                       int(self._step), float(self.grad_scale), stream_ptr()), "nb_adam_step")
        self.bump_versions()
        return loss

    # ---- torch.optim.Adam-compatible (de)serialisation -------------------------------------
    def state_dict(self):
        state = {}
        for i, (p, o) in enumerate(zip(self._params, self._offsets)):
            n = p.numel()
            state[i] = {"step": torch.tensor(float(self._step)),
                        "exp_avg": self.exp_avg[o:o + n].view(p.shape).clone(),
                        "exp_avg_sq": self.exp_avg_sq[o:o + n].view(p.shape).clone()}
        groups, k = [], 0
        for g in self.param_groups:
            d = {key: val for key, val in g.items() if key != "params"}
            d["params"] = list(range(k, k + len(g["params"])))
            k += len(g["params"])
            groups.append(d)
        return {"state": state if self._step > 0 else {}, "param_groups": groups}

    def load_state_dict(self, sd):
        for g, saved in zip(self.param_groups, sd["param_groups"]):
            for key, val in saved.items():
                if key != "params":
                    g[key] = val
        st = sd.get("state", {})
        steps = []
        with torch.no_grad():
            for i, (p, o) in enumerate(zip(self._params, self._offsets)):
                s = st.get(i, st.get(str(i)))
                if s is None:
                    continue
                n = p.numel()
                self.exp_avg[o:o + n].copy_(torch.as_tensor(s["exp_avg"]).reshape(-1).to(self.exp_avg.device))
                self.exp_avg_sq[o:o + n].copy_(torch.as_tensor(s["exp_avg_sq"]).reshape(-1).to(self.exp_avg.device))
                steps.append(int(float(s["step"])))
        self._step = max(steps) if steps else 0
