#!/usr/bin/env python
"""bench.py -- the reference's headline workload on B200 (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W            # this framework (sm_100a kernels)
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (oracle port)

A "step" is one training step of BASELINE.json configs[1]: lego-shaped synthetic Blender data,
N_rand = 1024 rays per GPU, 64 coarse + 128 fine samples, coarse + fine forward/backward + Adam
(reference main.py:72-112).  `value` is rays/s with the step's rays and targets already in HBM;
`e2e` is the same step through the public API (utils.sample_random_ray_batch -> Renderer ->
loss -> backward -> optimizer.step -> loss read-back), pinned-host -> device copies and the loss
read-back inside the timed region.  One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

# torchrun exports OMP_NUM_THREADS=1 to every rank; rank 0 times the CPU arm and needs the host cores
if os.environ.get("RANK", "0") == "0" and os.environ.get("OMP_NUM_THREADS") == "1":
    os.environ["OMP_NUM_THREADS"] = os.environ["MKL_NUM_THREADS"] = str(os.cpu_count() or 1)

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

_real_stdout = sys.stdout
FLOP_FWD = 1186816        # per sample, SURVEY.md section 8d
FLOP_TRAIN = 3489024      # fwd + wgrad + needed dgrad per sample
H = W = 400
N_IMAGES = 8
S_COARSE, S_FINE = 64, 192


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rays", type=int, default=1024, help="N_rand per GPU (weak scaling)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-render", action="store_true")
    ap.add_argument("--no-sustained", action="store_true")
    ap.add_argument("--large-rays", type=int, default=65536,
                    help="global batch of the configs[2] leg (65,536 rays per step, sharded over the GPUs); 0 = skip")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops", 1590.0), d.get("bf16_tflops_sustained", 1400.0), d.get("hbm_gbs", 6650.0), "measured"
    return 1590.0, 1400.0, 6650.0, "fallback"


def workload_config(rays, world):
    """`config` of the JSON line, identical for the B200 arm and the reference arm."""
    return {"workload": "lego-shaped synthetic Blender training step (BASELINE.json configs[1]): N_rand=%d rays/GPU, "
                        "64 coarse + 128 fine samples, D8 W256 coarse+fine fwd/bwd + Adam" % rays,
            "global_rays_per_step": world * rays, "parallelism": "dp%d (ray-sharded, flat-gradient all-reduce)" % world}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.lines, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: [self.lines.append(l) for l in self.proc.stdout], daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except Exception:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def synth_scene(seed=0):
    """Blender-shaped synthetic data: poses on the standard orbit, random targets."""
    from nerf_shared_b200 import datasets
    images, poses, _, K = datasets.synthetic_blender(H, W, N_IMAGES, seed)
    return K, poses, images


class Args:  # the subset of config_parser flags the API helpers read (configs/lego.txt)
    N_importance, N_samples, N_rand = 128, 64, 1024
    netdepth = netdepth_fine = 8
    netwidth = netwidth_fine = 256
    use_viewdirs, multires, multires_views, i_embed = True, 10, 4, 0
    perturb, raw_noise_std, white_bkgd, lindisp = 1.0, 0.0, True, False
    dataset_type, no_ndc, lrate, no_batching = "blender", False, 5e-4, True
    precrop_iters, precrop_frac, chunk = 0, 0.5, 32768


def large_batch_leg(a, args, renderer, coarse, fine, opt, K, poses_d, images, dev, world, flush, steps=3):
    """BASELINE.json configs[2] as written: ONE optimizer step on a fixed global batch of 65,536 rays, ray-sharded over the
    GPUs of the box (strong scaling: 65,536 / N rays per GPU).  The saved activations of a whole shard do not fit in HBM
    at N <= 2 (168 GB at 65,536 rays), so every rank accumulates the gradient over chunks of 8,192 rays
    (train.ChunkedBatchStep: main.py's calls per chunk, one all-reduce + Adam per step)."""
    import torch.distributed as dist
    from nerf_shared_b200 import train, utils
    total = int(a.large_rays)
    if total % world != 0:
        return None
    n = total // world
    step = train.ChunkedBatchStep(renderer, coarse, fine, opt, H, W, K, chunk_rays=min(8192, n))
    pix = torch.from_numpy(np.random.choice(H * W, n, replace=False)).to(dev)
    ro, rd = utils.get_rays_at(H, W, K, poses_d[0, :3, :4], pix)
    tgt = torch.from_numpy(images[0].reshape(-1, 3)[pix.cpu().numpy()]).to(dev)
    rays = torch.stack([ro, rd], 0)
    for _ in range(2):
        step(rays, tgt)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for i in range(steps):
        flush.zero_()
        ev[i][0].record()
        loss = step(rays, tgt)
        ev[i][1].record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    assert np.isfinite(float(loss))
    ms = sum(s_.elapsed_time(e) for s_, e in ev) / steps
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    samples = n * (args.N_samples + args.N_samples + args.N_importance)
    renderer.__dict__.pop("_nb_graphs", None)        # release the chunk-sized graphs
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    return {"workload": "BASELINE.json configs[2]: one training step on a fixed global batch of %d rays, %d rays per GPU on %d GPU(s), "
                        "gradient accumulated over %d-ray chunks, one all-reduce + Adam" % (total, n, world, min(8192, n)),
            "scaling": "strong", "rays_per_s": total / (ms * 1e-3), "ms_per_step": ms, "steps": steps,
            "mlp_train_tflops_per_gpu": samples * FLOP_TRAIN / (ms * 1e-3) / 1e12}


def run_b200(a):
    import torch.distributed as dist
    from nerf_shared_b200 import _lib, _ops, utils

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    utils.device = dev
    args = Args()
    args.N_rand = a.rays
    K, poses, images = synth_scene()
    torch.manual_seed(0)                       # identical random-init weights (nn.Linear default init) on every rank
    coarse, fine = utils.create_nerf_models(args)
    torch.manual_seed(1234 + rank)             # per-rank ray batches / jitter
    np.random.seed(1234 + rank)
    opt = utils.get_optimizer(coarse, fine, args)
    opt.grad_scale = 1.0 / world
    renderer = utils.get_renderer(args, {"near": 2.0, "far": 6.0})
    renderer.train()
    poses_d = torch.tensor(poses, device=dev)
    images_pinned = torch.from_numpy(images).pin_memory()
    hwf = [H, W, float(K[0][0])]
    i_train = np.arange(N_IMAGES)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    from nerf_shared_b200 import dist as nd_, train
    step = train.GraphedTrainStep(renderer, coarse, fine, opt, H, W, K, a.rays, chunk=args.chunk)

    def eager_step(batch_rays, target):
        """main.py:77-104 call by call (what the unmodified main.py executes)."""
        rgb, _, _, extras = renderer.render_from_rays(H, W, K, chunk=args.chunk, rays=batch_rays, coarse_model=coarse,
                                                      fine_model=fine, retraw=True)
        opt.zero_grad()
        loss = utils.img2mse(rgb, target) + utils.img2mse(extras["rgb0"], target)
        loss.backward()
        if world > 1:
            nd_.step_with_allreduce(opt)      # fused all-reduce + Adam over peer memory when available, else NCCL + step
        else:
            opt.step()
        return loss

    # ---- device-resident batches (value) ------------------------------------------------------
    n_pool = 8
    pool = []
    for i in range(n_pool):
        pix = torch.from_numpy(np.random.choice(H * W, a.rays, replace=False)).to(dev)
        ro, rd = utils.get_rays_at(H, W, K, poses_d[i % N_IMAGES, :3, :4], pix)
        tgt = torch.from_numpy(images[i % N_IMAGES].reshape(-1, 3)[pix.cpu().numpy()]).to(dev)
        pool.append((torch.stack([ro, rd], 0), tgt))

    # The clock sampler (rank 0 only) starts BEFORE the warm-up: nvidia-smi's start-up (NVML initialisation) stalls
    # kernel submission for tens of milliseconds -- inside the timed region it doubled the step time of an 8-GPU run --
    # so the warm-up keeps stepping (untimed) until the first sample has arrived; after that it is one light query
    # every 100 ms.
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for i in range(max(a.warmup, 3)):
        step(*pool[i % n_pool])
    t_wait = time.perf_counter()
    ready = torch.zeros(1, device=dev)
    while True:                               # all ranks leave together (the steps contain a collective)
        ready.fill_(1.0 if (rank != 0 or sampler.proc is None or sampler.lines or time.perf_counter() - t_wait > 3.0) else 0.0)
        if world > 1:
            dist.all_reduce(ready, op=dist.ReduceOp.MIN)
        if float(ready.item()) == 1.0:
            break
        for i in range(20):
            step(*pool[i % n_pool])
    sync_all()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    sync_all()
    for i in range(a.steps):
        flush.zero_()                        # L2 flush between timed iterations (outside the events)
        ev[i][0].record()
        step(*pool[i % n_pool])
        ev[i][1].record()
    sync_all()
    launches = step.launches_per_replay * a.steps
    clocks = sampler.stop()
    ms = sum(s.elapsed_time(e) for s, e in ev) / a.steps
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item())
    value = world * a.rays / (ms_step * 1e-3)

    # ---- sustained leg: the rate a 200k-iteration run lives at (power-capped clocks), not the 26 ms burst above -------
    # >= 2 s of back-to-back steps (no flush: a step streams ~5 GB, 40x the L2), per-step CUDA events, median
    sustained = None
    if not a.no_sustained:
        n_sus = max(200, int(2.5 / (ms_step * 1e-3)))
        sampler2 = ClockSampler(local)
        if rank == 0:
            sampler2.start()
        for i in range(50):
            step(*pool[i % n_pool])
        sync_all()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(n_sus + 1)]
        t0 = time.perf_counter()
        evs[0].record()
        for i in range(n_sus):
            step(*pool[i % n_pool])
            evs[i + 1].record()
        sync_all()
        wall = time.perf_counter() - t0
        per = np.array([evs[i].elapsed_time(evs[i + 1]) for i in range(n_sus)])
        t = torch.tensor([float(np.median(per)), float(per.mean())], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        med, mean = float(t[0].item()), float(t[1].item())
        sustained = {"steps": n_sus, "seconds": wall, "ms_per_step_median": med, "ms_per_step_mean": mean,
                     "rays_per_s": world * a.rays / (med * 1e-3), "clocks": sampler2.stop() if rank == 0 else None,
                     "mlp_tflops_per_gpu": a.rays * (S_COARSE + S_FINE) * FLOP_TRAIN / (med * 1e-3) / 1e12}

    # ---- end to end through the public API (e2e) -----------------------------------------------
    # host images (pinned) -> utils.sample_random_ray_batch (pixel ids + targets H2D) -> graphed step ->
    # loss.item() (D2H) every step
    def sample(i):
        return utils.sample_random_ray_batch(args, images_pinned, poses_d, None, a.rays, False, None, i_train, hwf, K, 0, i + 1)

    def time_e2e(fn):
        """Per step: host batch sampling (pixel ids + targets gathered on the host, H2D from pinned memory), the
        step, and a D2H read of that step's loss into pinned memory.  Software-pipelined by one step, as a training
        loop that logs its loss does: while the GPU runs step i the host samples batch i+1, queues step i+1 and only
        then waits for the loss of step i.  The loss of EVERY step is read back inside the timed region."""
        host_loss = [torch.zeros(1).pin_memory() for _ in range(2)]
        done = [torch.cuda.Event() for _ in range(2)]

        def loop(k):
            batch_rays, target, _, _ = sample(0)
            prev, losses = None, []
            for i in range(k):
                out = fn(batch_rays, target)
                loss = out[0] if isinstance(out, tuple) else out
                host_loss[i & 1].copy_(loss.detach().reshape(1), non_blocking=True)   # D2H of this step's loss
                done[i & 1].record()
                batch_rays, target, _, _ = sample(i + 1)          # overlaps the step on the GPU
                if prev is not None:
                    done[prev].synchronize()
                    losses.append(float(host_loss[prev][0]))
                prev = i & 1
            done[prev].synchronize()
            losses.append(float(host_loss[prev][0]))
            assert len(losses) == k and all(np.isfinite(losses)), losses
        loop(3)
        sync_all()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        loop(a.steps)
        e1.record()
        sync_all()
        ms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / a.steps
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    graph_ms = time_e2e(step)
    eager_ms = time_e2e(eager_step)
    h2d = a.rays * 8 + a.rays * 3 * 4      # int64 pixel ids + fp32 targets per step
    # The headline e2e is the loop the reference's main.py executes, call by call through the reference API
    # (utils.sample_random_ray_batch -> Renderer.render_from_rays -> utils.img2mse x2 -> backward -> optimizer.step);
    # train.GraphedTrainStep (not a reference API: the whole step as one CUDA graph) is reported beside it.
    e2e = {"value": world * a.rays / (eager_ms * 1e-3), "unit": "rays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
           "ms_per_step": eager_ms,
           "api": "main.py:72-104 unchanged: utils.sample_random_ray_batch (host pixel sampling, pinned H2D) -> render_from_rays -> "
                  "img2mse x2 -> loss.backward() -> optimizer.step() -> loss D2H (pinned, read one step later)",
           "graphed_step": {"value": world * a.rays / (graph_ms * 1e-3), "ms_per_step": graph_ms,
                            "api": "utils.sample_random_ray_batch -> train.GraphedTrainStep -> loss D2H"}}

    cfg = workload_config(a.rays, world)          # identical in the reference arm (the driver compares the two)
    how = {"l2": "256 MiB flush between timed iterations", "timing": "CUDA events per step, max over ranks",
           "launch": "one CUDA graph per step (train.GraphedTrainStep, mode %s%s)" % (
               step.mode, ", coarse backward captured as a parallel branch beside the fine pass" if step.overlap else "")}
    out = {"metric": "train_rays_per_s", "value": value, "unit": "rays/s", "n_gpus": world, "steps": a.steps,
           "warmup": max(a.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": cfg,
           "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "measurement": how}
    if sustained is not None:
        out["sustained"] = sustained

    large = None if (a.no_render or a.large_rays <= 0) else large_batch_leg(a, args, renderer, coarse, fine, opt, K, poses_d,
                                                                             images, dev, world, flush)
    if large is not None:
        out["large_batch"] = large
        out["e2e"]["configs2_strong_scaling"] = large         # BASELINE.json configs[2] as written, in a driver-parsed field
    render = None if a.no_render else render_leg(renderer, coarse, fine, K, poses_d, dev, world)   # all ranks (row-sharded)
    if rank == 0:
        out.update(kernel_breakdown(a, coarse, fine, pool[0], dev))
        burst, sus_peak, _, _ = peaks()
        step_flop = a.rays * (S_COARSE + S_FINE) * FLOP_TRAIN
        out["roofline"]["step"] = {"algorithmic_tflop": step_flop / 1e12, "tflops": step_flop / (ms_step * 1e-3) / 1e12,
                                   "frac_of_burst_peak": step_flop / (ms_step * 1e-3) / 1e12 / burst,
                                   "frac_of_sustained_peak": step_flop / (ms_step * 1e-3) / 1e12 / sus_peak,
                                   "sustained_leg_frac_of_sustained_peak": (None if sustained is None else
                                                                            sustained["mlp_tflops_per_gpu"] / sus_peak)}
        if render is not None:
            out["render"] = render
        if world == 1 and not a.no_render:
            out["llff_pose"] = pose_leg(dev)
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import bench_stream
            out["stream_kernels"] = bench_stream.run(quiet=True)
        if not a.no_cpu_baseline and world == 1:     # the CPU baseline is reported at N=1 only
            out["cpu_baseline"] = cpu_train_baseline()
        print(json.dumps(out), file=_real_stdout, flush=True)
    # teardown: a CUDA graph that captured NCCL kernels must be released before its communicator;
    # if anything in the teardown still blocks, dump the stacks and leave (the result is already printed)
    import faulthandler
    import gc
    faulthandler.dump_traceback_later(90, exit=True)
    step.graph = None
    gc.collect()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    faulthandler.cancel_dump_traceback_later()


def kernel_breakdown(a, coarse, fine, batch, dev):
    """Times the three tensor-core kernels of the fine pass alone (CUDA events on the launching
    stream) and reports the roofline of the dominant one."""
    import ctypes as C

    from nerf_shared_b200 import _bwd, _lib, _ops
    burst, sustained, hbm, src = peaks()
    M = a.rays * S_FINE
    pts = torch.rand(M, 3, device=dev) * 6 - 3
    vd = torch.nn.functional.normalize(torch.randn(a.rays, 3, device=dev), dim=-1)
    d_raw = torch.randn(M, 4, device=dev) * 1e-3
    lib = _lib.load()

    def timed(fn, reps=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    t_fwd = timed(lambda: _ops.mlp_forward_tc(fine, pts, vd, S_FINE, save_acts=True))
    raw, acts = _ops.mlp_forward_tc(fine, pts, vd, S_FINE, save_acts=True)
    offs, total = _bwd.grad_layout(fine)
    grads = torch.zeros(total, device=dev)
    dA = torch.empty(int(lib.nb_mlp_da_bytes(M)), dtype=torch.uint8, device=dev)
    pk = _ops.packed_weights(fine, backward=True)
    off_arr = (C.c_int64 * 24)(*offs)

    def bwd(with_w):
        _lib.check(lib.nb_mlp_bwd_tc(_lib.ptr(pk), 10, 4, _lib.ptr(acts), _lib.ptr(d_raw), M, _lib.ptr(dA), off_arr,
                                     _lib.ptr(grads) if with_w else None, _lib.stream_ptr()))
    t_dgrad = timed(lambda: bwd(False))
    t_bwd = timed(lambda: bwd(True))
    t_wgrad = max(t_bwd - t_dgrad, 1e-6)
    # inference-sized forward (no saved activations): the "MLP tensor-pipe utilisation" of BASELINE.json
    Mi = 32768 * S_FINE
    pts_i = torch.rand(Mi, 3, device=dev) * 6 - 3
    vd_i = torch.nn.functional.normalize(torch.randn(32768, 3, device=dev), dim=-1)
    t_inf = timed(lambda: _ops.mlp_forward_tc(fine, pts_i, vd_i, S_FINE, save_acts=False), reps=5)
    del pts_i, vd_i
    dgrad_flop = FLOP_FWD - 2 * (63 * 256 * 2 + 27 * 128 + 256 + 384)
    # design traffic per sample (NOT the algorithmic 20 B/sample of SURVEY 8d): saved activations 40 slabs,
    # dA 38 slabs of 128 B per row, relu' bit masks 288 B
    act_b, da_b, mask_b = 40 * 128 + 288, 38 * 128, 288
    kern = {
        "mlp_fwd_tc(fine, saves activations)": (t_fwd, FLOP_FWD * M, (12 + 16 + act_b) * M),
        "mlp_dgrad(fine)": (t_dgrad, dgrad_flop * M, (16 + mask_b + da_b) * M),
        "mlp_wgrad+heads(fine)": (t_wgrad, FLOP_FWD * M, (40 * 128 + da_b + 768 + 16) * M),
        "mlp_fwd_tc(inference 32768x192)": (t_inf, FLOP_FWD * Mi, 28 * Mi),
    }
    traffic = ncu_traffic()
    per = {}
    for k, (t, fl, by) in kern.items():
        tf, gb = fl / t / 1e9, by / t / 1e6
        per[k] = {"ms": round(t, 4), "algorithmic_flop": int(fl), "tflops": round(tf, 1), "frac_tensor_burst": round(tf / burst, 3),
                  "design_traffic_bytes": int(by), "design_traffic_GBps": round(gb, 1), "design_traffic_frac_hbm": round(gb / hbm, 3)}
    # The path is tensor-bound by its algorithm (the MLP is the only dense contraction, its algorithmic
    # HBM I/O is 20 B/sample): the roofline of the dominant kernel is ALGORITHMIC FLOP / launch time / measured bf16 peak.
    train_k = [k for k in kern if "inference" not in k]
    top = max(train_k, key=lambda k: kern[k][0])
    t, fl, by = kern[top]
    roof = {"bound": "tensor", "kernel": top, "achieved": fl / t / 1e9, "peak": burst, "unit": "TFLOP/s", "frac": fl / t / 1e9 / burst,
            "peak_source": src + " bf16 burst (kernel timed alone, CUDA events on the launching stream)",
            "algorithmic_flop_per_launch": int(fl), "launch_ms": t}
    roof["traffic"] = traffic.get(top.split("(")[0].split("+")[0] + "_kernel")
    roof["traffic_source"] = "profiles/ (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch)" if roof["traffic"] else None
    roof["algorithmic_bytes_per_launch"] = int(20 * M)
    # secondary view: why the kernel sits where it does -- the DESIGN moves saved activations and activation gradients
    # through HBM, and that traffic (not the algorithm's) is what the weight-gradient kernel is limited by
    roof["hbm_design_traffic"] = {"bytes_per_launch": int(by), "GBps": by / t / 1e6, "frac_of_copy_peak": by / t / 1e6 / hbm,
                                  "peak": hbm, "note": "design traffic (saved bf16 activations + activation gradients), not algorithmic bytes"}
    # the weight-gradient kernel by itself (the lumped entry above also holds the heads kernel, the unpad pass, a memset and
    # the gaps between them): first CTA start to last CTA end of its last launch, from the kernel's own globaltimer trace
    try:
        buf = (C.c_uint64 * 296)()
        bwd(True)
        torch.cuda.synchronize()
        if lib.nb_debug_wgrad_trace(buf, 296) == 0:
            tt = np.array(buf[:], dtype=np.int64).reshape(148, 2)
            t_k = float(tt[:, 1].max() - tt[:, 0].min()) / 1e6
            w_flop = (FLOP_FWD - 2 * (256 + 384)) * M            # dW GEMMs: the forward's MACs without the two CUDA-core heads
            w_by = 80 * 128 * M                                  # 80 operand slabs per tile: 76 distinct ones, dA_5 twice
            per["mlp_wgrad_kernel alone (device globaltimer)"] = {
                "ms": round(t_k, 4), "algorithmic_flop": int(w_flop), "tflops": round(w_flop / t_k / 1e9, 1),
                "frac_tensor_burst": round(w_flop / t_k / 1e9 / burst, 3), "design_traffic_bytes": int(w_by),
                "design_traffic_GBps": round(w_by / t_k / 1e6, 1), "design_traffic_frac_hbm": round(w_by / t_k / 1e6 / hbm, 3)}
    except Exception as e:      # a debugging aid: never fail the bench on it
        per["mlp_wgrad_kernel alone (device globaltimer)"] = {"error": str(e)}
    roof["per_kernel"] = per
    roof["mlp_tensor_pipe"] = {"kernel": "mlp_fwd_tc(inference 32768x192)", "tflops": per["mlp_fwd_tc(inference 32768x192)"]["tflops"],
                               "frac_of_burst_peak": round(FLOP_FWD * Mi / t_inf / 1e9 / burst, 3),
                               "frac_of_sustained_peak": round(FLOP_FWD * Mi / t_inf / 1e9 / sustained, 3)}
    return {"roofline": roof}


def ncu_traffic():
    """DRAM bytes per launch of the MLP kernels from the committed ncu summary (same sizes as kernel_breakdown:
    tools/prof_kernels.py, 1024 rays x 192 samples); {} if the file is missing."""
    import csv
    import glob
    out = {}
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_mlp_kernels.csv")))
    if not files:
        return out
    rows = list(csv.reader(open(files[-1])))
    if len(rows) < 3:
        return out
    h = rows[0]
    try:
        ik, ir, iw = h.index("Kernel Name"), h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
    except ValueError:
        return out
    unit = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
    ur, uw = unit.get(rows[1][ir], 1.0), unit.get(rows[1][iw], 1.0)
    for r in rows[2:]:
        name = r[ik].replace("void ", "").replace("nb::", "").split("<")[0].split("(")[0].strip()
        if name not in out:     # first launch of each kernel = the training-sized one
            out[name] = float(r[ir]) * ur + float(r[iw]) * uw
    return out


def render_leg(renderer, coarse, fine, K, poses_d, dev, world=1):
    """800x800 full-res render (BASELINE.json configs[3]): frames/s with the pose resident; image rows are
    split across the ranks (dist.render_sharded) and the tiles all-gathered, time = max over ranks."""
    import torch.distributed as dist
    from nerf_shared_b200 import datasets, dist as nd
    K8 = datasets.synthetic_blender(800, 800, 1)[3]
    renderer.eval()
    perturb = renderer.perturb
    renderer.perturb = 0.0

    def frame(i):
        return nd.render_sharded(renderer, 800, 800, K8, 32768, poses_d[i % N_IMAGES, :3, :4], coarse, fine)

    with torch.no_grad():
        frame(0)
        frame(1)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 4
        for i in range(n):
            frame(i + 2)
        e1.record()
        torch.cuda.synchronize()
        # end to end: host pose in (H2D, 64 B), uint8 frame out (to8b on the device, D2H into pinned memory, 1.92 MB);
        # frame i's copy runs under frame i+1's kernels, the host reads every frame
        from nerf_shared_b200 import utils
        poses_h = poses_d.cpu().pin_memory()          # a pageable source would make the H2D copy block the host
        host = [torch.empty((800, 800, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
        done = [torch.cuda.Event() for _ in range(2)]
        copy_stream = torch.cuda.Stream(device=dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        checksum = 0
        n2 = 8
        for i in range(n2):
            c2w = poses_h[(i + 2) % N_IMAGES].to(dev, non_blocking=True)[:3, :4]
            rgb8 = utils.to8b_device(nd.render_sharded(renderer, 800, 800, K8, 32768, c2w, coarse, fine)[0])
            copy_stream.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(copy_stream):
                host[i & 1].copy_(rgb8, non_blocking=True)
                rgb8.record_stream(copy_stream)
                done[i & 1].record(copy_stream)
            if i > 0:
                done[(i - 1) & 1].synchronize()
                checksum += int(host[(i - 1) & 1][400, 400, 0])
        done[(n2 - 1) & 1].synchronize()
        checksum += int(host[(n2 - 1) & 1][400, 400, 0])
        e2e_ms = (time.perf_counter() - t0) * 1e3 / n2
        extra = {}
        if world == 1:
            # configs[0] on the GPU: the 400x400 view the CPU arm renders (bench.py --impl reference, configs0_frame)
            K4 = datasets.synthetic_blender(400, 400, 1)[3]
            for _ in range(2):
                renderer.render_from_pose(400, 400, K4, 32768, poses_d[0, :3, :4], coarse, fine, retraw=False)
            torch.cuda.synchronize()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record()
            for i in range(4):
                renderer.render_from_pose(400, 400, K4, 32768, poses_d[i % N_IMAGES, :3, :4], coarse, fine, retraw=False)
            f1.record()
            torch.cuda.synchronize()
            ms4 = f0.elapsed_time(f1) / 4
            extra["configs0_400x400"] = {"ms_per_frame": ms4, "fps": 1e3 / ms4, "rays_per_s": 160000 / (ms4 * 1e-3)}
            # the product API with its asynchronous frame tail: Renderer.render_from_batch_poses writes PNG files
            # (render_utils.py:293-319); frame i's D2H + PNG encode overlap frame i+1's kernels
            import shutil
            import tempfile
            tmp = tempfile.mkdtemp(prefix="nb_frames_")
            try:
                batch = poses_d[:6, :3, :4]
                renderer.render_from_batch_poses(800, 800, K8, 32768, batch[:1], coarse, fine, False, tmp)
                torch.cuda.synchronize()
                tb = time.perf_counter()
                renderer.render_from_batch_poses(800, 800, K8, 32768, batch, coarse, fine, False, tmp)
                torch.cuda.synchronize()
                msb = (time.perf_counter() - tb) * 1e3 / batch.shape[0]
                extra["render_from_batch_poses_800x800_png"] = {"ms_per_frame": msb, "fps": 1e3 / msb, "frames": int(batch.shape[0]),
                                                                 "what": "wall clock incl. uint8 conversion, pinned D2H and PNG encoding of every frame"}
            finally:
                shutil.rmtree(tmp, ignore_errors=True)
    renderer.perturb = perturb
    t = torch.tensor([e0.elapsed_time(e1) / n, e2e_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t[0].item()), float(t[1].item())
    burst, sustained, _, _ = peaks()
    tf = 640000 * (64 + 192) * FLOP_FWD / ms / 1e9
    return {"workload": "800x800 Blender render, 64+128 samples, rows split over %d GPU(s)" % world, "fps": 1e3 / ms,
            "ms_per_frame": ms, "mlp_tflops": tf, "frac_of_sustained_peak": tf / (sustained * world),
            "e2e": {"fps": 1e3 / e2e_ms, "ms_per_frame": e2e_ms, "h2d_bytes_per_frame": 64, "d2h_bytes_per_frame": 800 * 800 * 3,
                    "api": "host pose -> dist.render_sharded -> utils.to8b_device -> pinned uint8 frame (wall clock)"},
            **extra}


def pose_leg(dev, n_rays=512, steps=20):
    """BASELINE.json configs[4]: fern-shaped LLFF scene (504x378, NDC, 64 + 64 samples): one full-frame render and
    the pose-estimation step of examples/relative_pose_estimation_demo/demo_est_rel_pose.py:74-98 -- rays of the whole
    grid from a pose that requires grad, a 512-pixel batch, render with frozen weights, MSE, backward to the pose."""
    from nerf_shared_b200 import datasets, render_utils, utils

    class A(Args):
        N_importance = 64
    Hf, Wf = 378, 504
    images, poses, _, Kf = datasets.synthetic_llff(Hf, Wf, 2)
    torch.manual_seed(0)
    mc, mf = utils.create_nerf_models(A())
    for m in (mc, mf):
        for p in m.parameters():
            p.requires_grad_(False)
    r = render_utils.Renderer(perturb=0.0, N_importance=64, N_samples=64, use_viewdirs=True, white_bkgd=False,
                              raw_noise_std=0.0, ndc=True, lindisp=False, near=0.0, far=1.0)
    pose0 = torch.tensor(poses[0, :3, :4], device=dev)
    with torch.no_grad():
        for _ in range(2):
            r.render_from_pose(Hf, Wf, Kf, 32768, pose0, mc, mf, retraw=False)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            r.render_from_pose(Hf, Wf, Kf, 32768, pose0, mc, mf, retraw=False)
        e1.record()
        torch.cuda.synchronize()
    frame_ms = e0.elapsed_time(e1) / 3
    g = torch.Generator().manual_seed(5)
    target = torch.from_numpy(images[1].reshape(-1, 3)).to(dev)
    pose = pose0.clone().requires_grad_(True)

    target_np = images[1]

    # (the demo draws its pixel batch with np.random.choice(..., replace=False): a 190 k-element permutation per step, 3 ms of
    #  NumPy that is the caller's, not this library's -- the batches are drawn before the timed loop)
    batches = [np.random.choice(Hf * Wf, size=n_rays, replace=False) for _ in range(16)]
    counter = [0]

    def step():       # demo_est_rel_pose.py:78-97 line by line (full-grid get_rays, [ys, xs] indexing, render, img2mse, backward)
        sel = batches[counter[0] % len(batches)]
        counter[0] += 1
        ys, xs = sel // Wf, sel % Wf
        target_s = torch.Tensor(target_np[ys, xs]).to(dev)
        rays_o, rays_d = utils.get_rays(Hf, Wf, Kf, pose)
        ro, rd = rays_o[ys, xs], rays_d[ys, xs]
        rgb = r.render_from_rays(Hf, Wf, Kf, chunk=32768, rays=torch.stack([ro, rd], 0), coarse_model=mc, fine_model=mf,
                                 retraw=True)[0]
        loss = utils.img2mse(rgb, target_s)
        pose.grad = None
        loss.backward()
        return loss

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        loss = step()
    gn = float(pose.grad.norm())            # D2H read of the result
    ms = (time.perf_counter() - t0) * 1e3 / steps
    assert np.isfinite(gn) and np.isfinite(float(loss))

    # the same step as one CUDA graph (train.GraphedPoseStep): rays of the 512 selected pixels only, render with frozen
    # weights, MSE, backward to the pose; per step the host copies (pose, pixel ids, targets) into the static inputs,
    # replays, and applies a gradient step to the pose (the demo's optimizer acts on a 7-parameter twist instead)
    from nerf_shared_b200 import train
    gstep = train.GraphedPoseStep(r, mc, mf, Hf, Wf, Kf, n_rays)
    cur = pose0.clone()
    cur4 = torch.eye(4, device=dev)

    def gstep_once():
        pix = torch.randint(0, Hf * Wf, (n_rays,), generator=g).to(dev, non_blocking=True)
        cur4[:3] = cur
        loss_g, grad = gstep(cur4, pix, target[pix])
        cur.sub_(1e-4 * grad[:3])
        return loss_g

    for _ in range(5):
        gstep_once()
    torch.cuda.synchronize()
    gsteps = 100
    t0 = time.perf_counter()
    for _ in range(gsteps):
        loss_g = gstep_once()
    lg = float(loss_g)                      # D2H read of the result
    gms = (time.perf_counter() - t0) * 1e3 / gsteps
    assert np.isfinite(lg)
    return {"workload": "fern-shaped LLFF NDC scene 504x378, 64+64 samples (configs[4])",
            "render_fps": 1e3 / frame_ms, "render_ms_per_frame": frame_ms,
            "render_mlp_tflops": Hf * Wf * (64 + 128) * FLOP_FWD / frame_ms / 1e9,
            "pose_step": {"rays": n_rays, "ms_per_step": gms, "steps_per_s": 1e3 / gms, "launches_per_replay": gstep.launches_per_replay,
                          "what": "train.GraphedPoseStep: get_rays_at(512 pixels, pose requires grad) -> render (frozen weights) -> "
                                  "MSE -> backward to the pose as one CUDA graph + a gradient step on the pose; wall clock incl. the host",
                          "unchanged_demo_loop": {"ms_per_step": ms, "steps_per_s": 1e3 / ms,
                                                  "what": "demo_est_rel_pose.py:74-97 call by call: NumPy pixel batch -> get_rays(504x378 grid, pose requires grad) "
                                                          "[ys, xs] -> render -> img2mse -> backward; wall clock"}}}


def cpu_step_fn():
    """The reference's CPU path for the same step (oracle port: torch CPU fp32 autograd + Adam,
    main.py:72-112), all host threads.  Returns step(n_rays) -> seconds."""
    from oracle import nerf_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    pc, pf = O.make_params(11), O.make_params(12)
    K = O.blender_intrinsics(H, W)
    c2w = torch.from_numpy(O.pose_spherical(30.0, -30.0, 4.0))[:3, :4]
    ro, rd = O.get_rays(H, W, K, c2w)
    ro, rd = ro.reshape(-1, 3), rd.reshape(-1, 3)
    state = {k: [torch.zeros_like(v), torch.zeros_like(v)] for p in (pc, pf) for k, v in p.items()}

    def step(n, it=[0]):
        it[0] += 1
        idx = torch.from_numpy(np.random.default_rng(it[0]).choice(H * W, n, replace=False))
        tgt = torch.rand(n, 3)
        t0 = time.perf_counter()
        loss, gc, gf, _ = O.train_step_grads(pc, pf, ro[idx].contiguous(), rd[idx].contiguous(), tgt, 2.0, 6.0,
                                             t_rand=torch.rand(n, 64), u=torch.rand(n, 128))
        for p, g in ((pc, gc), (pf, gf)):
            for k in p:
                p[k], state[k][0], state[k][1] = O.adam_step(p[k], g[k], state[k][0], state[k][1], it[0], 5e-4)
        return time.perf_counter() - t0

    return step


CPU_SAMPLE_RAYS = 256      # fixed: the CPU arm times this many rays of the 1,024-ray step (rays/s is batch-flat on the CPU)


def cpu_train_baseline():
    step = cpu_step_fn()
    n = CPU_SAMPLE_RAYS
    step(n)                                             # discarded: thread pool / allocator warm-up
    best = min(step(n) for _ in range(3))
    return {"value": n / best, "unit": "rays/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": "%d rays of the 1,024-ray training step (coarse+fine fwd/bwd + Adam), 1 discarded + best of 3; oracle port, "
                      "torch %s CPU fp32" % (n, torch.__version__),
            "s_per_step": best}


def cpu_frame_leg(budget_s=240.0):
    """BASELINE.json configs[0] / BASELINE.md section 3: the 400 x 400 lego-shaped view (64 + 128 samples, random-init weights,
    perturb = 0, chunk = 32,768) rendered forward by the reference's CPU path (oracle port), one warm-up chunk and then the
    frame.  If the warm-up chunk projects the frame beyond `budget_s`, whole chunks are rendered up to the budget and the
    sample is stated."""
    from oracle import nerf_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    pc, pf = O.make_params(11), O.make_params(12)
    K = O.blender_intrinsics(H, W)
    c2w = torch.from_numpy(O.pose_spherical(40.0, -30.0, 4.0))[:3, :4]
    ro, rd = O.get_rays(H, W, K, c2w)
    vd = O.unit_viewdirs(rd)
    ro, rd = ro.reshape(-1, 3).float(), rd.reshape(-1, 3).float()
    chunk, n = 32768, H * W

    def render(b, e):
        with torch.no_grad():
            return O.render_rays(pc, pf, ro[b:e], rd[b:e], vd[b:e], 2.0, 6.0)["rgb_map"]

    t0 = time.perf_counter()
    render(0, chunk)                                    # warm-up chunk
    warm = time.perf_counter() - t0
    n_chunks = (n + chunk - 1) // chunk
    do = n_chunks if warm * n_chunks <= budget_s else max(1, int(budget_s / warm))
    t0 = time.perf_counter()
    done = 0
    for c in range(do):
        done += render(c * chunk, min((c + 1) * chunk, n)).shape[0]
    dt = time.perf_counter() - t0
    return {"workload": "BASELINE.json configs[0]: 400x400 lego-shaped view, 64+128 samples, forward render on the host cores",
            "rays_per_s": done / dt, "s_per_frame": dt * n / done, "rays_timed": done, "full_frame": done == n,
            "cores": torch.get_num_threads(), "kind": "port", "torch": torch.__version__}


def run_reference(a):
    """--impl reference: the reference's own CPU implementation of the step (oracle port) timed on
    the host cores; rank 0 only.  Every step is a bounded sample (256 rays) of the 1024-ray step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    step = cpu_step_fn()
    n = CPU_SAMPLE_RAYS
    step(n)                                             # discarded (cold thread pool)
    for _ in range(a.warmup):
        step(n)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step(n)
    per = (time.perf_counter() - t0) / a.steps
    v = n / per
    sample = "%d rays of the 1,024-ray step per timed step x %d steps (1 discarded + %d warm-up); oracle port of main.py:72-112, torch %s CPU fp32" % (
        n, a.steps, a.warmup, torch.__version__)
    out = {"impl": "reference", "metric": "train_rays_per_s", "value": v, "unit": "rays/s", "n_gpus": world, "steps": a.steps,
           "warmup": a.warmup, "ms_per_step": per * 1e3, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": workload_config(a.rays, world),
           "cpu_baseline": {"kind": "port", "cores": torch.get_num_threads(), "sample": sample, "value": v, "unit": "rays/s"},
           "e2e": {"value": v, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if world == 1 and os.environ.get("NB_REF_FRAME", "1") != "0":
        out["configs0_frame"] = cpu_frame_leg()
    print(json.dumps(out), file=_real_stdout, flush=True)


if __name__ == "__main__":
    a = parse()
    # stdout carries ONE JSON line: Python-level prints (the library prints its configuration) AND C-level writes to fd 1
    # (NCCL prints its version banner there at N > 1) go to stderr; the JSON line goes to the original stdout
    sys.stdout.flush()
    _json_fd = os.dup(1)
    os.dup2(2, 1)
    _real_stdout = os.fdopen(_json_fd, "w")
    sys.stdout = sys.stderr
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
