"""Read-only / write-only / copy bandwidth of this GPU's HBM (torch kernels, CUDA events, best of 10): the
denominators behind the per-kernel HBM rooflines (the forward / dgrad kernels only write, wgrad only reads)."""
import json
import torch

dev = torch.device("cuda:0")
n = 1 << 29            # 2 GiB of fp32
a = torch.empty(n, device=dev)
b = torch.empty(n, device=dev)


def best(fn, reps=10):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts)


out = {}
t = best(lambda: a.zero_());            out["write_only_GBps"] = 4 * n / t / 1e6
t = best(lambda: a.fill_(1.5));         out["fill_GBps"] = 4 * n / t / 1e6
t = best(lambda: b.copy_(a));           out["copy_GBps"] = 8 * n / t / 1e6
t = best(lambda: torch.sum(a));         out["read_only_GBps"] = 4 * n / t / 1e6
t = best(lambda: torch.add(a, b, out=b)); out["read2_write1_GBps"] = 12 * n / t / 1e6
print(json.dumps({k: round(v, 1) for k, v in out.items()}))
