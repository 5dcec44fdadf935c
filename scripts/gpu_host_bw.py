"""Platform limit for host-array executes at N GPUs: aggregate pinned-memory copy bandwidth with all visible GPUs copying at
once (one process, one stream pair per GPU, no engine involved).  H2D alone, D2H alone, both directions together.
Usage: python scripts/gpu_host_bw.py [MiB per GPU]   -> prints one line per (number of GPUs, mode)"""
import sys
import time

import torch

mib = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
nd = torch.cuda.device_count()
host_in = [torch.empty(mib << 20, dtype=torch.uint8).pin_memory() for _ in range(nd)]
host_out = [torch.empty(mib << 20, dtype=torch.uint8).pin_memory() for _ in range(nd)]
dev_a, dev_b, s_in, s_out = [], [], [], []
for d in range(nd):
    with torch.cuda.device(d):
        dev_a.append(torch.empty(mib << 20, dtype=torch.uint8, device=f"cuda:{d}"))
        dev_b.append(torch.empty(mib << 20, dtype=torch.uint8, device=f"cuda:{d}"))
        s_in.append(torch.cuda.Stream(d))
        s_out.append(torch.cuda.Stream(d))


def sync(k):
    for d in range(k):
        torch.cuda.synchronize(d)


def run(k, h2d, d2h, reps=3):
    best = 1e9
    for _ in range(reps + 1):
        sync(k)
        t0 = time.perf_counter()
        for d in range(k):
            if h2d:
                with torch.cuda.stream(s_in[d]):
                    dev_a[d].copy_(host_in[d], non_blocking=True)
            if d2h:
                with torch.cuda.stream(s_out[d]):
                    host_out[d].copy_(dev_b[d], non_blocking=True)
        sync(k)
        best = min(best, time.perf_counter() - t0)
    return best


for k in sorted({1, 2, 4, nd} & set(range(1, nd + 1))):
    for name, h, o in (("h2d", 1, 0), ("d2h", 0, 1), ("both", 1, 1)):
        t = run(k, h, o)
        per_dir = mib / 1024 * 1.073741824 / t  # GB/s per GPU per direction
        print(f"gpus={k} {name:5s} {t * 1e3:8.2f} ms  per GPU per direction {per_dir:6.1f} GB/s  aggregate per direction {per_dir * k:7.1f} GB/s")
