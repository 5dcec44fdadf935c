"""Run the slab leg's correctness checks (fftw3_b200/bench_slab.py) for one cube size; under torchrun for > 1 rank.
Usage: python scripts/gpu_slab_parity.py N [steps]"""
import json
import os
import sys

sys.path.insert(0, ".")
import torch

import bench
from fftw3_b200 import bench_slab

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(sys.argv[1])
lib = bench.engine_lib("fp32")
lib.lib.b2fft_set_device(local)
rec, st = bench_slab.run(lib, (n, n, n), world, rank, steps=int(sys.argv[2]) if len(sys.argv) > 2 else 2, warmup=3, peak_gbs=6549.8)
par = bench_slab.checks(st)
if rank == 0:
    print(json.dumps({"n": n, "world": world, "ms": rec["ms_per_step"], "stage_ms": rec["stage_ms"],
                      "env": {k: v for k, v in os.environ.items() if k.startswith(("B2FFT_", "FFTW_B200_"))}, "parity": par}))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
