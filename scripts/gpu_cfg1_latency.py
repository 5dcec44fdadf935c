"""BASELINE configs[0]: one 1024-point complex double transform (fftw_plan_dft_1d + fftw_execute).  Where do the
microseconds go?  Prints a breakdown for profiles/r2_cfg1_latency.md.  Usage: python scripts/gpu_cfg1_latency.py"""
import ctypes as C
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import bench
from fftw3_b200 import capi

lib = bench.engine_lib("fp64")
L = lib.lib
x = torch.view_as_complex(torch.rand((1024, 2), device="cuda", dtype=torch.float64) - 0.5)
y = torch.empty_like(x)
p = lib.plan_dft_1d(1024, x, y, -1)
assert p
bp = bench.fftw_plan_bp(p)
st = torch.cuda.current_stream()
sp = C.c_void_p(st.cuda_stream)
REP = 2000


def wall(f, rep=REP):
    for _ in range(50):
        f()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(rep):
        f()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / rep * 1e6


sync_us = wall(lambda: lib.execute_dft(p, x, y))                        # fftw_execute_dft: enqueue + kernel + stream sync
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
for _ in range(50):
    L.b2fft_execute_async(bp, x.data_ptr(), y.data_ptr(), sp)
torch.cuda.synchronize()
t0 = time.perf_counter()
e0.record(st)
for _ in range(REP):
    L.b2fft_execute_async(bp, x.data_ptr(), y.data_ptr(), sp)
t_enq = (time.perf_counter() - t0) / REP * 1e6                           # host cost of one asynchronous enqueue
e1.record(st)
torch.cuda.synchronize()
gpu_us = e0.elapsed_time(e1) / REP * 1e3                                 # back-to-back launches: device time per transform
z = torch.zeros(1, device="cuda")
floor_us = wall(lambda: (z.add_(1), torch.cuda.synchronize()))          # any one-kernel launch + sync from this process
ref = bench.run_reference(bench.WORKLOADS["c2c_1d_1024_fp64"], steps=20000, warmup=100, threads=1, measure=False)
print(f"fftw_execute_dft, device arrays, synchronous : {sync_us:7.2f} us per call")
print(f"  host cost of the asynchronous enqueue       : {t_enq:7.2f} us (b2fft_execute_async: device scope, descriptor copy, cudaLaunchKernel)")
print(f"  device time per transform, back to back     : {gpu_us:7.2f} us (kernel + launch gap; 32 KiB of traffic)")
print(f"  launch + sync floor of a trivial kernel     : {floor_us:7.2f} us (torch add_ + cudaDeviceSynchronize, same process)")
print(f"reference on one host core, FFTW_ESTIMATE     : {ref['ms_per_step']*1e3:7.2f} us per fftw_execute")
print(lib.sprint_plan(p))
