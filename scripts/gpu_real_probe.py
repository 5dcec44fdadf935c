"""Time batched / multi-dimensional real transforms on device pointers.  Usage: python scripts/gpu_real_probe.py"""
import sys, os
import numpy as np, torch
sys.path.insert(0, ".")
from fftw3_b200 import build, capi
lib = capi.FFTW3Lib(build.lib_path("fp32"), True)

def timeit(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

def r2c_many(n, hm):
    x = torch.zeros((hm, n), dtype=torch.float32, device="cuda")
    y = torch.zeros((hm, n // 2 + 1), dtype=torch.complex64, device="cuda")
    p = lib.plan_many_dft_r2c([n], hm, x, None, 1, n, y, None, 1, n // 2 + 1)
    ms = timeit(lambda: lib.execute_dft_r2c(p, x, y))
    gb = (x.numel() * 4 + y.numel() * 8) / 1e9
    print(f"r2c n={n} x{hm}: {ms:.3f} ms  {gb/ms:.2f} TB/s algorithmic, {lib.sprint_plan(p).count('pass ')} pass(es)")
    q = lib.plan_many_dft_c2r([n], hm, y, None, 1, n // 2 + 1, x, None, 1, n)
    ms = timeit(lambda: lib.execute_dft_c2r(q, y, x))
    print(f"c2r n={n} x{hm}: {ms:.3f} ms  {gb/ms:.2f} TB/s algorithmic, {lib.sprint_plan(q).count('pass ')} pass(es)")
    lib.destroy_plan(p); lib.destroy_plan(q)

def r2c_nd(shape):
    x = torch.zeros(shape, dtype=torch.float32, device="cuda")
    y = torch.zeros(shape[:-1] + (shape[-1] // 2 + 1,), dtype=torch.complex64, device="cuda")
    p = lib.plan_dft_r2c(list(shape), x, y)
    ms = timeit(lambda: lib.execute_dft_r2c(p, x, y))
    gb = (x.numel() * 4 + y.numel() * 8) / 1e9
    print(f"r2c {shape}: {ms:.3f} ms  {gb/ms:.2f} TB/s algorithmic")
    print("   " + lib.sprint_plan(p).replace("\n", "\n   "))
    lib.destroy_plan(p)

for n, hm in ((256, 2 ** 20), (1024, 2 ** 18), (4096, 2 ** 16), (16384, 2 ** 14), (65536, 2 ** 12), (2 ** 20, 256)):
    r2c_many(n, hm)
r2c_nd((512, 512, 512))
r2c_nd((4096, 4096))
