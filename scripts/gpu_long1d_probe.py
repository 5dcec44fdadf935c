"""Time long 1-D c2c transforms (two-pass four-step and three-pass six-step plans) on the GPU.
Usage: python scripts/gpu_long1d_probe.py [fp32|fp64]"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from fftw3_b200 import build, capi

prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
lib = capi.FFTW3Lib(build.lib_path(prec), prec == "fp32")
ct = torch.complex64 if prec == "fp32" else torch.complex128
esz = 8 if prec == "fp32" else 16


def t(n, hm):
    x = torch.zeros((hm, n), dtype=ct, device="cuda")
    y = torch.empty_like(x)
    p = lib.plan_many_dft([n], hm, x, None, 1, n, y, None, 1, n, -1)
    assert p, n
    for _ in range(3):
        lib.execute_dft(p, x, y)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10):
        lib.execute_dft(p, x, y)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    npass = lib.sprint_plan(p).count("  pass ")
    print(f"c2c {prec} n=2^{int(np.log2(n))} x {hm}: {ms:.3f} ms  {5 * n * np.log2(n) * hm / ms / 1e6:.0f} GFLOP/s  "
          f"{2 * esz * n * hm / ms / 1e6:.0f} GB/s algorithmic  {npass} passes")
    print("   " + lib.sprint_plan(p).strip().replace("\n", "\n   "))
    lib.destroy_plan(p)


sizes = ((1 << 20, 128), (1 << 22, 32), (1 << 23, 16), (1 << 24, 8), (1 << 26, 2), (1 << 27, 1))
if prec == "fp64":
    sizes = ((1 << 20, 64), (1 << 21, 32), (1 << 24, 4), (1 << 26, 1))
for n, hm in sizes:
    t(n, hm)
