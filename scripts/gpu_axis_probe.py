"""Time single axis passes on the GPU: column transforms of length n over `cols` contiguous columns,
`planes` times (fftw_plan_many_dft with istride = cols), and row transforms for comparison.
Usage: python scripts/gpu_axis_probe.py [fp32|fp64]"""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from fftw3_b200 import build, capi

prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
import os
lib = capi.FFTW3Lib(os.environ.get("B2FFT_LIB") or build.lib_path(prec), prec == "fp32")
print("library:", os.environ.get("B2FFT_LIB") or build.lib_path(prec), "PF=", os.environ.get("FFTW_B200_PF"))
ct = torch.complex64 if prec == "fp32" else torch.complex128
esz = 8 if prec == "fp32" else 16

def timeit(p, x, y, reps=20):
    for _ in range(3): lib.execute_dft(p, x, y)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): lib.execute_dft(p, x, y)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

def col(n, cols, planes, inplace=True):
    x = torch.zeros((planes, n, cols), dtype=ct, device="cuda")
    y = x if inplace else torch.empty_like(x)
    # one plan: loop over planes (howmany_rank 2 via plan_many with 2-D embedding trick is not
    # available) -> use a 2-D plan with n0 = n and a length-1 ... simpler: rank-1 many over cols, per plane
    p = lib.plan_guru_dft([(n, cols, cols)], [(cols, 1, 1), (planes, n * cols, n * cols)], x, y, -1)
    assert p, "no plan"
    ms = timeit(p, x, y)
    gb = 2 * x.numel() * esz / 1e9
    print(f"col n={n} cols={cols} planes={planes}: {ms:.3f} ms  {gb/ms:.2f} TB/s   {lib.sprint_plan(p).count('pass ')} pass(es)")
    print("   " + lib.sprint_plan(p).replace("\n", "\n   "))
    lib.destroy_plan(p)

def row(n, rows):
    x = torch.zeros((rows, n), dtype=ct, device="cuda")
    p = lib.plan_many_dft([n], rows, x, None, 1, n, x, None, 1, n, -1)
    ms = timeit(p, x, x)
    gb = 2 * x.numel() * esz / 1e9
    print(f"row n={n} rows={rows}: {ms:.3f} ms  {gb/ms:.2f} TB/s")
    lib.destroy_plan(p)

GiB = 2 ** 30
tot = GiB // esz  # elements in 1 GiB
for n in (64, 128, 256, 512, 1024, 2048, 4096):
    row(n, tot // n)
for n, cols in ((512, 512), (1024, 1024), (2048, 2048), (512, 512 * 512), (2048, 2048 * 32), (256, 256), (4096, 4096)):
    planes = max(1, tot // (n * cols))
    col(n, cols, planes)
