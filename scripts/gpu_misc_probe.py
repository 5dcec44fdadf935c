"""Non-power-of-two and Bluestein lengths, batched rows, device pointers.  python scripts/gpu_misc_probe.py"""
import sys
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
tot = 2 ** 27  # complex elements = 1 GiB
for n in (1000, 1536, 3072, 3000, 4000, 6561, 243, 100, 1001, 1009, 4099, 2 ** 13 * 3, 10 ** 5):
    rows = tot // n
    x = torch.zeros((rows, n), dtype=torch.complex64, device="cuda")
    p = lib.plan_many_dft([n], rows, x, None, 1, n, x, None, 1, n, -1)
    if not p:
        print(f"n={n}: NULL plan"); continue
    ms = timeit(lambda: lib.execute_dft(p, x, x))
    gb = 2 * x.numel() * 8 / 1e9
    d = lib.sprint_plan(p).split("\n")
    print(f"n={n} rows={rows}: {ms:.3f} ms  {gb/ms:.2f} TB/s algorithmic  {len(d)-2} pass(es): {d[1][:100]}")
    lib.destroy_plan(p)
    del x
