"""First against second kernel shape (csrc/mixtile.cu, FFTW_B200_MIX_ALT=1) for the lengths that have two, fp32,
device arrays, 1 GiB per case.  python scripts/gpu_mix_alt_probe.py > gpurun_out/r2_mix_alt_probe.log"""
import os
import sys

import torch

sys.path.insert(0, ".")
from fftw3_b200 import build, capi

lib = capi.FFTW3Lib(build.lib_path("fp32"), True)
tot = 1 << 27


def timeit(fn, reps=8):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def run(n, col):
    res = []
    for alt in (0, 1):
        if alt:
            os.environ["FFTW_B200_MIX_ALT"] = "1"
        else:
            os.environ.pop("FFTW_B200_MIX_ALT", None)
        if col:
            w = (tot // n) // 64 * 64
            x = torch.zeros((n, w), dtype=torch.complex64, device="cuda")
            p = lib.plan_many_dft([n], w, x, None, w, 1, x, None, w, 1, -1)
        else:
            rows = tot // n
            x = torch.zeros((rows, n), dtype=torch.complex64, device="cuda")
            p = lib.plan_many_dft([n], rows, x, None, 1, n, x, None, 1, n, -1)
        d = lib.sprint_plan(p).split("\n")[1].strip()
        ms = timeit(lambda: lib.execute_dft(p, x, x))
        res.append((ms, d[d.index("E="):d.index("E=") + 60]))
        lib.destroy_plan(p)
        del x
    os.environ.pop("FFTW_B200_MIX_ALT", None)
    print(f"{'cols' if col else 'rows'} n={n}: first {res[0][0]:.3f} ms [{res[0][1]}] | second {res[1][0]:.3f} ms [{res[1][1]}]", flush=True)


for n in (250, 500, 1000, 243, 729):
    run(n, True)
for n in (5000, 6561, 625, 3125):
    run(n, False)
# long lengths whose passes use those column kernels
for n in (10 ** 6, 250000, 243 * 243):
    for alt in (0, 1):
        if alt:
            os.environ["FFTW_B200_MIX_ALT"] = "1"
        else:
            os.environ.pop("FFTW_B200_MIX_ALT", None)
        rows = max(1, tot // n)
        x = torch.zeros((rows, n), dtype=torch.complex64, device="cuda")
        p = lib.plan_many_dft([n], rows, x, None, 1, n, x, None, 1, n, -1)
        ms = timeit(lambda: lib.execute_dft(p, x, x))
        print(f"long n={n} {'second' if alt else 'first'}: {ms:.3f} ms | " + " | ".join(l.strip()[:95] for l in lib.sprint_plan(p).split("\n")[1:3]), flush=True)
        lib.destroy_plan(p)
        del x
os.environ.pop("FFTW_B200_MIX_ALT", None)
