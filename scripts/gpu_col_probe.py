"""Column (strided-axis) passes: correctness against torch.fft (probe only, not a parity test) and time,
round-1 kernels (FFTW_B200_NO_TMA=1) against the TMA column kernels, same process, same arrays.
Usage: python scripts/gpu_col_probe.py [quick]"""
import os
import sys

import torch

sys.path.insert(0, ".")
from fftw3_b200 import build, capi

lib = capi.FFTW3Lib(os.environ.get("B2FFT_LIB") or build.lib_path("fp32"), True)
quick = len(sys.argv) > 1 and sys.argv[1] == "quick"


def timeit(p, x, y, reps=20):
    for _ in range(3):
        lib.execute_dft(p, x, y)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps):
        lib.execute_dft(p, x, y)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def col(n, cols, planes, sign=-1, inplace=False):
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.view_as_complex(torch.rand((planes, n, cols, 2), device="cuda", generator=g) - 0.5)
    out = {}
    for mode in ("r1", "tma"):
        if mode == "r1":
            os.environ["FFTW_B200_NO_TMA"] = "1"
        else:
            os.environ.pop("FFTW_B200_NO_TMA", None)
        xi = x.clone()
        y = xi if inplace else torch.empty_like(xi)
        p = lib.plan_guru_dft([(n, cols, cols)], [(cols, 1, 1), (planes, n * cols, n * cols)], xi, y, sign)
        assert p, "no plan"
        lib.execute_dft(p, xi, y)
        torch.cuda.synchronize()
        sl = slice(0, min(planes, 2))
        xr = x[sl].to(torch.complex128)
        ref = torch.fft.fft(xr, dim=1) if sign < 0 else torch.fft.ifft(xr, dim=1) * n
        err = float(torch.linalg.norm(y[sl].to(torch.complex128) - ref) / torch.linalg.norm(ref))
        ms = timeit(p, xi, y)
        desc = [l for l in lib.sprint_plan(p).split("\n") if "pass 0" in l][0]
        out[mode] = (ms, err, desc)
        lib.destroy_plan(p)
    gb = 2 * x.numel() * 8 / 1e9
    a, b = out["r1"], out["tma"]
    print(f"col n={n} pitch={cols*8}B planes={planes} sign={sign:+d} {'in place' if inplace else 'out of place'}: "
          f"r1 {a[0]:.3f} ms ({gb/a[0]:.2f} TB/s, err {a[1]:.1e})  tma {b[0]:.3f} ms ({gb/b[0]:.2f} TB/s, err {b[1]:.1e})")
    if "TMA" not in b[2]:
        print("   NOTE: the TMA kernel was not selected:", b[2].strip())
    sys.stdout.flush()


tot = (1 << 30) // 8
col(1024, 1024, 4)           # small first: a wrong kernel shows here before the big ones run
col(512, 1000, 8)            # ragged last tile (1000 = 62 * 16 + 8)
col(2048, 2056, 2, sign=+1)  # ragged, backward
col(1024, 1024, 16, sign=+1, inplace=True)
cases = [(512, 512), (1024, 1024), (2048, 2048)]
if not quick:
    cases += [(512, 512 * 512), (1024, 1024 * 128), (2048, 2048 * 32), (1024, 512)]
for n, cols in cases:
    col(n, cols, max(1, tot // (n * cols)))
