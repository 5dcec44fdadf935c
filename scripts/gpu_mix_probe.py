"""Mixed-radix tile kernels (csrc/mixtile.cu) against the generic kernel they replace, fp32 and fp64, device arrays,
1 GiB of complex data per case.  python scripts/gpu_mix_probe.py > gpurun_out/r2_mix_probe.log
  rows : batched contiguous lines (the list of profiles/r1_nonpow2_probe.log and the other instantiated lengths)
  cols : the first axis of an [n][w] array, in place (column form)
  long : 1-D lengths whose two passes are mixed-radix kernels"""
import json
import os
import sys

import torch

sys.path.insert(0, ".")
from fftw3_b200 import build, capi

PEAK = 6549.8
try:
    PEAK = float(json.load(open("MEASURED_PEAKS.json")).get("hbm_gbs", PEAK))
except Exception:
    pass


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


def case(lib, dt, esz, n, rows, col_w=0):
    """returns {variant: (ms, passes, first pass line)}"""
    out = {}
    for variant in ("mix", "generic"):
        if variant == "generic":
            os.environ["FFTW_B200_NO_MIX"] = "1"
        else:
            os.environ.pop("FFTW_B200_NO_MIX", None)
        if col_w:
            x = torch.zeros((n, col_w), dtype=dt, device="cuda")
            p = lib.plan_many_dft([n], col_w, x, None, col_w, 1, x, None, col_w, 1, -1)
        else:
            x = torch.zeros((rows, n), dtype=dt, device="cuda")
            p = lib.plan_many_dft([n], rows, x, None, 1, n, x, None, 1, n, -1)
        if not p:
            out[variant] = None
            continue
        d = lib.sprint_plan(p).split("\n")
        ms = timeit(lambda: lib.execute_dft(p, x, x))
        npass = sum(1 for l in d if l.strip().startswith("pass "))
        out[variant] = (ms, npass, d[1].strip()[:110], 2 * x.numel() * esz / 1e6 / ms)
        lib.destroy_plan(p)
        del x
    os.environ.pop("FFTW_B200_NO_MIX", None)
    return out


def show(tag, n, r):
    m, g = r["mix"], r["generic"]
    if not m or not g:
        print(f"{tag} n={n}: NULL plan")
        return
    print(f"{tag} n={n}: mix {m[0]:.3f} ms = {m[3]:.0f} GB/s = {m[3] / PEAK:.2f} of HBM peak, {m[1]} pass(es) "
          f"(per pass {m[3] * m[1] / PEAK:.2f}) | generic {g[0]:.3f} ms ({g[1]} pass(es)) | x{g[0] / m[0]:.2f} | {m[2]}", flush=True)


for prec, dt, esz in (("fp32", torch.complex64, 8), ("fp64", torch.complex128, 16)):
    lib = capi.FFTW3Lib(build.lib_path(prec), prec == "fp32")
    tot = (1 << 30) // esz
    print(f"== {prec} rows (HBM peak {PEAK:.0f} GB/s; algorithmic bytes = read + write once)")
    rows_list = [1000, 1536, 3072, 3000, 4000, 6561, 243, 100, 1001, 96, 192, 384, 768, 200, 250, 400, 500, 2000, 5000, 6144,
                 81, 729, 2187, 125, 625, 3125, 49, 343, 2401, 1080]
    for n in rows_list:
        if prec == "fp64" and n > 4096:
            continue
        show("rows", n, case(lib, dt, esz, n, tot // n))
    print(f"== {prec} cols")
    for n in (96, 100, 192, 250, 384, 400, 500, 768, 1000, 1536, 243, 729, 343):
        if prec == "fp64" and n > 1024:
            continue
        w = (tot // n) // 64 * 64
        show("cols", n, case(lib, dt, esz, n, 0, col_w=w))
    print(f"== {prec} long")
    for n in (100000, 24576, 96000, 160000, 3 * 2 ** 17, 10 ** 6):
        show("long", n, case(lib, dt, esz, n, max(1, tot // n)))
