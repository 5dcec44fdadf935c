"""Does padding a power-of-two pitch fix the slow large-pitch column passes?  Same column pass, pitch = 2^k bytes
against 2^k + a few hundred bytes (only `cols` of each row are transformed; the row is longer).
Usage: python scripts/gpu_pitch_probe.py"""
import os
import sys

import torch

sys.path.insert(0, ".")
from fftw3_b200 import build, capi

lib = capi.FFTW3Lib(build.lib_path("fp32"), True)


def timeit(p, x, y, reps=10):
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


def col(n, cols, pitch, tag):
    """n-point transforms down `cols` adjacent columns of an [n][pitch] array, in place"""
    x = torch.zeros((n, pitch), dtype=torch.complex64, device="cuda")
    for mode in ("r1", "tma"):
        if mode == "r1":
            os.environ["FFTW_B200_NO_TMA"] = "1"
        else:
            os.environ.pop("FFTW_B200_NO_TMA", None)
            os.environ["FFTW_B200_TMA_ALL"] = "1"
        p = lib.plan_guru_dft([(n, pitch, pitch)], [(cols, 1, 1)], x, x, -1)
        assert p
        ms = timeit(p, x, x)
        gib = 2 * n * cols * 8 / 2 ** 30
        print(f"n={n} cols={cols} pitch={pitch*8:>9} B ({tag}) {mode}: {ms:.3f} ms = {ms/gib:.3f} ms per GiB moved*", flush=True)
        lib.destroy_plan(p)


print("(* per GiB of read+write, i.e. comparable with the 'ms per GiB pass' figures / 2)")
for n, cols in ((512, 262144), (2048, 65536), (1024, 131072)):
    for pad, tag in ((0, "2^k"), (32, "+256 B"), (16 * 17, "+2176 B"), (4096 + 32, "+32 KiB+256 B")):
        col(n, cols, cols + pad, tag)
