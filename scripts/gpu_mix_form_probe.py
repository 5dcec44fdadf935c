"""Per-form choice between the two shapes of a mixed-radix length (csrc/mixtile.cu): the column + twiddle form and the
row -> column form only occur as the two passes of a long 1-D transform, so each is isolated by pairing the length with a
power-of-two partner (whose kernel does not change): n = N * P runs col+twiddle(N) then row->col(P); n = P * N the reverse.
python scripts/gpu_mix_form_probe.py > gpurun_out/r2_mix_form_probe.log"""
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


for N, plo, phi in ((250, 128, 256), (500, 256, 512), (1000, 512, 1024), (243, 128, 256), (729, 512, 1024)):
    for n, what in ((N * plo, "col+twiddle"), (phi * N, "row->col")):
        res = []
        for alt in (0, 1):
            if alt:
                os.environ["FFTW_B200_MIX_ALT"] = "1"
            else:
                os.environ.pop("FFTW_B200_MIX_ALT", None)
            rows = max(1, tot // n)
            x = torch.zeros((rows, n), dtype=torch.complex64, device="cuda")
            p = lib.plan_many_dft([n], rows, x, None, 1, n, x, None, 1, n, -1)
            ms = timeit(lambda: lib.execute_dft(p, x, x))
            d = [l.strip() for l in lib.sprint_plan(p).split("\n")[1:3]]
            line = [l for l in d if f"n={N} " in l]
            res.append((ms, line[0][line[0].index("E="):][:44] if line else "?? " + " / ".join(l[:60] for l in d)))
            lib.destroy_plan(p)
            del x
        os.environ.pop("FFTW_B200_MIX_ALT", None)
        print(f"N={N} {what} (n={n}): listed first {res[0][0]:.3f} ms [{res[0][1]}] | listed last {res[1][0]:.3f} ms [{res[1][1]}]", flush=True)
