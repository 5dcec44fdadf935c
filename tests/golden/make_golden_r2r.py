"""Generate tests/golden/r2r_golden.npz from the UNMODIFIED reference build (oracle/_ref, see
make_golden.py): every r2r kind at several sizes, 1-D and one mixed 2-D case, double precision.
Run where /root/reference exists:  python tests/golden/make_golden_r2r.py"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402

SIZES = [2, 3, 4, 5, 8, 9, 16, 30, 64]


def main():
    pyoracle.build_reference()
    lib = pyoracle.reference_lib(False)
    L = lib.lib
    L.fftw_plan_r2r.restype = C.c_void_p
    L.fftw_plan_r2r.argtypes = [C.c_int, C.POINTER(C.c_int), C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.c_uint]
    rng = np.random.default_rng(20261105)
    out = {}

    def run(x, kinds):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.zeros_like(x)
        n = (C.c_int * x.ndim)(*x.shape)
        k = (C.c_int * x.ndim)(*kinds)
        p = L.fftw_plan_r2r(x.ndim, n, x.ctypes.data, y.ctypes.data, k, 1 << 6)
        assert p, (x.shape, kinds)
        xin = x.copy()
        lib.execute(p)
        lib.destroy_plan(p)
        return xin, y

    for name, kd in pyoracle.R2R_KINDS.items():
        for n in SIZES:
            x = rng.random(n) - 0.5
            xin, y = run(x, [kd])
            out[f"{name}_{n}_in"] = xin
            out[f"{name}_{n}_out"] = y
    x = rng.random((6, 10)) - 0.5
    xin, y = run(x, [pyoracle.R2R_KINDS["REDFT10"], pyoracle.R2R_KINDS["RODFT00"]])
    out["2d_REDFT10xRODFT00_in"] = xin
    out["2d_REDFT10xRODFT00_out"] = y
    x = rng.random((4, 5, 8)) - 0.5
    xin, y = run(x, [pyoracle.R2R_KINDS["DHT"], pyoracle.R2R_KINDS["REDFT00"], pyoracle.R2R_KINDS["R2HC"]])
    out["3d_DHTxREDFT00xR2HC_in"] = xin
    out["3d_DHTxREDFT00xR2HC_out"] = y
    path = os.path.join(ROOT, "tests", "golden", "r2r_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, len(out), "arrays", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
