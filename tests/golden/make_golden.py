"""Generate tests/golden/ref_golden.npz from the UNMODIFIED reference build.

Run here (the container that has /root/reference):
    make -C oracle/refbuild PREC=f64 && make -C oracle/refbuild PREC=f32
    python tests/golden/make_golden.py
The reference has no golden vectors of its own, so these are
outputs of the reference itself (oracle/_ref/libfftw3_ref_f64.so / _f32.so, built by
oracle/refbuild/Makefile from the sources under /root/reference) on seeded inputs:
uniform [-0.5, 0.5) like the reference verifier (fftw/libbench2/verify-lib.c:64-67).
Inputs are stored next to outputs so the fixtures do not depend on any RNG.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from fftw3_b200 import capi  # noqa: E402
from oracle import pyoracle  # noqa: E402

C2C_SHAPES = [(1,), (2,), (3,), (4,), (5,), (7,), (8,), (16,), (30,), (64,), (100,), (128,),
              (243,), (1024,), (4096,), (8, 12), (16, 16), (4, 6, 8), (16, 16, 16), (5, 3, 7)]
R2C_SHAPES = [(2,), (4,), (6,), (7,), (16,), (30,), (64,), (1024,), (6, 8), (5, 9), (4, 4, 6),
              (3, 5, 7)]


def main():
    pyoracle.build_reference()
    out = {}
    rng = np.random.default_rng(20261104)
    for single in (False, True):
        lib = pyoracle.reference_lib(single)
        assert lib is not None, "reference build missing: make -C oracle/refbuild"
        tag = "f32" if single else "f64"
        for shape in C2C_SHAPES:
            key = "c2c_" + "x".join(map(str, shape)) + "_" + tag
            x = ((rng.random(shape) - 0.5) + 1j * (rng.random(shape) - 0.5)).astype(lib.cplx)
            out[key + "_in"] = x
            out[key + "_fwd"] = lib.dft(x, capi.FFTW_FORWARD)
            out[key + "_bwd"] = lib.dft(x, capi.FFTW_BACKWARD)
        for shape in R2C_SHAPES:
            key = "r2c_" + "x".join(map(str, shape)) + "_" + tag
            x = (rng.random(shape) - 0.5).astype(lib.real)
            X = lib.r2c(x)
            out[key + "_in"] = x
            out[key + "_out"] = X
            out[key + "_c2r"] = lib.c2r(X, shape[-1])
        # advanced interface: strided / embedded batch (fftw_plan_many_dft, fftw_api.c:642-666)
        n, howmany, istride, idist, ostride, odist = 12, 5, 3, 40, 2, 30
        buf = ((rng.random(howmany * idist + n * istride) - 0.5)
               + 1j * (rng.random(howmany * idist + n * istride) - 0.5)).astype(lib.cplx)
        res = np.zeros(howmany * odist + n * ostride, dtype=lib.cplx)
        p = lib.plan_many_dft([n], howmany, buf, None, istride, idist, res, None, ostride, odist,
                              capi.FFTW_FORWARD)
        lib.execute(p)
        lib.destroy_plan(p)
        out[f"many_{tag}_in"] = buf
        out[f"many_{tag}_out"] = res
        out[f"many_{tag}_geom"] = np.array([n, howmany, istride, idist, ostride, odist])
        # 2-D transform embedded in a larger array (inembed/onembed)
        n2 = [4, 6]
        inembed, onembed = [5, 8], [4, 7]
        hm, idist2, odist2 = 3, 5 * 8 + 3, 4 * 7 + 1
        buf2 = ((rng.random(hm * idist2) - 0.5) + 1j * (rng.random(hm * idist2) - 0.5)).astype(lib.cplx)
        res2 = np.zeros(hm * odist2, dtype=lib.cplx)
        p = lib.plan_many_dft(n2, hm, buf2, inembed, 1, idist2, res2, onembed, 1, odist2,
                              capi.FFTW_BACKWARD)
        lib.execute(p)
        lib.destroy_plan(p)
        out[f"embed_{tag}_in"] = buf2
        out[f"embed_{tag}_out"] = res2
        # in-place r2c with the padded layout of fftw_rdft2_pad (fftw_api.c:774-788)
        shape = (3, 10)
        pad = 2 * (shape[1] // 2 + 1)
        a = np.zeros((shape[0], pad), dtype=lib.real)
        a[:, :shape[1]] = (rng.random(shape) - 0.5).astype(lib.real)
        out[f"r2c_inplace_{tag}_in"] = a.copy()
        p = lib.plan_dft_r2c(list(shape), a, a)
        lib.execute(p)
        lib.destroy_plan(p)
        out[f"r2c_inplace_{tag}_out"] = a.copy()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
