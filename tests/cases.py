"""Parity cases shared by the CPU (emulation) and GPU test suites.

Every case drives the engine through the public fftw3.h C ABI (fftw3_b200/capi.py is a plain
ctypes binding of it) and compares with the oracle (oracle/fft_oracle.c via oracle/pyoracle.py,
itself pinned to the unmodified reference by tests/test_oracle.py) or with the committed golden
vectors produced by the reference itself (tests/golden/ref_golden.npz).

Tolerance (BASELINE.json north_star): relative L2 error < 1e-6*log2(N) in fp32 and
< 1e-14*log2(N) in fp64, against double-precision reference output.
"""
from __future__ import annotations

import math

import ctypes as C

import numpy as np

from fftw3_b200 import capi
from oracle import pyoracle as orc


def rel_l2(a, b):
    a = np.asarray(a).astype(np.complex128).ravel()
    b = np.asarray(b).astype(np.complex128).ravel()
    den = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / (den if den > 0 else 1.0))


def tol(lib, N):
    return (1e-6 if lib.single else 1e-14) * max(1.0, math.log2(max(2, N)))


def rand_c(rng, shape, lib):
    return ((rng.random(shape) - 0.5) + 1j * (rng.random(shape) - 0.5)).astype(lib.cplx)


def rand_r(rng, shape, lib):
    return (rng.random(shape) - 0.5).astype(lib.real)


C2C_SHAPES = [(1,), (2,), (3,), (4,), (5,), (7,), (8,), (11,), (13,), (16,), (30,), (64,), (100,),
              (128,), (243,), (1000,), (1024,), (4096,), (8, 12), (16, 16), (4, 6, 8), (16, 16, 16),
              (5, 3, 7), (2, 3, 4, 5)]
C2C_BIG = [(8192,), (16384,), (3 * 2 ** 13,), (2 ** 17,), (64, 64, 64), (4, 2 ** 14),
           # tile-kernel modes with ragged tiles (line counts that are not multiples of the tile)
           (3, 128, 72), (20, 64), (64, 20), (7, 1024), (5, 2048), (2048, 5), (2 ** 18,), (3, 2 ** 16), (256, 256)]
# long strided axes: the persistent async-copy pipelined column kernels (several tiles per CTA stripe,
# ragged last tiles, plane dims, four-step twiddle epilogue)
C2C_PIPE = [(512, 80), (1024, 36), (3, 512, 24), (2048, 20), (2, 1024, 8), (2 ** 19,)]
R2C_SHAPES = [(2,), (4,), (6,), (7,), (9,), (16,), (30,), (64,), (1024,), (6, 8), (5, 9), (4, 4, 6),
              (3, 5, 7)]
R2C_BIG = [(4096,), (2 ** 15,), (2 ** 17,), (3, 2 ** 14), (32, 32, 32),
           (2 ** 19,), (2, 2 ** 20)]
# lengths with a prime factor > 13 (Bluestein path), alone and inside multi-dimensional transforms
C2C_PRIME = [(17,), (19,), (97,), (34,), (1009,), (4099,), (2 * 8191,), (3, 23), (29, 8), (5, 31, 6), (127, 127)]
R2C_PRIME = [(17,), (34,), (97,), (1009,), (2018,), (4, 46), (3, 5, 38), (37, 6), (10007,)]


def check_c2c(lib, shape, seed=0):
    rng = np.random.default_rng(seed)
    x = rand_c(rng, shape, lib)
    N = int(np.prod(shape))
    x0 = x.copy()
    for sign in (capi.FFTW_FORWARD, capi.FFTW_BACKWARD):
        y = lib.dft(x, sign)
        assert rel_l2(y, orc.dft(x0, sign)) < tol(lib, N), (shape, sign)
        assert np.array_equal(x, x0), "out-of-place c2c must preserve its input"
    # in place
    z = x0.copy()
    p = lib.plan_dft(list(shape), z, z, capi.FFTW_FORWARD)
    assert p, "in-place plan is NULL"
    lib.execute(p)
    lib.destroy_plan(p)
    assert rel_l2(z, orc.dft(x0, -1)) < tol(lib, N), (shape, "in place")


def check_r2c_c2r(lib, shape, seed=0):
    rng = np.random.default_rng(seed)
    x = rand_r(rng, shape, lib)
    N = int(np.prod(shape))
    X = lib.r2c(x)
    Xo = orc.r2c(x)
    assert rel_l2(X, Xo) < tol(lib, N), (shape, "r2c")
    Xin = Xo.astype(lib.cplx)
    y = lib.c2r(Xin, shape[-1])
    assert rel_l2(y, orc.c2r(Xin, shape[-1])) < tol(lib, N), (shape, "c2r")
    # unnormalised round trip (fftw3.h convention): c2r(r2c(x)) = N x
    assert rel_l2(lib.c2r(X, shape[-1]), N * x.astype(np.float64)) < 2 * tol(lib, N)


def check_r2c_inplace(lib, shape, seed=0):
    """in-place r2c / c2r with the padded rows of fftw_rdft2_pad (reference fftw_api.c:774-788)"""
    rng = np.random.default_rng(seed)
    n = shape[-1]
    pad = 2 * (n // 2 + 1)
    x = rand_r(rng, shape, lib)
    a = np.zeros(shape[:-1] + (pad,), dtype=lib.real)
    a[..., :n] = x
    p = lib.plan_dft_r2c(list(shape), a, a)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    N = int(np.prod(shape))
    got = a.view(lib.cplx)
    assert rel_l2(got, orc.r2c(x)) < tol(lib, N), (shape, "r2c in place")
    # and back
    Xo = orc.r2c(x).astype(lib.cplx)
    b = np.ascontiguousarray(Xo).view(lib.real).copy()
    p = lib.plan_dft_c2r(list(shape), b, b)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    assert rel_l2(b[..., :n], orc.c2r(Xo, n)) < tol(lib, N), (shape, "c2r in place")


def check_many_strided(lib, n=12, howmany=5, istride=3, idist=40, ostride=2, odist=30, seed=1):
    """advanced interface, element stride and dist (fftw_plan_many_dft, reference fftw_api.c:642-666)"""
    rng = np.random.default_rng(seed)
    buf = rand_c(rng, howmany * idist + n * istride, lib)
    res = np.zeros(howmany * odist + n * ostride, dtype=lib.cplx)
    p = lib.plan_many_dft([n], howmany, buf, None, istride, idist, res, None, ostride, odist,
                          capi.FFTW_FORWARD)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    ref = np.zeros(res.shape, dtype=np.complex128)
    orc.dft_many([n], howmany, buf.astype(np.complex128), None, istride, idist, ref, None, ostride,
                 odist, -1)
    assert rel_l2(res, ref) < tol(lib, n)
    assert np.array_equal(res == 0, ref == 0), "gaps of a strided output must stay untouched"


def check_embedded_2d(lib, seed=2):
    rng = np.random.default_rng(seed)
    n, inembed, onembed = [4, 6], [5, 8], [4, 7]
    hm, idist, odist = 3, 5 * 8 + 3, 4 * 7 + 1
    buf = rand_c(rng, hm * idist, lib)
    res = np.zeros(hm * odist, dtype=lib.cplx)
    p = lib.plan_many_dft(n, hm, buf, inembed, 1, idist, res, onembed, 1, odist, capi.FFTW_BACKWARD)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    ref = np.zeros(res.shape, dtype=np.complex128)
    orc.dft_many(n, hm, buf.astype(np.complex128), inembed, 1, idist, ref, onembed, 1, odist, +1)
    assert rel_l2(res, ref) < tol(lib, 24)


def check_batched_transposed(lib, n=64, howmany=48, seed=3):
    """batch along the contiguous dimension: istride = howmany, idist = 1 (column transforms)"""
    rng = np.random.default_rng(seed)
    x = rand_c(rng, (n, howmany), lib)
    y = np.zeros_like(x)
    p = lib.plan_many_dft([n], howmany, x, None, howmany, 1, y, None, howmany, 1, capi.FFTW_FORWARD)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    ref = np.fft.fft(x.astype(np.complex128), axis=0)
    assert rel_l2(y, ref) < tol(lib, n)


def check_many_r2c(lib, n=32, howmany=7, seed=4):
    rng = np.random.default_rng(seed)
    x = rand_r(rng, (howmany, n), lib)
    y = np.zeros((howmany, n // 2 + 1), dtype=lib.cplx)
    p = lib.plan_many_dft_r2c([n], howmany, x, None, 1, n, y, None, 1, n // 2 + 1)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    assert rel_l2(y, np.fft.rfft(x.astype(np.float64), axis=1)) < tol(lib, n)
    z = np.zeros((howmany, n), dtype=lib.real)
    yc = y.copy()
    p = lib.plan_many_dft_c2r([n], howmany, yc, None, 1, n // 2 + 1, z, None, 1, n)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    assert rel_l2(z, n * x.astype(np.float64)) < 2 * tol(lib, n)


def check_golden(lib, golden):
    """outputs of the unmodified reference on stored inputs (tests/golden/make_golden.py)"""
    tag = "f32" if lib.single else "f64"
    from tests.golden.make_golden import C2C_SHAPES as GC, R2C_SHAPES as GR
    for shape in GC:
        key = "c2c_" + "x".join(map(str, shape)) + "_" + tag
        x = golden[key + "_in"]
        N = int(np.prod(shape))
        t = 2 * tol(lib, N)  # both sides carry working-precision rounding
        assert rel_l2(lib.dft(x, -1), golden[key + "_fwd"]) < t, key
        assert rel_l2(lib.dft(x, +1), golden[key + "_bwd"]) < t, key
    for shape in GR:
        key = "r2c_" + "x".join(map(str, shape)) + "_" + tag
        x = golden[key + "_in"]
        X = golden[key + "_out"]
        N = int(np.prod(shape))
        t = 2 * tol(lib, N)
        assert rel_l2(lib.r2c(x), X) < t, key
        assert rel_l2(lib.c2r(X, shape[-1]), golden[key + "_c2r"]) < t, key


def check_edge_cases(lib):
    x = np.zeros(8, dtype=lib.cplx)
    # invalid sizes -> NULL plan, like the reference (fftw_api.c:863-877)
    assert not lib.plan_many_dft([0], 1, x, None, 1, 1, x, None, 1, 1, -1)
    assert not lib.plan_many_dft([4], -1, x, None, 1, 4, x, None, 1, 4, -1)
    # FFTW_WISDOM_ONLY without wisdom -> NULL (fftw_api.c:332-337); with it -> a plan; wisdom survives an
    # export / forget / import round trip (fftw_api.c:895-1168)
    L = lib.lib
    L.fftw_export_wisdom_to_string.restype = C.c_void_p
    L.fftw_import_wisdom_from_string.argtypes = [C.c_char_p]
    L.fftw_forget_wisdom()
    wo = capi.FFTW_ESTIMATE | capi.FFTW_WISDOM_ONLY
    assert not lib.plan_dft_1d(8, x, x, -1, wo)
    p = lib.plan_dft_1d(8, x, x, -1)
    assert p
    lib.destroy_plan(p)
    p = lib.plan_dft_1d(8, x, x, -1, wo)
    assert p, "the problem was planned before: it is in the accumulated wisdom"
    lib.destroy_plan(p)
    assert not lib.plan_dft_1d(16, np.zeros(16, lib.cplx), np.zeros(16, lib.cplx), -1, wo)
    sp = L.fftw_export_wisdom_to_string()
    text = C.string_at(sp)
    C.CDLL(None).free(C.c_void_p(sp))
    assert text.startswith(b"(b200fft_wisdom-1 ") and text.rstrip().endswith(b")")
    L.fftw_forget_wisdom()
    assert not lib.plan_dft_1d(8, x, x, -1, wo)
    assert L.fftw_import_wisdom_from_string(text) == 1
    p = lib.plan_dft_1d(8, x, x, -1, wo)
    assert p
    lib.destroy_plan(p)
    assert L.fftw_import_wisdom_from_string(b"(fftw-3.3.8 fftw_wisdom #x0 #x0)") == 0  # not ours: rejected
    # howmany = 0 is a legal no-op
    p = lib.plan_many_dft([4], 0, x, None, 1, 4, x, None, 1, 4, -1)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    # destroy_plan(NULL) is legal (fftw_api.c:410)
    lib.destroy_plan(None)
    # rank 0 = copy
    a = np.arange(5, dtype=np.float64).astype(lib.cplx) + 1j
    b = np.zeros_like(a)
    p = lib.plan_many_dft([], 5, a, None, 1, 1, b, None, 1, 1, -1)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    assert np.array_equal(a, b)
    # execute_dft on new arrays
    rng = np.random.default_rng(5)
    u, v = rand_c(rng, 64, lib), np.zeros(64, dtype=lib.cplx)
    p = lib.plan_dft_1d(64, u, v, -1)
    u2, v2 = rand_c(rng, 64, lib), np.zeros(64, dtype=lib.cplx)
    lib.execute_dft(p, u2, v2)
    lib.destroy_plan(p)
    assert rel_l2(v2, np.fft.fft(u2.astype(np.complex128))) < tol(lib, 64)
    # a length with a prime factor > 13 goes through the Bluestein path (reference: dft-bluestein,
    # fftw_api.c:1643-1688) -- a plan, and the same answer
    w, wo = rand_c(rng, 17, lib), np.zeros(17, dtype=lib.cplx)
    p = lib.plan_dft_1d(17, w, wo, -1)
    assert p and "bluestein" in lib.sprint_plan(p)
    lib.execute(p)
    lib.destroy_plan(p)
    assert rel_l2(wo, np.fft.fft(w.astype(np.complex128))) < 4 * tol(lib, 17)


def check_properties(lib, shape, seed=7):
    """size-independent checks in the style of the reference verifier
    (fftw/libbench2/verify-lib.c:293-436): impulse, linearity, shift theorem, Parseval,
    backward(forward(x)) = N x."""
    rng = np.random.default_rng(seed)
    N = int(np.prod(shape))
    t = 4 * tol(lib, N)
    imp = np.zeros(shape, dtype=lib.cplx)
    imp.flat[0] = 1
    assert rel_l2(lib.dft(imp, -1), np.ones(shape)) < t
    a, b = rand_c(rng, shape, lib), rand_c(rng, shape, lib)
    alpha, beta = 0.75 - 0.25j, -0.5 + 1.25j
    lhs = lib.dft((alpha * a + beta * b).astype(lib.cplx), -1)
    rhs = alpha * lib.dft(a, -1).astype(np.complex128) + beta * lib.dft(b, -1).astype(np.complex128)
    assert rel_l2(lhs, rhs) < t
    A = lib.dft(a, -1)
    assert abs(np.sum(np.abs(A.astype(np.complex128)) ** 2) / (N * np.sum(np.abs(a.astype(np.complex128)) ** 2)) - 1) < t
    assert rel_l2(lib.dft(A, +1), N * a.astype(np.complex128)) < t
    # time shift along the last axis <-> phase ramp
    n = shape[-1]
    sh = np.roll(a, 1, axis=-1)
    ramp = np.exp(-2j * np.pi * np.arange(n) / n)
    assert rel_l2(lib.dft(sh, -1), A.astype(np.complex128) * ramp) < t


def check_chained_launch(lib, device=False):
    """FFTW_B200_CHAIN=1: dependent passes fused into one launch (csrc/chain_kernel.cuh); the plan
    must say so and the results must not change.  Reference equivalents: rank_geq2_apply
    fftw_api.c:4435-4445 (3-D), ct_apply_dit :2079-2090 (four-step), ct_hc2c_apply_dit_dft :5579-5590."""
    import os
    old = os.environ.get("FFTW_B200_CHAIN")
    os.environ["FFTW_B200_CHAIN"] = "1"
    os.environ["FFTW_B200_NO_TWO_PASS_REAL"] = "1"  # the three-pass real plans are the ones with something to chain
    try:
        rng = np.random.default_rng(11)
        # 3-D: z and y passes of each plane group
        x = rand_c(rng, (64, 64, 64), lib)
        y = np.zeros_like(x)
        p = lib.plan_dft(list(x.shape), x, y, capi.FFTW_FORWARD)
        assert p and "chained launch" in lib.sprint_plan(p)
        lib.execute(p)
        lib.destroy_plan(p)
        assert rel_l2(y, np.fft.fftn(x.astype(np.complex128))) < tol(lib, x.size)
        # batched four-step c2c
        n, hm = 2 ** 14, 4
        x = rand_c(rng, (hm, n), lib)
        y = np.zeros_like(x)
        p = lib.plan_many_dft([n], hm, x, None, 1, n, y, None, 1, n, capi.FFTW_BACKWARD)
        assert p and "chained launch" in lib.sprint_plan(p)
        lib.execute(p)
        lib.destroy_plan(p)
        assert rel_l2(y, n * np.fft.ifft(x.astype(np.complex128), axis=1)) < tol(lib, n)
        # batched long r2c (two passes + Hermitian split) and c2r (merge + two passes)
        n, hm = 2 ** 15, 6
        xr = rand_r(rng, (hm, n), lib)
        yc = np.zeros((hm, n // 2 + 1), dtype=lib.cplx)
        p = lib.plan_many_dft_r2c([n], hm, xr, None, 1, n, yc, None, 1, n // 2 + 1)
        assert p and "chained launch" in lib.sprint_plan(p)
        lib.execute(p)
        lib.destroy_plan(p)
        assert rel_l2(yc, np.fft.rfft(xr.astype(np.float64), axis=1)) < tol(lib, n)
        z = np.zeros_like(xr)
        p = lib.plan_many_dft_c2r([n], hm, yc, None, 1, n // 2 + 1, z, None, 1, n)
        assert p and "chained launch" in lib.sprint_plan(p)
        lib.execute(p)
        lib.destroy_plan(p)
        assert rel_l2(z, n * xr.astype(np.float64)) < 2 * tol(lib, n)
    finally:
        os.environ.pop("FFTW_B200_NO_TWO_PASS_REAL", None)
        if old is None:
            os.environ.pop("FFTW_B200_CHAIN", None)
        else:
            os.environ["FFTW_B200_CHAIN"] = old


def check_two_pass_c2r(lib):
    """FFTW_B200_TWO_PASS_C2R=1: the mirror image of the two-pass r2c (unfold + column pass, pair merge
    + row pass); off by default because the three-pass plan is faster at 2^20, kept parity-tested."""
    import os
    os.environ["FFTW_B200_TWO_PASS_C2R"] = "1"
    try:
        rng = np.random.default_rng(5)
        for n, hm in ((2 ** 15, 3), (2 ** 17, 1)):
            x = rand_r(rng, (hm, n), lib)
            X = np.fft.rfft(x.astype(np.float64), axis=1).astype(lib.cplx)
            z = np.zeros_like(x)
            p = lib.plan_many_dft_c2r([n], hm, X, None, 1, n // 2 + 1, z, None, 1, n)
            assert p and "real-pair merge" in lib.sprint_plan(p)
            lib.execute(p)
            lib.destroy_plan(p)
            assert rel_l2(z, n * x.astype(np.float64)) < 2 * tol(lib, n)
    finally:
        os.environ.pop("FFTW_B200_TWO_PASS_C2R", None)


FUZZ_SIZES = [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 13, 15, 16, 17, 20, 23, 24, 25, 27, 30, 32, 36, 48, 49, 60, 64, 77,
              96, 100, 121, 127, 128, 130, 169, 200, 243, 256, 289, 384, 500, 512, 1000, 1024, 1536, 2048, 2187, 4096]


def check_fuzz(lib, ncases, seed, max_points=1 << 17):
    """Random geometries through the advanced interface (fftw_plan_many_dft / _r2c / _c2r, reference
    fftw_api.c:642-736): rank 1-3, sizes with and without large prime factors, random howmany, element
    strides, embeddings and dists, in place and out of place, both signs -- against numpy's pocketfft in
    double precision on the gathered data, and the gaps of strided outputs must stay untouched."""
    rng = np.random.default_rng(seed)
    done = 0
    while done < ncases:
        rank = int(rng.integers(1, 4))
        n = [int(rng.choice(FUZZ_SIZES)) for _ in range(rank)]
        hm = int(rng.integers(1, 6))
        if int(np.prod(n)) * hm > max_points:
            continue
        kind = rng.choice(["c2c", "r2c", "c2r"])
        inplace = bool(rng.integers(0, 2))
        N = int(np.prod(n))
        nc = n[:-1] + [n[-1] // 2 + 1]          # complex-side logical dims of a real transform
        if kind == "c2c":
            li, lo = n, n
        elif kind == "r2c":
            li, lo = n, nc
        else:
            li, lo = nc, n
        def embed(logical):
            e = list(logical)
            for d in range(1, rank):
                e[d] += int(rng.integers(0, 3))
            return e
        if inplace:
            if kind == "c2c":
                ie = oe = embed(n)
                istride = ostride = int(rng.integers(1, 3))
            else:  # padded rows: the real side has 2*(n/2+1) reals per row
                ce = embed(nc)
                re = ce[:-1] + [2 * ce[-1]]
                ie, oe = (re, ce) if kind == "r2c" else (ce, re)
                istride = ostride = 1
        else:
            ie, oe = embed(li), embed(lo)
            istride, ostride = int(rng.integers(1, 3)), int(rng.integers(1, 3))
        isz, osz = int(np.prod(ie)) * istride, int(np.prod(oe)) * ostride
        if inplace and kind != "c2c":
            idist = isz + (2 if kind == "r2c" else 1) * int(rng.integers(0, 2))
            odist = idist // 2 if kind == "r2c" else idist * 2
            if kind == "r2c" and idist % 2:
                continue
        elif inplace:
            idist = odist = isz + int(rng.integers(0, 4))
        else:
            idist, odist = isz + int(rng.integers(0, 4)), osz + int(rng.integers(0, 4))
        it = lib.real if kind == "r2c" else lib.cplx
        ot = lib.real if kind == "c2r" else lib.cplx
        ilen, olen = hm * idist + isz, hm * odist + osz

        def gather(buf, emb, logical, stride, dist):
            out = np.empty([hm] + list(logical), dtype=buf.dtype)
            idx = np.indices(logical).reshape(rank, -1)
            flat = np.zeros(idx.shape[1], dtype=np.int64)
            for d in range(rank):
                flat = flat * emb[d] + idx[d]
            for k in range(hm):
                out[k] = buf[k * dist + flat * stride].reshape(logical)
            return out, flat

        # input data
        if kind == "c2r":
            xr = rng.random([hm] + n) - 0.5
            spec = np.fft.rfftn(xr, axes=tuple(range(1, rank + 1)))
            src = np.zeros(ilen, dtype=np.complex128)
            _, flat = gather(src, ie, li, istride, idist)
            for k in range(hm):
                src[k * idist + flat * istride] = spec[k].reshape(-1)
            inbuf = src.astype(it)
        else:
            inbuf = (rand_r(rng, ilen, lib) if kind == "r2c" else rand_c(rng, ilen, lib))
        sign = int(rng.choice([-1, 1])) if kind == "c2c" else (-1 if kind == "r2c" else 1)
        if inplace:
            nbytes = max(inbuf.nbytes, olen * np.dtype(ot).itemsize)
            raw = np.zeros(nbytes + 16, dtype=np.uint8)
            off = (-raw.ctypes.data) % 16
            raw = raw[off:off + nbytes]
            raw[:inbuf.nbytes] = inbuf.view(np.uint8)
            a_in = raw[:inbuf.nbytes].view(it)
            a_out = raw[:olen * np.dtype(ot).itemsize].view(ot)
        else:
            a_in, a_out = inbuf.copy(), np.zeros(olen, dtype=ot)
        x_log, _ = gather(a_in.copy(), ie, li, istride, idist)
        args = (n, hm, a_in, ie, istride, idist, a_out, oe, ostride, odist)
        if kind == "c2c":
            p = lib.plan_many_dft(*args, sign)
        elif kind == "r2c":
            p = lib.plan_many_dft_r2c(*args)
        else:
            p = lib.plan_many_dft_c2r(*args)
        assert p, (kind, n, hm, ie, oe, istride, ostride, idist, odist, inplace)
        lib.execute(p)
        lib.destroy_plan(p)
        y_log, flat_o = gather(a_out, oe, lo, ostride, odist)
        axes = tuple(range(1, rank + 1))
        if kind == "c2c":
            ref = np.fft.fftn(x_log.astype(np.complex128), axes=axes) if sign < 0 else \
                N * np.fft.ifftn(x_log.astype(np.complex128), axes=axes)
        elif kind == "r2c":
            ref = np.fft.rfftn(x_log.astype(np.float64), axes=axes)
        else:
            ref = N * xr
        err = rel_l2(y_log, ref)
        assert err < 4 * tol(lib, max(N, 2)), (kind, n, hm, ie, oe, istride, ostride, idist, odist, inplace, err)
        if not inplace:
            mask = np.ones(olen, dtype=bool)
            for k in range(hm):
                mask[k * odist + flat_o * ostride] = False
            assert not a_out[mask].any(), "gaps of a strided output must stay untouched"
        done += 1


def check_split_arrays(lib, device=False):
    """guru split-array interfaces (reference fftw_api.c:1420-1517): c2c on separate re / im arrays
    (forward; backward by swapping the arrays), r2c with a split output, c2r with a split input;
    strided and batched loop nests; new-array execute."""
    rng = np.random.default_rng(11)
    if device:
        import torch
        up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        down = lambda t: t.cpu().numpy()
    else:
        up = lambda a: np.ascontiguousarray(a)
        down = lambda a: a
    # --- c2c, 2-D 12 x 16 with a batch of 3, arrays [batch][12][16], then backward by swapping
    shape = (3, 12, 16)
    x = rand_c(rng, shape, lib)
    ri, ii = up(x.real.astype(lib.real)), up(x.imag.astype(lib.real))
    ro, io = up(np.zeros(shape, lib.real)), up(np.zeros(shape, lib.real))
    dims = [(12, 16, 16), (16, 1, 1)]
    hd = [(3, 192, 192)]
    p = lib.plan_guru_split("c2c", dims, hd, ri, ii, ro, io)
    assert p, "split c2c plan is NULL"
    assert "split-array" in lib.sprint_plan(p)
    lib.execute(p)
    ref = np.stack([orc.dft(x[b], -1) for b in range(3)])
    got = down(ro) + 1j * down(io)
    assert rel_l2(got, ref) < tol(lib, 192)
    # backward through the same plan: swap real and imaginary arrays on both sides
    br, bi = up(np.zeros(shape, lib.real)), up(np.zeros(shape, lib.real))
    lib.execute_split("c2c", p, io, ro, bi, br)
    back = down(br) + 1j * down(bi)
    assert rel_l2(back / 192, x) < tol(lib, 192)
    lib.destroy_plan(p)
    # --- c2c 1-D, strided (every 3rd real), batch with distance, output transposed ([k][batch])
    n, hm = 64, 5
    rbuf, ibuf = rand_r(rng, (hm, 3 * n), lib), rand_r(rng, (hm, 3 * n), lib)
    robuf, iobuf = np.full((n, hm), 7, lib.real), np.full((n, hm), 7, lib.real)
    dri, dii, dro, dio = up(rbuf), up(ibuf), up(robuf), up(iobuf)
    p = lib.plan_guru_split("c2c", [(n, 3, hm)], [(hm, 3 * n, 1)], dri, dii, dro, dio)
    assert p
    lib.execute(p)
    xs = rbuf[:, ::3] + 1j * ibuf[:, ::3]
    ref = np.stack([orc.dft(xs[b].astype(lib.cplx), -1) for b in range(hm)], axis=1)
    assert rel_l2(down(dro) + 1j * down(dio), ref) < tol(lib, n)
    lib.destroy_plan(p)
    # --- r2c with a split output and c2r back, 2-D 10 x 18, batch 2
    shape = (2, 10, 18)
    hc = 18 // 2 + 1
    xr = rand_r(rng, shape, lib)
    din = up(xr)
    ro, io = up(np.zeros((2, 10, hc), lib.real)), up(np.zeros((2, 10, hc), lib.real))
    p = lib.plan_guru_split("r2c", [(10, 18, hc), (18, 1, 1)], [(2, 180, 10 * hc)], din, ro, io)
    assert p, "split r2c plan is NULL"
    lib.execute(p)
    ref = np.stack([orc.r2c(xr[b]) for b in range(2)])
    got = down(ro) + 1j * down(io)
    assert rel_l2(got, ref) < tol(lib, 180)
    lib.destroy_plan(p)
    dout = up(np.zeros(shape, lib.real))
    p = lib.plan_guru_split("c2r", [(10, hc, 18), (18, 1, 1)], [(2, 10 * hc, 180)], ro, io, dout)
    assert p, "split c2r plan is NULL"
    lib.execute(p)
    assert rel_l2(down(dout) / 180, xr) < tol(lib, 180)
    lib.destroy_plan(p)


def check_split_interleaved_host(lib):
    """the documented split usage on ONE interleaved host array: ri = in, ii = in + 1, ro = out, io = out + 1,
    all strides 2 (reference fftw_api.c:1420-1450).  The two output ranges overlap in memory: the write-back
    of one must not bring back stale values of the other.  Also the backward form (pointers swapped)."""
    rng = np.random.default_rng(17)
    n, hm = 48, 3
    x = rand_c(rng, (hm, n), lib)
    xin = np.ascontiguousarray(x).view(lib.real).reshape(hm, 2 * n).copy()
    out = np.full((hm, 2 * n), 9, lib.real)
    ri, ii = xin.reshape(-1)[0:], xin.reshape(-1)[1:]
    ro, io = out.reshape(-1)[0:], out.reshape(-1)[1:]
    p = lib.plan_guru_split("c2c", [(n, 2, 2)], [(hm, 2 * n, 2 * n)], ri, ii, ro, io)
    assert p, "split c2c plan on interleaved arrays is NULL"
    lib.execute(p)
    ref = np.stack([orc.dft(x[b], -1) for b in range(hm)])
    got = out.view(lib.cplx).reshape(hm, n)
    assert rel_l2(got, ref) < tol(lib, n), rel_l2(got, ref)
    # backward: swap real and imaginary pointers on both sides
    back = np.full((hm, 2 * n), 5, lib.real)
    lib.execute_split("c2c", p, io, ro, back.reshape(-1)[1:], back.reshape(-1)[0:])
    assert rel_l2(back.view(lib.cplx).reshape(hm, n) / n, x) < tol(lib, n)
    lib.destroy_plan(p)


def check_negative_strides(lib, device=False):
    """guru c2c with negative strides (the reference's tensors carry signed strides, fftw_api.h:472-487; guru entry
    fftw_api.c:1172-1517): the pointers address element 0, which is then not the lowest address of the array.
    Reversed input, reversed output rows, a batch walked backwards on both sides, in place with equal strides."""
    rng = np.random.default_rng(31)
    esz = np.dtype(lib.cplx).itemsize
    if device:
        import torch
        up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        down = lambda t: t.cpu().numpy()
        addr = lambda t, off: t.data_ptr() + off * esz
    else:
        up = lambda a: np.array(a, copy=True, order="C")  # (a copy: case 3 transforms in place)
        down = lambda a: a
        addr = lambda a, off: a.ctypes.data + off * esz
    # 1. batched 1-D, input read backwards (is = -1), output forwards
    n, hm = 48, 3
    x = rand_c(rng, (hm, n), lib)
    dx, dy = up(x), up(np.zeros((hm, n), lib.cplx))
    p = lib.plan_guru_dft([(n, -1, 1)], [(hm, n, n)], addr(dx, n - 1), addr(dy, 0), -1)
    assert p, "plan with a negative input stride is NULL"
    lib.execute(p)
    lib.destroy_plan(p)
    ref = np.stack([orc.dft(np.ascontiguousarray(x[b][::-1]), -1) for b in range(hm)])
    assert rel_l2(down(dy), ref) < tol(lib, n)
    # 2. 2-D 6 x 20, output rows stored in reverse order (os0 < 0), batch of 4 walked backwards on both sides
    n0, n1, hm = 6, 20, 4
    x = rand_c(rng, (hm, n0, n1), lib)
    dx, dy = up(x), up(np.full((hm, n0, n1), 3, lib.cplx))
    p = lib.plan_guru_dft([(n0, n1, -n1), (n1, 1, 1)], [(hm, -n0 * n1, -n0 * n1)],
                          addr(dx, (hm - 1) * n0 * n1), addr(dy, (hm - 1) * n0 * n1 + (n0 - 1) * n1), -1)
    assert p, "plan with negative output / batch strides is NULL"
    lib.execute(p)
    lib.destroy_plan(p)
    got = down(dy)
    for b in range(hm):  # batch element k sits at block hm-1-k on both sides; its output row r at row n0-1-r
        ref = orc.dft(np.ascontiguousarray(x[hm - 1 - b]), -1)
        assert rel_l2(got[hm - 1 - b][::-1], ref) < tol(lib, n0 * n1), b
    # 3. in place, the same negative stride on both sides (4096 points: two passes with a work array on the GPU)
    n = 4096 * 4
    x = rand_c(rng, (n,), lib)
    dz = up(x)
    p = lib.plan_guru_dft([(n, -1, -1)], [], addr(dz, n - 1), addr(dz, n - 1), +1)
    assert p, "in-place plan with negative strides is NULL"
    lib.execute(p)
    lib.destroy_plan(p)
    assert rel_l2(down(dz)[::-1], orc.dft(np.ascontiguousarray(x[::-1]), +1)) < tol(lib, n)
    # in place with DIFFERENT signs cannot address the same locations: NULL, as in the reference
    dz = up(x)
    assert not lib.plan_guru_dft([(n, -1, 1)], [], addr(dz, n - 1), addr(dz, n - 1), -1)


def check_measure_and_wisdom(lib, device=False):
    """FFTW_MEASURE (flags = 0) times the candidate configurations of the selector and keeps the fastest (reference:
    patience loop fftw_api.c:351-376, cost by time :15106-15153); the winner goes into wisdom, is exported with the
    problem's signature and is replayed, unmeasured, by later plans for the same problem (also FFTW_WISDOM_ONLY
    after forget + import).  A 3-D transform whose middle axis has 1024 points has at least one alternative."""
    import ctypes as C
    L = lib.lib
    L.fftw_export_wisdom_to_string.restype = C.c_void_p
    L.fftw_import_wisdom_from_string.argtypes = [C.c_char_p]
    L.fftw_forget_wisdom()
    rng = np.random.default_rng(41)
    shape = (3, 1024, 32)
    x = rand_c(rng, shape, lib)
    if device:
        import torch
        dx = torch.from_numpy(x).cuda()
        dy = torch.empty_like(dx)
        get = lambda: dy.cpu().numpy()
    else:
        dx, dy = x.copy(), np.zeros(shape, lib.cplx)
        get = lambda: dy
    ref = orc.dft(x, -1)
    p = lib.plan_dft(list(shape), dx, dy, -1, capi.FFTW_MEASURE)
    assert p, "FFTW_MEASURE plan is NULL"
    d1 = lib.sprint_plan(p)
    assert "measured (FFTW_MEASURE):" in d1 and "default:" in d1 and " ms]" in d1, d1
    if device:
        dx.copy_(torch.from_numpy(x))  # (planning with MEASURE may overwrite the arrays in the reference; here it does not)
    lib.execute(p)
    assert rel_l2(get(), ref) < tol(lib, int(np.prod(shape)))
    lib.destroy_plan(p)
    sp = L.fftw_export_wisdom_to_string()
    text = C.string_at(sp)
    lib.free_string(sp) if hasattr(lib, "free_string") else None
    assert b":" in text, text  # "<signature>:<configuration>" for the measured problem
    # the same problem again, FFTW_ESTIMATE: the remembered winner is replayed, nothing is timed
    p = lib.plan_dft(list(shape), dx, dy, -1, capi.FFTW_ESTIMATE)
    d2 = lib.sprint_plan(p)
    assert "measured" not in d2
    assert [l for l in d2.split("\n") if "pass " in l] == [l for l in d1.split("\n") if "pass " in l]
    lib.destroy_plan(p)
    # forget, import, WISDOM_ONLY
    L.fftw_forget_wisdom()
    assert not lib.plan_dft(list(shape), dx, dy, -1, capi.FFTW_ESTIMATE | capi.FFTW_WISDOM_ONLY)
    assert L.fftw_import_wisdom_from_string(text) == 1
    p = lib.plan_dft(list(shape), dx, dy, -1, capi.FFTW_ESTIMATE | capi.FFTW_WISDOM_ONLY)
    assert p, "imported wisdom must answer FFTW_WISDOM_ONLY"
    assert [l for l in lib.sprint_plan(p).split("\n") if "pass " in l] == [l for l in d1.split("\n") if "pass " in l]
    lib.execute(p)
    assert rel_l2(get(), ref) < tol(lib, int(np.prod(shape)))
    lib.destroy_plan(p)
    # a geometry without alternatives still plans under FFTW_MEASURE
    y = np.zeros(64, lib.cplx)
    p = lib.plan_dft([64], y.copy(), y, -1, capi.FFTW_MEASURE)
    assert p and "no alternative configuration" in lib.sprint_plan(p)
    lib.destroy_plan(p)
    L.fftw_forget_wisdom()


R2R_SIZES = [2, 3, 4, 5, 8, 9, 16, 30, 64, 100, 255, 1024]


def check_r2r(lib, r2r_golden, device=False):
    """every fftw_r2r_kind (fftw3.h:117-121): against the reference's own outputs (golden, fp64 inputs),
    against the oracle at sizes the golden file does not hold, multi-dimensional mixes, batched /
    strided `many` form, in place, new-array execute."""
    rng = np.random.default_rng(23)
    if device:
        import torch
        up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        down = lambda t: t.cpu().numpy()
    else:
        up = lambda a: np.ascontiguousarray(a)
        down = lambda a: a

    def run(x, kinds):
        xi = up(np.ascontiguousarray(x, dtype=lib.real))
        yo = up(np.zeros(x.shape, lib.real))
        p = lib.plan_r2r(list(x.shape), xi, yo, kinds)
        assert p, ("r2r plan is NULL", x.shape, kinds)
        lib.execute(p)
        lib.destroy_plan(p)
        return down(yo)

    scale = 40.0  # the zero-stuffed 8n extension adds 3 bits to log2 N; sums of 2n terms
    for name, kd in orc.R2R_KINDS.items():
        for n in R2R_SIZES:
            if name == "REDFT00" and n < 2:
                continue
            key = f"{name}_{n}_in"
            if key in r2r_golden.files:
                x, ref = r2r_golden[key], r2r_golden[f"{name}_{n}_out"]
            else:
                x = rng.random(n) - 0.5
                ref = orc.r2r_1d(x, kd)
            got = run(x, [kd])
            assert rel_l2(got, ref) < scale * tol(lib, 8 * n), (name, n, rel_l2(got, ref))
    for key, kinds in (("2d_REDFT10xRODFT00", [5, 7]), ("3d_DHTxREDFT00xR2HC", [2, 3, 0])):
        x, ref = r2r_golden[key + "_in"], r2r_golden[key + "_out"]
        assert rel_l2(run(x, kinds), ref) < scale * tol(lib, x.size * 8)
    # round trips: REDFT10 <-> REDFT01 and R2HC <-> HC2R are inverses up to 2n and n
    x = rng.random((3, 48)) - 0.5
    a = run(x[0], [orc.R2R_KINDS["REDFT10"]])
    assert rel_l2(run(a, [orc.R2R_KINDS["REDFT01"]]) / (2 * 48), x[0]) < scale * tol(lib, 8 * 48)
    h = run(x[1], [orc.R2R_KINDS["R2HC"]])
    assert rel_l2(run(h, [orc.R2R_KINDS["HC2R"]]) / 48, x[1]) < scale * tol(lib, 48)
    # many: 5 transforms of 12 x 10 (RODFT10 x REDFT11) embedded in 13 x 11 with distance, in place
    n, emb, hm = [12, 10], [13, 11], 5
    dist = 13 * 11 + 4
    buf = (rng.random(hm * dist) - 0.5).astype(lib.real)
    ref = buf.astype(np.float64).copy()
    for b in range(hm):
        blk = ref[b * dist:b * dist + 13 * 11].reshape(13, 11)
        blk[:12, :10] = orc.r2r(blk[:12, :10].copy(), [9, 6])
    d = up(buf)
    p = lib.plan_many_r2r(n, hm, d, emb, 1, dist, d, emb, 1, dist, [9, 6])
    assert p, "many_r2r plan is NULL"
    lib.execute(p)
    assert rel_l2(down(d), ref) < scale * tol(lib, 8 * 120)
    # new-array execute on the same plan
    buf2 = (rng.random(hm * dist) - 0.5).astype(lib.real)
    ref2 = buf2.astype(np.float64).copy()
    for b in range(hm):
        blk = ref2[b * dist:b * dist + 13 * 11].reshape(13, 11)
        blk[:12, :10] = orc.r2r(blk[:12, :10].copy(), [9, 6])
    d2 = up(buf2)
    lib.execute_r2r(p, d2, d2)
    assert rel_l2(down(d2), ref2) < scale * tol(lib, 8 * 120)
    lib.destroy_plan(p)


def check_threads(lib):
    """fftw_execute_dft from several threads on ONE plan with different arrays, and plan creation /
    destruction from several threads at once (reference contract: execute is re-entrant,
    fftw_api.c:428-454; the planner needs fftw_make_planner_thread_safe, threads/api.c:83-86)."""
    import threading
    lib.lib.fftw_make_planner_thread_safe()
    rng = np.random.default_rng(5)
    n, hm = 1 << 14, 3  # two kernel passes and a plan-owned scratch array
    x0 = rand_c(rng, (hm, n), lib)
    p = lib.plan_many_dft([n], hm, x0, None, 1, n, x0.copy(), None, 1, n, -1)
    assert p
    errs = []

    def work(seed):
        try:
            r = np.random.default_rng(seed)
            for it in range(4):
                x = rand_c(r, (hm, n), lib)
                y = np.empty_like(x)
                lib.execute_dft(p, x, y)
                ref = orc.dft_batched(x, -1)
                e = rel_l2(y, ref)
                if not e < tol(lib, n):
                    errs.append(("execute", seed, it, e))
                m = 64 << (seed % 3)
                a = rand_c(r, (m,), lib)
                q = lib.plan_dft_1d(m, a, a.copy(), -1)
                if not q:
                    errs.append(("plan", seed, it))
                lib.destroy_plan(q)
        except Exception as ex:  # pragma: no cover
            errs.append(("exception", seed, repr(ex)))

    th = [threading.Thread(target=work, args=(s,)) for s in range(6)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    lib.destroy_plan(p)
    assert not errs, errs


def check_inplace_relayout(lib, device=False):
    """in place with different input and output strides that address the same locations (reference:
    fftw_tensor_inplace_locations fftw_api.c:17298; e.g. a batch written back transposed); different
    locations -> NULL"""
    rng = np.random.default_rng(3)
    if device:
        import torch
        up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        down = lambda t: t.cpu().numpy()
    else:
        up = lambda a: np.ascontiguousarray(a)
        down = lambda a: a
    for n, hm in [(16, 5), (64, 48), (100, 7), (4096, 3), (2 ** 14, 2), (17, 4)]:
        x = rand_c(rng, (hm, n), lib)
        buf = up(x.copy().ravel())
        p = lib.plan_many_dft([n], hm, buf, None, 1, n, buf, None, hm, 1, capi.FFTW_FORWARD)
        assert p, (n, hm)
        lib.execute(p)
        lib.destroy_plan(p)
        assert rel_l2(down(buf).reshape(n, hm), orc.dft_batched(x, -1).T) < tol(lib, n), (n, hm)
    x = rand_c(rng, (12, 20), lib)
    buf = up(x.copy().ravel())
    p = lib.plan_guru_dft([(12, 20, 1), (20, 1, 12)], [], buf, buf, capi.FFTW_BACKWARD)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    assert rel_l2(down(buf).reshape(20, 12), orc.dft(x, +1).T) < tol(lib, 240)
    y = up(np.zeros(64, dtype=lib.cplx))
    assert not lib.plan_many_dft([8], 2, y, None, 1, 8, y, None, 2, 20, capi.FFTW_FORWARD)


# lengths with a mixed-radix tile kernel (csrc/mixtile.cu): (n, has the column forms)
MIX_LENGTHS = [(96, True), (192, True), (384, True), (768, True), (1536, True), (3072, False), (100, True), (200, True),
               (250, True), (400, True), (500, True), (1000, True), (2000, False), (3000, False), (4000, False),
               (81, True), (243, True), (729, True), (2187, False), (125, True), (625, False), (3125, False),
               (49, True), (343, True), (2401, False), (1080, False)]
MIX_LENGTHS_FP32_ONLY = [(5000, False), (6144, False), (6561, False)]


def check_mixed_radix(lib, n, cols, seed=0):
    """Compile-time mixed-radix tile kernels for lengths that are not powers of two: every form the selector can pick
    (row; column, column + four-step twiddle and row -> column for the lengths that have them), full and ragged
    tiles, both directions, against numpy fp64.  The plan text must name the kernel, so a silent fall back to the
    generic kernel fails the test.  Reference: the n1_* / t1_* codelets for radices 3 5 6 7 9 10 12 15 20
    (fftw/dft_scalar/codelets/codlist.c:57-109)."""
    rng = np.random.default_rng(seed + n)
    # ---- row form: batched contiguous lines, a batch that is not a multiple of any tile width
    hm = 67 if n <= 512 else (9 if n <= 2048 else 3)
    x = rand_c(rng, (hm, n), lib)
    for sign in (capi.FFTW_FORWARD, capi.FFTW_BACKWARD):
        y = np.zeros_like(x)
        p = lib.plan_many_dft([n], hm, x, None, 1, n, y, None, 1, n, sign)
        assert p
        txt = lib.sprint_plan(p)
        assert "mixed radix" in txt and "tile-kernel[row" in txt, txt
        lib.execute(p)
        lib.destroy_plan(p)
        ref = np.fft.fft(x.astype(np.complex128), axis=1) if sign < 0 else n * np.fft.ifft(x.astype(np.complex128), axis=1)
        assert rel_l2(y, ref) < tol(lib, n), (n, "row", sign)
    if not cols or n > (2048 if lib.single else 1024):  # (the column cap, csrc/selector.h caps())
        return
    # ---- column form: the first axis of a 2-D array (ragged: 20 columns; full: 64)
    for w in (20, 64):
        x = rand_c(rng, (n, w), lib)
        y = np.zeros_like(x)
        p = lib.plan_many_dft([n], w, x, None, w, 1, y, None, w, 1, capi.FFTW_FORWARD)
        assert p
        txt = lib.sprint_plan(p)
        assert "mixed radix" in txt and "tile-kernel[col" in txt, txt
        lib.execute(p)
        lib.destroy_plan(p)
        assert rel_l2(y, np.fft.fft(x.astype(np.complex128), axis=0)) < tol(lib, n), (n, "col", w)
    # ---- inside an N-D transform (column pass in place after the row pass), backward
    x = rand_c(rng, (n, 24), lib)
    y = lib.dft(x, capi.FFTW_BACKWARD)
    assert rel_l2(y, x.size * np.fft.ifft2(x.astype(np.complex128))) < tol(lib, x.size), (n, "2-D")


def check_mixed_radix_long(lib):
    """Long 1-D lengths whose two passes are mixed-radix kernels: column + four-step twiddle, then row -> column."""
    import os
    rng = np.random.default_rng(5)
    old = os.environ.get("FFTW_B200_ROW_CAP")
    os.environ["FFTW_B200_ROW_CAP"] = "1024"  # make the two-pass split reachable at test sizes
    try:
        for n, hm in ((100000, 2), (24576, 3), (1000 * 96, 1), (243 * 250, 2), (400 * 400, 1), (3 * 2 ** 17, 1)):
            x = rand_c(rng, (hm, n), lib)
            for sign in (capi.FFTW_FORWARD, capi.FFTW_BACKWARD):
                y = np.zeros_like(x)
                p = lib.plan_many_dft([n], hm, x, None, 1, n, y, None, 1, n, sign)
                assert p
                txt = lib.sprint_plan(p)
                assert "mixed radix" in txt, txt
                lib.execute(p)
                lib.destroy_plan(p)
                xd = x.astype(np.complex128)
                ref = np.fft.fft(xd, axis=1) if sign < 0 else n * np.fft.ifft(xd, axis=1)
                assert rel_l2(y, ref) < tol(lib, n), (n, sign, txt)
    finally:
        if old is None:
            os.environ.pop("FFTW_B200_ROW_CAP", None)
        else:
            os.environ["FFTW_B200_ROW_CAP"] = old


# lengths with a second kernel shape in the table (csrc/mixtile.cu, FFTW_B200_MIX_ALT=1 selects it)
MIX_SECOND_SHAPES = [(250, True), (500, True), (1000, True), (243, True), (729, True), (625, False), (3125, False)]
MIX_SECOND_SHAPES_FP32_ONLY = [(5000, False), (6561, False)]


def check_mixed_radix_second_shapes(lib, n, cols):
    import os
    os.environ["FFTW_B200_MIX_ALT"] = "1"
    try:
        check_mixed_radix(lib, n, cols, seed=3)
    finally:
        os.environ.pop("FFTW_B200_MIX_ALT", None)


# lengths N whose row shape has an even E: r2c / c2r of 2N reals run on the mixed-radix kernel with the Hermitian split / merge fused
MIX_REAL_HALF = [96, 192, 384, 768, 1536, 3072, 100, 200, 250, 400, 500, 1000, 2000, 3000, 4000, 1080]
MIX_REAL_HALF_FP32_ONLY = [5000, 6144]


def check_mixed_radix_real(lib, N, seed=0):
    """r2c / c2r rows of n = 2N reals for N with a mixed-radix row kernel (csrc/mixtile.cu B2_MIX_REAL): ragged batch, out of
    place and in place (padded rows), and as the last axis of a 2-D real transform; against numpy fp64.  Reference: rdft2 via
    hc2c (ct_hc2c_direct_apply fftw/fftw_api.c:5831-5845)."""
    rng = np.random.default_rng(seed + N)
    n = 2 * N
    hm = 37 if N <= 512 else (7 if N <= 2048 else 3)
    x = rand_r(rng, (hm, n), lib)
    X = np.zeros((hm, N + 1), dtype=lib.cplx)
    p = lib.plan_many_dft_r2c([n], hm, x, None, 1, n, X, None, 1, N + 1)
    assert p
    txt = lib.sprint_plan(p)
    assert "mixed radix" in txt and "r2c" in txt, txt
    lib.execute(p)
    lib.destroy_plan(p)
    ref = np.fft.rfft(x.astype(np.float64), axis=1)
    assert rel_l2(X, ref) < tol(lib, n), (N, "r2c")
    y = np.zeros_like(x)
    Xc = X.copy()
    p = lib.plan_many_dft_c2r([n], hm, Xc, None, 1, N + 1, y, None, 1, n)
    assert p
    txt = lib.sprint_plan(p)
    assert "mixed radix" in txt and "c2r" in txt, txt
    lib.execute(p)
    lib.destroy_plan(p)
    assert rel_l2(y, n * x.astype(np.float64)) < 2 * tol(lib, n), (N, "c2r")
    # in place, rows padded to 2(N+1) reals
    buf = np.zeros((hm, 2 * (N + 1)), dtype=lib.real)
    buf[:, :n] = x
    p = lib.plan_many_dft_r2c([n], hm, buf, None, 1, 2 * (N + 1), buf, None, 1, N + 1)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    assert rel_l2(buf.view(lib.cplx), ref) < tol(lib, n), (N, "r2c in place")
    # last axis of a 2-D real transform
    if N <= 1536:
        x2 = rand_r(rng, (12, n), lib)
        Y = lib.r2c(x2)
        assert rel_l2(Y, np.fft.rfft2(x2.astype(np.float64))) < tol(lib, x2.size), (N, "2-D r2c")


# second list (csrc/mixtile2.cu): row forms only; (N, False) in the format of MIX_LENGTHS
MIX_MORE = [(144, False), (288, False), (576, False), (1152, False), (2304, False), (1296, False), (1728, False), (216, False), (432, False), (864, False), (320, False), (640, False), (1280, False), (800, False), (1600, False), (3200, False), (162, False), (486, False), (1458, False), (150, False), (300, False), (450, False), (600, False), (750, False), (900, False), (1500, False), (1800, False), (2700, False), (1250, False), (2500, False), (360, False), (540, False), (720, False), (196, False), (392, False), (686, False), (1372, False), (2744, False), (448, False), (896, False), (1792, False), (441, False), (121, False), (1331, False), (169, False), (2197, False)]
MIX_MORE_FP32_ONLY = [(4608, False), (8000, False), (4500, False), (6000, False), (7500, False), (6250, False)]
MIX_MORE_REAL_HALF = [144, 288, 576, 1152, 2304, 1296, 1728, 216, 432, 864, 320, 640, 1280, 800, 1600, 3200, 300, 600, 900, 1500, 1800, 1250, 2500, 360, 720]

# 60 points per thread, fp32 library only (csrc/mixtile2.cu): video widths and multiples of 240; complex rows and r2c / c2r rows
MIX_E60 = [240, 480, 960, 1920, 3840, 1200, 2400, 1440, 2880, 3600, 4800, 5760, 7200, 7680]

# second-list lengths with column forms (csrc/mixtile3.cu); in double precision only those within the column cap of 1024
MIX_MORE_COLS = [144, 288, 576, 1152, 320, 640, 1280, 800, 1600, 300, 600, 900, 1500, 1800, 360, 720, 1080, 216, 432, 864, 1296, 1728]


def check_mixed_radix_frames(lib):
    """N-D transforms whose axes all have mixed-radix kernels: rows on the row form, the other axes on the column forms
    (a 1080 x 1920 frame in single precision, a 144 x 96 x 100 box, a 720 x 1280 real frame), and long 1-D products of the
    third list's lengths."""
    import os
    rng = np.random.default_rng(9)
    shapes = [(144, 96, 100), (360, 600)] + ([(1080, 1920)] if lib.single else [(720, 1000)])
    for shape in shapes:
        x = rand_c(rng, shape, lib)
        y = np.zeros_like(x)
        p = lib.plan_dft(list(shape), x, y, capi.FFTW_FORWARD)
        assert p
        txt = lib.sprint_plan(p)
        passes = [l for l in txt.split("\n") if l.strip().startswith("pass ")]
        assert len(passes) == len(shape) and all("mixed radix" in l for l in passes), txt
        lib.execute(p)
        lib.destroy_plan(p)
        assert rel_l2(y, np.fft.fftn(x.astype(np.complex128))) < tol(lib, x.size), shape
    # real frame: r2c rows of 1280 reals (N = 640), then 720-point columns over the 641 half-spectrum columns
    xr = rand_r(rng, (720, 1280), lib)
    Y = lib.r2c(xr)
    assert rel_l2(Y, np.fft.rfft2(xr.astype(np.float64))) < tol(lib, xr.size)
    old = os.environ.get("FFTW_B200_ROW_CAP")
    os.environ["FFTW_B200_ROW_CAP"] = "1024"
    try:
        for n in (144 * 144, 720 * 720, 600 * 360):
            x = rand_c(rng, (2, n), lib)
            y = np.zeros_like(x)
            p = lib.plan_many_dft([n], 2, x, None, 1, n, y, None, 1, n, capi.FFTW_BACKWARD)
            assert p
            txt = lib.sprint_plan(p)
            passes = [l for l in txt.split("\n") if l.strip().startswith("pass ")]
            assert len(passes) == 2 and all("mixed radix" in l for l in passes), txt
            lib.execute(p)
            lib.destroy_plan(p)
            assert rel_l2(y, n * np.fft.ifft(x.astype(np.complex128), axis=1)) < tol(lib, n), n
    finally:
        if old is None:
            os.environ.pop("FFTW_B200_ROW_CAP", None)
        else:
            os.environ["FFTW_B200_ROW_CAP"] = old


def check_negative_strides_real(lib, device=False):
    """guru r2c / c2r with negative strides, where the engine serves them (one real step + complex leading passes): the real
    input read backwards, the batch walked backwards on both sides, output rows of a 2-D r2c stored in reverse order, c2r
    writing its real output backwards; and the documented refusals (a length that needs several passes)."""
    rng = np.random.default_rng(37)
    rsz, csz = np.dtype(lib.real).itemsize, np.dtype(lib.cplx).itemsize
    if device:
        import torch
        up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        down = lambda t: t.cpu().numpy()
        addr = lambda t, off, sz: t.data_ptr() + off * sz
    else:
        up = lambda a: np.array(a, copy=True, order="C")
        down = lambda a: a
        addr = lambda a, off, sz: a.ctypes.data + off * sz
    # 1. batched 1-D r2c: input read backwards (is = -1), batch walked backwards on the output side
    n, hm = 60, 4
    hc = n // 2 + 1
    x = rand_r(rng, (hm, n), lib)
    dx, dy = up(x), up(np.zeros((hm, hc), lib.cplx))
    p = lib.plan_guru_real("r2c", [(n, -1, 1)], [(hm, n, -hc)], addr(dx, n - 1, rsz), addr(dy, (hm - 1) * hc, csz))
    assert p, "r2c plan with negative strides is NULL"
    lib.execute(p)
    lib.destroy_plan(p)
    got = down(dy)
    for b in range(hm):
        ref = np.fft.rfft(x[b][::-1].astype(np.float64))
        assert rel_l2(got[hm - 1 - b], ref) < tol(lib, n), b
    # 2. 2-D r2c 6 x 20, output rows in reverse order (os0 < 0)
    n0, n1 = 6, 20
    h1 = n1 // 2 + 1
    x = rand_r(rng, (n0, n1), lib)
    dx, dy = up(x), up(np.zeros((n0, h1), lib.cplx))
    p = lib.plan_guru_real("r2c", [(n0, n1, -h1), (n1, 1, 1)], [], addr(dx, 0, rsz), addr(dy, (n0 - 1) * h1, csz))
    assert p, "2-D r2c plan with a negative output stride is NULL"
    lib.execute(p)
    lib.destroy_plan(p)
    assert rel_l2(down(dy)[::-1], np.fft.rfft2(x.astype(np.float64))) < tol(lib, n0 * n1)
    # 3. batched c2r: real output written backwards (os = -1), input batch walked backwards
    n, hm = 64, 3
    hc = n // 2 + 1
    x = rand_r(rng, (hm, n), lib)
    X = np.fft.rfft(x.astype(np.float64), axis=1).astype(lib.cplx)
    dX, dz = up(X), up(np.zeros((hm, n), lib.real))
    p = lib.plan_guru_real("c2r", [(n, 1, -1)], [(hm, -hc, n)], addr(dX, (hm - 1) * hc, csz), addr(dz, n - 1, rsz))
    assert p, "c2r plan with negative strides is NULL"
    lib.execute(p)
    lib.destroy_plan(p)
    got = down(dz)
    for b in range(hm):  # batch element b reads X[hm-1-b] and writes row b backwards
        assert rel_l2(got[b][::-1], n * x[hm - 1 - b].astype(np.float64)) < 2 * tol(lib, n), b
    # refused: a real length that needs more than one pass with a negative stride
    big = 1 << 16
    dx, dy = up(np.zeros(big, lib.real)), up(np.zeros(big // 2 + 1, lib.cplx))
    assert not lib.plan_guru_real("r2c", [(big, -1, 1)], [], addr(dx, big - 1, rsz), addr(dy, 0, csz))
