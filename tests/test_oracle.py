"""Pins the CPU oracle (oracle/fft_oracle.c) to the unmodified reference.

Golden vectors = outputs of the reference itself (tests/golden/make_golden.py).
Tolerance: the oracle computes in float64; against the reference's float64 build the
bound is the north-star fp64 tolerance 1e-14*log2(N) (relative L2); against the float32
build it is 1e-6*log2(N).
"""
import math

import numpy as np
import pytest

from oracle import pyoracle as orc
from tests.golden.make_golden import C2C_SHAPES, R2C_SHAPES


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.complex128).ravel()
    b = np.asarray(b, dtype=np.complex128).ravel()
    den = np.linalg.norm(b)
    return np.linalg.norm(a - b) / (den if den > 0 else 1.0)


def tol(shape, tag):
    N = int(np.prod(shape))
    return (1e-6 if tag == "f32" else 1e-14) * max(1.0, math.log2(N))


@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("shape", C2C_SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_c2c_matches_reference(golden, shape, tag):
    key = "c2c_" + "x".join(map(str, shape)) + "_" + tag
    x = golden[key + "_in"]
    assert rel_l2(orc.dft(x, -1), golden[key + "_fwd"]) < tol(shape, tag)
    assert rel_l2(orc.dft(x, +1), golden[key + "_bwd"]) < tol(shape, tag)


@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("shape", R2C_SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_r2c_c2r_match_reference(golden, shape, tag):
    key = "r2c_" + "x".join(map(str, shape)) + "_" + tag
    x = golden[key + "_in"]
    X = golden[key + "_out"]
    assert rel_l2(orc.r2c(x), X) < tol(shape, tag)
    assert rel_l2(orc.c2r(X, shape[-1]), golden[key + "_c2r"]) < tol(shape, tag)


@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_many_strided_matches_reference(golden, tag):
    n, howmany, istride, idist, ostride, odist = (int(v) for v in golden[f"many_{tag}_geom"])
    buf = golden[f"many_{tag}_in"].astype(np.complex128)
    ref = golden[f"many_{tag}_out"]
    out = np.zeros(ref.shape, dtype=np.complex128)
    orc.dft_many([n], howmany, buf, None, istride, idist, out, None, ostride, odist, -1)
    assert rel_l2(out, ref) < tol((n,), tag)
    # untouched gaps stay zero in both
    assert np.array_equal(out == 0, ref == 0)


@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_embedded_2d_matches_reference(golden, tag):
    buf = golden[f"embed_{tag}_in"].astype(np.complex128)
    ref = golden[f"embed_{tag}_out"]
    out = np.zeros(ref.shape, dtype=np.complex128)
    orc.dft_many([4, 6], 3, buf, [5, 8], 1, 5 * 8 + 3, out, [4, 7], 1, 4 * 7 + 1, +1)
    assert rel_l2(out, ref) < tol((4, 6), tag)


@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_r2c_inplace_padded_layout(golden, tag):
    a = golden[f"r2c_inplace_{tag}_in"].astype(np.float64)
    ref = golden[f"r2c_inplace_{tag}_out"]
    rows, pad = a.shape
    n = 10
    out = np.zeros((rows, pad // 2), dtype=np.complex128)
    orc.r2c_many([rows, n], 1, a, [rows, pad], 1, 1, out, [rows, pad // 2], 1, 1)
    refc = ref.astype(np.float64).view(np.complex128)
    assert rel_l2(out, refc) < tol((rows, n), tag)


def test_edge_cases():
    # howmany = 0 is legal and a no-op (fftw_many_kosherp, fftw_api.c:875)
    x = np.zeros(4, dtype=np.complex128)
    orc.dft_many([4], 0, x, None, 1, 4, x, None, 1, 4, -1)
    # invalid sizes are rejected like the reference's NULL plan (fftw_api.c:863-877)
    with pytest.raises(ValueError):
        orc.dft_many([0], 1, x, None, 1, 1, x, None, 1, 1, -1)
    with pytest.raises(ValueError):
        orc.dft_many([4], -1, x, None, 1, 1, x, None, 1, 1, -1)
    # in place is allowed
    y = np.arange(8, dtype=np.complex128)
    ref = np.fft.fft(y)
    orc.dft_many([8], 1, y, None, 1, 8, y, None, 1, 8, -1)
    assert rel_l2(y, ref) < 1e-15 * 8
    # unnormalised round trip: backward(forward(x)) = N x  (fftw3.h:469-470 convention)
    z = np.random.default_rng(0).random(30) + 0j
    assert rel_l2(orc.dft(orc.dft(z, -1), +1), 30 * z) < 1e-14


def test_impulse_and_constant_closed_forms():
    # the closed forms the reference verifier uses (fftw/libbench2/verify-lib.c:293-323)
    n = 48
    imp = np.zeros(n, dtype=np.complex128)
    imp[0] = 1
    assert rel_l2(orc.dft(imp), np.ones(n)) < 1e-15
    const = np.ones(n, dtype=np.complex128)
    out = np.zeros(n)
    out[0] = n
    assert rel_l2(orc.dft(const), out) < 1e-14


@pytest.mark.parametrize("P", [1, 2, 3, 4, 8])
def test_mpi_slab_layout(P):
    # block distribution + transposed-output layout (fftw/mpi/block.c:39-50, doc/mpi.texi:470-509)
    n0, n1, n2 = 8, 6, 4
    rng = np.random.default_rng(P)
    x = rng.random((n0, n1, n2)) - 0.5 + 1j * (rng.random((n0, n1, n2)) - 0.5)
    full = np.fft.fftn(x)
    seen0 = seen1 = 0
    for p in range(P):
        N, ln0, ls0, ln1, ls1 = orc.mpi_local_size_3d_transposed(n0, n1, n2, P, p)
        b0, b1 = -(-n0 // P), -(-n1 // P)
        assert ln0 == max(0, min(b0, n0 - p * b0)) and ln1 == max(0, min(b1, n1 - p * b1))
        assert N >= max(ln0 * n1 * n2, ln1 * n0 * n2, 1)
        seen0 += ln0
        seen1 += ln1
        t = orc.mpi_dft_3d_local(x, P, p, transposed_out=True)
        assert t.shape == (ln1, n0, n2)
        if ln1:
            assert rel_l2(t, full[:, ls1:ls1 + ln1, :].transpose(1, 0, 2)) < 1e-14
        u = orc.mpi_dft_3d_local(x, P, p, transposed_out=False)
        if ln0:
            assert rel_l2(u, full[ls0:ls0 + ln0]) < 1e-14
        N2, a, b = orc.mpi_local_size_3d(n0, n1, n2, P, p)
        assert (a, b) == (ln0, ls0) and N2 >= ln0 * n1 * n2
    assert seen0 == n0 and seen1 == n1


def test_r2r_restatement_matches_the_reference(r2r_golden):
    """oracle.r2r_1d / r2r (direct sums from the manual's definitions) against outputs of the reference"""
    g = r2r_golden
    for name, kd in orc.R2R_KINDS.items():
        for n in (2, 3, 4, 5, 8, 9, 16, 30, 64):
            x, y = g[f"{name}_{n}_in"], g[f"{name}_{n}_out"]
            assert np.linalg.norm(orc.r2r_1d(x, kd) - y) / np.linalg.norm(y) < 1e-13, (name, n)
    assert np.linalg.norm(orc.r2r(g["2d_REDFT10xRODFT00_in"], [5, 7]) - g["2d_REDFT10xRODFT00_out"]) < 1e-12
    assert np.linalg.norm(orc.r2r(g["3d_DHTxREDFT00xR2HC_in"], [2, 3, 0]) - g["3d_DHTxREDFT00xR2HC_out"]) < 1e-12
