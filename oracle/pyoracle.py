"""Python loader for the CPU oracle (oracle/fft_oracle.c) and the reference build.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  Never imported by fftw3_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "liboracle.so")
REF_DIR = os.path.join(HERE, "_ref")
REFERENCE_ROOT = "/root/reference"

_ip = C.POINTER(C.c_int)
_pp = C.POINTER(C.c_ssize_t)


def build(force: bool = False) -> str:
    """gcc the C restatement into oracle/_build/liboracle.so (a few hundred ms)."""
    src = os.path.join(HERE, "fft_oracle.c")
    hdr = os.path.join(HERE, "fft_oracle.h")
    if (not force and os.path.exists(LIB)
            and os.path.getmtime(LIB) >= max(os.path.getmtime(src), os.path.getmtime(hdr))):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-o", LIB, src, "-lm"])
    return LIB


def build_reference(jobs: int = 8) -> bool:
    """Compile the unmodified reference into oracle/_ref/ when /root/reference exists.
    Returns True when both precision builds are present afterwards."""
    have = all(os.path.exists(os.path.join(REF_DIR, f"libfftw3_ref_{p}.so")) for p in ("f64", "f32"))
    if have or not os.path.isdir(REFERENCE_ROOT):
        return have
    for prec in ("f64", "f32"):
        subprocess.check_call(["make", "-s", f"-j{jobs}", f"PREC={prec}", f"REF={REFERENCE_ROOT}"],
                              cwd=os.path.join(HERE, "refbuild"))
    return True


def build_verifier(jobs: int = 8) -> bool:
    """Link the reference's own bench/verifier CLI (unmodified objects compiled from
    /root/reference) against the B200 engine and against the test-only emulation build:
    oracle/_ref/refbench_{b200,emu}_{f32,f64}.  Needs those libraries to exist already."""
    names = [f"refbench_{e}_{p}" for e in ("b200", "emu") for p in ("f32", "f64")]
    if not os.path.isdir(REFERENCE_ROOT):
        return all(os.path.exists(os.path.join(REF_DIR, n)) for n in names)
    for prec in ("f64", "f32"):
        subprocess.check_call(["make", "-s", f"-j{jobs}", f"PREC={prec}", f"REF={REFERENCE_ROOT}", "verifier"],
                              cwd=os.path.join(HERE, "refbuild"))
    return True


def reference_lib(single: bool):
    """FFTW3Lib bound to the unmodified reference build, or None when it is absent."""
    from fftw3_b200.capi import FFTW3Lib  # generic fftw3.h ctypes binding (no compute)
    path = os.path.join(REF_DIR, f"libfftw3_ref_{'f32' if single else 'f64'}.so")
    if not os.path.exists(path):
        return None
    return FFTW3Lib(path, single)


_lib = None


def _load():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        many = [C.c_int, _ip, C.c_int, C.c_void_p, _ip, C.c_int, C.c_int,
                C.c_void_p, _ip, C.c_int, C.c_int]
        L.oracle_dft_many.argtypes = many + [C.c_int]
        L.oracle_r2c_many.argtypes = many
        L.oracle_c2r_many.argtypes = many
        L.oracle_mpi_local_size_3d.restype = C.c_ssize_t
        L.oracle_mpi_local_size_3d.argtypes = [C.c_ssize_t] * 3 + [C.c_int, C.c_int, _pp, _pp]
        L.oracle_mpi_local_size_3d_transposed.restype = C.c_ssize_t
        L.oracle_mpi_local_size_3d_transposed.argtypes = ([C.c_ssize_t] * 3 + [C.c_int, C.c_int]
                                                          + [_pp] * 4)
        L.oracle_mpi_dft_3d_local.argtypes = ([C.c_ssize_t] * 3 + [C.c_int] * 4
                                              + [C.c_void_p, C.c_void_p])
        _lib = L
    return _lib


def _ints(seq):
    return None if seq is None else (C.c_int * len(seq))(*[int(v) for v in seq])


def dft_many(n, howmany, inp, inembed, istride, idist, out, onembed, ostride, odist, sign):
    """Raw strided call; inp/out are complex128 numpy buffers (may be the same array)."""
    assert inp.dtype == np.complex128 and out.dtype == np.complex128
    rc = _load().oracle_dft_many(len(n), _ints(n), howmany, inp.ctypes.data, _ints(inembed),
                                 istride, idist, out.ctypes.data, _ints(onembed), ostride,
                                 odist, sign)
    if rc:
        raise ValueError("oracle_dft_many rejected the geometry")
    return out


def dft(x, sign=-1):
    """Out-of-place DFT over all axes of x (any complex/real dtype), computed in float64."""
    x = np.ascontiguousarray(x, dtype=np.complex128)
    y = np.empty_like(x)
    return dft_many(list(x.shape), 1, x, None, 1, 1, y, None, 1, 1, sign)


def dft_batched(x, sign=-1):
    """1-D DFT along the last axis of a 2-D array [howmany, n]."""
    x = np.ascontiguousarray(x, dtype=np.complex128)
    y = np.empty_like(x)
    h, n = x.shape
    return dft_many([n], h, x, None, 1, n, y, None, 1, n, sign)


def r2c_many(n, howmany, inp, inembed, istride, idist, out, onembed, ostride, odist):
    assert inp.dtype == np.float64 and out.dtype == np.complex128
    rc = _load().oracle_r2c_many(len(n), _ints(n), howmany, inp.ctypes.data, _ints(inembed),
                                 istride, idist, out.ctypes.data, _ints(onembed), ostride, odist)
    if rc:
        raise ValueError("oracle_r2c_many rejected the geometry")
    return out


def c2r_many(n, howmany, inp, inembed, istride, idist, out, onembed, ostride, odist):
    assert inp.dtype == np.complex128 and out.dtype == np.float64
    rc = _load().oracle_c2r_many(len(n), _ints(n), howmany, inp.ctypes.data, _ints(inembed),
                                 istride, idist, out.ctypes.data, _ints(onembed), ostride, odist)
    if rc:
        raise ValueError("oracle_c2r_many rejected the geometry")
    return out


def r2c(x):
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.empty(x.shape[:-1] + (x.shape[-1] // 2 + 1,), dtype=np.complex128)
    return r2c_many(list(x.shape), 1, x, None, 1, 1, y, None, 1, 1)


def c2r(X, n_last):
    X = np.ascontiguousarray(X, dtype=np.complex128)
    shape = X.shape[:-1] + (n_last,)
    y = np.empty(shape, dtype=np.float64)
    return c2r_many(list(shape), 1, X, None, 1, 1, y, None, 1, 1)


def mpi_local_size_3d(n0, n1, n2, n_pes, my_pe):
    a, b = C.c_ssize_t(), C.c_ssize_t()
    N = _load().oracle_mpi_local_size_3d(n0, n1, n2, n_pes, my_pe, C.byref(a), C.byref(b))
    return N, a.value, b.value


def mpi_local_size_3d_transposed(n0, n1, n2, n_pes, my_pe):
    a, b, c, d = (C.c_ssize_t() for _ in range(4))
    N = _load().oracle_mpi_local_size_3d_transposed(n0, n1, n2, n_pes, my_pe, C.byref(a),
                                                    C.byref(b), C.byref(c), C.byref(d))
    return N, a.value, b.value, c.value, d.value


def mpi_dft_3d_local(x_global, n_pes, my_pe, transposed_out=True, sign=-1):
    x = np.ascontiguousarray(x_global, dtype=np.complex128)
    n0, n1, n2 = x.shape
    _, ln0, _, ln1, _ = mpi_local_size_3d_transposed(n0, n1, n2, n_pes, my_pe)
    shape = (ln1, n0, n2) if transposed_out else (ln0, n1, n2)
    out = np.empty(shape, dtype=np.complex128)
    rc = _load().oracle_mpi_dft_3d_local(n0, n1, n2, n_pes, my_pe, int(transposed_out), sign,
                                         x.ctypes.data, out.ctypes.data)
    if rc:
        raise ValueError("oracle_mpi_dft_3d_local failed")
    return out


# ---------------------------------------------------------------------------------------------------
# real-to-real kinds (TEST INFRASTRUCTURE like the rest of this module): direct O(n^2) restatement of
# the definitions the reference implements (fftw/doc/reference.texi "1d Real-even DFTs (DCTs)" and
# "1d Real-odd DFTs (DSTs)"; solvers fftw/fftw_api.c:11700-14160, kinds fftw/fftw3.h:117-121).  Pinned
# against outputs of the reference itself: tests/golden/r2r_golden.npz (tests/golden/make_golden_r2r.py).
R2R_KINDS = {"R2HC": 0, "HC2R": 1, "DHT": 2, "REDFT00": 3, "REDFT01": 4, "REDFT10": 5, "REDFT11": 6,
             "RODFT00": 7, "RODFT01": 8, "RODFT10": 9, "RODFT11": 10}


def r2r_1d(x, kind):
    """one-dimensional r2r transform of the last axis, float64"""
    x = np.asarray(x, dtype=np.float64)
    n = x.shape[-1]
    j = np.arange(n)[None, :]
    k = np.arange(n)[:, None]
    pi = np.pi
    if kind == 0:  # R2HC: r0, r1, ..., r_{n/2}, i_{(n+1)/2-1}, ..., i_1
        F = np.fft.fft(x, axis=-1)
        out = np.empty_like(x)
        out[..., :n // 2 + 1] = F.real[..., :n // 2 + 1]
        for kk in range(1, (n + 1) // 2):
            out[..., n - kk] = F.imag[..., kk]
        return out
    if kind == 1:  # HC2R: unnormalised inverse of R2HC
        F = np.zeros(x.shape[:-1] + (n // 2 + 1,), dtype=np.complex128)
        F[..., :n // 2 + 1] = x[..., :n // 2 + 1]
        for kk in range(1, (n + 1) // 2):
            F[..., kk] += 1j * x[..., n - kk]
        return np.fft.irfft(F, n=n, axis=-1) * n
    if kind == 2:  # DHT
        F = np.fft.fft(x, axis=-1)
        return F.real - F.imag
    if kind == 3:  # REDFT00
        M = 2 * np.cos(pi * j * k / (n - 1))
        M[:, 0] = 1
        M[:, n - 1] = (-1.0) ** np.arange(n)
    elif kind == 5:  # REDFT10
        M = 2 * np.cos(pi * (j + 0.5) * k / n)
    elif kind == 4:  # REDFT01
        M = 2 * np.cos(pi * j * (k + 0.5) / n)
        M[:, 0] = 1
    elif kind == 6:  # REDFT11
        M = 2 * np.cos(pi * (j + 0.5) * (k + 0.5) / n)
    elif kind == 7:  # RODFT00
        M = 2 * np.sin(pi * (j + 1) * (k + 1) / (n + 1))
    elif kind == 9:  # RODFT10
        M = 2 * np.sin(pi * (j + 0.5) * (k + 1) / n)
    elif kind == 8:  # RODFT01
        M = 2 * np.sin(pi * (j + 1) * (k + 0.5) / n)
        M[:, n - 1] = (-1.0) ** np.arange(n)
    elif kind == 10:  # RODFT11
        M = 2 * np.sin(pi * (j + 0.5) * (k + 0.5) / n)
    else:
        raise ValueError(kind)
    return x @ M.T


def r2r(x, kinds):
    """multi-dimensional: kinds[i] along axis -(len(kinds)) + i (the transform dims are the last ones)"""
    y = np.asarray(x, dtype=np.float64)
    r = len(kinds)
    for i, kd in enumerate(kinds):
        ax = y.ndim - r + i
        y = np.moveaxis(r2r_1d(np.moveaxis(y, ax, -1), kd), -1, ax)
    return y
