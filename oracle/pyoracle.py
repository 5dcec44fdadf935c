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
