"""Slab-distributed 3-D c2c (the fftw_mpi_plan_dft_3d form): geometry, message layout and the
two local FFT stages, checked on the CPU.

The product exchanges over NCCL, which needs GPUs; here the ranks run the TEST-ONLY emulation
build with "detached" communicators (rank, size only) and the test moves the messages itself:
in one process for P = 1..8 (ragged block sizes, empty ranks), and between two real processes
over torch.distributed/gloo.  The expected result is the reference's own definition of the
distributed transform: the serial DFT, redistributed (fftw/mpi/mpi-bench.c:236-257;
oracle_mpi_dft_3d_local)."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

from fftw3_b200 import build, capi
from oracle import pyoracle as orc

TRANSPOSED_OUT = 1 << 30


def load(prec="fp64"):
    lib = capi.FFTW3Lib(build.build_emu(prec), prec == "fp32")
    L = lib.lib
    L.b2fft_comm_detached.restype = C.c_void_p
    L.b2fft_comm_detached.argtypes = [C.c_int, C.c_int]
    L.b2fft_comm_destroy.argtypes = [C.c_void_p]
    L.fftw_mpi_plan_dft_3d.restype = C.c_void_p
    L.fftw_mpi_plan_dft_3d.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] * 3 + [C.c_int, C.c_uint]
    L.fftw_mpi_local_size_3d_transposed.restype = C.c_ssize_t
    L.fftw_mpi_local_size_3d_transposed.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] + [C.POINTER(C.c_ssize_t)] * 4
    L.b2fft_slab_stage.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    L.b2fft_slab_buffer.restype = C.c_void_p
    L.b2fft_slab_buffer.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_size_t)]
    return lib


def local_sizes(lib, shape, comm):
    v = [C.c_ssize_t() for _ in range(4)]
    N = lib.lib.fftw_mpi_local_size_3d_transposed(*shape, comm, *[C.byref(x) for x in v])
    return (N,) + tuple(x.value for x in v)


def buffer_view(lib, plan, which, dtype):
    nb = C.c_size_t()
    bp = C.cast(plan, C.POINTER(C.c_void_p))[0]
    ptr = lib.lib.b2fft_slab_buffer(bp, which, C.byref(nb))
    return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_byte)), shape=(nb.value,)).view(dtype), bp


@pytest.mark.parametrize("P", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("shape", [(8, 6, 4), (10, 6, 5), (16, 16, 16)], ids=lambda s: "x".join(map(str, s)))
def test_local_sizes_match_the_reference_rule(P, shape):
    lib = load()
    for pe in range(P):
        comm = lib.lib.b2fft_comm_detached(P, pe)
        assert local_sizes(lib, shape, comm) == orc.mpi_local_size_3d_transposed(*shape, P, pe)
        lib.lib.b2fft_comm_destroy(comm)


@pytest.mark.parametrize("transposed", [True, False], ids=["transposed_out", "natural_out"])
@pytest.mark.parametrize("P", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("shape", [(8, 6, 4), (10, 6, 5), (32, 16, 8)], ids=lambda s: "x".join(map(str, s)))
def test_ranks_simulated_in_one_process(shape, P, transposed):
    lib = load()
    n0, n1, n2 = shape
    rng = np.random.default_rng(P)
    x = ((rng.random(shape) - 0.5) + 1j * (rng.random(shape) - 0.5)).astype(lib.cplx)
    b0, b1 = -(-n0 // P), -(-n1 // P)
    ranks = []
    for pe in range(P):
        comm = lib.lib.b2fft_comm_detached(P, pe)
        N, ln0, s0, ln1, s1 = local_sizes(lib, shape, comm)
        xin = np.zeros(max(N, 1), dtype=lib.cplx)
        xin[:ln0 * n1 * n2] = x[s0:s0 + ln0].ravel()
        out = np.zeros(max(N, 1), dtype=lib.cplx)
        plan = lib.lib.fftw_mpi_plan_dft_3d(n0, n1, n2, xin.ctypes.data, out.ctypes.data, comm, -1,
                                            capi.FFTW_ESTIMATE | (TRANSPOSED_OUT if transposed else 0))
        assert plan, "slab plan is NULL"
        S, bp = buffer_view(lib, plan, 0, lib.cplx)
        R, _ = buffer_view(lib, plan, 1, lib.cplx)
        ranks.append(dict(comm=comm, ln0=ln0, s0=s0, ln1=ln1, s1=s1, xin=xin, out=out, plan=plan, bp=bp, S=S, R=R))
    for r in ranks:
        assert lib.lib.b2fft_slab_stage(r["bp"], 0, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    # the exchange, spelled out from the documented layout: S = [ln0][n1][n2], R = [n0][ln1][n2]
    for src in ranks:
        Sv = src["S"][:src["ln0"] * n1 * n2].reshape(src["ln0"], n1, n2)
        for dst in ranks:
            if dst["ln1"] == 0 or src["ln0"] == 0:
                continue
            Rv = dst["R"][:n0 * dst["ln1"] * n2].reshape(n0, dst["ln1"], n2)
            Rv[src["s0"]:src["s0"] + src["ln0"]] = Sv[:, dst["s1"]:dst["s1"] + dst["ln1"], :]
    for pe, r in enumerate(ranks):
        assert lib.lib.b2fft_slab_stage(r["bp"], 2, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    if not transposed:  # second exchange, back to [ln0][n1][n2]
        for src in ranks:
            if src["ln1"] == 0:
                continue
            Rv = src["R"][:n0 * src["ln1"] * n2].reshape(n0, src["ln1"], n2)
            for dst in ranks:
                if dst["ln0"] == 0:
                    continue
                Ov = dst["out"][:dst["ln0"] * n1 * n2].reshape(dst["ln0"], n1, n2)
                Ov[:, src["s1"]:src["s1"] + src["ln1"], :] = Rv[dst["s0"]:dst["s0"] + dst["ln0"]]
    for pe, r in enumerate(ranks):
        ref = orc.mpi_dft_3d_local(x, P, pe, transposed_out=transposed)
        got = r["out"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13, (pe, P)
        lib.destroy_plan(r["plan"])
        lib.lib.b2fft_comm_destroy(r["comm"])


@pytest.mark.parametrize("prec", ["fp64", "fp32"])
@pytest.mark.parametrize("P", [1, 2, 3, 8])
@pytest.mark.parametrize("shape", [(12, 32, 6), (9, 64, 16), (20, 128, 3)], ids=lambda s: "x".join(map(str, s)))
def test_fused_exchange_scatters_into_the_owners_receive_buffers(shape, P, prec):
    """Stage 0 with the exchange fused into the n1 pass (tile kernel EPI 8): every rank stores row k1 of
    its planes straight into the receive buffer of rank k1 / block -- here the 'peer memory' is the
    other simulated ranks' buffers in this process.  No message is moved by the test."""
    lib = load(prec)
    lib.lib.b2fft_slab_set_peer_recv.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_int]
    n0, n1, n2 = shape
    rng = np.random.default_rng(P + n1)
    x = ((rng.random(shape) - 0.5) + 1j * (rng.random(shape) - 0.5)).astype(lib.cplx)
    ranks = []
    for pe in range(P):
        comm = lib.lib.b2fft_comm_detached(P, pe)
        N, ln0, s0, ln1, s1 = local_sizes(lib, shape, comm)
        xin = np.zeros(max(N, 1), dtype=lib.cplx)
        xin[:ln0 * n1 * n2] = x[s0:s0 + ln0].ravel()
        out = np.zeros(max(N, 1), dtype=lib.cplx)
        plan = lib.lib.fftw_mpi_plan_dft_3d(n0, n1, n2, xin.ctypes.data, out.ctypes.data, comm, -1,
                                            capi.FFTW_ESTIMATE | TRANSPOSED_OUT)
        assert plan
        R, bp = buffer_view(lib, plan, 1, lib.cplx)
        ranks.append(dict(comm=comm, xin=xin, out=out, plan=plan, bp=bp, R=R))
    recv = (C.c_void_p * P)(*[r["R"].ctypes.data for r in ranks])
    for r in ranks:
        assert lib.lib.b2fft_slab_set_peer_recv(r["bp"], recv, P) == 1, "fused path not taken"
        assert "fused into the n1 pass" in lib.sprint_plan(r["plan"])
    for r in ranks:
        assert lib.lib.b2fft_slab_stage(r["bp"], 0, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
        assert lib.lib.b2fft_slab_stage(r["bp"], 1, r["xin"].ctypes.data, r["out"].ctypes.data) == 0  # no-op
    tol = (1e-6 if lib.single else 1e-14) * np.log2(n0 * n1 * n2)
    for pe, r in enumerate(ranks):
        assert lib.lib.b2fft_slab_stage(r["bp"], 2, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
        ref = orc.mpi_dft_3d_local(x, P, pe, transposed_out=True)
        got = r["out"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < tol, (pe, P)
    for r in ranks:
        lib.destroy_plan(r["plan"])
        lib.lib.b2fft_comm_destroy(r["comm"])


TRANSPOSED_IN = 1 << 29


def _exchange(ranks, n_a, n_b, rl):
    """S = [la][n_b][rl] on the sender, R = [n_a][lb][rl] on the receiver (blocks over a, b)"""
    for src in ranks:
        if src["la"] == 0:
            continue
        Sv = src["S"][:src["la"] * n_b * rl].reshape(src["la"], n_b, rl)
        for dst in ranks:
            if dst["lb"] == 0:
                continue
            Rv = dst["R"][:n_a * dst["lb"] * rl].reshape(n_a, dst["lb"], rl)
            Rv[src["sa"]:src["sa"] + src["la"]] = Sv[:, dst["sb"]:dst["sb"] + dst["lb"], :]


@pytest.mark.parametrize("flags", [TRANSPOSED_IN, TRANSPOSED_IN | TRANSPOSED_OUT], ids=["tin", "tin_tout"])
@pytest.mark.parametrize("P", [1, 2, 3])
@pytest.mark.parametrize("shape", [(8, 6, 4), (10, 7, 5), (12, 32, 6)], ids=lambda s: "x".join(map(str, s)))
def test_transposed_input(shape, P, flags):
    """FFTW_MPI_TRANSPOSED_IN (fftw/doc/mpi.texi:470-509): the input is [local_n1][n0][n2], distributed
    over n1; the output is natural ([local_n0][n1][n2]) or, with TRANSPOSED_OUT too, stays transposed."""
    lib = load()
    n0, n1, n2 = shape
    rng = np.random.default_rng(17 + P)
    x = ((rng.random(shape) - 0.5) + 1j * (rng.random(shape) - 0.5)).astype(lib.cplx)
    xT = np.ascontiguousarray(x.transpose(1, 0, 2))
    Y = orc.dft(x, -1)
    tout = bool(flags & TRANSPOSED_OUT)
    ranks = []
    for pe in range(P):
        comm = lib.lib.b2fft_comm_detached(P, pe)
        N, ln0, s0, ln1, s1 = local_sizes(lib, shape, comm)
        xin = np.zeros(max(N, 1), dtype=lib.cplx)
        xin[:ln1 * n0 * n2] = xT[s1:s1 + ln1].ravel()
        out = np.zeros(max(N, 1), dtype=lib.cplx)
        plan = lib.lib.fftw_mpi_plan_dft_3d(n0, n1, n2, xin.ctypes.data, out.ctypes.data, comm, -1,
                                            capi.FFTW_ESTIMATE | flags)
        assert plan, "slab plan is NULL"
        S, bp = buffer_view(lib, plan, 0, lib.cplx)
        R, _ = buffer_view(lib, plan, 1, lib.cplx)
        # inside the plan the first two dims are exchanged: its "a" axis is n1, its "b" axis n0
        ranks.append(dict(comm=comm, la=ln1, sa=s1, lb=ln0, sb=s0, xin=xin, out=out, plan=plan, bp=bp, S=S, R=R))
    for r in ranks:
        assert lib.lib.b2fft_slab_stage(r["bp"], 0, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    if "fused into" not in lib.sprint_plan(ranks[0]["plan"]):
        _exchange(ranks, n1, n0, n2)
    for r in ranks:
        assert lib.lib.b2fft_slab_stage(r["bp"], 2, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    if tout:  # the plan's natural output = this problem's transposed output: second exchange, back
        for src in ranks:
            if src["lb"] == 0:
                continue
            Rv = src["R"][:n1 * src["lb"] * n2].reshape(n1, src["lb"], n2)
            for dst in ranks:
                if dst["la"] == 0:
                    continue
                Ov = dst["out"][:dst["la"] * n0 * n2].reshape(dst["la"], n0, n2)
                Ov[:, src["sb"]:src["sb"] + src["lb"], :] = Rv[dst["sa"]:dst["sa"] + dst["la"]]
    for r in ranks:
        ref = Y.transpose(1, 0, 2)[r["sa"]:r["sa"] + r["la"]] if tout else Y[r["sb"]:r["sb"] + r["lb"]]
        got = r["out"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13
        lib.destroy_plan(r["plan"])
        lib.lib.b2fft_comm_destroy(r["comm"])


@pytest.mark.parametrize("P", [1, 2, 3])
@pytest.mark.parametrize("howmany", [2, 3])
@pytest.mark.parametrize("shape", [(8, 6, 4), (9, 32, 5)], ids=lambda s: "x".join(map(str, s)))
def test_many_interleaved_transforms(shape, howmany, P):
    """fftw_mpi_plan_many_dft with howmany > 1: every element is a tuple of `howmany` complex numbers
    (array n0 x n1 x n2 x howmany, last index contiguous; fftw/doc/mpi.texi 'Advanced distributed-transpose
    interface' / mpi/api.c:592-636)."""
    lib = load()
    L = lib.lib
    L.fftw_mpi_plan_many_dft.restype = C.c_void_p
    L.fftw_mpi_plan_many_dft.argtypes = [C.c_int, C.POINTER(C.c_ssize_t), C.c_ssize_t, C.c_ssize_t, C.c_ssize_t,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint]
    L.fftw_mpi_local_size_many_transposed.restype = C.c_ssize_t
    L.fftw_mpi_local_size_many_transposed.argtypes = [C.c_int, C.POINTER(C.c_ssize_t), C.c_ssize_t, C.c_ssize_t,
                                                      C.c_ssize_t, C.c_void_p] + [C.POINTER(C.c_ssize_t)] * 4
    n0, n1, n2 = shape
    rng = np.random.default_rng(howmany)
    x = ((rng.random(shape + (howmany,)) - 0.5) + 1j * (rng.random(shape + (howmany,)) - 0.5)).astype(lib.cplx)
    Y = np.stack([orc.dft(np.ascontiguousarray(x[..., h]), -1) for h in range(howmany)], axis=-1)
    nn = (C.c_ssize_t * 3)(*shape)
    rl = n2 * howmany
    ranks = []
    for pe in range(P):
        comm = L.b2fft_comm_detached(P, pe)
        v = [C.c_ssize_t() for _ in range(4)]
        N = L.fftw_mpi_local_size_many_transposed(3, nn, howmany, 0, 0, comm, *[C.byref(t) for t in v])
        ln0, s0, ln1, s1 = (t.value for t in v)
        assert N == howmany * local_sizes(lib, shape, comm)[0]
        xin = np.zeros(max(N, 1), dtype=lib.cplx)
        xin[:ln0 * n1 * rl] = x[s0:s0 + ln0].ravel()
        out = np.zeros(max(N, 1), dtype=lib.cplx)
        plan = L.fftw_mpi_plan_many_dft(3, nn, howmany, 0, 0, xin.ctypes.data, out.ctypes.data, comm, -1,
                                        capi.FFTW_ESTIMATE | TRANSPOSED_OUT)
        assert plan, "slab plan is NULL"
        S, bp = buffer_view(lib, plan, 0, lib.cplx)
        R, _ = buffer_view(lib, plan, 1, lib.cplx)
        ranks.append(dict(comm=comm, la=ln0, sa=s0, lb=ln1, sb=s1, xin=xin, out=out, plan=plan, bp=bp, S=S, R=R))
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 0, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    if "fused into" not in lib.sprint_plan(ranks[0]["plan"]):
        _exchange(ranks, n0, n1, rl)
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 2, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
        ref = Y.transpose(1, 0, 2, 3)[r["sb"]:r["sb"] + r["lb"]]
        got = r["out"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13
        lib.destroy_plan(r["plan"])
        L.b2fft_comm_destroy(r["comm"])


@pytest.mark.parametrize("P", [2, 3, 8])
@pytest.mark.parametrize("shape", [(12, 32, 6), (9, 64, 16)], ids=lambda s: "x".join(map(str, s)))
def test_hybrid_exchange_stores_the_local_block_itself(shape, P, monkeypatch):
    """hybrid exchange (default for n1 > 1024, forced here): the n1 pass writes the rows this rank owns
    straight into its own receive buffer and leaves the rest in the send buffer; the test moves only the
    REMOTE blocks, so a local block that the kernel did not place would stay zero."""
    monkeypatch.setenv("B2FFT_SLAB_HYBRID", "1")
    lib = load()
    lib.lib.b2fft_slab_set_peer_recv.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_int]
    n0, n1, n2 = shape
    rng = np.random.default_rng(3 * P + n1)
    x = ((rng.random(shape) - 0.5) + 1j * (rng.random(shape) - 0.5)).astype(lib.cplx)
    ranks = []
    for pe in range(P):
        comm = lib.lib.b2fft_comm_detached(P, pe)
        N, ln0, s0, ln1, s1 = local_sizes(lib, shape, comm)
        xin = np.zeros(max(N, 1), dtype=lib.cplx)
        xin[:ln0 * n1 * n2] = x[s0:s0 + ln0].ravel()
        out = np.zeros(max(N, 1), dtype=lib.cplx)
        plan = lib.lib.fftw_mpi_plan_dft_3d(n0, n1, n2, xin.ctypes.data, out.ctypes.data, comm, -1,
                                            capi.FFTW_ESTIMATE | TRANSPOSED_OUT)
        assert plan
        S, bp = buffer_view(lib, plan, 0, lib.cplx)
        R, _ = buffer_view(lib, plan, 1, lib.cplx)
        R[:] = 0
        ranks.append(dict(comm=comm, pe=pe, la=ln0, sa=s0, lb=ln1, sb=s1, xin=xin, out=out, plan=plan, bp=bp, S=S, R=R))
    recv = (C.c_void_p * P)(*[r["R"].ctypes.data for r in ranks])
    for r in ranks:
        assert lib.lib.b2fft_slab_set_peer_recv(r["bp"], recv, P) == 0  # not the fully fused form ...
        assert "local block stored by the n1 pass" in lib.sprint_plan(r["plan"])  # ... but the hybrid one
    for r in ranks:
        assert lib.lib.b2fft_slab_stage(r["bp"], 0, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    for src in ranks:  # remote blocks only
        if src["la"] == 0:
            continue
        Sv = src["S"][:src["la"] * n1 * n2].reshape(src["la"], n1, n2)
        for dst in ranks:
            if dst["lb"] == 0 or dst["pe"] == src["pe"]:
                continue
            Rv = dst["R"][:n0 * dst["lb"] * n2].reshape(n0, dst["lb"], n2)
            Rv[src["sa"]:src["sa"] + src["la"]] = Sv[:, dst["sb"]:dst["sb"] + dst["lb"], :]
    for pe, r in enumerate(ranks):
        assert lib.lib.b2fft_slab_stage(r["bp"], 2, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
        ref = orc.mpi_dft_3d_local(x, P, pe, transposed_out=True)
        got = r["out"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13, (pe, P)
    for r in ranks:
        lib.destroy_plan(r["plan"])
        lib.lib.b2fft_comm_destroy(r["comm"])


@pytest.mark.parametrize("flags", [0, TRANSPOSED_OUT, TRANSPOSED_IN], ids=["natural", "tout", "tin"])
@pytest.mark.parametrize("P", [1, 2, 3])
@pytest.mark.parametrize("shape", [(8, 6), (10, 32), (64, 9)], ids=lambda s: "x".join(map(str, s)))
def test_two_dimensional_slab(shape, P, flags):
    """fftw_mpi_plan_dft_2d (fftw/mpi/api.c:631-636): the 3-D machinery with rows of one element"""
    lib = load()
    L = lib.lib
    L.fftw_mpi_plan_dft_2d.restype = C.c_void_p
    L.fftw_mpi_plan_dft_2d.argtypes = [C.c_ssize_t] * 2 + [C.c_void_p] * 3 + [C.c_int, C.c_uint]
    L.fftw_mpi_local_size_2d_transposed.restype = C.c_ssize_t
    L.fftw_mpi_local_size_2d_transposed.argtypes = [C.c_ssize_t] * 2 + [C.c_void_p] + [C.POINTER(C.c_ssize_t)] * 4
    n0, n1 = shape
    rng = np.random.default_rng(7 + P)
    x = ((rng.random(shape) - 0.5) + 1j * (rng.random(shape) - 0.5)).astype(lib.cplx)
    Y = orc.dft(x, -1)
    tin, tout = bool(flags & TRANSPOSED_IN), bool(flags & TRANSPOSED_OUT)
    ranks = []
    for pe in range(P):
        comm = L.b2fft_comm_detached(P, pe)
        v = [C.c_ssize_t() for _ in range(4)]
        N = L.fftw_mpi_local_size_2d_transposed(n0, n1, comm, *[C.byref(t) for t in v])
        ln0, s0, ln1, s1 = (t.value for t in v)
        assert (N, ln0, s0, ln1, s1) == orc.mpi_local_size_3d_transposed(n0, n1, 1, P, pe)
        xin = np.zeros(max(N, 1), dtype=lib.cplx)
        if tin:
            xin[:ln1 * n0] = x.T[s1:s1 + ln1].ravel()
        else:
            xin[:ln0 * n1] = x[s0:s0 + ln0].ravel()
        out = np.zeros(max(N, 1), dtype=lib.cplx)
        plan = L.fftw_mpi_plan_dft_2d(n0, n1, xin.ctypes.data, out.ctypes.data, comm, -1, capi.FFTW_ESTIMATE | flags)
        assert plan, "2-D slab plan is NULL"
        S, bp = buffer_view(lib, plan, 0, lib.cplx)
        R, _ = buffer_view(lib, plan, 1, lib.cplx)
        # inside a TRANSPOSED_IN plan the two dims are exchanged
        a = dict(la=ln1, sa=s1, lb=ln0, sb=s0) if tin else dict(la=ln0, sa=s0, lb=ln1, sb=s1)
        ranks.append(dict(comm=comm, xin=xin, out=out, plan=plan, bp=bp, S=S, R=R, ln0=ln0, s0=s0, ln1=ln1, s1=s1, **a))
    na, nb = (n1, n0) if tin else (n0, n1)
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 0, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    if "fused into" not in lib.sprint_plan(ranks[0]["plan"]):
        _exchange(ranks, na, nb, 1)
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 2, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    plan_natural = (tin == tout)  # the plan's own output is natural: second exchange, back
    if plan_natural:
        for src in ranks:
            if src["lb"] == 0:
                continue
            Rv = src["R"][:na * src["lb"]].reshape(na, src["lb"])
            for dst in ranks:
                if dst["la"] == 0:
                    continue
                Ov = dst["out"][:dst["la"] * nb].reshape(dst["la"], nb)
                Ov[:, src["sb"]:src["sb"] + src["lb"]] = Rv[dst["sa"]:dst["sa"] + dst["la"]]
    for r in ranks:
        ref = Y.T[r["s1"]:r["s1"] + r["ln1"]] if tout else Y[r["s0"]:r["s0"] + r["ln0"]]
        got = r["out"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13
        lib.destroy_plan(r["plan"])
        L.b2fft_comm_destroy(r["comm"])


@pytest.mark.parametrize("flags", [TRANSPOSED_OUT, TRANSPOSED_IN], ids=["tout", "tin"])
@pytest.mark.parametrize("shape,P,b_n0,b_n1", [((10, 12, 4), 4, 4, 5), ((9, 32, 3), 3, 5, 16), ((8, 6, 5), 2, 6, 3)],
                         ids=["10x12x4", "9x32x3", "8x6x5"])
def test_explicit_block_sizes(shape, P, b_n0, b_n1, flags):
    """fftw_mpi_plan_many_dft with block / tblock other than ceil(n/P) (fftw/mpi/api.c:592-636, block.c:46-50):
    uneven and empty ranks; more blocks than ranks -> NULL"""
    lib = load()
    L = lib.lib
    L.fftw_mpi_plan_many_dft.restype = C.c_void_p
    L.fftw_mpi_plan_many_dft.argtypes = [C.c_int, C.POINTER(C.c_ssize_t), C.c_ssize_t, C.c_ssize_t, C.c_ssize_t,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint]
    L.fftw_mpi_local_size_many_transposed.restype = C.c_ssize_t
    L.fftw_mpi_local_size_many_transposed.argtypes = [C.c_int, C.POINTER(C.c_ssize_t), C.c_ssize_t, C.c_ssize_t,
                                                      C.c_ssize_t, C.c_void_p] + [C.POINTER(C.c_ssize_t)] * 4
    n0, n1, n2 = shape
    tin = bool(flags & TRANSPOSED_IN)
    blk, tblk = (b_n1, b_n0) if tin else (b_n0, b_n1)
    rng = np.random.default_rng(P + blk)
    x = ((rng.random(shape) - 0.5) + 1j * (rng.random(shape) - 0.5)).astype(lib.cplx)
    Y = orc.dft(x, -1)
    nn = (C.c_ssize_t * 3)(*shape)
    # local_size_many_transposed: block0 over n0, block1 over n1; the plan: block over the INPUT dim, tblock over the OUTPUT dim
    ranks = []
    for pe in range(P):
        comm = L.b2fft_comm_detached(P, pe)
        v = [C.c_ssize_t() for _ in range(4)]
        N = L.fftw_mpi_local_size_many_transposed(3, nn, 1, b_n0, b_n1, comm, *[C.byref(t) for t in v])
        ln0, s0, ln1, s1 = (t.value for t in v)
        assert ln0 == max(0, min(b_n0, n0 - pe * b_n0)) and ln1 == max(0, min(b_n1, n1 - pe * b_n1))
        xin = np.zeros(max(N, 1), dtype=lib.cplx)
        if tin:
            xin[:ln1 * n0 * n2] = x.transpose(1, 0, 2)[s1:s1 + ln1].ravel()
        else:
            xin[:ln0 * n1 * n2] = x[s0:s0 + ln0].ravel()
        out = np.zeros(max(N, 1), dtype=lib.cplx)
        plan = L.fftw_mpi_plan_many_dft(3, nn, 1, blk, tblk, xin.ctypes.data, out.ctypes.data, comm, -1,
                                        capi.FFTW_ESTIMATE | flags)
        assert plan, "slab plan with explicit blocks is NULL"
        S, bp = buffer_view(lib, plan, 0, lib.cplx)
        R, _ = buffer_view(lib, plan, 1, lib.cplx)
        a = dict(la=ln1, sa=s1, lb=ln0, sb=s0) if tin else dict(la=ln0, sa=s0, lb=ln1, sb=s1)
        ranks.append(dict(comm=comm, xin=xin, out=out, plan=plan, bp=bp, S=S, R=R, ln0=ln0, s0=s0, ln1=ln1, s1=s1, **a))
    na, nb = (n1, n0) if tin else (n0, n1)
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 0, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    _exchange(ranks, na, nb, n2)
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 2, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
        ref = Y[r["s0"]:r["s0"] + r["ln0"]] if tin else Y.transpose(1, 0, 2)[r["s1"]:r["s1"] + r["ln1"]]
        got = r["out"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13
        lib.destroy_plan(r["plan"])
    # a block size that needs more blocks than ranks cannot be distributed
    assert not L.fftw_mpi_plan_many_dft(3, nn, 1, 1, 0, ranks[0]["xin"].ctypes.data, ranks[0]["out"].ctypes.data,
                                        ranks[0]["comm"], -1, capi.FFTW_ESTIMATE | flags) or max(n0, n1) <= P
    for r in ranks:
        L.b2fft_comm_destroy(r["comm"])


def _real_api(lib):
    L = lib.lib
    for nm in ("fftw_mpi_plan_dft_r2c_3d", "fftw_mpi_plan_dft_c2r_3d"):
        f = getattr(L, nm)
        f.restype = C.c_void_p
        f.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] * 3 + [C.c_uint]
    return L


@pytest.mark.parametrize("transposed", [True, False], ids=["transposed_out", "natural_out"])
@pytest.mark.parametrize("P", [1, 2, 3])
@pytest.mark.parametrize("shape", [(8, 6, 10), (10, 7, 9), (6, 32, 12)], ids=lambda s: "x".join(map(str, s)))
def test_real_to_complex_slab(shape, P, transposed):
    """fftw_mpi_plan_dft_r2c_3d (fftw/mpi/api.c:853-858; layout fftw/doc/mpi.texi 'Multi-dimensional MPI
    DFTs of Real Data'): real [local_n0][n1][2(n2/2+1)] -> complex [local_n1][n0][n2/2+1] (transposed)
    or [local_n0][n1][n2/2+1]."""
    lib = load()
    L = _real_api(lib)
    n0, n1, n2 = shape
    n2c = n2 // 2 + 1
    rng = np.random.default_rng(31 + P)
    x = (rng.random(shape) - 0.5).astype(lib.real)
    Y = orc.r2c(x)
    ranks = []
    for pe in range(P):
        comm = L.b2fft_comm_detached(P, pe)
        N, ln0, s0, ln1, s1 = local_sizes(lib, (n0, n1, n2c), comm)
        xin = np.zeros(2 * max(N, 1), dtype=lib.real)
        pad = np.zeros((ln0, n1, 2 * n2c), dtype=lib.real)
        pad[:, :, :n2] = x[s0:s0 + ln0]
        xin[:pad.size] = pad.ravel()
        out = np.zeros(max(N, 1), dtype=lib.cplx)
        plan = L.fftw_mpi_plan_dft_r2c_3d(n0, n1, n2, xin.ctypes.data, out.ctypes.data, comm,
                                          capi.FFTW_ESTIMATE | (TRANSPOSED_OUT if transposed else 0))
        assert plan, "slab r2c plan is NULL"
        S, bp = buffer_view(lib, plan, 0, lib.cplx)
        R, _ = buffer_view(lib, plan, 1, lib.cplx)
        ranks.append(dict(comm=comm, la=ln0, sa=s0, lb=ln1, sb=s1, xin=xin, out=out, plan=plan, bp=bp, S=S, R=R))
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 0, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    if "fused into" not in lib.sprint_plan(ranks[0]["plan"]):
        _exchange(ranks, n0, n1, n2c)
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 2, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    if not transposed:
        for src in ranks:
            if src["lb"] == 0:
                continue
            Rv = src["R"][:n0 * src["lb"] * n2c].reshape(n0, src["lb"], n2c)
            for dst in ranks:
                if dst["la"] == 0:
                    continue
                Ov = dst["out"][:dst["la"] * n1 * n2c].reshape(dst["la"], n1, n2c)
                Ov[:, src["sb"]:src["sb"] + src["lb"], :] = Rv[dst["sa"]:dst["sa"] + dst["la"]]
    for r in ranks:
        ref = Y.transpose(1, 0, 2)[r["sb"]:r["sb"] + r["lb"]] if transposed else Y[r["sa"]:r["sa"] + r["la"]]
        got = r["out"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13
        lib.destroy_plan(r["plan"])
        L.b2fft_comm_destroy(r["comm"])


@pytest.mark.parametrize("tin", [True, False], ids=["transposed_in", "natural_in"])
@pytest.mark.parametrize("P", [1, 2, 3])
@pytest.mark.parametrize("shape", [(8, 6, 10), (10, 7, 9), (6, 32, 12)], ids=lambda s: "x".join(map(str, s)))
def test_complex_to_real_slab(shape, P, tin):
    """fftw_mpi_plan_dft_c2r_3d: the pipeline backwards -- n0 FFTs, reverse exchange, local 2-D c2r"""
    lib = load()
    L = _real_api(lib)
    n0, n1, n2 = shape
    n2c = n2 // 2 + 1
    rng = np.random.default_rng(41 + P)
    x = (rng.random(shape) - 0.5).astype(lib.real)
    Y = orc.r2c(x)
    YT = np.ascontiguousarray(Y.transpose(1, 0, 2))
    ranks = []
    for pe in range(P):
        comm = L.b2fft_comm_detached(P, pe)
        N, ln0, s0, ln1, s1 = local_sizes(lib, (n0, n1, n2c), comm)
        cin = np.zeros(max(N, 1), dtype=lib.cplx)
        if tin:
            cin[:ln1 * n0 * n2c] = YT[s1:s1 + ln1].ravel()
        else:
            cin[:ln0 * n1 * n2c] = Y[s0:s0 + ln0].ravel()
        out = np.zeros(2 * max(N, 1), dtype=lib.real)
        plan = L.fftw_mpi_plan_dft_c2r_3d(n0, n1, n2, cin.ctypes.data, out.ctypes.data, comm,
                                          capi.FFTW_ESTIMATE | (TRANSPOSED_IN if tin else 0))
        assert plan, "slab c2r plan is NULL"
        S, bp = buffer_view(lib, plan, 0, lib.cplx)
        R, _ = buffer_view(lib, plan, 1, lib.cplx)
        ranks.append(dict(comm=comm, la=ln0, sa=s0, lb=ln1, sb=s1, cin=cin, out=out, plan=plan, bp=bp, S=S, R=R))
    if P == 1:  # one rank: the whole pipeline through fftw_execute
        r = ranks[0]
        lib.execute(r["plan"])
    else:
        if not tin:  # forward exchange of the natural input straight into the receive layout
            for src in ranks:
                if src["la"] == 0:
                    continue
                Iv = src["cin"][:src["la"] * n1 * n2c].reshape(src["la"], n1, n2c)
                for dst in ranks:
                    if dst["lb"] == 0:
                        continue
                    Rv = dst["R"][:n0 * dst["lb"] * n2c].reshape(n0, dst["lb"], n2c)
                    Rv[src["sa"]:src["sa"] + src["la"]] = Iv[:, dst["sb"]:dst["sb"] + dst["lb"], :]
        for r in ranks:
            assert L.b2fft_slab_stage(r["bp"], 1, r["cin"].ctypes.data, r["out"].ctypes.data) == 0
        for src in ranks:  # reverse exchange: receive layout -> [la][n1][n2c]
            if src["lb"] == 0:
                continue
            Rv = src["R"][:n0 * src["lb"] * n2c].reshape(n0, src["lb"], n2c)
            for dst in ranks:
                if dst["la"] == 0:
                    continue
                Sv = dst["S"][:dst["la"] * n1 * n2c].reshape(dst["la"], n1, n2c)
                Sv[:, src["sb"]:src["sb"] + src["lb"], :] = Rv[dst["sa"]:dst["sa"] + dst["la"]]
        for r in ranks:
            assert L.b2fft_slab_stage(r["bp"], 3, r["cin"].ctypes.data, r["out"].ctypes.data) == 0
    for r in ranks:
        ref = x[r["sa"]:r["sa"] + r["la"]] * (n0 * n1 * n2)
        got = r["out"][:r["la"] * n1 * 2 * n2c].reshape(r["la"], n1, 2 * n2c)[:, :, :n2]
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13
        lib.destroy_plan(r["plan"])
        L.b2fft_comm_destroy(r["comm"])


@pytest.mark.parametrize("transposed", [True, False], ids=["transposed", "natural"])
@pytest.mark.parametrize("P", [1, 2, 3])
@pytest.mark.parametrize("shape", [(8, 10), (12, 9), (6, 64)], ids=lambda s: "x".join(map(str, s)))
def test_two_dimensional_real_slab(shape, P, transposed):
    """fftw_mpi_plan_dft_r2c_2d and _c2r_2d back (fftw/mpi/api.c:845-851, 867-873): the n1/2+1 complex columns
    are the dimension the exchange distributes"""
    lib = load()
    L = lib.lib
    for nm in ("fftw_mpi_plan_dft_r2c_2d", "fftw_mpi_plan_dft_c2r_2d"):
        f = getattr(L, nm)
        f.restype = C.c_void_p
        f.argtypes = [C.c_ssize_t] * 2 + [C.c_void_p] * 3 + [C.c_uint]
    n0, n1 = shape
    n1c = n1 // 2 + 1
    rng = np.random.default_rng(51 + P)
    x = (rng.random(shape) - 0.5).astype(lib.real)
    Y = orc.r2c(x)
    fwd, bwd = [], []
    for pe in range(P):
        comm = L.b2fft_comm_detached(P, pe)
        N, ln0, s0, ln1, s1 = local_sizes(lib, (n0, n1c, 1), comm)
        rin = np.zeros(2 * max(N, 1), dtype=lib.real)
        pad = np.zeros((ln0, 2 * n1c), dtype=lib.real)
        pad[:, :n1] = x[s0:s0 + ln0]
        rin[:pad.size] = pad.ravel()
        cbuf = np.zeros(max(N, 1), dtype=lib.cplx)
        rout = np.zeros(2 * max(N, 1), dtype=lib.real)
        pf = L.fftw_mpi_plan_dft_r2c_2d(n0, n1, rin.ctypes.data, cbuf.ctypes.data, comm,
                                        capi.FFTW_ESTIMATE | (TRANSPOSED_OUT if transposed else 0))
        pb = L.fftw_mpi_plan_dft_c2r_2d(n0, n1, cbuf.ctypes.data, rout.ctypes.data, comm,
                                        capi.FFTW_ESTIMATE | (TRANSPOSED_IN if transposed else 0))
        assert pf and pb, "2-D real slab plan is NULL"
        geo = dict(comm=comm, la=ln0, sa=s0, lb=ln1, sb=s1, rin=rin, cbuf=cbuf, rout=rout)
        for plan, lst in ((pf, fwd), (pb, bwd)):
            S, bp = buffer_view(lib, plan, 0, lib.cplx)
            R, _ = buffer_view(lib, plan, 1, lib.cplx)
            lst.append(dict(plan=plan, bp=bp, S=S, R=R, **geo))

    def reverse_exchange(ranks, dst_key):  # receive layout [n0][lb] -> [la][n1c] rows
        for src in ranks:
            if src["lb"] == 0:
                continue
            Rv = src["R"][:n0 * src["lb"]].reshape(n0, src["lb"])
            for dst in ranks:
                if dst["la"] == 0:
                    continue
                Dv = dst[dst_key][:dst["la"] * n1c].reshape(dst["la"], n1c)
                Dv[:, src["sb"]:src["sb"] + src["lb"]] = Rv[dst["sa"]:dst["sa"] + dst["la"]]

    # ---- r2c
    for r in fwd:
        assert L.b2fft_slab_stage(r["bp"], 0, r["rin"].ctypes.data, r["cbuf"].ctypes.data) == 0
    _exchange(fwd, n0, n1c, 1)
    for r in fwd:
        assert L.b2fft_slab_stage(r["bp"], 2, r["rin"].ctypes.data, r["cbuf"].ctypes.data) == 0
    if not transposed:
        reverse_exchange(fwd, "cbuf")
    for r in fwd:
        ref = Y.T[r["sb"]:r["sb"] + r["lb"]] if transposed else Y[r["sa"]:r["sa"] + r["la"]]
        got = r["cbuf"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13
    # ---- c2r from that layout
    if not transposed:  # forward exchange of the natural input straight into the receive layout
        for src in bwd:
            if src["la"] == 0:
                continue
            Iv = src["cbuf"][:src["la"] * n1c].reshape(src["la"], n1c)
            for dst in bwd:
                if dst["lb"] == 0:
                    continue
                Rv = dst["R"][:n0 * dst["lb"]].reshape(n0, dst["lb"])
                Rv[src["sa"]:src["sa"] + src["la"]] = Iv[:, dst["sb"]:dst["sb"] + dst["lb"]]
    for r in bwd:
        assert L.b2fft_slab_stage(r["bp"], 1, r["cbuf"].ctypes.data, r["rout"].ctypes.data) == 0
    reverse_exchange(bwd, "S")
    for r in bwd:
        assert L.b2fft_slab_stage(r["bp"], 3, r["cbuf"].ctypes.data, r["rout"].ctypes.data) == 0
        want = x[r["sa"]:r["sa"] + r["la"]] * (n0 * n1)
        got = r["rout"][:r["la"] * 2 * n1c].reshape(r["la"], 2 * n1c)[:, :n1]
        if want.size:
            assert np.linalg.norm(got - want) / np.linalg.norm(want) < 1e-13
    for r in fwd + bwd:
        lib.destroy_plan(r["plan"])
    for r in fwd:
        L.b2fft_comm_destroy(r["comm"])


@pytest.mark.parametrize("P", [1, 2, 3])
@pytest.mark.parametrize("howmany", [2, 3])
def test_real_slab_with_interleaved_transforms(howmany, P):
    """fftw_mpi_plan_many_dft_r2c / _c2r with howmany > 1: tuples of howmany reals (complex) per element"""
    lib = load()
    L = lib.lib
    for nm in ("fftw_mpi_plan_many_dft_r2c", "fftw_mpi_plan_many_dft_c2r"):
        f = getattr(L, nm)
        f.restype = C.c_void_p
        f.argtypes = [C.c_int, C.POINTER(C.c_ssize_t), C.c_ssize_t, C.c_ssize_t, C.c_ssize_t, C.c_void_p, C.c_void_p,
                      C.c_void_p, C.c_uint]
    n0, n1, n2 = 6, 8, 10
    n2c, hm = n2 // 2 + 1, howmany
    rl = n2c * hm
    rng = np.random.default_rng(61 + P)
    x = (rng.random((n0, n1, n2, hm)) - 0.5).astype(lib.real)
    Y = np.stack([orc.r2c(np.ascontiguousarray(x[..., h])) for h in range(hm)], axis=-1)  # [n0][n1][n2c][hm]
    nn = (C.c_ssize_t * 3)(n0, n1, n2)
    fwd, bwd = [], []
    for pe in range(P):
        comm = L.b2fft_comm_detached(P, pe)
        N, ln0, s0, ln1, s1 = local_sizes(lib, (n0, n1, n2c), comm)
        N *= hm
        pad = np.zeros((ln0, n1, 2 * n2c, hm), dtype=lib.real)
        pad[:, :, :n2, :] = x[s0:s0 + ln0]
        rin = np.zeros(2 * max(N, 1), dtype=lib.real)
        rin[:pad.size] = pad.ravel()
        cbuf = np.zeros(max(N, 1), dtype=lib.cplx)
        rout = np.zeros(2 * max(N, 1), dtype=lib.real)
        pf = L.fftw_mpi_plan_many_dft_r2c(3, nn, hm, 0, 0, rin.ctypes.data, cbuf.ctypes.data, comm,
                                          capi.FFTW_ESTIMATE | TRANSPOSED_OUT)
        pb = L.fftw_mpi_plan_many_dft_c2r(3, nn, hm, 0, 0, cbuf.ctypes.data, rout.ctypes.data, comm,
                                          capi.FFTW_ESTIMATE | TRANSPOSED_IN)
        assert pf and pb, "slab real plan with howmany is NULL"
        geo = dict(comm=comm, la=ln0, sa=s0, lb=ln1, sb=s1, rin=rin, cbuf=cbuf, rout=rout)
        for plan, lst in ((pf, fwd), (pb, bwd)):
            S, bp = buffer_view(lib, plan, 0, lib.cplx)
            R, _ = buffer_view(lib, plan, 1, lib.cplx)
            lst.append(dict(plan=plan, bp=bp, S=S, R=R, **geo))
    for r in fwd:
        assert L.b2fft_slab_stage(r["bp"], 0, r["rin"].ctypes.data, r["cbuf"].ctypes.data) == 0
    if "fused into" not in lib.sprint_plan(fwd[0]["plan"]):
        _exchange(fwd, n0, n1, rl)
    for r in fwd:
        assert L.b2fft_slab_stage(r["bp"], 2, r["rin"].ctypes.data, r["cbuf"].ctypes.data) == 0
        ref = Y.transpose(1, 0, 2, 3)[r["sb"]:r["sb"] + r["lb"]]
        got = r["cbuf"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13
    for r in bwd:  # transposed input -> n0 FFTs -> reverse exchange -> local 2-D c2r
        assert L.b2fft_slab_stage(r["bp"], 1, r["cbuf"].ctypes.data, r["rout"].ctypes.data) == 0
    for src in bwd:
        if src["lb"] == 0:
            continue
        Rv = src["R"][:n0 * src["lb"] * rl].reshape(n0, src["lb"], rl)
        for dst in bwd:
            if dst["la"] == 0:
                continue
            Sv = dst["S"][:dst["la"] * n1 * rl].reshape(dst["la"], n1, rl)
            Sv[:, src["sb"]:src["sb"] + src["lb"], :] = Rv[dst["sa"]:dst["sa"] + dst["la"]]
    for r in bwd:
        assert L.b2fft_slab_stage(r["bp"], 3, r["cbuf"].ctypes.data, r["rout"].ctypes.data) == 0
        want = x[r["sa"]:r["sa"] + r["la"]] * (n0 * n1 * n2)
        got = r["rout"][:r["la"] * n1 * 2 * rl].reshape(r["la"], n1, 2 * n2c, hm)[:, :, :n2, :]
        if want.size:
            assert np.linalg.norm(got - want) / np.linalg.norm(want) < 1e-13
    for r in fwd + bwd:
        lib.destroy_plan(r["plan"])
    for r in fwd:
        L.b2fft_comm_destroy(r["comm"])


def _gloo_worker(rank, world, port, shape, q):
    try:
        import torch
        import torch.distributed as dist
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        lib = load()
        n0, n1, n2 = shape
        rng = np.random.default_rng(5)  # every rank can regenerate the global array
        x = ((rng.random(shape) - 0.5) + 1j * (rng.random(shape) - 0.5)).astype(lib.cplx)
        comm = lib.lib.b2fft_comm_detached(world, rank)
        N, ln0, s0, ln1, s1 = local_sizes(lib, shape, comm)
        xin = np.zeros(max(N, 1), dtype=lib.cplx)
        xin[:ln0 * n1 * n2] = x[s0:s0 + ln0].ravel()
        out = np.zeros(max(N, 1), dtype=lib.cplx)
        plan = lib.lib.fftw_mpi_plan_dft_3d(n0, n1, n2, xin.ctypes.data, out.ctypes.data, comm, -1,
                                            capi.FFTW_ESTIMATE | TRANSPOSED_OUT)
        assert plan
        S, bp = buffer_view(lib, plan, 0, lib.cplx)
        R, _ = buffer_view(lib, plan, 1, lib.cplx)
        assert lib.lib.b2fft_slab_stage(bp, 0, xin.ctypes.data, out.ctypes.data) == 0
        b0, b1 = -(-n0 // world), -(-n1 // world)
        Sv = S[:ln0 * n1 * n2].reshape(ln0, n1, n2)
        sends = [torch.from_numpy(np.ascontiguousarray(Sv[:, p * b1:min(n1, (p + 1) * b1), :]).view(np.float64).copy())
                 for p in range(world)]
        recvs = [torch.empty((max(0, min(b0, n0 - p * b0)), ln1, n2 * 2), dtype=torch.float64).reshape(-1)
                 for p in range(world)]
        sends = [t.reshape(-1) for t in sends]
        # gloo has no all_to_all: pairwise exchange
        for p in range(world):
            if p == rank:
                recvs[p].copy_(sends[p])
            elif rank < p:
                dist.send(sends[p], p)
                dist.recv(recvs[p], p)
            else:
                dist.recv(recvs[p], p)
                dist.send(sends[p], p)
        Rv = R[:n0 * ln1 * n2].reshape(n0, ln1, n2)
        for p in range(world):
            l0p = max(0, min(b0, n0 - p * b0))
            if l0p and ln1:
                Rv[p * b0:p * b0 + l0p] = recvs[p].numpy().view(np.complex128).reshape(l0p, ln1, n2)
        assert lib.lib.b2fft_slab_stage(bp, 2, xin.ctypes.data, out.ctypes.data) == 0
        ref = orc.mpi_dft_3d_local(x, world, rank, transposed_out=True)
        err = np.linalg.norm(out[:ref.size].reshape(ref.shape) - ref) / np.linalg.norm(ref)
        # and the batch-sharding rule of bench.py: per-rank howmany, whole-job value by max time
        t = torch.tensor([1.0 + rank])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.destroy_process_group()
        q.put((rank, float(err), float(t.item())))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "ERR " + traceback.format_exc(), 0.0))


def test_two_processes_over_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, (12, 10, 6), q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, err, tmax in res:
        assert not isinstance(err, str), err
        assert err < 1e-13
        assert tmax == 2.0


def _many_real_api(lib):
    L = lib.lib
    for nm in ("fftw_mpi_plan_many_dft_r2c", "fftw_mpi_plan_many_dft_c2r"):
        f = getattr(L, nm)
        f.restype = C.c_void_p
        f.argtypes = [C.c_int, C.POINTER(C.c_ssize_t), C.c_ssize_t, C.c_ssize_t, C.c_ssize_t, C.c_void_p, C.c_void_p,
                      C.c_void_p, C.c_uint]
    L.fftw_mpi_local_size_many_transposed.restype = C.c_ssize_t
    L.fftw_mpi_local_size_many_transposed.argtypes = [C.c_int, C.POINTER(C.c_ssize_t), C.c_ssize_t, C.c_ssize_t,
                                                      C.c_ssize_t, C.c_void_p] + [C.POINTER(C.c_ssize_t)] * 4
    return L


def _block_sizes(L, shape_c, b_n0, b_n1, comm):
    v = [C.c_ssize_t() for _ in range(4)]
    nn = (C.c_ssize_t * 3)(*shape_c)
    N = L.fftw_mpi_local_size_many_transposed(3, nn, 1, b_n0, b_n1, comm, *[C.byref(t) for t in v])
    return (N,) + tuple(t.value for t in v)


@pytest.mark.parametrize("transposed", [True, False], ids=["transposed_out", "natural_out"])
@pytest.mark.parametrize("shape,P,b_n0,b_n1", [((10, 12, 6), 4, 4, 5), ((9, 32, 7), 3, 5, 16), ((8, 6, 10), 2, 6, 3)],
                         ids=["10x12x6", "9x32x7", "8x6x10"])
def test_real_to_complex_slab_with_explicit_block_sizes(shape, P, b_n0, b_n1, transposed):
    """fftw_mpi_plan_many_dft_r2c with iblock / oblock other than ceil(n/P) (fftw/mpi/api.c:790-812: iblock over the input's
    distributed dimension n0, oblock over the output's -- n1 with FFTW_MPI_TRANSPOSED_OUT, else n0 again): uneven and empty
    ranks; a natural-output plan whose oblock differs from its iblock is refused (NULL)."""
    lib = load()
    L = _many_real_api(lib)
    n0, n1, n2 = shape
    n2c = n2 // 2 + 1
    rng = np.random.default_rng(51 + P)
    x = (rng.random(shape) - 0.5).astype(lib.real)
    Y = orc.r2c(x)
    nn = (C.c_ssize_t * 3)(*shape)
    ranks = []
    for pe in range(P):
        comm = L.b2fft_comm_detached(P, pe)
        # (natural output: the distribution over n1 in the middle of the plan is the plan's own business -- default blocks)
        N, ln0, s0, ln1, s1 = _block_sizes(L, (n0, n1, n2c), b_n0, b_n1 if transposed else 0, comm)
        assert ln0 == max(0, min(b_n0, n0 - pe * b_n0)) and (not transposed or ln1 == max(0, min(b_n1, n1 - pe * b_n1)))
        xin = np.zeros(2 * max(N, 1), dtype=lib.real)
        pad = np.zeros((ln0, n1, 2 * n2c), dtype=lib.real)
        pad[:, :, :n2] = x[s0:s0 + ln0]
        xin[:pad.size] = pad.ravel()
        out = np.zeros(max(N, 1), dtype=lib.cplx)
        plan = L.fftw_mpi_plan_many_dft_r2c(3, nn, 1, b_n0, b_n1 if transposed else b_n0, xin.ctypes.data, out.ctypes.data,
                                            comm, capi.FFTW_ESTIMATE | (TRANSPOSED_OUT if transposed else 0))
        assert plan, "slab r2c plan with explicit blocks is NULL"
        S, bp = buffer_view(lib, plan, 0, lib.cplx)
        R, _ = buffer_view(lib, plan, 1, lib.cplx)
        ranks.append(dict(comm=comm, la=ln0, sa=s0, lb=ln1, sb=s1, xin=xin, out=out, plan=plan, bp=bp, S=S, R=R))
    if not transposed:  # different blocks on the two sides of a natural-output plan cannot be served
        assert not L.fftw_mpi_plan_many_dft_r2c(3, nn, 1, b_n0, b_n0 + 1, ranks[0]["xin"].ctypes.data,
                                                ranks[0]["out"].ctypes.data, ranks[0]["comm"], capi.FFTW_ESTIMATE)
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 0, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    if "fused into" not in lib.sprint_plan(ranks[0]["plan"]):
        _exchange(ranks, n0, n1, n2c)
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 2, r["xin"].ctypes.data, r["out"].ctypes.data) == 0
    if not transposed:
        for src in ranks:
            if src["lb"] == 0:
                continue
            Rv = src["R"][:n0 * src["lb"] * n2c].reshape(n0, src["lb"], n2c)
            for dst in ranks:
                if dst["la"] == 0:
                    continue
                Ov = dst["out"][:dst["la"] * n1 * n2c].reshape(dst["la"], n1, n2c)
                Ov[:, src["sb"]:src["sb"] + src["lb"], :] = Rv[dst["sa"]:dst["sa"] + dst["la"]]
    for r in ranks:
        ref = Y.transpose(1, 0, 2)[r["sb"]:r["sb"] + r["lb"]] if transposed else Y[r["sa"]:r["sa"] + r["la"]]
        got = r["out"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13
        lib.destroy_plan(r["plan"])
        L.b2fft_comm_destroy(r["comm"])


@pytest.mark.parametrize("tin", [True, False], ids=["transposed_in", "natural_in"])
@pytest.mark.parametrize("shape,P,b_n0,b_n1", [((10, 12, 6), 4, 4, 5), ((9, 32, 7), 3, 5, 16), ((8, 6, 10), 2, 6, 3)],
                         ids=["10x12x6", "9x32x7", "8x6x10"])
def test_complex_to_real_slab_with_explicit_block_sizes(shape, P, b_n0, b_n1, tin):
    """fftw_mpi_plan_many_dft_c2r with explicit blocks (fftw/mpi/api.c:814-836): iblock over the complex input's distributed
    dimension (n1 with FFTW_MPI_TRANSPOSED_IN, else n0), oblock over n0 of the real output."""
    lib = load()
    L = _many_real_api(lib)
    n0, n1, n2 = shape
    n2c = n2 // 2 + 1
    rng = np.random.default_rng(61 + P)
    x = (rng.random(shape) - 0.5).astype(lib.real)
    Y = orc.r2c(x)
    YT = np.ascontiguousarray(Y.transpose(1, 0, 2))
    nn = (C.c_ssize_t * 3)(*shape)
    ranks = []
    for pe in range(P):
        comm = L.b2fft_comm_detached(P, pe)
        N, ln0, s0, ln1, s1 = _block_sizes(L, (n0, n1, n2c), b_n0, b_n1 if tin else 0, comm)
        cin = np.zeros(max(N, 1), dtype=lib.cplx)
        if tin:
            cin[:ln1 * n0 * n2c] = YT[s1:s1 + ln1].ravel()
        else:
            cin[:ln0 * n1 * n2c] = Y[s0:s0 + ln0].ravel()
        out = np.zeros(2 * max(N, 1), dtype=lib.real)
        plan = L.fftw_mpi_plan_many_dft_c2r(3, nn, 1, b_n1 if tin else b_n0, b_n0, cin.ctypes.data, out.ctypes.data, comm,
                                            capi.FFTW_ESTIMATE | (TRANSPOSED_IN if tin else 0))
        assert plan, "slab c2r plan with explicit blocks is NULL"
        S, bp = buffer_view(lib, plan, 0, lib.cplx)
        R, _ = buffer_view(lib, plan, 1, lib.cplx)
        ranks.append(dict(comm=comm, la=ln0, sa=s0, lb=ln1, sb=s1, cin=cin, out=out, plan=plan, bp=bp, S=S, R=R))
    if not tin:  # forward exchange of the natural input straight into the receive layout
        for src in ranks:
            if src["la"] == 0:
                continue
            Iv = src["cin"][:src["la"] * n1 * n2c].reshape(src["la"], n1, n2c)
            for dst in ranks:
                if dst["lb"] == 0:
                    continue
                Rv = dst["R"][:n0 * dst["lb"] * n2c].reshape(n0, dst["lb"], n2c)
                Rv[src["sa"]:src["sa"] + src["la"]] = Iv[:, dst["sb"]:dst["sb"] + dst["lb"], :]
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 1, r["cin"].ctypes.data, r["out"].ctypes.data) == 0
    for src in ranks:  # reverse exchange: receive layout -> [la][n1][n2c]
        if src["lb"] == 0:
            continue
        Rv = src["R"][:n0 * src["lb"] * n2c].reshape(n0, src["lb"], n2c)
        for dst in ranks:
            if dst["la"] == 0:
                continue
            Sv = dst["S"][:dst["la"] * n1 * n2c].reshape(dst["la"], n1, n2c)
            Sv[:, src["sb"]:src["sb"] + src["lb"], :] = Rv[dst["sa"]:dst["sa"] + dst["la"]]
    for r in ranks:
        assert L.b2fft_slab_stage(r["bp"], 3, r["cin"].ctypes.data, r["out"].ctypes.data) == 0
    for r in ranks:
        ref = x[r["sa"]:r["sa"] + r["la"]] * (n0 * n1 * n2)
        got = r["out"][:r["la"] * n1 * 2 * n2c].reshape(r["la"], n1, 2 * n2c)[:, :, :n2]
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13
        lib.destroy_plan(r["plan"])
        L.b2fft_comm_destroy(r["comm"])


@pytest.mark.parametrize("transposed", [True, False], ids=["transposed", "natural"])
@pytest.mark.parametrize("shape,P,b_n0,b_n1c", [((10, 22), 4, 4, 5), ((9, 62), 3, 5, 16), ((8, 10), 2, 6, 3)],
                         ids=["10x22", "9x62", "8x10"])
def test_two_dimensional_real_slab_with_explicit_block_sizes(shape, P, b_n0, b_n1c, transposed):
    """fftw_mpi_plan_many_dft_r2c / _c2r with rnk = 2 and explicit iblock / oblock: block over n0 on the real side, over the
    n1/2+1 complex columns on a transposed complex side (fftw/mpi/api.c:790-836); r2c forward, then c2r back from that layout."""
    lib = load()
    L = _many_real_api(lib)
    n0, n1 = shape
    n1c = n1 // 2 + 1
    assert (n1c + b_n1c - 1) // b_n1c <= P
    rng = np.random.default_rng(71 + P)
    x = (rng.random(shape) - 0.5).astype(lib.real)
    Y = orc.r2c(x)
    nn = (C.c_ssize_t * 2)(n0, n1)
    nc = (C.c_ssize_t * 2)(n0, n1c)
    fwd, bwd = [], []
    for pe in range(P):
        comm = L.b2fft_comm_detached(P, pe)
        v = [C.c_ssize_t() for _ in range(4)]
        N = L.fftw_mpi_local_size_many_transposed(2, nc, 1, b_n0, b_n1c if transposed else 0, comm, *[C.byref(t) for t in v])
        ln0, s0, ln1, s1 = (t.value for t in v)
        assert ln0 == max(0, min(b_n0, n0 - pe * b_n0))
        rin = np.zeros(2 * max(N, 1), dtype=lib.real)
        pad = np.zeros((ln0, 2 * n1c), dtype=lib.real)
        pad[:, :n1] = x[s0:s0 + ln0]
        rin[:pad.size] = pad.ravel()
        cbuf = np.zeros(max(N, 1), dtype=lib.cplx)
        rout = np.zeros(2 * max(N, 1), dtype=lib.real)
        cb = b_n1c if transposed else b_n0
        pf = L.fftw_mpi_plan_many_dft_r2c(2, nn, 1, b_n0, cb, rin.ctypes.data, cbuf.ctypes.data, comm,
                                          capi.FFTW_ESTIMATE | (TRANSPOSED_OUT if transposed else 0))
        pb = L.fftw_mpi_plan_many_dft_c2r(2, nn, 1, cb, b_n0, cbuf.ctypes.data, rout.ctypes.data, comm,
                                          capi.FFTW_ESTIMATE | (TRANSPOSED_IN if transposed else 0))
        assert pf and pb, "2-D real slab plan with explicit blocks is NULL"
        geo = dict(comm=comm, la=ln0, sa=s0, lb=ln1, sb=s1, rin=rin, cbuf=cbuf, rout=rout)
        for plan, lst in ((pf, fwd), (pb, bwd)):
            S, bp = buffer_view(lib, plan, 0, lib.cplx)
            R, _ = buffer_view(lib, plan, 1, lib.cplx)
            lst.append(dict(plan=plan, bp=bp, S=S, R=R, **geo))

    def reverse_exchange(ranks, dst_key):  # receive layout [n0][lb] -> [la][n1c] rows
        for src in ranks:
            if src["lb"] == 0:
                continue
            Rv = src["R"][:n0 * src["lb"]].reshape(n0, src["lb"])
            for dst in ranks:
                if dst["la"] == 0:
                    continue
                Dv = dst[dst_key][:dst["la"] * n1c].reshape(dst["la"], n1c)
                Dv[:, src["sb"]:src["sb"] + src["lb"]] = Rv[dst["sa"]:dst["sa"] + dst["la"]]

    for r in fwd:
        assert L.b2fft_slab_stage(r["bp"], 0, r["rin"].ctypes.data, r["cbuf"].ctypes.data) == 0
    _exchange(fwd, n0, n1c, 1)
    for r in fwd:
        assert L.b2fft_slab_stage(r["bp"], 2, r["rin"].ctypes.data, r["cbuf"].ctypes.data) == 0
    if not transposed:
        reverse_exchange(fwd, "cbuf")
    for r in fwd:
        ref = Y.T[r["sb"]:r["sb"] + r["lb"]] if transposed else Y[r["sa"]:r["sa"] + r["la"]]
        got = r["cbuf"][:ref.size].reshape(ref.shape)
        if ref.size:
            assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-13
    if not transposed:
        for src in bwd:
            if src["la"] == 0:
                continue
            Iv = src["cbuf"][:src["la"] * n1c].reshape(src["la"], n1c)
            for dst in bwd:
                if dst["lb"] == 0:
                    continue
                Rv = dst["R"][:n0 * dst["lb"]].reshape(n0, dst["lb"])
                Rv[src["sa"]:src["sa"] + src["la"]] = Iv[:, dst["sb"]:dst["sb"] + dst["lb"]]
    for r in bwd:
        assert L.b2fft_slab_stage(r["bp"], 1, r["cbuf"].ctypes.data, r["rout"].ctypes.data) == 0
    reverse_exchange(bwd, "S")
    for r in bwd:
        assert L.b2fft_slab_stage(r["bp"], 3, r["cbuf"].ctypes.data, r["rout"].ctypes.data) == 0
        want = x[r["sa"]:r["sa"] + r["la"]] * (n0 * n1)
        got = r["rout"][:r["la"] * 2 * n1c].reshape(r["la"], 2 * n1c)[:, :n1]
        if want.size:
            assert np.linalg.norm(got - want) / np.linalg.norm(want) < 1e-13
    for lst in (fwd, bwd):
        for r in lst:
            lib.destroy_plan(r["plan"])
    for r in fwd:
        L.b2fft_comm_destroy(r["comm"])
