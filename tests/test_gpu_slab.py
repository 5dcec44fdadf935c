"""GPU tests of the slab-distributed 3-D transform: one rank (self communicator, device arrays)
and, when the box has >= 2 GPUs, 2 ranks over NCCL (tests/slab_nccl_worker.py under torchrun)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from fftw3_b200 import build, capi
from oracle import pyoracle as orc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("prec", ["fp32", "fp64"])
@pytest.mark.parametrize("transposed", [True, False])
@pytest.mark.parametrize("shape", [(64, 48, 32), (20, 64, 24), (6, 1024, 40), (8, 12, 512)], ids=lambda s: "x".join(map(str, s)))
def test_single_rank_device_arrays(prec, transposed, shape):
    """n1 = 64 / 1024: the n1 pass scatters straight into this rank's own receive buffer (fused exchange);
    n1 = 48: staged copy"""
    import torch
    lib = capi.FFTW3Lib(build.lib_path(prec), prec == "fp32")
    L = lib.lib
    L.fftw_mpi_comm_self.restype = C.c_void_p
    L.fftw_mpi_plan_dft_3d.restype = C.c_void_p
    L.fftw_mpi_plan_dft_3d.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] * 3 + [C.c_int, C.c_uint]
    comm = L.fftw_mpi_comm_self()
    rng = np.random.default_rng(3)
    xh = ((rng.random(shape) - 0.5) + 1j * (rng.random(shape) - 0.5)).astype(lib.cplx)
    x = torch.from_numpy(xh).cuda()
    y = torch.zeros_like(x)
    plan = L.fftw_mpi_plan_dft_3d(*shape, x.data_ptr(), y.data_ptr(), comm, -1,
                                  capi.FFTW_ESTIMATE | ((1 << 30) if transposed else 0))
    assert plan
    lib.execute(plan)
    torch.cuda.synchronize()
    ref = orc.mpi_dft_3d_local(xh, 1, 0, transposed_out=transposed)
    got = y.cpu().numpy().reshape(ref.shape)
    tol = (1e-6 if lib.single else 1e-14) * np.log2(np.prod(shape))
    assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < tol
    lib.destroy_plan(plan)


@pytest.mark.parametrize("prec", ["fp32", "fp64"])
@pytest.mark.parametrize("flags", [1 << 29, (1 << 29) | (1 << 30)], ids=["tin", "tin_tout"])
@pytest.mark.parametrize("shape,howmany", [((12, 10, 8), 1), ((16, 64, 6), 3), ((64, 20, 5), 2)],
                         ids=lambda v: "x".join(map(str, v)) if isinstance(v, tuple) else str(v))
def test_single_rank_transposed_in_and_howmany(prec, flags, shape, howmany):
    """FFTW_MPI_TRANSPOSED_IN and interleaved howmany > 1 through fftw_mpi_plan_many_dft + fftw_execute"""
    import torch
    lib = capi.FFTW3Lib(build.lib_path(prec), prec == "fp32")
    L = lib.lib
    L.fftw_mpi_comm_self.restype = C.c_void_p
    L.fftw_mpi_plan_many_dft.restype = C.c_void_p
    L.fftw_mpi_plan_many_dft.argtypes = [C.c_int, C.POINTER(C.c_ssize_t), C.c_ssize_t, C.c_ssize_t, C.c_ssize_t,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint]
    comm = L.fftw_mpi_comm_self()
    n0, n1, n2 = shape
    rng = np.random.default_rng(13)
    xh = ((rng.random(shape + (howmany,)) - 0.5) + 1j * (rng.random(shape + (howmany,)) - 0.5)).astype(lib.cplx)
    Y = np.stack([orc.dft(np.ascontiguousarray(xh[..., h]), -1) for h in range(howmany)], axis=-1)
    x = torch.from_numpy(np.ascontiguousarray(xh.transpose(1, 0, 2, 3))).cuda()   # [n1][n0][n2][howmany]
    y = torch.zeros(xh.size, dtype=x.dtype, device="cuda")
    nn = (C.c_ssize_t * 3)(*shape)
    plan = L.fftw_mpi_plan_many_dft(3, nn, howmany, 0, 0, x.data_ptr(), y.data_ptr(), comm, -1, capi.FFTW_ESTIMATE | flags)
    assert plan
    lib.execute(plan)
    torch.cuda.synchronize()
    ref = Y.transpose(1, 0, 2, 3) if flags & (1 << 30) else Y
    got = y.cpu().numpy().reshape(ref.shape)
    tol = (1e-6 if lib.single else 1e-14) * np.log2(np.prod(shape))
    assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < tol
    lib.destroy_plan(plan)


@pytest.mark.parametrize("prec", ["fp32", "fp64"])
@pytest.mark.parametrize("shape", [(16, 12, 20), (10, 64, 50), (5, 1024, 18)], ids=lambda s: "x".join(map(str, s)))
def test_single_rank_real_data(prec, shape):
    """fftw_mpi_plan_dft_r2c_3d / _c2r_3d on one rank, device arrays, all four layouts"""
    import torch
    lib = capi.FFTW3Lib(build.lib_path(prec), prec == "fp32")
    L = lib.lib
    L.fftw_mpi_comm_self.restype = C.c_void_p
    for nm in ("fftw_mpi_plan_dft_r2c_3d", "fftw_mpi_plan_dft_c2r_3d"):
        f = getattr(L, nm)
        f.restype = C.c_void_p
        f.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] * 3 + [C.c_uint]
    comm = L.fftw_mpi_comm_self()
    n0, n1, n2 = shape
    n2c = n2 // 2 + 1
    rng = np.random.default_rng(9)
    xr = (rng.random(shape) - 0.5).astype(lib.real)
    Y = orc.r2c(xr)
    tol = (1e-6 if lib.single else 1e-14) * np.log2(np.prod(shape))
    pad = np.zeros((n0, n1, 2 * n2c), dtype=lib.real)
    pad[:, :, :n2] = xr
    for transposed in (True, False):
        x = torch.from_numpy(pad).cuda()
        y = torch.zeros((n0 * n1 * n2c,), dtype=torch.complex64 if lib.single else torch.complex128, device="cuda")
        plan = L.fftw_mpi_plan_dft_r2c_3d(n0, n1, n2, x.data_ptr(), y.data_ptr(), comm,
                                          capi.FFTW_ESTIMATE | ((1 << 30) if transposed else 0))
        assert plan
        lib.execute(plan)
        torch.cuda.synchronize()
        ref = Y.transpose(1, 0, 2) if transposed else Y
        got = y.cpu().numpy().reshape(ref.shape)
        assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < tol
        lib.destroy_plan(plan)
        z = torch.zeros_like(x)
        plan = L.fftw_mpi_plan_dft_c2r_3d(n0, n1, n2, y.data_ptr(), z.data_ptr(), comm,
                                          capi.FFTW_ESTIMATE | ((1 << 29) if transposed else 0))
        assert plan
        lib.execute(plan)
        torch.cuda.synchronize()
        back = z.cpu().numpy()[:, :, :n2] / np.prod(shape)
        assert np.linalg.norm(back - xr) / np.linalg.norm(xr) < tol
        lib.destroy_plan(plan)


@pytest.mark.parametrize("shape", [(48, 40, 32), (50, 30, 16), (24, 64, 32), (21, 512, 48), (10, 12, 512), (512, 16, 256)],
                         ids=lambda s: "x".join(map(str, s)))
def test_two_ranks_over_nccl(shape):
    """transposed output: the round-2 schedule (chunks along the contiguous dimension: 1 chunk for n2 < 256, 2 at 256,
    4 at 512; uneven blocks at n0 = 21, 50); natural output: n1 = 64, 512 fuse the exchange into the n1 pass
    (stores into peer memory), 40, 30 take the DMA path"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(29700 + os.getpid() % 200),
           os.path.join(ROOT, "tests", "slab_nccl_worker.py"), *map(str, shape)]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "SLAB_OK" in r.stdout, r.stdout[-3000:]


def test_two_ranks_hybrid_exchange():
    """the exchange taken above n1 = 1024 (local block stored by the n1 pass, remote blocks by DMA), forced
    at a size the oracle can check"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(29900 + os.getpid() % 90),
           os.path.join(ROOT, "tests", "slab_nccl_worker.py"), "24", "64", "32"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600,
                       env=dict(os.environ, B2FFT_SLAB_HYBRID="1"))
    assert r.returncode == 0 and "SLAB_OK" in r.stdout, r.stdout[-3000:]
