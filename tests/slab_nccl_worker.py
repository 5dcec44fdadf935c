"""torchrun worker: slab-distributed 3-D c2c over NCCL on WORLD_SIZE GPUs, checked against the
oracle (small cube) -- launched by tests/test_gpu_slab.py and usable by hand:
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/slab_nccl_worker.py 48 40 32
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    from fftw3_b200 import build, capi
    from oracle import pyoracle as orc
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    shape = tuple(int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (48, 40, 32)
    prec = sys.argv[4] if len(sys.argv) > 4 else "fp32"
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = capi.FFTW3Lib(build.lib_path(prec), prec == "fp32")
    L = lib.lib
    L.b2fft_set_device(local)
    L.fftw_mpi_comm_from_nccl.restype = C.c_void_p
    L.fftw_mpi_comm_from_nccl.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.fftw_mpi_plan_dft_3d.restype = C.c_void_p
    L.fftw_mpi_plan_dft_3d.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] * 3 + [C.c_int, C.c_uint]
    L.fftw_mpi_local_size_3d_transposed.restype = C.c_ssize_t
    L.fftw_mpi_local_size_3d_transposed.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] + [C.POINTER(C.c_ssize_t)] * 4
    L.b2fft_last_error.restype = C.c_char_p
    idt = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        buf = (C.c_char * 128)()
        assert L.fftw_mpi_nccl_unique_id(buf) == 0, L.b2fft_last_error()
        idt = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
    idt = idt.cuda()
    dist.broadcast(idt, 0)
    comm = L.fftw_mpi_comm_from_nccl(bytes(idt.cpu().numpy().tobytes()), world, rank)
    assert comm, L.b2fft_last_error()
    n0, n1, n2 = shape
    v = [C.c_ssize_t() for _ in range(4)]
    nalloc = L.fftw_mpi_local_size_3d_transposed(n0, n1, n2, comm, *[C.byref(t) for t in v])
    ln0, s0, ln1, s1 = (t.value for t in v)
    rng = np.random.default_rng(7)
    xg = ((rng.random(shape) - 0.5) + 1j * (rng.random(shape) - 0.5)).astype(lib.cplx)
    tol = (1e-6 if lib.single else 1e-14) * np.log2(n0 * n1 * n2)
    worst = 0.0
    for transposed in (True, False):
        for sign in (-1, +1):
            xin = np.zeros(nalloc, dtype=lib.cplx)
            xin[:ln0 * n1 * n2] = xg[s0:s0 + ln0].ravel()
            x = torch.from_numpy(xin).cuda()
            y = torch.zeros_like(x)
            plan = L.fftw_mpi_plan_dft_3d(n0, n1, n2, x.data_ptr(), y.data_ptr(), comm, sign,
                                          capi.FFTW_ESTIMATE | ((1 << 30) if transposed else 0))
            assert plan, L.b2fft_last_error()
            lib.execute(plan)
            lib.execute(plan)  # plans are re-executable
            torch.cuda.synchronize()
            ref = orc.mpi_dft_3d_local(xg, world, rank, transposed_out=transposed, sign=sign)
            got = y.cpu().numpy()[:ref.size].reshape(ref.shape)
            err = float(np.linalg.norm(got - ref) / np.linalg.norm(ref)) if ref.size else 0.0
            worst = max(worst, err)
            assert err < tol, (rank, transposed, sign, err)
            assert np.array_equal(x.cpu().numpy(), xin), "out-of-place slab transform must preserve its input"
            lib.destroy_plan(plan)
            if transposed:  # in place on one alloc_local-sized array (how BASELINE config 5 is quoted)
                z = torch.from_numpy(xin).cuda()
                plan = L.fftw_mpi_plan_dft_3d(n0, n1, n2, z.data_ptr(), z.data_ptr(), comm, sign,
                                              capi.FFTW_ESTIMATE | (1 << 30))
                assert plan, L.b2fft_last_error()
                lib.execute(plan)
                torch.cuda.synchronize()
                got = z.cpu().numpy()[:ref.size].reshape(ref.shape)
                err = float(np.linalg.norm(got - ref) / np.linalg.norm(ref)) if ref.size else 0.0
                worst = max(worst, err)
                assert err < tol, (rank, "in place", sign, err)
                lib.destroy_plan(plan)
    # ---- real data: r2c (natural / transposed output) and c2r back (natural / transposed input)
    for nm in ("fftw_mpi_plan_dft_r2c_3d", "fftw_mpi_plan_dft_c2r_3d"):
        f = getattr(L, nm)
        f.restype = C.c_void_p
        f.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] * 3 + [C.c_uint]
    n2c = n2 // 2 + 1
    nalloc_r = L.fftw_mpi_local_size_3d_transposed(n0, n1, n2c, comm, *[C.byref(t) for t in v])
    ln0, s0, ln1, s1 = (t.value for t in v)
    xr = (rng.random(shape) - 0.5).astype(lib.real)
    Y = orc.r2c(xr)
    for transposed in (True, False):
        pad = np.zeros((ln0, n1, 2 * n2c), dtype=lib.real)
        pad[:, :, :n2] = xr[s0:s0 + ln0]
        rin = np.zeros(2 * nalloc_r, dtype=lib.real)
        rin[:pad.size] = pad.ravel()
        x = torch.from_numpy(rin).cuda()
        y = torch.zeros(nalloc_r, dtype=torch.complex64 if lib.single else torch.complex128, device="cuda")
        plan = L.fftw_mpi_plan_dft_r2c_3d(n0, n1, n2, x.data_ptr(), y.data_ptr(), comm,
                                          capi.FFTW_ESTIMATE | ((1 << 30) if transposed else 0))
        assert plan, L.b2fft_last_error()
        lib.execute(plan)
        lib.execute(plan)
        torch.cuda.synchronize()
        ref = Y.transpose(1, 0, 2)[s1:s1 + ln1] if transposed else Y[s0:s0 + ln0]
        got = y.cpu().numpy()[:ref.size].reshape(ref.shape)
        err = float(np.linalg.norm(got - ref) / np.linalg.norm(ref)) if ref.size else 0.0
        worst = max(worst, err)
        assert err < tol, (rank, "r2c", transposed, err)
        lib.destroy_plan(plan)
        # and back: c2r from that layout
        z = torch.zeros(2 * nalloc_r, dtype=x.dtype, device="cuda")
        plan = L.fftw_mpi_plan_dft_c2r_3d(n0, n1, n2, y.data_ptr(), z.data_ptr(), comm,
                                          capi.FFTW_ESTIMATE | ((1 << 29) if transposed else 0))
        assert plan, L.b2fft_last_error()
        lib.execute(plan)
        lib.execute(plan)
        torch.cuda.synchronize()
        back = z.cpu().numpy()[:ln0 * n1 * 2 * n2c].reshape(ln0, n1, 2 * n2c)[:, :, :n2] / (n0 * n1 * n2)
        want = xr[s0:s0 + ln0]
        err = float(np.linalg.norm(back - want) / np.linalg.norm(want)) if want.size else 0.0
        worst = max(worst, err)
        assert err < tol, (rank, "c2r", transposed, err)
        lib.destroy_plan(plan)
    t = torch.tensor([worst], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"SLAB_OK world={world} shape={shape} {prec} worst_rel_l2={float(t.item()):.3e} tol={tol:.1e}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
