"""The slab-distributed 3-D c2c leg of bench.py (BASELINE config 5): fftw_mpi_plan_dft_3d + fftw_execute over the
ranks of the run (one process per GPU; one rank = fftw_mpi_comm_self), in place on one alloc_local-sized array with
FFTW_MPI_TRANSPOSED_OUT -- how the reference quotes the distributed transform (fftw/mpi/dft-rank-geq2-transposed.c:47-70)
-- followed, outside the timed region, by the correctness checks SURVEY.md section 8(c) prescribes for a cube no CPU
oracle can hold (scatter -> distributed plan -> gather -> serial verifier is the reference's own pattern,
fftw/mpi/mpi-bench.c:213-257):

  * a small cube through the same ranks against numpy's double-precision FFT (full array, relative L2);
  * at the timed size: K plane waves -> K spikes of height N; >= 1000 output bins of a random input against direct
    double-precision summation on the GPU; Parseval; backward(forward(x)) = N x through a FFTW_MPI_TRANSPOSED_IN plan.

torch is used as device memory, random numbers, reductions for the CHECKS and torch.distributed plumbing only; every
transform goes through the C ABI.  Peak memory at 2048^3 on one GPU: the array (64 GiB) + the plan's receive buffer
(64 GiB) + chunk-sized temporaries.
"""
from __future__ import annotations

import ctypes as C
import math
import time

import numpy as np

FFTW_FORWARD, FFTW_BACKWARD = -1, +1
FFTW_ESTIMATE = 1 << 6
TRANSPOSED_IN, TRANSPOSED_OUT = 1 << 29, 1 << 30
GEN_CHUNK = 1 << 26  # complex elements per random-number call: regeneration repeats the same call sequence


def bind(L):
    L.fftw_mpi_comm_from_nccl.restype = C.c_void_p
    L.fftw_mpi_comm_from_nccl.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.fftw_mpi_comm_self.restype = C.c_void_p
    L.fftw_mpi_plan_dft_3d.restype = C.c_void_p
    L.fftw_mpi_plan_dft_3d.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] * 3 + [C.c_int, C.c_uint]
    L.fftw_mpi_local_size_3d_transposed.restype = C.c_ssize_t
    L.fftw_mpi_local_size_3d_transposed.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] + [C.POINTER(C.c_ssize_t)] * 4
    L.b2fft_slab_last_times.argtypes = [C.c_void_p] + [C.POINTER(C.c_float)] * 3
    L.b2fft_last_error.restype = C.c_char_p
    L.b2fft_launch_count.restype = C.c_ulonglong


def make_comm(L, world, rank):
    """MPI_Comm stand-in from an NCCL unique id passed through torch.distributed (include/fftw3-mpi.h)"""
    import torch
    import torch.distributed as dist
    if world == 1:
        return L.fftw_mpi_comm_self()
    idt = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        buf = (C.c_char * 128)()
        assert L.fftw_mpi_nccl_unique_id(buf) == 0, L.b2fft_last_error()
        idt = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
    idt = idt.cuda()
    dist.broadcast(idt, 0)
    comm = L.fftw_mpi_comm_from_nccl(bytes(idt.cpu().numpy().tobytes()), world, rank)
    assert comm, L.b2fft_last_error()
    return comm


def local_sizes(L, comm, n):
    v = [C.c_ssize_t() for _ in range(4)]
    nalloc = L.fftw_mpi_local_size_3d_transposed(n[0], n[1], n[2], comm, *[C.byref(t) for t in v])
    return (nalloc,) + tuple(t.value for t in v)  # nalloc, ln0, s0, ln1, s1


def fill_random(x, seed):
    """uniform [-0.5, 0.5) per real component (the reference verifier's distribution, libbench2/verify-lib.c:64-67),
    generated chunk by chunk so that regenerate-and-compare needs no second array"""
    import torch
    g = torch.Generator(device="cuda").manual_seed(seed)
    xr = torch.view_as_real(x)
    for a in range(0, x.numel(), GEN_CHUNK):
        xr[a:a + GEN_CHUNK].uniform_(-0.5, 0.5, generator=g)


def sumsq(x, sub=None, scale=1.0):
    """sum |scale * x - sub|^2 in double precision, chunk by chunk; sub: callable(a, b) -> chunk to subtract"""
    import torch
    tot = torch.zeros((), dtype=torch.float64, device="cuda")
    step = 1 << 25
    for a in range(0, x.numel(), step):
        c = x[a:a + step].to(torch.complex128)
        if scale != 1.0:
            c = c * scale
        if sub is not None:
            c = c - sub(a, min(a + step, x.numel()))
        tot += (c.real * c.real + c.imag * c.imag).sum()
    return tot


def allsum(t, world):
    import torch.distributed as dist
    if world > 1:
        dist.all_reduce(t)
    return t


def run(lib, n, world, rank, steps, warmup, peak_gbs):
    """times `steps` in-place executes of the slab plan; returns (record, state for the checks)"""
    import torch
    import torch.distributed as dist
    L = lib.lib
    bind(L)
    comm = make_comm(L, world, rank)
    nalloc, ln0, s0, ln1, s1 = local_sizes(L, comm, n)
    N = n[0] * n[1] * n[2]
    x = torch.empty(nalloc, dtype=torch.complex64, device="cuda")
    plan = L.fftw_mpi_plan_dft_3d(n[0], n[1], n[2], x.data_ptr(), x.data_ptr(), comm, FFTW_FORWARD,
                                  FFTW_ESTIMATE | TRANSPOSED_OUT)
    assert plan, L.b2fft_last_error()
    bp = C.cast(plan, C.POINTER(C.c_void_p))[0]
    fill_random(x, 1 + rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # the transform is in place and unnormalised: rescale between steps (outside the timed calls) or the data
    # overflows fp32 after a few steps
    inv = 1.0 / math.sqrt(N)
    for _ in range(max(3, warmup)):
        lib.execute(plan)
        x.mul_(inv)
    l0 = L.b2fft_launch_count()
    times, stage = [], np.zeros(3)
    for _ in range(steps):
        barrier()
        t0 = time.perf_counter()
        lib.execute(plan)  # synchronous: returns when every stream of the plan has drained
        dt = time.perf_counter() - t0
        a, b, c = C.c_float(), C.c_float(), C.c_float()
        L.b2fft_slab_last_times(bp, C.byref(a), C.byref(b), C.byref(c))
        stage += (a.value, b.value, c.value)
        times.append(dt)
        x.mul_(inv)
    launches = int(L.b2fft_launch_count() - l0)
    t = torch.tensor([*times, *(stage / steps)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)  # per step: the slowest rank
    t = t.cpu().numpy()
    ms = float(t[:steps].mean() * 1e3)
    stage = t[steps:]
    flops = 5.0 * N * math.log2(N)
    slab_bytes = 2.0 * N * 8 / world                      # algorithmic: read the slab once + write it once
    sent = (N * 8 / world) * (world - 1) / world          # bytes each GPU sends over NVLink
    rec = {
        "workload": "c2c_3d_%d_slab_fp32" % n[0], "n": list(n), "api": "fftw_mpi_plan_dft_3d + fftw_execute",
        "layout": "in place, FFTW_MPI_TRANSPOSED_OUT, slabs over n0, alloc_local = %d complex per rank" % nalloc,
        "n_gpus": world, "scaling": "strong", "steps": steps, "warmup": max(3, warmup),
        "ms_per_step": ms, "min_step_ms": float(t[:steps].min() * 1e3), "value": flops / (ms * 1e-3) / 1e9, "unit": "GFLOP/s",
        "timing": "host clock around the synchronous fftw_execute, barrier + device sync before, max over ranks per step",
        "stage_ms": {"row_pass": float(stage[0]), "pipelined_n1_exchange_n0": float(stage[1]), "tail_n0": float(stage[2])},
        "nvlink_bytes_sent_per_gpu": sent,
        "nvlink_frac_of_900GBs": (sent / (ms * 1e-3) / 900e9) if world > 1 else None,
        "nvlink_frac_of_770GBs_measured_peer_copy": (sent / (ms * 1e-3) / 770e9) if world > 1 else None,
        "hbm_frac": slab_bytes / (ms * 1e-3) / 1e9 / peak_gbs,
        "hbm_note": "algorithmic bytes of this GPU's slab (read once + write once) / step time / measured HBM peak; 3 passes => <= 1/3",
        "gpu_launches": launches,
        "plan": lib.sprint_plan(plan).strip().split("\n"),
    }
    state = dict(L=L, lib=lib, comm=comm, plan=plan, x=x, n=n, N=N, world=world, rank=rank,
                 nalloc=nalloc, ln0=ln0, s0=s0, ln1=ln1, s1=s1)
    return rec, state


def _ks(n):
    """ten sampled frequencies per axis: both ends, the middle, thirds"""
    c = [0, 1, 2, n // 3, n // 2 - 1, n // 2, n // 2 + 1, (2 * n) // 3, n - 2, n - 1]
    return sorted(set(int(k) % n for k in c))


def checks(st, tol_per_log2=1e-6):
    """correctness of the slab path at the timed size and on a small cube; consumes (destroys) the state's plan"""
    import torch
    L, lib, comm, x, n, N, world, rank = (st[k] for k in ("L", "lib", "comm", "x", "n", "N", "world", "rank"))
    ln0, s0, ln1, s1 = st["ln0"], st["s0"], st["ln1"], st["s1"]
    n0, n1, n2 = n
    plan = st["plan"]
    dev = x.device
    tol = tol_per_log2 * math.log2(N)
    out = {"tolerance": tol, "tolerance_rule": "relative L2 < 1e-6 * log2(N) (north star, fp32)"}
    nin, nout = ln0 * n1 * n2, ln1 * n0 * n2
    x3 = x[:nin].view(ln0, n1, n2) if ln0 else None
    i0g = torch.arange(s0, s0 + ln0, device=dev, dtype=torch.float64)

    def root(k, idx, nn, sign):  # exp(sign * 2 pi i k idx / nn), phases reduced exactly before the trig call
        ph = torch.remainder(idx * float(k), float(nn)) / float(nn)
        return torch.polar(torch.ones_like(ph), sign * 2.0 * math.pi * ph)

    # ---- 1. K plane waves -> K spikes of height N ----
    waves = [(3 % n0, 5 % n1, 7 % n2), ((n0 - 2) % n0, 11 % n1, n2 // 2), (17 % n0, n1 - 1, 1 % n2)]
    i1 = torch.arange(n1, device=dev, dtype=torch.float64)
    i2 = torch.arange(n2, device=dev, dtype=torch.float64)
    B = max(1, (1 << 25) // (n1 * n2))
    for a0 in range(0, ln0, B):
        blk = None
        for (a, b, c) in waves:
            e0 = root(a, i0g[a0:a0 + B], n0, +1).to(torch.complex64)
            e1 = root(b, i1, n1, +1).to(torch.complex64)
            e2 = root(c, i2, n2, +1).to(torch.complex64)
            t = (e0[:, None, None] * e1[None, :, None]) * e2[None, None, :]
            blk = t if blk is None else blk.add_(t)
        x3[a0:a0 + B] = blk
    del blk, t
    torch.cuda.synchronize()  # (the fill above is asynchronous torch work)
    lib.execute(plan)
    torch.cuda.synchronize()
    y = x[:nout]
    err = torch.zeros((), dtype=torch.float64, device=dev)
    saved = []
    for (a, b, c) in waves:
        if s1 <= b < s1 + ln1:
            idx = ((b - s1) * n0 + a) * n2 + c
            v = y[idx].to(torch.complex128)
            err += (v - N).abs() ** 2
            saved.append((idx, y[idx].clone()))
            y[idx] = 0
    err += sumsq(y)
    for idx, v in saved:
        y[idx] = v
    out["planewave_spikes"] = len(waves)
    out["planewave_rel_l2"] = float(torch.sqrt(allsum(err, world)).item() / (N * math.sqrt(len(waves))))

    # ---- 2. random input: sampled bins by direct fp64 summation, Parseval ----
    fill_random(x, 101 + rank)
    ssx = allsum(sumsq(x[:nin]), world)
    k0s, k1s, k2s = _ks(n0), _ks(n1), _ks(n2)
    W2 = torch.stack([root(k, i2, n2, -1) for k in k2s], dim=1)      # [n2][K2]
    W1 = torch.stack([root(k, i1, n1, -1) for k in k1s], dim=1)      # [n1][K1]
    W0 = torch.stack([root(k, i0g, n0, -1) for k in k0s], dim=1) if ln0 else None  # [ln0][K0]
    acc = torch.zeros((len(k0s), len(k1s), len(k2s)), dtype=torch.complex128, device=dev)
    B = max(1, (1 << 24) // (n1 * n2))
    for a0 in range(0, ln0, B):
        xb = x3[a0:a0 + B].to(torch.complex128)
        t2 = (xb.reshape(-1, n2) @ W2).view(xb.shape[0], n1, len(k2s))
        t1 = torch.einsum("bjc,ja->bac", t2, W1)
        acc += torch.einsum("bac,bz->zac", t1, W0[a0:a0 + B])
    del xb, t2, t1
    acc = allsum(acc, world)
    torch.cuda.synchronize()
    lib.execute(plan)
    torch.cuda.synchronize()
    got = torch.zeros_like(acc)
    y3 = x[:nout].view(ln1, n0, n2) if ln1 else None
    for ia, b in enumerate(k1s):
        if s1 <= b < s1 + ln1:
            row = y3[b - s1]
            got[:, ia, :] = row[k0s][:, k2s].to(torch.complex128)
    got = allsum(got, world)
    out["sampled_bins"] = int(acc.numel())
    out["sampled_bins_rel_l2"] = float((torch.linalg.norm(got - acc) / torch.linalg.norm(acc)).item())
    ssy = allsum(sumsq(x[:nout]), world)
    out["parseval_rel"] = float(abs(ssy.item() / (N * ssx.item()) - 1.0))

    # ---- 3. backward(forward(x)) = N x, through a FFTW_MPI_TRANSPOSED_IN plan (natural output) ----
    lib.destroy_plan(plan)  # (its receive buffer goes with it: one GPU cannot hold two at 2048^3)
    st["plan"] = None
    pb = L.fftw_mpi_plan_dft_3d(n0, n1, n2, x.data_ptr(), x.data_ptr(), comm, FFTW_BACKWARD, FFTW_ESTIMATE | TRANSPOSED_IN)
    assert pb, L.b2fft_last_error()
    torch.cuda.synchronize()
    lib.execute(pb)
    torch.cuda.synchronize()
    lib.destroy_plan(pb)
    g = torch.Generator(device="cuda").manual_seed(101 + rank)
    d2 = torch.zeros((), dtype=torch.float64, device=dev)
    for a in range(0, x.numel(), GEN_CHUNK):
        m = min(GEN_CHUNK, x.numel() - a)
        ref = torch.empty((m, 2), dtype=torch.float32, device=dev).uniform_(-0.5, 0.5, generator=g)
        hi = min(a + m, nin)
        if hi > a:
            d = x[a:hi].to(torch.complex128) / N - torch.view_as_complex(ref)[:hi - a].to(torch.complex128)
            d2 += (d.real * d.real + d.imag * d.imag).sum()
    out["roundtrip_rel_l2"] = float(math.sqrt(allsum(d2, world).item() / ssx.item()))

    # ---- 4. a small cube through the same ranks against numpy's double-precision FFT (full array) ----
    m = min(256, n0)
    ns = (m, m, m)
    nalloc, l0, t0, l1, t1 = local_sizes(L, comm, ns)
    rng = np.random.default_rng(11)
    xg = ((rng.random(ns, dtype=np.float32) - 0.5) + 1j * (rng.random(ns, dtype=np.float32) - 0.5)).astype(np.complex64)
    xin = np.zeros(nalloc, dtype=np.complex64)
    xin[:l0 * m * m] = xg[t0:t0 + l0].ravel()
    xs = torch.from_numpy(xin).cuda()
    ps = L.fftw_mpi_plan_dft_3d(m, m, m, xs.data_ptr(), xs.data_ptr(), comm, FFTW_FORWARD, FFTW_ESTIMATE | TRANSPOSED_OUT)
    assert ps, L.b2fft_last_error()
    lib.execute(ps)
    torch.cuda.synchronize()
    lib.destroy_plan(ps)
    ref = np.fft.fftn(xg.astype(np.complex128)).transpose(1, 0, 2)[t1:t1 + l1]
    gotc = xs.cpu().numpy()[:l1 * m * m].reshape(l1, m, m).astype(np.complex128)
    e = torch.tensor([float(np.sum(np.abs(gotc - ref) ** 2)), float(np.sum(np.abs(ref) ** 2))], dtype=torch.float64, device=dev)
    e = allsum(e, world)
    out["small_cube"] = [m, m, m]
    out["small_cube_rel_l2_vs_numpy_fp64"] = float(math.sqrt(e[0].item() / e[1].item()))
    out["small_cube_tolerance"] = tol_per_log2 * math.log2(m ** 3)

    worst = max(out["planewave_rel_l2"], out["sampled_bins_rel_l2"], out["parseval_rel"], out["roundtrip_rel_l2"])
    out["worst"] = worst
    out["pass"] = bool(worst < tol and out["small_cube_rel_l2_vs_numpy_fp64"] < out["small_cube_tolerance"])
    return out
