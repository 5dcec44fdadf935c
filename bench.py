#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 FFT engine (contract: see DESIGN.md section 6).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A "step" is one fftw_execute of the workload's plan.  Default workload = BASELINE.json
configs[1]: batched 1-D c2c fp32, n = 4096 x 65536 via fftw_plan_many_dft (4 GiB of
algorithmic traffic per step: larger than the 126 MB L2, so no flush is needed between steps).

  value     : FFTW-convention GFLOP/s = 5 N log2 N * howmany / t (fftw/libbench2/mflops.c:6-31),
              arrays resident in HBM, timed with CUDA events on the launching stream
  e2e       : same metric through fftw_execute with HOST (pinned, fftw_malloc) arrays:
              host->device and device->host copies inside the timed region
  roofline  : algorithmic bytes / kernel time against the measured HBM copy bandwidth
  cpu_baseline / --impl reference : the UNMODIFIED reference (oracle/_ref, built from
              /root/reference by oracle/refbuild/Makefile) with threads + AVX2 on the host cores
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import math
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from fftw3_b200 import capi  # noqa: E402
from fftw3_b200 import build as b2build  # noqa: E402

GIB = float(1 << 30)

WORKLOADS = {
    # name: (kind, dims, howmany, dtype)
    "c2c_1d_4096x65536_fp32": dict(kind="c2c", n=[4096], howmany=65536, prec="fp32"),
    "c2c_1d_1024_fp64": dict(kind="c2c", n=[1024], howmany=1, prec="fp64"),
    "r2c_1d_1048576x256_fp32": dict(kind="r2c", n=[1 << 20], howmany=256, prec="fp32"),
    "c2r_1d_1048576x256_fp32": dict(kind="c2r", n=[1 << 20], howmany=256, prec="fp32"),
    "c2c_3d_512_fp32": dict(kind="c2c", n=[512, 512, 512], howmany=1, prec="fp32"),
    "c2c_3d_512_fp32_inplace": dict(kind="c2c", n=[512, 512, 512], howmany=1, prec="fp32", inplace=True),
    # slab-distributed 3-D (fftw_mpi_plan_dft_3d, FFTW_MPI_TRANSPOSED_OUT): the WHOLE cube is split
    # over the ranks (strong scaling); 2048^3 needs >= 2 GPUs
    "c2c_3d_2048_slab_fp32": dict(kind="slab", n=[2048, 2048, 2048], howmany=1, prec="fp32"),
    "c2c_3d_1024_slab_fp32": dict(kind="slab", n=[1024, 1024, 1024], howmany=1, prec="fp32"),
    "c2c_3d_512_slab_fp32": dict(kind="slab", n=[512, 512, 512], howmany=1, prec="fp32"),
}
DEFAULT_WORKLOAD = "c2c_1d_4096x65536_fp32"


def workload_numbers(w):
    """flops by the FFTW convention and algorithmic bytes (read input once + write output once)"""
    N = int(np.prod(w["n"]))
    hm = w["howmany"]
    rb = 4 if w["prec"] == "fp32" else 8
    if w["kind"] in ("c2c", "slab"):
        flops = 5.0 * N * math.log2(N) * hm
        nbytes = 2.0 * N * hm * 2 * rb
    else:
        flops = 2.5 * N * math.log2(N) * hm
        last = w["n"][-1]
        rows = N // last * hm
        nbytes = rows * (last * rb + (last // 2 + 1) * 2 * rb)
    return flops, nbytes


class ClockSampler:
    """samples SM clock and throttle reasons during the timed region (pynvml)"""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.002)  # the timed region of the default workload is ~15 ms

    def __enter__(self):
        if self.nv:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------------------- reference arm
def reference_lib(prec):
    """the UNMODIFIED reference build (oracle/_ref), else None"""
    path = os.path.join(ROOT, "oracle", "_ref", f"libfftw3_ref_{'f32' if prec == 'fp32' else 'f64'}.so")
    if not os.path.exists(path):
        return None
    return capi.FFTW3Lib(path, prec == "fp32")


def cpu_sample_of(w, budget_s=20.0):
    """a bounded sample of the workload: same transform, fewer of them"""
    s = dict(w)
    N = int(np.prod(w["n"]))
    if w["howmany"] > 1:
        # ~1.5 GFLOP-conv per repetition keeps a rep well under a second on a few cores
        s["howmany"] = max(1, min(w["howmany"], int(1.5e9 / (5.0 * N * math.log2(N))) or 1))
    elif len(w["n"]) == 3 and N > 256 ** 3:
        s["n"] = [256, 256, 256]
    return s


def run_reference(w, steps, warmup, threads=None, measure=True):
    """times the reference's own CPU implementation on this host: threads + AVX2 + FFTW_MEASURE
    (oracle/_ref is built with -DWITH_SLOW_TIMER so MEASURE really measures)"""
    lib = reference_lib(w["prec"])
    kind = "reference"
    cores = threads or (os.cpu_count() or 1)
    if lib is None:
        return None
    lib.lib.fftw_init_threads()
    lib.plan_with_nthreads(cores)
    n, hm = w["n"], w["howmany"]
    N = int(np.prod(n))
    rng = np.random.default_rng(1)
    flags = capi.FFTW_MEASURE if measure else capi.FFTW_ESTIMATE
    lib.lib.fftw_set_timelimit.argtypes = [C.c_double]
    lib.lib.fftw_set_timelimit(20.0)
    if w["kind"] in ("c2c", "slab"):
        x = np.zeros((hm, N), dtype=lib.cplx)
        y = x if w.get("inplace") else np.zeros_like(x)
        p = lib.plan_many_dft(n, hm, x, None, 1, N, y, None, 1, N, capi.FFTW_FORWARD, flags)
        x[:] = ((rng.random(x.shape) - 0.5) + 1j * (rng.random(x.shape) - 0.5)).astype(lib.cplx)
    elif w["kind"] == "r2c":
        x = np.zeros((hm, N), dtype=lib.real)
        y = np.zeros((hm, N // n[-1] * (n[-1] // 2 + 1)), dtype=lib.cplx)
        p = lib.plan_many_dft_r2c(n, hm, x, None, 1, N, y, None, 1, y.shape[1], flags)
        x[:] = (rng.random(x.shape) - 0.5).astype(lib.real)
    else:
        y = np.zeros((hm, N), dtype=lib.real)
        x = np.zeros((hm, N // n[-1] * (n[-1] // 2 + 1)), dtype=lib.cplx)
        p = lib.plan_many_dft_c2r(n, hm, x, None, 1, x.shape[1], y, None, 1, N, flags)
        x[:] = ((rng.random(x.shape) - 0.5) + 1j * (rng.random(x.shape) - 0.5)).astype(lib.cplx)
    assert p, "reference plan is NULL"
    for _ in range(warmup):
        lib.execute(p)
    t0 = time.perf_counter()
    for _ in range(steps):
        lib.execute(p)
    dt = (time.perf_counter() - t0) / steps
    lib.destroy_plan(p)
    flops, _ = workload_numbers(w)
    return {"value": flops / dt / 1e9, "unit": "GFLOP/s", "cores": cores, "kind": kind,
            "sample": f"{w['kind']} n={'x'.join(map(str, n))} howmany={hm} {w['prec']}, "
                      f"{'FFTW_MEASURE' if measure else 'FFTW_ESTIMATE'}, {cores} threads, AVX2",
            "ms_per_step": dt * 1e3}


# ------------------------------------------------------------------------------- GPU arm
def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def engine_lib(prec):
    path = b2build.lib_path(prec)
    if not os.path.exists(path):
        path = b2build.build_product(prec)
    L = capi.FFTW3Lib(path, prec == "fp32")
    L.lib.b2fft_launch_count.restype = C.c_ulonglong
    L.lib.b2fft_execute_async.restype = C.c_int
    L.lib.b2fft_execute_async.argtypes = [C.c_void_p] * 4
    L.lib.b2fft_last_error.restype = C.c_char_p
    L.lib.b2fft_plan_num_passes.argtypes = [C.c_void_p]
    return L


def plan_for(lib, w, x, y):
    n, hm = w["n"], w["howmany"]
    N = int(np.prod(n))
    if w["kind"] == "c2c":
        return lib.plan_many_dft(n, hm, x, None, 1, N, y, None, 1, N, capi.FFTW_FORWARD)
    hc = N // n[-1] * (n[-1] // 2 + 1)
    if w["kind"] == "r2c":
        return lib.plan_many_dft_r2c(n, hm, x, None, 1, N, y, None, 1, hc)
    return lib.plan_many_dft_c2r(n, hm, x, None, 1, hc, y, None, 1, N)


def io_shapes(lib, w):
    n, hm = w["n"], w["howmany"]
    N = int(np.prod(n))
    hc = N // n[-1] * (n[-1] // 2 + 1)
    if w["kind"] == "c2c":
        return ((hm, N), lib.cplx), ((hm, N), lib.cplx)
    if w["kind"] == "r2c":
        return ((hm, N), lib.real), ((hm, hc), lib.cplx)
    return ((hm, hc), lib.cplx), ((hm, N), lib.real)


def fftw_plan_bp(plan):
    """struct fftw_plan_s { b2fft_plan *bp; ... }: first member (csrc/fftw_api.c)"""
    return C.cast(plan, C.POINTER(C.c_void_p))[0]


def run_gpu(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    w = WORKLOADS[args.workload]
    lib = engine_lib(w["prec"])
    lib.lib.b2fft_set_device(local)
    flops, nbytes = workload_numbers(w)
    (ishape, idt), (oshape, odt) = io_shapes(lib, w)
    tmap = {np.complex64: torch.complex64, np.complex128: torch.complex128,
            np.float32: torch.float32, np.float64: torch.float64}

    # ---- device-resident arrays: synthetic uniform [-0.5, 0.5) like the reference verifier
    g = torch.Generator(device="cuda").manual_seed(1 + rank)
    rdt = torch.float32 if w["prec"] == "fp32" else torch.float64
    if idt in (np.complex64, np.complex128):
        x = torch.view_as_complex(torch.rand(ishape + (2,), device="cuda", dtype=rdt, generator=g) - 0.5)
    else:
        x = torch.rand(ishape, device="cuda", dtype=rdt, generator=g) - 0.5
    inplace = bool(w.get("inplace"))
    y = x if inplace else torch.empty(oshape, device="cuda", dtype=tmap[odt])
    plan = plan_for(lib, w, x, y)
    assert plan, "plan is NULL: " + lib.lib.b2fft_last_error().decode()
    bp = fftw_plan_bp(plan)
    stream = torch.cuda.current_stream()
    sp = C.c_void_p(stream.cuda_stream)

    def step_dev():
        rc = lib.lib.b2fft_execute_async(bp, x.data_ptr(), y.data_ptr(), sp)
        assert rc == 0, lib.lib.b2fft_last_error()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step_dev()
    barrier()
    l0 = lib.lib.b2fft_launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    with ClockSampler(local) as clk:
        ev[0].record(stream)
        for i in range(args.steps):
            step_dev()
            ev[i + 1].record(stream)
        barrier()
    launches = int(lib.lib.b2fft_launch_count() - l0)
    total_ms = ev[0].elapsed_time(ev[-1])
    per = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps))
    if world > 1:
        t = torch.tensor([total_ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_step = total_ms / args.steps
    value = world * flops / (ms_step * 1e-3) / 1e9
    npasses = lib.lib.b2fft_plan_num_passes(bp)
    peak, peak_src = measured_peak_gbs()
    achieved = nbytes / (ms_step * 1e-3) / 1e9
    plan_desc = lib.sprint_plan(plan)

    # ---- end to end: host (pinned) arrays through fftw_execute, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        ib = int(np.prod(ishape)) * np.dtype(idt).itemsize
        ob = int(np.prod(oshape)) * np.dtype(odt).itemsize
        hin = lib.malloc(ib)
        hout = hin if inplace else lib.malloc(ob)
        assert hin and hout, "fftw_malloc failed"
        hx = np.ctypeslib.as_array(C.cast(hin, C.POINTER(C.c_byte)), shape=(ib,)).view(idt).reshape(ishape)
        hx[...] = x.cpu().numpy()
        hplan = plan_for(lib, w, hin, hout)
        assert hplan
        esteps = max(1, min(args.steps, args.e2e_steps))
        for _ in range(2):
            lib.execute(hplan)
        barrier()
        t0 = time.perf_counter()
        for _ in range(esteps):
            lib.execute(hplan)  # synchronous: returns when the result is in host memory
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / esteps
        if world > 1:
            t = torch.tensor([dt], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e = {"value": world * flops / dt / 1e9, "unit": "GFLOP/s", "h2d_bytes_per_step": ib,
               "d2h_bytes_per_step": ob, "ms_per_step": dt * 1e3, "steps": esteps}
        lib.destroy_plan(hplan)
        lib.free(hin)
        if not inplace:
            lib.free(hout)
    lib.destroy_plan(plan)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        s = cpu_sample_of(w)
        cpu = run_reference(s, steps=3, warmup=1, measure=True)
        if cpu:
            cpu.pop("ms_per_step", None)
    if rank == 0:
        traffic = traffic_note = None
        try:  # measured DRAM bytes per launch of the dominant kernel (one ncu --set full capture, profiles/)
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                t = json.load(f).get(args.workload)
            if t:
                traffic, traffic_note = t["dram_bytes_per_launch"], "%s, %s" % (t["kernel"], t["source"])
        except (OSError, ValueError, KeyError):
            pass
        line = {
            "metric": "FFTW-convention GFLOP/s (5 N log2 N / t), fftw_execute",
            "value": value, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if w["prec"] == "fp32" else "f64", "data": "synthetic uniform[-0.5,0.5), seed 1",
            "config": {"workload": args.workload, "api": "fftw_plan_many_dft + fftw_execute" if w["kind"] == "c2c"
                       else f"fftw_plan_many_dft_{w['kind']} + fftw_execute",
                       "n": w["n"], "howmany_per_gpu": w["howmany"], "in_place": inplace,
                       "l2": "inputs+outputs %.2f GiB per step >> 126 MB L2, no flush needed" % (nbytes / GIB)
                       if nbytes > 512e6 else "working set fits L2 (latency-bound config)",
                       "global_passes": int(npasses), "sharding": "batch split across ranks, no collective"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_note,
                         "peak_source": peak_src, "algorithmic_bytes_per_step": nbytes, "kernels_per_step": launches / args.steps,
                         "median_step_ms": per[len(per) // 2], "min_step_ms": per[0]},
            "clocks": clk.summary(),
            "gpu_launches": launches,
            "plan": plan_desc.strip().split("\n"),
        }
        if e2e:
            line["e2e"] = e2e
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_slab(args):
    """3-D c2c split over the ranks by slabs: local 2-D FFTs, NCCL all-to-all, 1-D FFTs"""
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    w = WORKLOADS[args.workload]
    lib = engine_lib(w["prec"])
    L = lib.lib
    L.b2fft_set_device(local)
    L.fftw_mpi_comm_from_nccl.restype = C.c_void_p
    L.fftw_mpi_comm_from_nccl.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.fftw_mpi_comm_self.restype = C.c_void_p
    L.fftw_mpi_plan_dft_3d.restype = C.c_void_p
    L.fftw_mpi_plan_dft_3d.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] * 3 + [C.c_int, C.c_uint]
    L.fftw_mpi_local_size_3d_transposed.restype = C.c_ssize_t
    L.fftw_mpi_local_size_3d_transposed.argtypes = [C.c_ssize_t] * 3 + [C.c_void_p] + [C.POINTER(C.c_ssize_t)] * 4
    L.b2fft_slab_last_times.argtypes = [C.c_void_p] + [C.POINTER(C.c_float)] * 3
    if world > 1:
        idt = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            buf = (C.c_char * 128)()
            assert L.fftw_mpi_nccl_unique_id(buf) == 0, L.b2fft_last_error()
            idt = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
        idt = idt.cuda()
        dist.broadcast(idt, 0)
        idb = bytes(idt.cpu().numpy().tobytes())
        comm = L.fftw_mpi_comm_from_nccl(idb, world, rank)
    else:
        comm = L.fftw_mpi_comm_self()
    assert comm, L.b2fft_last_error()
    n0, n1, n2 = w["n"]
    v = [C.c_ssize_t() for _ in range(4)]
    nalloc = L.fftw_mpi_local_size_3d_transposed(n0, n1, n2, comm, *[C.byref(t) for t in v])
    ln0, s0, ln1, s1 = (t.value for t in v)
    g = torch.Generator(device="cuda").manual_seed(1 + rank)
    x = torch.view_as_complex(torch.rand((nalloc, 2), device="cuda", dtype=torch.float32, generator=g) - 0.5)
    y = torch.empty_like(x)
    plan = L.fftw_mpi_plan_dft_3d(n0, n1, n2, x.data_ptr(), y.data_ptr(), comm, capi.FFTW_FORWARD,
                                  capi.FFTW_ESTIMATE | (1 << 30))
    assert plan, L.b2fft_last_error()
    bp = fftw_plan_bp(plan)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        lib.execute(plan)
    barrier()
    l0 = L.b2fft_launch_count()
    stage = np.zeros(3)
    with ClockSampler(local) as clk:
        t0 = time.perf_counter()
        for _ in range(args.steps):
            lib.execute(plan)  # synchronous on the plan's streams
            a, b, c = C.c_float(), C.c_float(), C.c_float()
            L.b2fft_slab_last_times(bp, C.byref(a), C.byref(b), C.byref(c))
            stage += (a.value, b.value, c.value)
        barrier()
        dt = (time.perf_counter() - t0) / args.steps
    launches = int(L.b2fft_launch_count() - l0)
    stage /= args.steps
    if world > 1:
        t = torch.tensor([dt, *stage], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt, stage = float(t[0]), t[1:].cpu().numpy()
    flops, nbytes = workload_numbers(w)
    peak, peak_src = measured_peak_gbs()
    sent = (n0 * n1 * n2 * 8 / world) * (world - 1) / world  # bytes each GPU sends
    if rank == 0:
        line = {
            "metric": "FFTW-convention GFLOP/s (5 N log2 N / t), fftw_mpi_plan_dft_3d + fftw_execute",
            "value": flops / dt / 1e9, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic uniform[-0.5,0.5)",
            "config": {"workload": args.workload, "n": w["n"], "layout": "FFTW_MPI_TRANSPOSED_OUT, slabs over n0",
                       "timing": "host clock around synchronous fftw_execute, barrier + max over ranks",
                       "stage_ms": {"local_2d_ffts": float(stage[0]), "exposed_exchange": float(stage[1]),
                                    "final_1d_ffts": float(stage[2])}},
            "roofline": {"bound": "hbm", "achieved": nbytes / world / dt / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": nbytes / world / dt / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                         "note": "per GPU: algorithmic bytes of its slab / whole step",
                         "nvlink_bytes_sent_per_gpu": sent,
                         "nvlink_min_ms_at_770GBs": sent / 770e9 * 1e3},
            "clocks": clk.summary(), "gpu_launches": launches,
            "plan": lib.sprint_plan(plan).strip().split("\n"),
        }
        print(json.dumps(line))
    lib.destroy_plan(plan)
    if world > 1:
        dist.destroy_process_group()


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    w = WORKLOADS[args.workload]
    s = cpu_sample_of(w)
    r = run_reference(s, steps=max(1, args.steps), warmup=max(1, min(args.warmup, 3)), measure=True)
    if r is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built (run __graft_entry__.build() where /root/reference exists)"}))
        return
    line = {"impl": "reference", "metric": "FFTW-convention GFLOP/s (5 N log2 N / t), fftw_execute",
            "value": r["value"], "unit": "GFLOP/s", "n_gpus": int(os.environ.get("WORLD_SIZE", 1)),
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if w["prec"] == "fp32" else "f64", "data": "synthetic uniform[-0.5,0.5), seed 1",
            "config": {"workload": args.workload, "n": w["n"], "howmany_per_gpu": w["howmany"]},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=5)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    elif WORKLOADS[args.workload]["kind"] == "slab":
        run_slab(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
