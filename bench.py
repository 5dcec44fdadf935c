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
    # lengths that are not powers of two (mixed-radix tile kernels, csrc/mixtile*.cu): not BASELINE configs -- side records
    # ("smooth_lengths" in the JSON line), about 1 GiB of data each
    "c2c_1d_1000x134217_fp32": dict(kind="c2c", n=[1000], howmany=134217, prec="fp32"),
    "c2c_1d_3072x43690_fp32": dict(kind="c2c", n=[3072], howmany=43690, prec="fp32"),
    "c2c_1d_1152x116508_fp32": dict(kind="c2c", n=[1152], howmany=116508, prec="fp32"),
    "c2c_1d_1920x69905_fp32": dict(kind="c2c", n=[1920], howmany=69905, prec="fp32"),
    "c2c_1d_100000x1342_fp32": dict(kind="c2c", n=[100000], howmany=1342, prec="fp32"),
    "c2c_1d_1000x67108_fp64": dict(kind="c2c", n=[1000], howmany=67108, prec="fp64"),
    "r2c_1d_2000x134217_fp32": dict(kind="r2c", n=[2000], howmany=134217, prec="fp32"),
    "r2c_1d_3840x69905_fp32": dict(kind="r2c", n=[3840], howmany=69905, prec="fp32"),
}
DEFAULT_WORKLOAD = "c2c_1d_4096x65536_fp32"
SMOOTH_SET = ["c2c_1d_1000x134217_fp32", "c2c_1d_3072x43690_fp32", "c2c_1d_1152x116508_fp32", "c2c_1d_1920x69905_fp32",
              "c2c_1d_100000x1342_fp32", "c2c_1d_1000x67108_fp64", "r2c_1d_2000x134217_fp32", "r2c_1d_3840x69905_fp32"]


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


def cpu_sample_of(w, name=None):
    """what the reference arm runs: the FULL workload where the host can hold it and a step stays around 0.1 s
    (the headline config: 65536 transforms of 4096 points, 2 GiB in + 2 GiB out), else a bounded sample of it
    (same transform, fewer of them); the record says which"""
    s = dict(w)
    N = int(np.prod(w["n"]))
    if name == DEFAULT_WORKLOAD:
        return s
    if w["howmany"] > 1:
        # ~1.5 GFLOP-conv per repetition keeps a rep well under a second on a few cores
        s["howmany"] = max(1, min(w["howmany"], int(1.5e9 / (5.0 * N * math.log2(N))) or 1))
    elif len(w["n"]) == 3 and N > 256 ** 3:
        s["n"] = [256, 256, 256]
    return s


def fill_uniform(a, rng, chunk=1 << 24):
    """uniform [-0.5, 0.5) per real component, in place, in chunks (no array-sized temporaries)"""
    flat = a.reshape(-1).view(np.float32 if a.dtype in (np.complex64, np.float32) else np.float64)
    for i in range(0, flat.size, chunk):
        flat[i:i + chunk] = rng.random(min(chunk, flat.size - i), dtype=np.float32 if flat.dtype == np.float32 else np.float64) - 0.5


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
        fill_uniform(x, rng)
    elif w["kind"] == "r2c":
        x = np.zeros((hm, N), dtype=lib.real)
        y = np.zeros((hm, N // n[-1] * (n[-1] // 2 + 1)), dtype=lib.cplx)
        p = lib.plan_many_dft_r2c(n, hm, x, None, 1, N, y, None, 1, y.shape[1], flags)
        fill_uniform(x, rng)
    else:
        y = np.zeros((hm, N), dtype=lib.real)
        x = np.zeros((hm, N // n[-1] * (n[-1] // 2 + 1)), dtype=lib.cplx)
        p = lib.plan_many_dft_c2r(n, hm, x, None, 1, x.shape[1], y, None, 1, N, flags)
        fill_uniform(x, rng)
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
            "ms_per_step": dt * 1e3, "howmany": hm, "n": list(n)}


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


def device_arrays(lib, w, seed):
    """device-resident synthetic input (uniform [-0.5, 0.5), like the reference verifier) and output array"""
    import torch
    (ishape, idt), (oshape, odt) = io_shapes(lib, w)
    tmap = {np.complex64: torch.complex64, np.complex128: torch.complex128, np.float32: torch.float32, np.float64: torch.float64}
    g = torch.Generator(device="cuda").manual_seed(seed)
    rdt = torch.float32 if w["prec"] == "fp32" else torch.float64
    if idt in (np.complex64, np.complex128):
        x = torch.view_as_complex(torch.empty(ishape + (2,), device="cuda", dtype=rdt).uniform_(-0.5, 0.5, generator=g))
    else:
        x = torch.empty(ishape, device="cuda", dtype=rdt).uniform_(-0.5, 0.5, generator=g)
    if w["kind"] == "c2r":  # a half-spectrum of REAL data: the DC and Nyquist bins are real
        x[..., 0] = x[..., 0].real
        x[..., -1] = x[..., -1].real
    y = x if w.get("inplace") else torch.empty(oshape, device="cuda", dtype=tmap[odt])
    return x, y


def validate(lib, w, x, y, plan):
    """the bench refuses to post a number for garbage: after the timed loop, the plan's output is compared with
    numpy's double-precision FFT on sampled rows (batched 1-D) or, for a 3-D cube, with direct double-precision
    summation at 64 sampled bins plus Parseval (the checker is numpy / torch arithmetic, never the timed path)"""
    import torch
    n, hm = w["n"], w["howmany"]
    N = int(np.prod(n))
    tol = (1e-6 if w["prec"] == "fp32" else 1e-14) * math.log2(N)
    if w.get("inplace"):  # re-seed: the timed loop transformed the array over and over
        x2, _ = device_arrays(lib, w, 1)
        x.copy_(x2)
        del x2
    rows = sorted({0, hm // 2, hm - 1}) if len(n) == 1 else []
    rows_in = {r: x[r].cpu().numpy() for r in rows}  # (before the execute: out-of-place c2r may destroy its input)
    xin = x.clone() if (w.get("inplace") and len(n) > 1) else x
    lib.execute(plan)
    torch.cuda.synchronize()
    errs = []
    if len(n) == 1:
        for r in rows:
            a = rows_in[r]
            got = y[r].cpu().numpy().astype(np.complex128 if w["kind"] != "c2r" else np.float64)
            if w["kind"] == "c2c":
                ref = np.fft.fft(a.astype(np.complex128))
            elif w["kind"] == "r2c":
                ref = np.fft.rfft(a.astype(np.float64))
            else:
                ref = np.fft.irfft(a.astype(np.complex128), n[0]) * n[0]
            errs.append(float(np.linalg.norm(got - ref) / np.linalg.norm(ref)))
    else:
        xd = xin.reshape(n).to(torch.complex128)
        ks = [sorted({0, 1, m // 2, m - 1}) for m in n]
        Ws = [torch.polar(torch.ones(m, len(k), dtype=torch.float64, device="cuda"),
                          -2.0 * math.pi * torch.remainder(torch.arange(m, device="cuda", dtype=torch.float64)[:, None]
                                                           * torch.tensor(k, device="cuda", dtype=torch.float64)[None, :], m) / m)
              for m, k in zip(n, ks)]
        t = torch.einsum("abc,cz->abz", xd, Ws[2])
        t = torch.einsum("abz,by->ayz", t, Ws[1])
        ref = torch.einsum("ayz,ax->xyz", t, Ws[0])
        got = y.reshape(n)[ks[0]][:, ks[1]][:, :, ks[2]].to(torch.complex128)
        errs.append(float((torch.linalg.norm(got - ref) / torch.linalg.norm(ref)).item()))
        sx = float((xd.real ** 2 + xd.imag ** 2).sum().item())
        yd = y.reshape(n).to(torch.complex128)
        errs.append(abs(float((yd.real ** 2 + yd.imag ** 2).sum().item()) / (N * sx) - 1.0))
        del xd, yd, t
    worst = max(errs)
    assert worst < tol, f"bench output check failed for {w}: {errs} (tolerance {tol})"
    return {"worst_rel_err": worst, "tolerance": tol,
            "against": "numpy fp64 FFT of sampled rows" if len(n) == 1 else "fp64 direct summation at 64 sampled bins + Parseval"}


def time_device(lib, w, plan, x, y, steps, warmup, world=1):
    """device-resident timing with CUDA events on the launching stream.  Out of place: ONE bracket around exactly
    `steps` steps.  In place (unnormalised transform applied to its own output): per-step event pairs with a
    rescale between them, outside the pairs, so that the data stays finite."""
    import torch
    import torch.distributed as dist
    bp = fftw_plan_bp(plan)
    stream = torch.cuda.current_stream()
    sp = C.c_void_p(stream.cuda_stream)
    inplace = bool(w.get("inplace"))
    inv = 1.0 / math.sqrt(float(np.prod(w["n"])))

    def step():
        rc = lib.lib.b2fft_execute_async(bp, x.data_ptr(), y.data_ptr(), sp)
        assert rc == 0, lib.lib.b2fft_last_error()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, warmup)):
        step()
        if inplace:
            x.mul_(inv)
    barrier()
    l0 = lib.lib.b2fft_launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * steps)]
    with ClockSampler(torch.cuda.current_device()) as clk:
        for i in range(steps):
            ev[2 * i].record(stream)
            step()
            ev[2 * i + 1].record(stream)
            if inplace:
                x.mul_(inv)
        barrier()
    launches = int(lib.lib.b2fft_launch_count() - l0)
    per = sorted(ev[2 * i].elapsed_time(ev[2 * i + 1]) for i in range(steps))
    total_ms = sum(per) if inplace else ev[0].elapsed_time(ev[-1])
    if world > 1:
        t = torch.tensor([total_ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    return total_ms / steps, per, launches, clk.summary()


def traffic_of(name):
    """measured DRAM bytes per step of the workload's kernels (ncu --set full captures, profiles/traffic.json)"""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get(name)
        if t:
            return t["dram_bytes_per_launch"], "%s, %s" % (t["kernel"], t["source"])
    except (OSError, ValueError, KeyError):
        pass
    return None, None


def config_record(name, steps, warmup):
    """one compact, validated record per BASELINE config (device-resident arrays, this GPU)"""
    import torch
    w = WORKLOADS[name]
    lib = engine_lib(w["prec"])
    flops, nbytes = workload_numbers(w)
    x, y = device_arrays(lib, w, 1)
    plan = plan_for(lib, w, x, y)
    assert plan, "plan is NULL: " + lib.lib.b2fft_last_error().decode()
    ms, per, launches, _ = time_device(lib, w, plan, x, y, steps, warmup)
    peak, _ = measured_peak_gbs()
    rec = {"workload": name, "n": w["n"], "howmany": w["howmany"], "in_place": bool(w.get("inplace")),
           "dtype": "f32" if w["prec"] == "fp32" else "f64", "ms_per_step": ms, "min_step_ms": per[0],
           "gflops": flops / (ms * 1e-3) / 1e9, "global_passes": int(lib.lib.b2fft_plan_num_passes(fftw_plan_bp(plan))),
           "algorithmic_bytes": nbytes, "hbm_frac": nbytes / (ms * 1e-3) / 1e9 / peak,
           "timed_call": "b2fft_execute_async on torch's current stream, CUDA events",
           "kernels_per_step": launches / steps, "traffic": traffic_of(name)[0]}
    if name == "c2c_1d_1024_fp64":  # BASELINE configs[0]: one transform, latency-bound -- also through fftw_execute itself
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(200):
            lib.execute_dft(plan, x, y)  # synchronous fftw_execute_dft on device arrays
        rec["fftw_execute_sync_us"] = (time.perf_counter() - t0) / 200 * 1e6
        rec["hbm_frac"] = None
        rec["note"] = "32 KiB of traffic: launch-latency bound; fftw_execute_sync_us = wall time of one synchronous fftw_execute_dft call"
    rec["check"] = validate(lib, w, x, y, plan)
    rec["plan"] = [l.strip()[:160] for l in lib.sprint_plan(plan).strip().split("\n")[1:]]
    lib.destroy_plan(plan)
    del x, y
    torch.cuda.empty_cache()
    return rec


CONFIG_SET = ["c2c_1d_1024_fp64", "r2c_1d_1048576x256_fp32", "c2r_1d_1048576x256_fp32", "c2c_3d_512_fp32", "c2c_3d_512_fp32_inplace"]


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
    inplace = bool(w.get("inplace"))

    # ---- headline: device-resident arrays, K steps in one CUDA-event bracket, max over ranks
    x, y = device_arrays(lib, w, 1 + rank)
    plan = plan_for(lib, w, x, y)
    assert plan, "plan is NULL: " + lib.lib.b2fft_last_error().decode()
    bp = fftw_plan_bp(plan)
    ms_step, per, launches, clocks = time_device(lib, w, plan, x, y, args.steps, args.warmup, world)
    value = world * flops / (ms_step * 1e-3) / 1e9
    npasses = lib.lib.b2fft_plan_num_passes(bp)
    peak, peak_src = measured_peak_gbs()
    achieved = nbytes / (ms_step * 1e-3) / 1e9
    plan_desc = lib.sprint_plan(plan)
    check = validate(lib, w, x, y, plan)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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
               "d2h_bytes_per_step": ob, "ms_per_step": dt * 1e3, "steps": esteps,
               "h2d_GBs_per_rank": ib / dt / 1e9, "d2h_GBs_per_rank": ob / dt / 1e9,
               "timed_call": "fftw_execute on fftw_malloc (pinned host) arrays, host clock, max over ranks"}
        lib.destroy_plan(hplan)
        lib.free(hin)
        if not inplace:
            lib.free(hout)
    lib.destroy_plan(plan)
    del x, y
    torch.cuda.empty_cache()

    # ---- the other single-GPU BASELINE configs, one compact validated record each (rank 0's GPU)
    configs = None
    if not args.no_configs and args.workload == DEFAULT_WORKLOAD:
        if rank == 0:
            configs = []
            for name in CONFIG_SET:
                try:
                    configs.append(config_record(name, steps=10, warmup=3))
                except Exception as e:  # a failing side record must not take the headline line with it
                    configs.append({"workload": name, "error": repr(e)[:300]})
        barrier()

    # ---- BASELINE config 5: the slab-distributed cube over all ranks of this run (strong scaling)
    slab = None
    if not args.no_slab and args.workload == DEFAULT_WORKLOAD:
        from fftw3_b200 import bench_slab
        try:
            free_b, _ = torch.cuda.mem_get_info()
            if world > 1:
                t = torch.tensor([float(free_b)], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MIN)  # every rank must pick the same cube
                free_b = int(t.item())
            cube = args.slab_n
            # in place: the array + the plan's receive buffer, 2 x 8 N^3 / world bytes per GPU (+ check temporaries)
            while cube > 256 and 2 * 8 * cube ** 3 / world + (6 << 30) > free_b:
                cube //= 2
            slab, st = bench_slab.run(lib, (cube, cube, cube), world, rank, steps=args.slab_steps, warmup=3, peak_gbs=peak)
            if cube != args.slab_n:
                slab["note"] = "%d^3 does not fit %d GPU(s) with %.0f GiB free each: ran %d^3" % (args.slab_n, world, free_b / GIB, cube)
            slab["parity"] = None if args.no_slab_checks else bench_slab.checks(st)
            del st
        except Exception as e:
            slab = {"workload": "c2c_3d_%d_slab_fp32" % args.slab_n, "error": repr(e)[:400]}
        torch.cuda.empty_cache()

    # ---- lengths that are not powers of two: side records, LAST of the GPU work so that whatever happens here cannot touch
    # the headline, the BASELINE configs or the slab record above
    smooth = None
    if not args.no_configs and args.workload == DEFAULT_WORKLOAD:
        if rank == 0:
            smooth = []
            for name in SMOOTH_SET:
                try:
                    smooth.append(config_record(name, steps=10, warmup=3))
                except Exception as e:
                    smooth.append({"workload": name, "error": repr(e)[:300]})
            torch.cuda.empty_cache()
        barrier()

    # ---- the reference's own CPU path on this box's host cores (rank 0; the other ranks wait)
    cpu = None
    if rank == 0 and not args.no_cpu:
        cpu = run_reference(cpu_sample_of(w, args.workload), steps=3, warmup=1, measure=True)
        if cpu:
            for k in ("ms_per_step", "howmany", "n"):
                cpu.pop(k, None)
    if rank == 0:
        traffic, traffic_note = traffic_of(args.workload)
        line = {
            "metric": "FFTW-convention GFLOP/s (5 N log2 N / t), fftw_plan_many_dft + execute",
            "value": value, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if w["prec"] == "fp32" else "f64", "data": "synthetic uniform[-0.5,0.5), seed 1",
            "timed_call": "value: b2fft_execute_async (the engine entry fftw_execute calls) on torch's current stream, device "
                          "arrays, CUDA events; e2e: fftw_execute itself on pinned host arrays",
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
            "output_check": check,
            "clocks": clocks,
            "gpu_launches": launches,
            "plan": plan_desc.strip().split("\n"),
        }
        if e2e:
            line["e2e"] = e2e
        if cpu:
            line["cpu_baseline"] = cpu
        if configs is not None:
            line["configs"] = configs
        if slab is not None:
            line["slab"] = slab
        if smooth is not None:
            line["smooth_lengths"] = smooth
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_slab(args):
    """a slab workload on its own (--workload c2c_3d_*_slab_fp32): the slab record is the line"""
    import torch
    import torch.distributed as dist
    from fftw3_b200 import bench_slab

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    w = WORKLOADS[args.workload]
    lib = engine_lib(w["prec"])
    lib.lib.b2fft_set_device(local)
    peak, peak_src = measured_peak_gbs()
    with ClockSampler(local) as clk:
        rec, st = bench_slab.run(lib, tuple(w["n"]), world, rank, steps=args.steps, warmup=args.warmup, peak_gbs=peak)
    parity = None if args.no_slab_checks else bench_slab.checks(st)
    if rank == 0:
        line = {"metric": "FFTW-convention GFLOP/s (5 N log2 N / t), fftw_mpi_plan_dft_3d + fftw_execute",
                "value": rec["value"], "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
                "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic uniform[-0.5,0.5)",
                "config": {"workload": args.workload, "n": w["n"], "layout": rec["layout"], "timing": rec["timing"],
                           "stage_ms": rec["stage_ms"]},
                "roofline": {"bound": "hbm", "achieved": rec["hbm_frac"] * peak, "peak": peak, "unit": "GB/s",
                             "frac": rec["hbm_frac"], "traffic": None, "peak_source": peak_src, "note": rec["hbm_note"],
                             "nvlink_bytes_sent_per_gpu": rec["nvlink_bytes_sent_per_gpu"],
                             "nvlink_frac_of_900GBs": rec["nvlink_frac_of_900GBs"]},
                "slab": dict(rec, parity=parity), "clocks": clk.summary(), "gpu_launches": rec["gpu_launches"], "plan": rec["plan"]}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    w = WORKLOADS[args.workload]
    s = cpu_sample_of(w, args.workload)
    same = s["howmany"] == w["howmany"] and list(s["n"]) == list(w["n"])
    r = run_reference(s, steps=max(1, args.steps), warmup=max(1, min(args.warmup, 3)), measure=True)
    if r is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built (run __graft_entry__.build() where /root/reference exists)"}))
        return
    line = {"impl": "reference", "metric": "FFTW-convention GFLOP/s (5 N log2 N / t), fftw_execute",
            "value": r["value"], "unit": "GFLOP/s", "n_gpus": int(os.environ.get("WORLD_SIZE", 1)),
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if w["prec"] == "fp32" else "f64", "data": "synthetic uniform[-0.5,0.5), seed 1",
            "config": {"workload": args.workload, "n": r["n"], "howmany": r["howmany"], "howmany_of_the_config": w["howmany"],
                       "same_config": same,
                       "note": "the reference's own fftw_plan_many_dft + fftw_execute on the host cores, on the "
                               + ("full config" if same else "bounded sample named in cpu_baseline.sample")},
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
    ap.add_argument("--no-configs", action="store_true", help="skip the compact records of the other single-GPU configs")
    ap.add_argument("--no-slab", action="store_true", help="skip the slab-distributed cube (BASELINE config 5)")
    ap.add_argument("--no-slab-checks", action="store_true")
    ap.add_argument("--slab-n", type=int, default=2048)
    ap.add_argument("--slab-steps", type=int, default=6)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    elif WORKLOADS[args.workload]["kind"] == "slab":
        run_slab(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
