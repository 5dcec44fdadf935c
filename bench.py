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
}
DEFAULT_WORKLOAD = "c2c_1d_4096x65536_fp32"


def workload_numbers(w):
    """flops by the FFTW convention and algorithmic bytes (read input once + write output once)"""
    N = int(np.prod(w["n"]))
    hm = w["howmany"]
    rb = 4 if w["prec"] == "fp32" else 8
    if w["kind"] == "c2c":
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
            self._stop.wait(0.02)

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
