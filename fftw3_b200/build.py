"""Builds the shared objects of the B200 FFT engine, in tree.

  product : fftw3_b200/lib/{fp32,fp64}/libfftw3.so   nvcc, sm_100a only, -lineinfo
  emu     : tests/_emu/libfftw3_emu_{fp32,fp64}.so   g++ -DB2_EMULATE (TEST-ONLY: executes the
            kernels' CTA bodies in host loops so the selector and index arithmetic can be
            checked without a GPU; never loaded by the product, bench.py or smoke()).
One library per precision because the reference ABI has no name mangling (FFTW_SINGLE is a
compile-time switch, reference fftw/fftw3.h:65-82).
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
INC = os.path.join(ROOT, "include")
LIBDIR = os.path.join(HERE, "lib")
EMUDIR = os.path.join(ROOT, "tests", "_emu")
OBJDIR = os.path.join(HERE, "lib", "_obj")

CU_SOURCES = ["engine.cu", "mixtile.cu", "mixtile2.cu", "mixtile3.cu", "slab.cu", "split.cu", "r2r.cu", "noalloc.cu"]
C_SOURCES = ["fftw_api.c", "fftw_mpi_api.c"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _newest_source() -> float:
    t = 0.0
    for d in (CSRC, INC):
        for f in os.listdir(d):
            t = max(t, os.path.getmtime(os.path.join(d, f)))
    return max(t, os.path.getmtime(__file__))


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: " + " ".join(cmd) + "\n" + r.stdout)
    return r.stdout


def lib_path(prec: str) -> str:
    return os.path.join(LIBDIR, prec, "libfftw3.so")


def emu_path(prec: str) -> str:
    return os.path.join(EMUDIR, f"libfftw3_emu_{prec}.so")


def build_product(prec: str, force: bool = False, verbose: bool = False) -> str:
    out = lib_path(prec)
    if not force and os.path.exists(out) and os.path.getmtime(out) >= _newest_source():
        return out
    os.makedirs(os.path.dirname(out), exist_ok=True)
    od = os.path.join(OBJDIR, prec)
    os.makedirs(od, exist_ok=True)
    defs = ["-DFFTW_SINGLE=1"] if prec == "fp32" else []
    csrc = [s for s in C_SOURCES if os.path.exists(os.path.join(CSRC, s))]
    jobs = []
    for s in CU_SOURCES:
        o = os.path.join(od, s + ".o")
        jobs.append((["nvcc", "-O3", "-std=c++17", "-lineinfo", "--expt-relaxed-constexpr", *ARCH, "-Xcompiler", "-fPIC,-O3,-fno-gnu-unique",
                      *(["-Xptxas", "-v"] if verbose else []), *defs, "-I", INC, "-I", CSRC,
                      "-c", os.path.join(CSRC, s), "-o", o], o))
    for s in csrc:
        o = os.path.join(od, s + ".o")
        jobs.append((["gcc", "-O2", "-fPIC", "-std=c99", "-Wall", *defs, "-I", INC, "-I", CSRC,
                      "-c", os.path.join(CSRC, s), "-o", o], o))
    with ThreadPoolExecutor(max_workers=8) as ex:
        logs = list(ex.map(lambda j: _run(j[0]), jobs))
    if verbose:
        print("\n".join(logs))
    _run(["nvcc", "-shared", *ARCH, "-o", out, *[j[1] for j in jobs], "-Xlinker", "-Bsymbolic",
          "-lcudart_static", "-ldl", "-lpthread", "-lrt"])
    return out


def build_emu(prec: str, force: bool = False) -> str:
    out = emu_path(prec)
    if not force and os.path.exists(out) and os.path.getmtime(out) >= _newest_source():
        return out
    os.makedirs(EMUDIR, exist_ok=True)
    od = os.path.join(EMUDIR, "_obj_" + prec)
    os.makedirs(od, exist_ok=True)
    defs = ["-DB2_EMULATE=1"] + (["-DFFTW_SINGLE=1"] if prec == "fp32" else [])
    csrc = [s for s in C_SOURCES if os.path.exists(os.path.join(CSRC, s))]
    jobs = []
    for s in CU_SOURCES:
        o = os.path.join(od, s + ".o")
        jobs.append((["g++", "-O2", "-std=c++17", "-fPIC", "-fno-gnu-unique", "-x", "c++", *defs, "-I", INC, "-I", CSRC,
                      "-c", os.path.join(CSRC, s), "-o", o], o))
    for s in csrc:
        o = os.path.join(od, s + ".o")
        jobs.append((["gcc", "-O2", "-fPIC", "-std=c99", "-Wall", *defs, "-I", INC, "-I", CSRC,
                      "-c", os.path.join(CSRC, s), "-o", o], o))
    with ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(lambda j: _run(j[0]), jobs))
    _run(["g++", "-shared", "-o", out, *[j[1] for j in jobs], "-Wl,-Bsymbolic", "-lpthread"])
    return out


def build_all(force: bool = False, verbose: bool = False):
    # the two precisions are independent nvcc runs dominated by one translation unit each
    with ThreadPoolExecutor(max_workers=2) as ex:
        return list(ex.map(lambda p: build_product(p, force, verbose), ("fp32", "fp64")))


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    force = "--force" in sys.argv
    verbose = "-v" in sys.argv
    if what in ("all", "product"):
        print("\n".join(build_all(force, verbose)))
    if what in ("all", "emu"):
        for p in ("fp32", "fp64"):
            print(build_emu(p, force))
