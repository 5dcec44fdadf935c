"""The C-ABI shared objects load and export every symbol that include/*.h declares
(no compute calls: there is no GPU here), and the product fails LOUDLY without a CUDA device
instead of falling back to the CPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from fftw3_b200 import build, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for hdr, pat in (("fftw3.h", r"\b(fftw_[a-z0-9_]+)\s*(?:\(|\[)"),
                     ("b200fft.h", r"\b(b2fft_[a-z0-9_]+)\s*\("),
                     ("fftw3-mpi.h", r"\b(fftw_mpi_[a-z0-9_]+)\s*\(")):
        path = os.path.join(ROOT, "include", hdr)
        if not os.path.exists(path):
            continue
        text = open(path).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        for line in text.splitlines():
            if line.lstrip().startswith("#") or "typedef" in line:
                continue
            for m in re.finditer(pat, line):
                nm = m.group(1)
                if nm.endswith("_do_not_use_me") or nm in ("fftw_plan", "fftw_iodim", "fftw_iodim64",
                                                           "fftw_r2r_kind", "fftw_plan_s", "fftw_mpi_ddim"):
                    continue
                names.add(nm)
    return sorted(names)


@pytest.fixture(scope="module", params=["fp32", "fp64"])
def product(request):
    return build.build_product(request.param), request.param


def test_every_declared_symbol_is_exported(product):
    path, prec = product
    lib = C.CDLL(path, mode=C.RTLD_LOCAL)
    syms = declared_symbols()
    assert len(syms) > 80
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing
    lib.b2fft_precision.restype = C.c_char_p
    assert lib.b2fft_precision().decode() == prec
    ver = (C.c_char * 7).in_dll(lib, "fftw_version")  # data symbol (a char array, fftw_api.c:28)
    assert ver.raw == b"b200fft"


def test_product_has_no_cpu_fallback(product):
    path, prec = product
    lib = capi.FFTW3Lib(path, prec == "fp32")
    lib.lib.b2fft_device_count.restype = C.c_int
    if lib.lib.b2fft_device_count() > 0:
        pytest.skip("a CUDA device is present")
    x = np.zeros(16, dtype=lib.cplx)
    assert not lib.plan_dft_1d(16, x, x.copy(), -1), "plan must be NULL without a GPU"
    lib.lib.b2fft_last_error.restype = C.c_char_p
    assert b"no CUDA device" in lib.lib.b2fft_last_error()


def test_product_contains_only_sm100a_code(product):
    import subprocess
    path, _ = product
    out = subprocess.run(["cuobjdump", "-lelf", path], stdout=subprocess.PIPE, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def _build_dropin(prec, lib_path, tmp_path):
    """examples/dropin.c: a C99 program written against include/fftw3.h only (-pedantic -Werror)"""
    import subprocess
    exe = str(tmp_path / f"dropin_{prec}")
    cmd = ["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", *(["-DFFTW_SINGLE"] if prec == "fp32" else []),
           "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "examples", "dropin.c"), "-o", exe, lib_path, "-lm",
           "-Wl,-rpath," + os.path.dirname(lib_path)]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    return exe


@pytest.mark.parametrize("prec", ["fp32", "fp64"])
def test_c_program_against_the_public_header_runs_on_the_emulation_build(prec, tmp_path):
    import subprocess
    from fftw3_b200 import build
    exe = _build_dropin(prec, build.build_emu(prec), tmp_path)
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 0 and "dropin: ok" in r.stdout, r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("prec", ["fp32", "fp64"])
def test_c_program_against_the_public_header_runs_on_b200(prec, tmp_path):
    import subprocess
    from fftw3_b200 import build
    exe = _build_dropin(prec, build.lib_path(prec), tmp_path)
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 0 and "dropin: ok" in r.stdout, r.stdout
