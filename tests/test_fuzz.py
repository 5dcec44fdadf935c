"""Seeded random geometries (tests/cases.check_fuzz) on the CPU emulation build and, marked gpu, on the
product library.  The reference's own random tests do the same with sizes of factors <= 13
(fftw/tests/check.pl:186-196); here sizes with larger prime factors are mixed in as well."""
import pytest

from fftw3_b200 import build, capi
from tests import cases


@pytest.mark.parametrize("prec", ["fp32", "fp64"])
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_fuzz_emulation(prec, seed):
    lib = capi.FFTW3Lib(build.build_emu(prec), prec == "fp32")
    cases.check_fuzz(lib, 40, seed, max_points=1 << 15)


@pytest.mark.gpu
@pytest.mark.parametrize("prec", ["fp32", "fp64"])
@pytest.mark.parametrize("seed", [11, 12])
def test_fuzz_b200(prec, seed):
    lib = capi.FFTW3Lib(build.lib_path(prec), prec == "fp32")
    cases.check_fuzz(lib, 120, seed)
