"""Extra seeded random geometries on the product library (tests/cases.check_fuzz): python scripts/gpu_fuzz_more.py A B"""
import sys
sys.path.insert(0, ".")
from fftw3_b200 import build, capi
from tests import cases
a, b = (int(v) for v in sys.argv[1:3])
for prec in ("fp32", "fp64"):
    lib = capi.FFTW3Lib(build.lib_path(prec), prec == "fp32")
    for seed in range(a, b):
        cases.check_fuzz(lib, 150, seed)
    print(prec, "fuzz ok", a, b)
