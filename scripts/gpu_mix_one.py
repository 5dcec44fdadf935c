"""Three mixed-radix kernels, two launches each (1 GiB of fp32 complex data, in place), no torch: the process ncu captures for
profiles/r2_ncu_full_mix.txt.  Launch order: 1000-point rows x2, 1536-point rows x2, 1000-point columns x2."""
import ctypes as C
import sys

sys.path.insert(0, ".")
from fftw3_b200 import build, capi

lib = capi.FFTW3Lib(build.lib_path("fp32"), True)
L = lib.lib
L.b2fft_device_malloc.restype = C.c_void_p
L.b2fft_device_malloc.argtypes = [C.c_size_t]
L.b2fft_device_free.argtypes = [C.c_void_p]
tot = 1 << 27
buf = L.b2fft_device_malloc(tot * 8)
assert buf, "device allocation failed"
for n, col in ((1000, False), (1536, False), (1000, True)):
    if col:
        w = (tot // n) // 64 * 64
        p = lib.plan_many_dft([n], w, buf, None, w, 1, buf, None, w, 1, -1)
    else:
        rows = tot // n
        p = lib.plan_many_dft([n], rows, buf, None, 1, n, buf, None, 1, n, -1)
    assert p
    print(lib.sprint_plan(p).split("\n")[1].strip())
    lib.execute(p)
    lib.execute(p)
    lib.destroy_plan(p)
L.b2fft_device_free(buf)
print("ok")
