"""b2fft_c2c_1d_static_scratch (include/b200fft.h, csrc/noalloc.cu): the entry point that allocates nothing
-- its intermediate array and twiddle table live in a static __device__ array of the library.  Checked
through the C ABI against the existing entry point (fftw_plan_many_dft + fftw_execute on device arrays),
within the tolerance of the other parity tests, on two calls in a row that share the static array.
Collected last (file name) so that a failure here cannot hide another test under pytest -x."""
import ctypes as C
import os

import numpy as np
import pytest

from fftw3_b200 import build, capi
from tests import cases

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["fp32", "fp64"])
def lib(request):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    path = build.lib_path(request.param)
    assert os.path.exists(path), "product library missing: run __graft_entry__.build()"
    L = capi.FFTW3Lib(path, request.param == "fp32")
    f = L.lib.b2fft_c2c_1d_static_scratch
    f.restype = C.c_int
    f.argtypes = [C.c_ssize_t, C.c_ssize_t, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.lib.b2fft_last_error.restype = C.c_char_p
    return L


def _existing(lib, x, sign):
    """the existing entry point on device arrays"""
    import torch
    hm, n = x.shape
    y = torch.empty_like(x)
    p = lib.plan_many_dft([n], hm, x, None, 1, n, y, None, 1, n, sign)
    assert p, "plan is NULL"
    lib.execute(p)
    torch.cuda.synchronize()
    lib.destroy_plan(p)
    return y.cpu().numpy()


def test_static_scratch_entry_point_matches_the_plan_path_on_two_calls_in_a_row(lib):
    import torch
    rng = np.random.default_rng(5)
    f = lib.lib.b2fft_c2c_1d_static_scratch
    stream = torch.cuda.current_stream().cuda_stream
    # call 1: forward, out of place, the largest n (256 x 256), fills a quarter of the work area
    # call 2: backward, IN PLACE, a shorter odd-log2 n (64 x 32) over a different part of the array, right behind call 1
    xs, outs, refs = [], [], []
    for (n, hm, sign, inplace) in ((65536, 4, -1, False), (2048, 37, +1, True)):
        xh = cases.rand_c(rng, (hm, n), lib)
        x = torch.from_numpy(xh).cuda()
        refs.append((_existing(lib, x, sign), n))
        y = x.clone() if inplace else torch.empty_like(x)
        xs.append((x if not inplace else y, y, n, hm, sign))
    for (x, y, n, hm, sign) in xs:  # back to back on one stream: the second call reuses the first one's array
        rc = f(n, hm, x.data_ptr(), y.data_ptr(), sign, stream)
        assert rc == 0, lib.lib.b2fft_last_error()
    torch.cuda.synchronize()
    for (x, y, n, hm, sign), (ref, _) in zip(xs, refs):
        err = cases.rel_l2(y.cpu().numpy(), ref)
        assert err < cases.tol(lib, n), (n, hm, sign, err)
    # and once more with the first geometry: the array now holds call 2's leftovers
    x, y, n, hm, sign = xs[0]
    y.zero_()
    assert f(n, hm, x.data_ptr(), y.data_ptr(), sign, stream) == 0
    torch.cuda.synchronize()
    assert cases.rel_l2(y.cpu().numpy(), refs[0][0]) < cases.tol(lib, n)


def test_static_scratch_entry_point_rejects_sizes_beyond_its_array(lib):
    import torch
    f = lib.lib.b2fft_c2c_1d_static_scratch
    x = torch.zeros(4096, dtype=torch.complex64 if lib.single else torch.complex128, device="cuda")
    assert f(1 << 17, 1, x.data_ptr(), x.data_ptr(), -1, None) != 0       # n beyond the table
    assert b"power of two" in lib.lib.b2fft_last_error()
    assert f(65536, 17, x.data_ptr(), x.data_ptr(), -1, None) != 0        # n * howmany beyond the work area
    assert b"exceeds the static scratch" in lib.lib.b2fft_last_error()
    assert f(1000, 1, x.data_ptr(), x.data_ptr(), -1, None) != 0          # not a power of two
    assert f(4096, 1, x.data_ptr(), x.data_ptr(), -1, None) == 0          # and a legal call still works
    torch.cuda.synchronize()
