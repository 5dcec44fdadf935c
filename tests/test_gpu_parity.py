"""GPU parity tests proper: the product library (fftw3_b200/lib/*/libfftw3.so, sm_100a CUDA
kernels) driven through the fftw3.h C ABI, compared with the oracle and with the golden
vectors of the reference.  Host-pointer cases go through the PCIe staging path; device-pointer
cases (torch tensors, used only as device memory) run in HBM with no copies."""
import ctypes as C
import math

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
    import os
    assert os.path.exists(path), "product library missing: run __graft_entry__.build()"
    L = capi.FFTW3Lib(path, request.param == "fp32")
    L.lib.b2fft_launch_count.restype = C.c_ulonglong
    return L


@pytest.mark.parametrize("shape", cases.C2C_SHAPES + cases.C2C_BIG + cases.C2C_PRIME + cases.C2C_PIPE, ids=lambda s: "x".join(map(str, s)))
def test_c2c(lib, shape):
    n0 = lib.lib.b2fft_launch_count()
    cases.check_c2c(lib, shape)
    assert lib.lib.b2fft_launch_count() > n0, "no CUDA kernel was launched"


@pytest.mark.parametrize("shape", [(2 ** 23,), (3 * 2 ** 22,), (2, 2 ** 22)], ids=lambda s: "x".join(map(str, s)))
def test_c2c_three_pass_lengths(lib, shape):
    """beyond the four-step limit (2^22 fp32 / 2^20 fp64): n = n1*n2*n3, three HBM passes"""
    cases.check_c2c(lib, shape)
    x = np.zeros(shape, dtype=lib.cplx)
    p = lib.plan_dft(list(shape), x, x.copy(), -1)
    assert lib.sprint_plan(p).count("pass ") >= 3
    lib.destroy_plan(p)


def test_r2c_c2r_three_pass_length(lib):
    """n/2 beyond the four-step limit: 2^24 reals (fp32) / 2^22 (fp64)"""
    cases.check_r2c_c2r(lib, (2 ** 24,) if lib.single else (2 ** 22,))


@pytest.mark.parametrize("shape", cases.R2C_SHAPES + cases.R2C_BIG + cases.R2C_PRIME, ids=lambda s: "x".join(map(str, s)))
def test_r2c_c2r(lib, shape):
    cases.check_r2c_c2r(lib, shape)


@pytest.mark.parametrize("shape", [(10,), (16,), (3, 10), (4, 6, 8), (7,), (2 ** 15,), (2 ** 19,)],
                         ids=lambda s: "x".join(map(str, s)))
def test_r2c_inplace(lib, shape):
    cases.check_r2c_inplace(lib, shape)


def test_many_strided(lib):
    cases.check_many_strided(lib)
    cases.check_many_strided(lib, n=64, howmany=9, istride=1, idist=70, ostride=1, odist=64)


def test_embedded_2d(lib):
    cases.check_embedded_2d(lib)


def test_batched_transposed(lib):
    cases.check_batched_transposed(lib)
    cases.check_batched_transposed(lib, n=4096, howmany=6)


def test_many_r2c(lib):
    cases.check_many_r2c(lib)


def test_golden_vectors_of_the_reference(lib, golden):
    cases.check_golden(lib, golden)


def test_split_array_guru_interfaces(lib):
    cases.check_split_arrays(lib)
    cases.check_split_arrays(lib, device=True)
    cases.check_split_interleaved_host(lib)


@pytest.mark.parametrize("n,cols", cases.MIX_LENGTHS + cases.MIX_LENGTHS_FP32_ONLY, ids=lambda v: str(v))
def test_mixed_radix_tile_kernels(lib, n, cols):
    if not lib.single and (n, cols) in cases.MIX_LENGTHS_FP32_ONLY:
        pytest.skip("beyond the fp64 row cap")
    cases.check_mixed_radix(lib, n, cols)


@pytest.mark.parametrize("n,cols", cases.MIX_SECOND_SHAPES + cases.MIX_SECOND_SHAPES_FP32_ONLY, ids=lambda v: str(v))
def test_mixed_radix_second_shapes(lib, n, cols):
    if not lib.single:
        pytest.skip("second shapes exist in the fp32 library only")
    cases.check_mixed_radix_second_shapes(lib, n, cols)


@pytest.mark.parametrize("N", cases.MIX_REAL_HALF + cases.MIX_REAL_HALF_FP32_ONLY)
def test_mixed_radix_real_rows(lib, N):
    if not lib.single and N in cases.MIX_REAL_HALF_FP32_ONLY:
        pytest.skip("beyond the fp64 row cap")
    cases.check_mixed_radix_real(lib, N)


@pytest.mark.parametrize("n,cols", cases.MIX_MORE + cases.MIX_MORE_FP32_ONLY, ids=lambda v: str(v))
def test_mixed_radix_second_list(lib, n, cols):
    if not lib.single and (n, cols) in cases.MIX_MORE_FP32_ONLY:
        pytest.skip("beyond the fp64 row cap")
    cases.check_mixed_radix(lib, n, cols)


@pytest.mark.parametrize("N", cases.MIX_MORE_REAL_HALF)
def test_mixed_radix_second_list_real_rows(lib, N):
    cases.check_mixed_radix_real(lib, N)


@pytest.mark.parametrize("n", cases.MIX_E60)
def test_mixed_radix_sixty_points_per_thread(lib, n):
    if not lib.single:
        pytest.skip("60 points per thread do not fit the register file in double precision")
    cases.check_mixed_radix(lib, n, False)
    if 2 * n <= 8192 * 2:
        cases.check_mixed_radix_real(lib, n)


@pytest.mark.parametrize("n", cases.MIX_MORE_COLS)
def test_mixed_radix_third_list_column_forms(lib, n):
    cases.check_mixed_radix(lib, n, True)


def test_mixed_radix_frames_and_boxes(lib):
    cases.check_mixed_radix_frames(lib)


def test_mixed_radix_long_lengths(lib):
    cases.check_mixed_radix_long(lib)


def test_negative_strides_real_transforms(lib):
    cases.check_negative_strides_real(lib)
    cases.check_negative_strides_real(lib, device=True)


def test_measure_and_wisdom(lib):
    cases.check_measure_and_wisdom(lib)
    cases.check_measure_and_wisdom(lib, device=True)


def test_negative_strides(lib):
    cases.check_negative_strides(lib)
    cases.check_negative_strides(lib, device=True)


def test_r2r_kinds(lib, r2r_golden):
    cases.check_r2r(lib, r2r_golden)
    cases.check_r2r(lib, r2r_golden, device=True)


def test_threads_share_a_plan_and_the_planner(lib):
    cases.check_threads(lib)


def test_in_place_with_different_strides(lib):
    cases.check_inplace_relayout(lib)
    cases.check_inplace_relayout(lib, device=True)


def test_edge_cases(lib):
    cases.check_edge_cases(lib)


@pytest.mark.parametrize("shape", [(96,), (16, 20), (2 ** 14,), (2 ** 20,), (128, 128, 128)],
                         ids=lambda s: "x".join(map(str, s)))
def test_verifier_properties(lib, shape):
    cases.check_properties(lib, shape)


# ---------------------------------------------------------------- device-resident arrays
def _tdtype(lib):
    import torch
    return (torch.complex64, torch.float32) if lib.single else (torch.complex128, torch.float64)


def test_device_pointers_batched_4096(lib):
    """config 2 shape at reduced batch: fftw_plan_many_dft n=4096, contiguous batch, in HBM"""
    import torch
    ct, _ = _tdtype(lib)
    howmany, n = 256, 4096
    rng = np.random.default_rng(11)
    xh = cases.rand_c(rng, (howmany, n), lib)
    x = torch.from_numpy(xh).cuda()
    y = torch.empty_like(x)
    p = lib.plan_many_dft([n], howmany, x, None, 1, n, y, None, 1, n, capi.FFTW_FORWARD)
    assert p
    lib.execute(p)
    torch.cuda.synchronize()
    got = y.cpu().numpy()
    from oracle import pyoracle as orc
    assert cases.rel_l2(got, orc.dft_batched(xh, -1)) < cases.tol(lib, n)
    assert torch.equal(x.cpu(), torch.from_numpy(xh)), "input must be preserved"
    # new-array execute with the same plan (fftw_execute_dft)
    x2 = torch.from_numpy(cases.rand_c(rng, (howmany, n), lib)).cuda()
    y2 = torch.empty_like(x2)
    lib.execute_dft(p, x2, y2)
    torch.cuda.synchronize()
    assert cases.rel_l2(y2.cpu().numpy(), np.fft.fft(x2.cpu().numpy().astype(np.complex128), axis=1)) < cases.tol(lib, n)
    lib.destroy_plan(p)


def test_full_size_config2_4096x65536_fp32(lib):
    """BASELINE.json configs[1] at its full size (2^28 points, 2 GiB in + 2 GiB out): 64 sampled transforms
    against the oracle, Parseval over the whole batch, backward(forward(x)) = N x."""
    import torch
    if not lib.single:
        pytest.skip("config 2 is single precision")
    n, howmany = 4096, 65536
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.view_as_complex(torch.rand((howmany, n, 2), device="cuda", dtype=torch.float32, generator=g) - 0.5)
    y = torch.empty_like(x)
    p = lib.plan_many_dft([n], howmany, x, None, 1, n, y, None, 1, n, capi.FFTW_FORWARD)
    assert p
    lib.execute(p)
    torch.cuda.synchronize()
    rows = np.random.default_rng(0).choice(howmany, 64, replace=False)
    rows.sort()
    idx = torch.from_numpy(rows).cuda()
    from oracle import pyoracle as orc
    ref = orc.dft_batched(x[idx].cpu().numpy(), -1)
    assert cases.rel_l2(y[idx].cpu().numpy(), ref) < cases.tol(lib, n)
    ex = float((x.real.double() ** 2 + x.imag.double() ** 2).sum())
    ey = float((y.real.double() ** 2 + y.imag.double() ** 2).sum())
    assert abs(ey / (n * ex) - 1) < 1e-5, "Parseval"
    pb = lib.plan_many_dft([n], howmany, y, None, 1, n, y, None, 1, n, capi.FFTW_BACKWARD)
    assert pb
    lib.execute(pb)
    torch.cuda.synchronize()
    y /= n
    num = float(((y - x).real.double() ** 2 + (y - x).imag.double() ** 2).sum())
    assert (num / ex) ** 0.5 < cases.tol(lib, n), "round trip"
    lib.destroy_plan(p)
    lib.destroy_plan(pb)


def test_device_3d_in_place_and_out_of_place(lib):
    import torch
    rng = np.random.default_rng(12)
    shape = (64, 64, 64)
    xh = cases.rand_c(rng, shape, lib)
    ref = np.fft.fftn(xh.astype(np.complex128))
    x = torch.from_numpy(xh).cuda()
    y = torch.empty_like(x)
    p = lib.plan_dft_3d(*shape, x, y, capi.FFTW_FORWARD)
    lib.execute(p)
    lib.destroy_plan(p)
    torch.cuda.synchronize()
    assert cases.rel_l2(y.cpu().numpy(), ref) < cases.tol(lib, 64 ** 3)
    p = lib.plan_dft_3d(*shape, x, x, capi.FFTW_FORWARD)
    lib.execute(p)
    lib.destroy_plan(p)
    torch.cuda.synchronize()
    assert cases.rel_l2(x.cpu().numpy(), ref) < cases.tol(lib, 64 ** 3)


def test_device_r2c_c2r_large_1d(lib):
    """config 3 shape at reduced batch: n = 2^20 real, four-step half-length path"""
    import torch
    rng = np.random.default_rng(13)
    n, howmany = 2 ** 20, 3
    xh = cases.rand_r(rng, (howmany, n), lib)
    x = torch.from_numpy(xh).cuda()
    X = torch.empty((howmany, n // 2 + 1), dtype=_tdtype(lib)[0], device="cuda")
    p = lib.plan_many_dft_r2c([n], howmany, x, None, 1, n, X, None, 1, n // 2 + 1)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    torch.cuda.synchronize()
    ref = np.fft.rfft(xh.astype(np.float64), axis=1)
    assert cases.rel_l2(X.cpu().numpy(), ref) < cases.tol(lib, n)
    z = torch.empty_like(x)
    p = lib.plan_many_dft_c2r([n], howmany, X, None, 1, n // 2 + 1, z, None, 1, n)
    assert p
    lib.execute(p)
    lib.destroy_plan(p)
    torch.cuda.synchronize()
    assert cases.rel_l2(z.cpu().numpy(), n * xh.astype(np.float64)) < 2 * cases.tol(lib, n)


def test_full_size_roundtrip_512_cubed_fp32(lib):
    """BASELINE config 4 at full size, through size-independent properties: closed-form
    spectrum of plane waves + backward(forward(x)) = N x + Parseval."""
    import torch
    if not lib.single:
        pytest.skip("config 4 is fp32")
    n = 512
    g = torch.Generator(device="cuda").manual_seed(3)
    x = (torch.rand((n, n, n, 2), device="cuda", generator=g) - 0.5)
    x = torch.view_as_complex(x).contiguous()
    # add a plane wave with integer wave numbers: a spike of height N at (5, 7, 11)
    i = torch.arange(n, device="cuda", dtype=torch.float32)
    ph = (5 * i[:, None, None] + 7 * i[None, :, None] + 11 * i[None, None, :]) * (2 * math.pi / n)
    x0 = x.clone()
    x = x + 3.0 * torch.polar(torch.ones_like(ph), ph)
    del ph
    y = torch.empty_like(x)
    p = lib.plan_dft_3d(n, n, n, x, y, capi.FFTW_FORWARD)
    assert p
    lib.execute(p)
    torch.cuda.synchronize()
    N = n ** 3
    e_in = float((x.abs() ** 2).sum(dtype=torch.float64))
    e_out = float((y.abs() ** 2).sum(dtype=torch.float64))
    assert abs(e_out / (N * e_in) - 1) < 1e-5
    # linearity: F(x0 + 3 wave) - F(x0) is a single spike 3N at (5,7,11)
    y0 = torch.empty_like(x0)
    lib.execute_dft(p, x0, y0)
    torch.cuda.synchronize()
    d = y - y0
    spike = d[5, 7, 11].item()
    assert abs(spike - 3.0 * N) / (3.0 * N) < 1e-5
    d[5, 7, 11] = 0
    assert float(d.abs().max()) / (3.0 * N) < 1e-4
    del d, y0, x0
    lib.destroy_plan(p)
    # in place, backward: N x
    p = lib.plan_dft_3d(n, n, n, y, y, capi.FFTW_BACKWARD)
    lib.execute(p)
    lib.destroy_plan(p)
    torch.cuda.synchronize()
    err = float(((y / N - x).abs() ** 2).sum(dtype=torch.float64).sqrt() / (x.abs() ** 2).sum(dtype=torch.float64).sqrt())
    assert err < 2e-6 * math.log2(N)


@pytest.mark.parametrize("howmany", [4, 37, 256])
def test_host_arrays_go_through_the_chunked_pipeline(lib, howmany):
    """host arrays + batch: H2D / kernels / D2H of different chunks overlap (engine.cu pipe_*);
    ragged last chunk, out of place, in place, and r2c"""
    rng = np.random.default_rng(howmany)
    n = 1024
    x = cases.rand_c(rng, (howmany, n), lib)
    y = np.zeros_like(x)
    p = lib.plan_many_dft([n], howmany, x, None, 1, n, y, None, 1, n, capi.FFTW_FORWARD)
    assert p
    lib.execute(p)
    lib.execute(p)
    lib.destroy_plan(p)
    ref = np.fft.fft(x.astype(np.complex128), axis=1)
    assert cases.rel_l2(y, ref) < cases.tol(lib, n)
    z = x.copy()
    p = lib.plan_many_dft([n], howmany, z, None, 1, n, z, None, 1, n, capi.FFTW_BACKWARD)
    lib.execute(p)
    lib.destroy_plan(p)
    assert cases.rel_l2(z, n * np.fft.ifft(x.astype(np.complex128), axis=1)) < cases.tol(lib, n)
    xr = cases.rand_r(rng, (howmany, n), lib)
    yr = np.zeros((howmany, n // 2 + 1), dtype=lib.cplx)
    p = lib.plan_many_dft_r2c([n], howmany, xr, None, 1, n, yr, None, 1, n // 2 + 1)
    lib.execute(p)
    lib.destroy_plan(p)
    assert cases.rel_l2(yr, np.fft.rfft(xr.astype(np.float64), axis=1)) < cases.tol(lib, n)


def test_chained_launch_opt_in(lib):
    """the single-launch chain (acquire/release between CTAs of one grid) gives the same answers"""
    cases.check_chained_launch(lib)


def test_two_pass_c2r_opt_in(lib):
    cases.check_two_pass_c2r(lib)


def test_engine_calls_leave_the_callers_current_device_alone(lib):
    """the engine made its own device current on every call and never restored it.  Planning and
    executing must leave the caller's current device as it was; with two GPUs a plan keeps running on the device it
    was created on after b2fft_set_device selects another one, and the table caches are per device."""
    import torch
    L = lib.lib
    ndev = torch.cuda.device_count()
    caller = ndev - 1            # the caller works on the LAST device, the engine (so far) on device 0
    rng = np.random.default_rng(21)
    xh = cases.rand_c(rng, (4, 1000), lib)   # 1000 = 8 x 125: twiddle tables; 1009 below: Bluestein tables
    ref = np.stack([np.fft.fft(r.astype(np.complex128)) for r in xh])
    try:
        L.b2fft_set_device(0)
        torch.cuda.set_device(caller)
        x0 = torch.from_numpy(xh).to("cuda:0")
        y0 = torch.empty_like(x0)
        p0 = lib.plan_many_dft([1000], 4, x0, None, 1, 1000, y0, None, 1, 1000, -1)
        assert p0 and torch.cuda.current_device() == caller
        lib.execute(p0)
        assert torch.cuda.current_device() == caller
        torch.cuda.synchronize(0)
        assert cases.rel_l2(y0.cpu().numpy(), ref) < cases.tol(lib, 1000)
        if ndev >= 2:
            assert L.b2fft_set_device(1) == 0 and torch.cuda.current_device() == caller
            x1 = torch.from_numpy(xh).to("cuda:1")
            y1 = torch.empty_like(x1)
            p1 = lib.plan_many_dft([1000], 4, x1, None, 1, 1000, y1, None, 1, 1000, -1)   # same length: its own tables on device 1
            assert p1
            lib.execute(p1)
            y0.zero_()
            lib.execute(p0)   # still on device 0
            torch.cuda.synchronize(0)
            torch.cuda.synchronize(1)
            assert torch.cuda.current_device() == caller
            assert cases.rel_l2(y1.cpu().numpy(), ref) < cases.tol(lib, 1000)
            assert cases.rel_l2(y0.cpu().numpy(), ref) < cases.tol(lib, 1000)
            xb = torch.from_numpy(cases.rand_c(rng, (2, 1009), lib)).to("cuda:1")
            yb = torch.empty_like(xb)
            pb = lib.plan_many_dft([1009], 2, xb, None, 1, 1009, yb, None, 1, 1009, -1)
            lib.execute(pb)
            torch.cuda.synchronize(1)
            refb = np.stack([np.fft.fft(r.astype(np.complex128)) for r in xb.cpu().numpy()])
            assert cases.rel_l2(yb.cpu().numpy(), refb) < cases.tol(lib, 1009)
            for p in (p1, pb):
                lib.destroy_plan(p)
        lib.destroy_plan(p0)
    finally:
        L.b2fft_set_device(0)
        torch.cuda.set_device(0)
