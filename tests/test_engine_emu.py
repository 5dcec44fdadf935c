"""CPU suite for the host logic of the engine: C API layer (csrc/fftw_api.c), geometry
canonicalisation, selector, buffer assignment, and every index computation of the kernels.

It runs the TEST-ONLY emulation build (tests/_emu/libfftw3_emu_*.so: the same sources compiled
with -DB2_EMULATE, each CTA executed by a host loop) -- there is no GPU in this container.
The product library has no such path; the real parity tests are tests/test_gpu_parity.py.
"""
import numpy as np
import pytest

from fftw3_b200 import build, capi
from tests import cases


@pytest.fixture(scope="module", params=["fp32", "fp64"])
def lib(request):
    path = build.build_emu(request.param)
    return capi.FFTW3Lib(path, request.param == "fp32")


@pytest.mark.parametrize("shape", cases.C2C_SHAPES + cases.C2C_BIG + cases.C2C_PRIME + cases.C2C_PIPE, ids=lambda s: "x".join(map(str, s)))
def test_c2c(lib, shape):
    cases.check_c2c(lib, shape)


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


def test_edge_cases(lib):
    cases.check_edge_cases(lib)


@pytest.mark.parametrize("shape", [(96,), (16, 20), (2 ** 14,)], ids=lambda s: "x".join(map(str, s)))
def test_verifier_properties(lib, shape):
    cases.check_properties(lib, shape)


def test_plan_description_and_flops(lib):
    x = np.zeros((4, 4096), dtype=lib.cplx)
    p = lib.plan_many_dft([4096], 4, x, None, 1, 4096, x.copy(), None, 1, 4096, -1)
    assert p
    s = lib.sprint_plan(p)
    assert "n=4096" in s and "pass 0" in s
    add, mul, fma = lib.flops(p)
    assert abs((add + mul + fma) - 5 * 4096 * 12 * 4) < 1e-6
    lib.destroy_plan(p)


def test_cost_is_the_flop_count_for_every_kind_of_plan(lib):
    """fftw_cost / fftw_estimate_cost (reference fftw_api.c: cost of the plan's operation counts): c2c, r2r and split-array
    plans all report the flop count fftw_flops reports -- r2r and split plans keep theirs in their own objects"""
    import ctypes as C
    L = lib.lib
    L.fftw_cost.restype = C.c_double
    L.fftw_cost.argtypes = [C.c_void_p]
    L.fftw_estimate_cost.restype = C.c_double
    L.fftw_estimate_cost.argtypes = [C.c_void_p]
    x = np.zeros((4, 96), dtype=lib.cplx)
    y = x.copy()
    plans = [lib.plan_many_dft([96], 4, x, None, 1, 96, y, None, 1, 96, -1)]
    r = np.zeros(64, dtype=lib.real)
    r2 = r.copy()
    plans.append(lib.plan_r2r([64], r, r2, [3]))  # REDFT00
    ri, ii, ro, io = (np.zeros(48, dtype=lib.real) for _ in range(4))
    plans.append(lib.plan_guru_split("c2c", [(48, 1, 1)], [], ri, ii, ro, io))
    for p in plans:
        assert p
        f = sum(lib.flops(p))
        assert f > 0 and abs(L.fftw_cost(p) - f) < 1e-6 and abs(L.fftw_estimate_cost(p) - f) < 1e-6
        lib.destroy_plan(p)


def test_long_column_axes_take_the_pipelined_kernel(lib):
    x = np.zeros((1024, 64), dtype=lib.cplx)
    p = lib.plan_dft([1024, 64], x, x.copy(), -1)
    assert p
    assert "async-copy pipelined" in lib.sprint_plan(p), lib.sprint_plan(p)
    lib.destroy_plan(p)


@pytest.mark.parametrize("shape", [(2 ** 15,), (3 * 2 ** 13,), (2, 2 ** 14), (2 ** 13, 3), (5 * 7 * 2 ** 9,)],
                         ids=lambda s: "x".join(map(str, s)))
def test_three_pass_split_beyond_the_four_step_limit(lib, shape, monkeypatch):
    """n = n1*n2*n3 (six-step): with the per-kernel limits lowered to 32 / 64 points the split that
    serves n > 2^22 (fp32) / 2^20 (fp64) on the GPU is reached at 2^15 points"""
    monkeypatch.setenv("FFTW_B200_COL_CAP", "32")
    monkeypatch.setenv("FFTW_B200_ROW_CAP", "64")
    cases.check_c2c(lib, shape)
    x = np.zeros(shape, dtype=lib.cplx)
    p = lib.plan_dft(list(shape), x, x.copy(), -1)
    assert lib.sprint_plan(p).count("pass ") >= 3
    lib.destroy_plan(p)


@pytest.mark.parametrize("shape", [(2 ** 16,), (3, 2 ** 16), (2, 3 * 2 ** 14)], ids=lambda s: "x".join(map(str, s)))
def test_real_transforms_beyond_the_four_step_limit(lib, shape, monkeypatch):
    monkeypatch.setenv("FFTW_B200_COL_CAP", "32")
    monkeypatch.setenv("FFTW_B200_ROW_CAP", "64")
    cases.check_r2c_c2r(lib, shape)
    if len(shape) == 1:
        cases.check_r2c_inplace(lib, shape)


def test_split_array_guru_interfaces(lib):
    cases.check_split_arrays(lib)
    cases.check_split_interleaved_host(lib)


def test_r2r_kinds(lib, r2r_golden):
    cases.check_r2r(lib, r2r_golden)


def test_negative_strides(lib):
    cases.check_negative_strides(lib)


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


def test_measure_and_wisdom(lib):
    cases.check_measure_and_wisdom(lib)


def test_threads_share_a_plan_and_the_planner(lib):
    cases.check_threads(lib)


@pytest.mark.parametrize("shape", [(2 ** 19,), (2, 2 ** 20), (3, 2 ** 19)], ids=lambda s: "x".join(map(str, s)))
def test_three_pass_real_transform(lib, shape, monkeypatch):
    """pair-split column pass, twiddled column pass inside every row, folded row pass (default above the
    two-pass limit, forced here at sizes the emulation can run)"""
    monkeypatch.setenv("FFTW_B200_THREE_PASS_REAL", "1")
    cases.check_r2c_c2r(lib, shape)
    x = np.zeros(shape[-1], dtype=lib.real)
    p = lib.plan_dft_r2c([shape[-1]], x, np.zeros(shape[-1] // 2 + 1, dtype=lib.cplx))
    assert "hermitian fold" in lib.sprint_plan(p) and lib.sprint_plan(p).count("  pass ") == 3
    lib.destroy_plan(p)
    if len(shape) == 1:
        cases.check_r2c_inplace(lib, shape)


def test_in_place_with_different_strides(lib):
    cases.check_inplace_relayout(lib)


def test_chained_launch_opt_in(lib):
    cases.check_chained_launch(lib)


def test_two_pass_c2r_opt_in(lib):
    cases.check_two_pass_c2r(lib)


@pytest.mark.parametrize("shape", [(6, 32, 128), (5, 3, 64, 256)], ids=lambda s: "x".join(map(str, s)))
def test_grouped_execution_of_the_first_two_axes(lib, shape, monkeypatch):
    """N-D c2c: the passes over the two innermost axes run plane group by plane group (engine.cu run_steps; opt-in), forced
    here at a size the emulation can run: groups of two planes, every plane larger than the 'fits the L2' threshold"""
    monkeypatch.setenv("FFTW_B200_GROUPS", "1")
    monkeypatch.setenv("FFTW_B200_GROUP_MIN_KIB", "1")
    monkeypatch.setenv("FFTW_B200_GROUP_KIB", str(2 * shape[-1] * shape[-2] * 2 * np.dtype(lib.real).itemsize // 1024))
    cases.check_c2c(lib, shape)
    monkeypatch.setenv("FFTW_B200_NO_GROUPS", "1")   # and the same plan without grouping agrees bit for bit
    rng = np.random.default_rng(4)
    x = cases.rand_c(rng, shape, lib)
    a = lib.dft(x, -1)
    monkeypatch.delenv("FFTW_B200_NO_GROUPS")
    assert np.array_equal(a, lib.dft(x, -1))


@pytest.mark.parametrize("shape", [(7, 2 ** 16), (4, 2 ** 17)], ids=lambda s: "x".join(map(str, s)))
def test_grouped_execution_of_the_two_pass_real_transform(lib, shape, monkeypatch):
    """batched two-pass r2c: both passes run transform group by transform group through the START of the scratch
    array (three transforms per group here; 7 = 3 + 3 + 1 also covers the short last group)"""
    monkeypatch.setenv("FFTW_B200_GROUPS", "1")
    monkeypatch.setenv("FFTW_B200_GROUP_MIN_KIB", "1")
    monkeypatch.setenv("FFTW_B200_GROUP_KIB", str(3 * (shape[-1] + 2 * int(np.sqrt(shape[-1]))) * np.dtype(lib.real).itemsize // 1024))
    x = np.zeros(shape, dtype=lib.real)
    p = lib.plan_many_dft_r2c([shape[-1]], shape[0], x, None, 1, shape[-1], np.zeros((shape[0], shape[-1] // 2 + 1), dtype=lib.cplx),
                              None, 1, shape[-1] // 2 + 1)
    two_pass = "real-pair split" in lib.sprint_plan(p)
    lib.destroy_plan(p)
    if not two_pass:
        pytest.skip("this precision does not take the two-pass real path at this length")
    cases.check_r2c_c2r(lib, shape)
