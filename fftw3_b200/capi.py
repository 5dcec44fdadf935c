"""ctypes binding of the public fftw3.h C API (reference: fftw/fftw3.h:143-463).

The binding is deliberately library-agnostic: it binds *any* shared object that exports
the reference's `fftw_*` symbols.  The product uses it on the B200 engine
(`fftw3_b200/lib/fp32|fp64/libfftw3.so`); the tests additionally point it at the
unmodified reference build (`oracle/_ref/libfftw3_ref_*.so`) to generate golden vectors.
Nothing here computes anything: every call goes straight through the C ABI.

Precision is a property of the shared object (the reference has no `fftwf_` prefix:
`FFTW_SINGLE` is a compile-time switch, fftw/fftw3.h:65-82), so each `FFTW3Lib`
instance carries its real dtype.
"""
from __future__ import annotations

import ctypes as C
import numpy as np

FFTW_FORWARD = -1   # fftw/fftw3.h:469
FFTW_BACKWARD = +1  # fftw/fftw3.h:470

# planner flags, fftw/fftw3.h:475-499
FFTW_MEASURE = 0
FFTW_DESTROY_INPUT = 1 << 0
FFTW_UNALIGNED = 1 << 1
FFTW_CONSERVE_MEMORY = 1 << 2
FFTW_EXHAUSTIVE = 1 << 3
FFTW_PRESERVE_INPUT = 1 << 4
FFTW_PATIENT = 1 << 5
FFTW_ESTIMATE = 1 << 6
FFTW_WISDOM_ONLY = 1 << 21

_vp = C.c_void_p
_ip = C.POINTER(C.c_int)


def _ptr(x):
    """numpy array -> host address; int -> raw (host or device) address; None -> NULL."""
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    if hasattr(x, "data_ptr"):  # torch tensor (device or host)
        return x.data_ptr()
    return int(x)


def _ints(seq):
    if seq is None:
        return None
    arr = (C.c_int * len(seq))(*[int(v) for v in seq])
    return arr


class FFTW3Lib:
    """One loaded fftw3-ABI shared object."""

    def __init__(self, path: str, single: bool):
        self.path = path
        self.single = bool(single)
        self.real = np.float32 if single else np.float64
        self.cplx = np.complex64 if single else np.complex128
        self.lib = C.CDLL(path, mode=C.RTLD_LOCAL)
        L = self.lib
        many = [C.c_int, _ip, C.c_int, _vp, _ip, C.c_int, C.c_int, _vp, _ip, C.c_int, C.c_int]
        L.fftw_plan_many_dft.restype = _vp
        L.fftw_plan_many_dft.argtypes = many + [C.c_int, C.c_uint]
        L.fftw_plan_many_dft_r2c.restype = _vp
        L.fftw_plan_many_dft_r2c.argtypes = many + [C.c_uint]
        L.fftw_plan_many_dft_c2r.restype = _vp
        L.fftw_plan_many_dft_c2r.argtypes = many + [C.c_uint]
        L.fftw_plan_dft.restype = _vp
        L.fftw_plan_dft.argtypes = [C.c_int, _ip, _vp, _vp, C.c_int, C.c_uint]
        L.fftw_plan_dft_1d.restype = _vp
        L.fftw_plan_dft_1d.argtypes = [C.c_int, _vp, _vp, C.c_int, C.c_uint]
        L.fftw_plan_dft_2d.restype = _vp
        L.fftw_plan_dft_2d.argtypes = [C.c_int, C.c_int, _vp, _vp, C.c_int, C.c_uint]
        L.fftw_plan_dft_3d.restype = _vp
        L.fftw_plan_dft_3d.argtypes = [C.c_int, C.c_int, C.c_int, _vp, _vp, C.c_int, C.c_uint]
        for nm in ("r2c", "c2r"):
            f = getattr(L, f"fftw_plan_dft_{nm}")
            f.restype = _vp
            f.argtypes = [C.c_int, _ip, _vp, _vp, C.c_uint]
            f = getattr(L, f"fftw_plan_dft_{nm}_1d")
            f.restype = _vp
            f.argtypes = [C.c_int, _vp, _vp, C.c_uint]
            f = getattr(L, f"fftw_plan_dft_{nm}_2d")
            f.restype = _vp
            f.argtypes = [C.c_int, C.c_int, _vp, _vp, C.c_uint]
            f = getattr(L, f"fftw_plan_dft_{nm}_3d")
            f.restype = _vp
            f.argtypes = [C.c_int, C.c_int, C.c_int, _vp, _vp, C.c_uint]
        L.fftw_execute.restype = None
        L.fftw_execute.argtypes = [_vp]
        for nm in ("fftw_execute_dft", "fftw_execute_dft_r2c", "fftw_execute_dft_c2r"):
            f = getattr(L, nm)
            f.restype = None
            f.argtypes = [_vp, _vp, _vp]
        L.fftw_destroy_plan.restype = None
        L.fftw_destroy_plan.argtypes = [_vp]
        L.fftw_malloc.restype = _vp
        L.fftw_malloc.argtypes = [C.c_size_t]
        L.fftw_free.restype = None
        L.fftw_free.argtypes = [_vp]
        L.fftw_cleanup.restype = None
        L.fftw_cleanup.argtypes = []
        L.fftw_sprint_plan.restype = _vp
        L.fftw_sprint_plan.argtypes = [_vp]
        L.fftw_flops.restype = None
        L.fftw_flops.argtypes = [_vp] + [C.POINTER(C.c_double)] * 3
        if hasattr(L, "fftw_init_threads"):
            L.fftw_init_threads.restype = C.c_int
            L.fftw_plan_with_nthreads.restype = None
            L.fftw_plan_with_nthreads.argtypes = [C.c_int]

    # ---- thin, argument-for-argument mirrors of the C prototypes -------------------
    def plan_many_dft(self, n, howmany, inp, inembed, istride, idist,
                      out, onembed, ostride, odist, sign, flags=FFTW_ESTIMATE):
        return self.lib.fftw_plan_many_dft(len(n), _ints(n), howmany, _ptr(inp), _ints(inembed),
                                           istride, idist, _ptr(out), _ints(onembed), ostride,
                                           odist, sign, flags)

    def plan_many_dft_r2c(self, n, howmany, inp, inembed, istride, idist,
                          out, onembed, ostride, odist, flags=FFTW_ESTIMATE):
        return self.lib.fftw_plan_many_dft_r2c(len(n), _ints(n), howmany, _ptr(inp),
                                               _ints(inembed), istride, idist, _ptr(out),
                                               _ints(onembed), ostride, odist, flags)

    def plan_many_dft_c2r(self, n, howmany, inp, inembed, istride, idist,
                          out, onembed, ostride, odist, flags=FFTW_ESTIMATE):
        return self.lib.fftw_plan_many_dft_c2r(len(n), _ints(n), howmany, _ptr(inp),
                                               _ints(inembed), istride, idist, _ptr(out),
                                               _ints(onembed), ostride, odist, flags)

    def plan_guru_dft(self, dims, howmany_dims, inp, out, sign, flags=FFTW_ESTIMATE):
        """dims / howmany_dims: lists of (n, is, os) in complex elements (fftw_iodim, fftw3.h)"""
        class IoDim(C.Structure):
            _fields_ = [("n", C.c_int), ("is_", C.c_int), ("os", C.c_int)]
        def arr(ds):
            a = (IoDim * max(1, len(ds)))()
            for i, (n, is_, os_) in enumerate(ds):
                a[i].n, a[i].is_, a[i].os = n, is_, os_
            return a
        f = self.lib.fftw_plan_guru_dft
        f.restype = _vp
        f.argtypes = [C.c_int, _vp, C.c_int, _vp, _vp, _vp, C.c_int, C.c_uint]
        d, h = arr(dims), arr(howmany_dims)
        return f(len(dims), C.cast(d, _vp), len(howmany_dims), C.cast(h, _vp), _ptr(inp), _ptr(out), sign, flags)

    def plan_guru_real(self, kind, dims, howmany_dims, inp, out, flags=FFTW_ESTIMATE):
        """fftw_plan_guru_dft_r2c / _c2r: dims / howmany_dims = lists of (n, is, os), strides in elements of the respective
        array (reals on the real side, complex on the other); n of the last dim is the logical real length"""
        f = self.lib.fftw_plan_guru_dft_r2c if kind == "r2c" else self.lib.fftw_plan_guru_dft_c2r
        f.restype = _vp
        f.argtypes = [C.c_int, _vp, C.c_int, _vp, _vp, _vp, C.c_uint]
        d, h = self._iodims(dims), self._iodims(howmany_dims)
        return f(len(dims), C.cast(d, _vp), len(howmany_dims), C.cast(h, _vp), _ptr(inp), _ptr(out), flags)

    def _iodims(self, ds):
        class IoDim(C.Structure):
            _fields_ = [("n", C.c_int), ("is_", C.c_int), ("os", C.c_int)]
        a = (IoDim * max(1, len(ds)))()
        for i, (n, is_, os_) in enumerate(ds):
            a[i].n, a[i].is_, a[i].os = n, is_, os_
        return a

    def plan_guru_split(self, kind, dims, howmany_dims, a, b, c, d=None, flags=FFTW_ESTIMATE):
        """split-array guru plans (fftw3.h: fftw_plan_guru_split_dft / _split_dft_r2c / _split_dft_c2r);
        strides in reals.  kind 'c2c': (ri, ii, ro, io); 'r2c': (in, ro, io); 'c2r': (ri, ii, out)."""
        dd, hh = self._iodims(dims), self._iodims(howmany_dims)
        head = [len(dims), C.cast(dd, _vp), len(howmany_dims), C.cast(hh, _vp)]
        if kind == "c2c":
            f = self.lib.fftw_plan_guru_split_dft
            f.restype, f.argtypes = _vp, [C.c_int, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp, C.c_uint]
            return f(*head, _ptr(a), _ptr(b), _ptr(c), _ptr(d), flags)
        f = self.lib.fftw_plan_guru_split_dft_r2c if kind == "r2c" else self.lib.fftw_plan_guru_split_dft_c2r
        f.restype, f.argtypes = _vp, [C.c_int, _vp, C.c_int, _vp, _vp, _vp, _vp, C.c_uint]
        return f(*head, _ptr(a), _ptr(b), _ptr(c), flags)

    def execute_split(self, kind, plan, a, b, c, d=None):
        if kind == "c2c":
            f = self.lib.fftw_execute_split_dft
            f.restype, f.argtypes = None, [_vp] * 5
            f(plan, _ptr(a), _ptr(b), _ptr(c), _ptr(d))
        else:
            f = self.lib.fftw_execute_split_dft_r2c if kind == "r2c" else self.lib.fftw_execute_split_dft_c2r
            f.restype, f.argtypes = None, [_vp] * 4
            f(plan, _ptr(a), _ptr(b), _ptr(c))

    def plan_r2r(self, n, inp, out, kinds, flags=FFTW_ESTIMATE):
        f = self.lib.fftw_plan_r2r
        f.restype, f.argtypes = _vp, [C.c_int, _ip, _vp, _vp, _ip, C.c_uint]
        return f(len(n), _ints(n), _ptr(inp), _ptr(out), _ints(kinds), flags)

    def plan_many_r2r(self, n, howmany, inp, inembed, istride, idist, out, onembed, ostride, odist, kinds,
                      flags=FFTW_ESTIMATE):
        f = self.lib.fftw_plan_many_r2r
        f.restype = _vp
        f.argtypes = [C.c_int, _ip, C.c_int, _vp, _ip, C.c_int, C.c_int, _vp, _ip, C.c_int, C.c_int, _ip, C.c_uint]
        return f(len(n), _ints(n), howmany, _ptr(inp), _ints(inembed), istride, idist, _ptr(out), _ints(onembed),
                 ostride, odist, _ints(kinds), flags)

    def execute_r2r(self, plan, inp, out):
        f = self.lib.fftw_execute_r2r
        f.restype, f.argtypes = None, [_vp, _vp, _vp]
        f(plan, _ptr(inp), _ptr(out))

    def plan_dft(self, n, inp, out, sign, flags=FFTW_ESTIMATE):
        return self.lib.fftw_plan_dft(len(n), _ints(n), _ptr(inp), _ptr(out), sign, flags)

    def plan_dft_1d(self, n, inp, out, sign, flags=FFTW_ESTIMATE):
        return self.lib.fftw_plan_dft_1d(n, _ptr(inp), _ptr(out), sign, flags)

    def plan_dft_2d(self, n0, n1, inp, out, sign, flags=FFTW_ESTIMATE):
        return self.lib.fftw_plan_dft_2d(n0, n1, _ptr(inp), _ptr(out), sign, flags)

    def plan_dft_3d(self, n0, n1, n2, inp, out, sign, flags=FFTW_ESTIMATE):
        return self.lib.fftw_plan_dft_3d(n0, n1, n2, _ptr(inp), _ptr(out), sign, flags)

    def plan_dft_r2c(self, n, inp, out, flags=FFTW_ESTIMATE):
        return self.lib.fftw_plan_dft_r2c(len(n), _ints(n), _ptr(inp), _ptr(out), flags)

    def plan_dft_c2r(self, n, inp, out, flags=FFTW_ESTIMATE):
        return self.lib.fftw_plan_dft_c2r(len(n), _ints(n), _ptr(inp), _ptr(out), flags)

    def execute(self, plan):
        self.lib.fftw_execute(plan)

    def execute_dft(self, plan, inp, out):
        self.lib.fftw_execute_dft(plan, _ptr(inp), _ptr(out))

    def execute_dft_r2c(self, plan, inp, out):
        self.lib.fftw_execute_dft_r2c(plan, _ptr(inp), _ptr(out))

    def execute_dft_c2r(self, plan, inp, out):
        self.lib.fftw_execute_dft_c2r(plan, _ptr(inp), _ptr(out))

    def sprint_plan(self, plan) -> str:
        p = self.lib.fftw_sprint_plan(plan)
        if not p:
            return ""
        s = C.string_at(p).decode()
        C.CDLL(None).free(C.c_void_p(p))
        return s

    def flops(self, plan):
        a, m, f = C.c_double(), C.c_double(), C.c_double()
        self.lib.fftw_flops(plan, C.byref(a), C.byref(m), C.byref(f))
        return a.value, m.value, f.value

    def plan_with_nthreads(self, n):
        self.lib.fftw_plan_with_nthreads(int(n))

    def destroy_plan(self, plan):
        self.lib.fftw_destroy_plan(plan)

    def malloc(self, nbytes):
        return self.lib.fftw_malloc(nbytes)

    def free(self, p):
        self.lib.fftw_free(p)

    def cleanup(self):
        self.lib.fftw_cleanup()

    # ---- conveniences for host numpy arrays (tests / golden generation) ------------
    def dft(self, x: np.ndarray, sign=FFTW_FORWARD, flags=FFTW_ESTIMATE) -> np.ndarray:
        """rank-x.ndim out-of-place complex DFT of a contiguous host array."""
        x = np.ascontiguousarray(x, dtype=self.cplx)
        y = np.empty_like(x)
        p = self.plan_dft(list(x.shape), x, y, sign, flags)
        if not p:
            raise RuntimeError("fftw_plan_dft returned NULL")
        self.execute(p)
        self.destroy_plan(p)
        return y

    def r2c(self, x: np.ndarray, flags=FFTW_ESTIMATE) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=self.real)
        y = np.empty(x.shape[:-1] + (x.shape[-1] // 2 + 1,), dtype=self.cplx)
        p = self.plan_dft_r2c(list(x.shape), x, y, flags)
        if not p:
            raise RuntimeError("fftw_plan_dft_r2c returned NULL")
        self.execute(p)
        self.destroy_plan(p)
        return y

    def c2r(self, X: np.ndarray, n_last: int, flags=FFTW_ESTIMATE) -> np.ndarray:
        X = np.array(X, dtype=self.cplx, order="C", copy=True)  # c2r may destroy its input
        shape = X.shape[:-1] + (n_last,)
        y = np.empty(shape, dtype=self.real)
        p = self.plan_dft_c2r(list(shape), X, y, flags)
        if not p:
            raise RuntimeError("fftw_plan_dft_c2r returned NULL")
        self.execute(p)
        self.destroy_plan(p)
        return y
