// csrc/split.cu -- split-array (separate real / imaginary arrays) transforms: the guru "split"
// interfaces of the reference (fftw_plan_guru_split_dft fftw/fftw_api.c:1420-1450,
// _split_dft_r2c / _split_dft_c2r :1452-1517, fftw_execute_split_dft* :444-454).
//
// The reference runs its codelets directly on (ri, ii) pointer pairs.  Here the hot kernels want
// interleaved complex elements (one 64/128-bit access per point), so a split transform is
//   gather (ri, ii) -> dense interleaved work array   [one sweep, coalesced along the unit-stride dim]
//   the ordinary plan, in place on the work array     [tile kernels at full speed]
//   scatter work array -> (ro, io)                    [one sweep]
// r2c reads the caller's real array directly and only scatters; c2r only gathers.  Strides of every
// split / real array are in reals, as in the reference's guru interface.
#include <algorithm>
#include <vector>

#include "b200fft.h"
#include "device_rt.h"
#include "plan_struct.h"

using namespace b2;

#define B2_SPLIT_MAXD (2 * B2FFT_MAXRANK)
struct SplitCopyParams {
    int nd;
    long long n[B2_SPLIT_MAXD], st[B2_SPLIT_MAXD];  // innermost dim last; strides in reals
    long long total;
};

B2_HD long long b2_split_offset(const SplitCopyParams &p, long long i) {
    long long off = 0;
    for (int d = p.nd - 1; d >= 0; --d) {
        const long long q = i / p.n[d];
        off += (i - q * p.n[d]) * p.st[d];
        i = q;
    }
    return off;
}

#if B2_CUDA
__global__ void __launch_bounds__(256) b2_split_gather_kernel(const SplitCopyParams p, const real_t *B2_RESTRICT re,
                                                              const real_t *B2_RESTRICT im, cpx_t *B2_RESTRICT dense) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.total; i += (long long)gridDim.x * blockDim.x) {
        const long long off = b2_split_offset(p, i);
        dense[i] = cpx_t{re[off], im[off]};
    }
}
__global__ void __launch_bounds__(256) b2_split_scatter_kernel(const SplitCopyParams p, const cpx_t *B2_RESTRICT dense,
                                                               real_t *B2_RESTRICT re, real_t *B2_RESTRICT im) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.total; i += (long long)gridDim.x * blockDim.x) {
        const long long off = b2_split_offset(p, i);
        const cpx_t v = dense[i];
        re[off] = v.x;
        im[off] = v.y;
    }
}
#endif

struct b2fft_split_s {
    std::mutex exec_mutex;  // executes of one plan share its work arrays: serialised
    int kind = B2FFT_C2C;
    SplitCopyParams in_cp, out_cp;  // complex-side loop nests (c2c: both; r2c: out; c2r: in)
    long long in_extent = 0, out_extent = 0;  // reals spanned by one split array
    b2fft_plan *inner = nullptr;
    cpx_t *dense = nullptr;
    real_t *stage[4] = {nullptr, nullptr, nullptr, nullptr};  // device copies of host ri, ii, ro, io
    long long stage_reals[4] = {0, 0, 0, 0};
    std::string desc;
};

static int split_copy(const SplitCopyParams &p, bool gather, real_t *re, real_t *im, cpx_t *dense) {
    if (p.total <= 0) return 0;
#if B2_CUDA
    const long long want = (p.total + 255) / 256;
    const unsigned grid = (unsigned)std::min<long long>(want, (long long)sm_count() * 32);
    if (gather)
        b2_split_gather_kernel<<<grid, 256>>>(p, re, im, dense);
    else
        b2_split_scatter_kernel<<<grid, 256>>>(p, dense, re, im);
    ++launch_counter();
    B2_CK(cudaGetLastError());
#else
    for (long long i = 0; i < p.total; ++i) {
        const long long off = b2_split_offset(p, i);
        if (gather)
            dense[i] = cpx_t{re[off], im[off]};
        else {
            re[off] = dense[i].x;
            im[off] = dense[i].y;
        }
    }
#endif
    return 0;
}

// loop nest of the complex side: batch dims outermost, then the transform dims; `half`: the last
// transform dim holds n/2+1 elements.  Dense strides (in complex elements) go to ds[].
static bool split_nest(const b2fft_geom &g, bool input_side, bool half, SplitCopyParams &cp, long long &extent,
                       b2fft_dim *dense_dims, b2fft_dim *dense_vdims, bool dense_is_input) {
    cp.nd = 0;
    cp.total = 1;
    extent = 1;
    std::vector<std::pair<const b2fft_dim *, b2fft_dim *>> order;
    for (int i = 0; i < g.vrank; ++i) order.push_back({&g.vdims[i], &dense_vdims[i]});
    for (int i = 0; i < g.rank; ++i) order.push_back({&g.dims[i], &dense_dims[i]});
    std::vector<long long> n(order.size());
    for (size_t i = 0; i < order.size(); ++i) {
        n[i] = order[i].first->n;
        if (half && i + 1 == order.size()) n[i] = n[i] / 2 + 1;
    }
    long long dstride = 1;
    for (int i = (int)order.size() - 1; i >= 0; --i) {
        const long long st = input_side ? order[i].first->is : order[i].first->os;
        if (st < 0) {
            set_error("split arrays with negative strides are not supported");
            return false;
        }
        (dense_is_input ? order[i].second->is : order[i].second->os) = dstride;
        dstride *= n[i];
    }
    for (size_t i = 0; i < order.size(); ++i) {
        const long long st = input_side ? order[i].first->is : order[i].first->os;
        cp.n[cp.nd] = n[i];
        cp.st[cp.nd] = st;
        ++cp.nd;
        cp.total *= n[i];
        extent += (n[i] - 1) * st;
    }
    if (cp.total == 0) extent = 0;
    return true;
}

extern "C" b2fft_split *b2fft_split_plan_create(const b2fft_geom *gu) {
    if (!gu || gu->rank < 0 || gu->rank > B2FFT_MAXRANK || gu->vrank < 0 || gu->vrank > B2FFT_MAXRANK) return nullptr;
    DeviceScope caller;
    if (ensure_device() != 0) return nullptr;
    for (int i = 0; i < gu->rank; ++i)
        if (gu->dims[i].n <= 0) return nullptr;
    for (int i = 0; i < gu->vrank; ++i)
        if (gu->vdims[i].n < 0) return nullptr;
    if (gu->kind != B2FFT_C2C && gu->rank < 1) return nullptr;
    b2fft_split *s = new b2fft_split_s();
    s->kind = gu->kind;
    b2fft_geom gi = *gu;  // the inner (interleaved) problem
    bool ok = true;
    if (gu->kind == B2FFT_C2C) {
        ok = split_nest(*gu, true, false, s->in_cp, s->in_extent, gi.dims, gi.vdims, true) &&
             split_nest(*gu, false, false, s->out_cp, s->out_extent, gi.dims, gi.vdims, false);
        gi.in_place = 1;
        gi.sign = -1;  // split plans are forward; callers swap the arrays for backward (fftw_api.c:1420)
    } else if (gu->kind == B2FFT_R2C) {
        // the real input keeps the caller's strides; the interleaved output is dense
        ok = split_nest(*gu, false, true, s->out_cp, s->out_extent, gi.dims, gi.vdims, false);
        gi.in_place = 0;
    } else {
        ok = split_nest(*gu, true, true, s->in_cp, s->in_extent, gi.dims, gi.vdims, true);
        gi.in_place = 0;
    }
    const long long total = gu->kind == B2FFT_R2C ? s->out_cp.total : s->in_cp.total;
    if (ok) {
        s->inner = b2fft_plan_create(&gi);
        ok = s->inner != nullptr;
    }
    if (ok) {
        s->dense = (cpx_t *)dev_malloc((size_t)std::max<long long>(total, 1) * sizeof(cpx_t));
        ok = s->dense != nullptr;
    }
    if (!ok) {
        b2fft_split_plan_destroy(s);
        return nullptr;
    }
    char buf[256];
    snprintf(buf, sizeof buf, "split-array %s: %s%lld interleaved elements in a dense work array%s\n",
             gu->kind == B2FFT_C2C ? "c2c" : gu->kind == B2FFT_R2C ? "r2c" : "c2r",
             gu->kind == B2FFT_R2C ? "" : "gather (re, im) -> ", total, gu->kind == B2FFT_C2R ? "" : " -> scatter (re, im)");
    s->desc = buf;
    return s;
}

extern "C" void b2fft_split_plan_destroy(b2fft_split *s) {
    if (!s) return;
    b2fft_plan_destroy(s->inner);
    dev_free(s->dense);
    for (real_t *p : s->stage) dev_free(p);
    delete s;
}

// device-visible version of a split array: the pointer itself, or a staged copy of its extent
static real_t *split_stage(b2fft_split *s, int slot, real_t *user, long long extent, bool upload) {
    if (!user || extent <= 0 || is_device_ptr(user)) return user;
#if B2_CUDA
    if (s->stage[slot] && s->stage_reals[slot] < extent) {  // the union span of overlapping arrays depends on the pointers
        dev_free(s->stage[slot]);
        s->stage[slot] = nullptr;
    }
    if (!s->stage[slot]) {
        s->stage[slot] = (real_t *)dev_malloc((size_t)extent * sizeof(real_t));
        s->stage_reals[slot] = extent;
    }
    if (!s->stage[slot]) return nullptr;
    if (upload && cudaMemcpy(s->stage[slot], user, (size_t)extent * sizeof(real_t), cudaMemcpyHostToDevice) != cudaSuccess) {
        set_error("split execute: host -> device copy failed");
        return nullptr;
    }
    return s->stage[slot];
#else
    (void)s, (void)slot, (void)upload;
    return user;
#endif
}
static int split_unstage(real_t *dev, real_t *user, long long extent) {
#if B2_CUDA
    if (dev != user && extent > 0) B2_CK(cudaMemcpy(user, dev, (size_t)extent * sizeof(real_t), cudaMemcpyDeviceToHost));
#else
    (void)dev, (void)user, (void)extent;
#endif
    return 0;
}

// c2c: (ri, ii) -> (ro, io).  r2c: `ri` is the real input, ii unused.  c2r: `ro` is the real output, io unused.
extern "C" int b2fft_split_execute(b2fft_split *s, void *ri_, void *ii_, void *ro_, void *io_) {
    if (!s) return -1;
    std::lock_guard<std::mutex> guard(s->exec_mutex);
    DeviceScope caller;
    if (ensure_device() != 0) return -1;
    real_t *ri = (real_t *)ri_, *ii = (real_t *)ii_, *ro = (real_t *)ro_, *io = (real_t *)io_;
    int rc = 0;
    if (s->kind != B2FFT_R2C) {
        real_t *dr = split_stage(s, 0, ri, s->in_extent, true), *di = split_stage(s, 1, ii, s->in_extent, true);
        if ((s->in_extent > 0) && (!dr || !di)) return -3;
        rc = split_copy(s->in_cp, true, dr, di, s->dense);
        if (rc) return rc;
    }
    if (s->kind == B2FFT_C2C)
        rc = b2fft_execute_async(s->inner, s->dense, s->dense, nullptr);
    else if (s->kind == B2FFT_R2C)
        rc = b2fft_execute(s->inner, ri, s->dense);
    else
        rc = b2fft_execute(s->inner, s->dense, ro);
    if (rc) return rc;
    if (s->kind != B2FFT_C2R) {
        // a strided host output keeps whatever lives in its gaps: stage its current content first
        const bool dense_out = s->out_extent == s->out_cp.total;
        real_t *dr, *di;
        real_t *lo = ro < io ? ro : io, *hi = ro < io ? io : ro;
        const bool overlap = ro && io && s->out_extent > 0 && (hi - lo) < s->out_extent && !is_device_ptr(ro) && !is_device_ptr(io);
        if (overlap) {
            // host (ro, io) ranges that interleave or overlap -- the documented usage ro = out, io = out + 1
            // with stride 2: ONE staged copy of the union span, both arrays addressed inside it, so that the
            // write-back of one array cannot bring back stale values of the other
            const long long span = (long long)(hi - lo) + s->out_extent;
            real_t *base = split_stage(s, 2, lo, span, true);
            if (!base) return -3;
            dr = base + (ro - lo);
            di = base + (io - lo);
            rc = split_copy(s->out_cp, false, dr, di, s->dense);
            if (rc) return rc;
            rc = split_unstage(base, lo, span);
            if (rc) return rc;
        } else {
            dr = split_stage(s, 2, ro, s->out_extent, !dense_out);
            di = split_stage(s, 3, io, s->out_extent, !dense_out);
            if ((s->out_extent > 0) && (!dr || !di)) return -3;
            rc = split_copy(s->out_cp, false, dr, di, s->dense);
            if (rc) return rc;
            rc = split_unstage(dr, ro, s->out_extent);
            if (!rc) rc = split_unstage(di, io, s->out_extent);
            if (rc) return rc;
        }
    }
#if B2_CUDA
    B2_CK(cudaDeviceSynchronize());
#endif
    return 0;
}

extern "C" int b2fft_split_describe(const b2fft_split *s, char *buf, size_t n) {
    if (!s) return 0;
    const int k = b2fft_plan_describe(s->inner, nullptr, 0);
    std::string inner((size_t)k + 1, '\0');
    b2fft_plan_describe(s->inner, &inner[0], (size_t)k + 1);
    inner.resize((size_t)k);
    return snprintf(buf, n, "%s%s", s->desc.c_str(), inner.c_str());
}
extern "C" double b2fft_split_flops(const b2fft_split *s) { return s ? b2fft_plan_flops(s->inner) : 0.0; }
