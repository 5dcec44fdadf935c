/* csrc/fftw_api.c -- plain-C host side of the B200 FFT engine: the public fftw3.h entry
 * points.  It is the drop-in counterpart of the API layer of the reference
 * (fftw/fftw_api.c:21-1517): same names, same argument meaning, same NULL-plan error
 * behaviour; but instead of building problem/tensor objects for a recursive planner
 * (fftw_mkproblem_dft_d :4110, fftw_mkapiplan :317-407) it fills a flat geometry descriptor
 * and hands it to the CUDA engine through the C-ABI of include/b200fft.h.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "b200fft.h"
#include "fftw3.h"

#include "fftw_plan_struct.h"

const char fftw_version[] = "b200fft-0.1 (fftw3.h ABI, sm_100a)";
const char fftw_cc[] = "nvcc -gencode arch=compute_100a,code=sm_100a";
const char fftw_codelet_optim[] = "";

/* same validity rule as fftw_many_kosherp (reference fftw_api.c:863-877) */
static int kosher(int rank, const int *n, int howmany) {
    int i;
    if (rank < 0 || rank > B2FFT_MAXRANK || howmany < 0) return 0;
    for (i = 0; i < rank; ++i)
        if (n[i] <= 0) return 0;
    return 1;
}

/* ---- planner lock.  In the reference every planner / destroy / wisdom call shares one global planner
 * and is NOT thread-safe unless fftw_make_planner_thread_safe() installed mutex hooks
 * (fftw_api.c:258-263, threads/api.c:83-86).  Here the shared state is the twiddle / chirp table cache
 * and the wisdom set; the lock is always taken (it costs nothing next to a plan), so the call that
 * installs it in the reference is a no-op. */
static pthread_mutex_t planner_lock = PTHREAD_MUTEX_INITIALIZER;
#define PLANNER_LOCK() pthread_mutex_lock(&planner_lock)
#define PLANNER_UNLOCK() pthread_mutex_unlock(&planner_lock)

/* ---- wisdom: the set of problems planned (or imported) so far: the problem's signature and, once a plan for
 * it was made with FFTW_MEASURE or more, the configuration that won the timing (b2fft_plan_create_tuned), which
 * later plans for the same problem replay without measuring again -- the reference's "wisdom = remembered winner"
 * (fftw_api.c:332-337, 15300-15426).  FFTW_WISDOM_ONLY gives NULL unless the problem is in the accumulated wisdom;
 * export / import / forget round-trip (fftw_api.c:895-1168).  All accesses under the planner lock. */
#define WIS_UNMEASURED 0xffffffffu
static uint64_t *wis_key = 0;
static unsigned *wis_cfg = 0;
static int wis_n = 0, wis_cap = 0;
static uint64_t wis_hash(const b2fft_geom *g) {
    uint64_t h = 1469598103934665603ull;
    int i;
#define WIS_MIX(v) h = (h ^ (uint64_t)(v)) * 1099511628211ull
    WIS_MIX(g->kind);
    WIS_MIX(g->sign + 2);
    WIS_MIX(g->rank);
    WIS_MIX(g->vrank);
    WIS_MIX(g->in_place);
    WIS_MIX(sizeof(FFTW_REAL_TYPE));
    for (i = 0; i < g->rank; ++i) {
        WIS_MIX(g->dims[i].n);
        WIS_MIX(g->dims[i].is);
        WIS_MIX(g->dims[i].os);
    }
    for (i = 0; i < g->vrank; ++i) {
        WIS_MIX(g->vdims[i].n);
        WIS_MIX(g->vdims[i].is);
        WIS_MIX(g->vdims[i].os);
    }
#undef WIS_MIX
    return h;
}
static int wis_find(uint64_t k) {
    int i;
    for (i = 0; i < wis_n; ++i)
        if (wis_key[i] == k) return i;
    return -1;
}
static int wis_has(uint64_t k) { return wis_find(k) >= 0; }
/* cfg: the measured configuration, or WIS_UNMEASURED (which never replaces a measured one) */
static void wis_add(uint64_t k, unsigned cfg) {
    int i = wis_find(k);
    if (i < 0) {
        if (wis_n == wis_cap) {
            int cap = wis_cap ? 2 * wis_cap : 256;
            uint64_t *nk = (uint64_t *)realloc(wis_key, (size_t)cap * sizeof *nk);
            unsigned *nc;
            if (!nk) return;
            wis_key = nk;
            nc = (unsigned *)realloc(wis_cfg, (size_t)cap * sizeof *nc);
            if (!nc) return;
            wis_cfg = nc;
            wis_cap = cap;
        }
        i = wis_n++;
        wis_key[i] = k;
        wis_cfg[i] = WIS_UNMEASURED;
    }
    if (cfg != WIS_UNMEASURED) wis_cfg[i] = cfg;
}

static fftw_plan finish_locked(b2fft_geom *g, void *in, void *out, unsigned flags);
static fftw_plan finish(b2fft_geom *g, void *in, void *out, unsigned flags) {
    fftw_plan p;
    PLANNER_LOCK();
    p = finish_locked(g, in, out, flags);
    PLANNER_UNLOCK();
    return p;
}
static fftw_plan finish_locked(b2fft_geom *g, void *in, void *out, unsigned flags) {
    fftw_plan p;
    b2fft_plan *bp;
    int wi;
    g->in_place = (in == out);
    /* the reference answers WISDOM_ONLY with NULL when it has no wisdom for the problem
     * (fftw_api.c:332-337) */
    wi = wis_find(wis_hash(g));
    if ((flags & FFTW_WISDOM_ONLY) && wi < 0) return 0;
    g->flags = flags;
    g->in_place = (in == out);
    if (wi >= 0 && wis_cfg[wi] != WIS_UNMEASURED)
        bp = b2fft_plan_create_tuned(g, (int)wis_cfg[wi]);           /* replay the winner remembered in wisdom */
    else if (!(flags & (FFTW_ESTIMATE | FFTW_WISDOM_ONLY)))
        bp = b2fft_plan_create_tuned(g, B2FFT_CFG_MEASURE);          /* FFTW_MEASURE (0), PATIENT, EXHAUSTIVE: time the candidates */
    else
        bp = b2fft_plan_create(g);
    if (!bp) return 0;
    p = (fftw_plan)malloc(sizeof *p);
    if (!p) {
        b2fft_plan_destroy(bp);
        return 0;
    }
    memset(p, 0, sizeof *p);
    p->bp = bp;
    p->in = in;
    p->out = out;
    p->kind = g->kind;
    p->g = *g;
    wis_add(wis_hash(g), b2fft_plan_measured(bp) ? (unsigned)b2fft_plan_config(bp) : WIS_UNMEASURED);
    return p;
}

static fftw_plan finish_split(b2fft_geom *g, void *ri, void *ii, void *ro, void *io, unsigned flags) {
    fftw_plan p;
    b2fft_split *sp;
    if (flags & FFTW_WISDOM_ONLY) return 0;
    g->flags = flags;
    PLANNER_LOCK();
    sp = b2fft_split_plan_create(g);
    PLANNER_UNLOCK();
    if (!sp) return 0;
    p = (fftw_plan)malloc(sizeof *p);
    if (!p) {
        b2fft_split_plan_destroy(sp);
        return 0;
    }
    memset(p, 0, sizeof *p);
    p->sp = sp;
    p->ri = ri;
    p->ii = ii;
    p->ro = ro;
    p->io = io;
    p->kind = g->kind;
    p->g = *g;
    return p;
}

/* row-major strides inside an embedding array, like fftw_mktensor_rowmajor (fftw_api.c:842-861)
 * but in elements rather than in reals */
static void rowmajor(b2fft_geom *g, int rank, const int *n, const int *ie, const int *oe, ptrdiff_t is,
                     ptrdiff_t os) {
    int i;
    g->rank = rank;
    if (rank <= 0) return;
    g->dims[rank - 1].n = n[rank - 1];
    g->dims[rank - 1].is = is;
    g->dims[rank - 1].os = os;
    for (i = rank - 1; i > 0; --i) {
        g->dims[i - 1].n = n[i - 1];
        g->dims[i - 1].is = g->dims[i].is * ie[i];
        g->dims[i - 1].os = g->dims[i].os * oe[i];
    }
}

/* ------------------------------------------------------------------ complex DFT */
fftw_plan fftw_plan_many_dft(int rank, const int *n, int howmany, FFTW_COMPLEX *in, const int *inembed,
                             int istride, int idist, FFTW_COMPLEX *out, const int *onembed, int ostride,
                             int odist, int sign, unsigned flags) {
    b2fft_geom g;
    if (!kosher(rank, n, howmany)) return 0;
    memset(&g, 0, sizeof g);
    g.kind = B2FFT_C2C;
    g.sign = sign < 0 ? -1 : +1;
    rowmajor(&g, rank, n, inembed ? inembed : n, onembed ? onembed : n, istride, ostride);
    g.vrank = 1;
    g.vdims[0].n = howmany;
    g.vdims[0].is = idist;
    g.vdims[0].os = odist;
    return finish(&g, in, out, flags);
}

fftw_plan fftw_plan_dft(int rank, const int *n, FFTW_COMPLEX *in, FFTW_COMPLEX *out, int sign, unsigned flags) {
    return fftw_plan_many_dft(rank, n, 1, in, 0, 1, 1, out, 0, 1, 1, sign, flags);
}
fftw_plan fftw_plan_dft_1d(int n, FFTW_COMPLEX *in, FFTW_COMPLEX *out, int sign, unsigned flags) {
    return fftw_plan_dft(1, &n, in, out, sign, flags);
}
fftw_plan fftw_plan_dft_2d(int n0, int n1, FFTW_COMPLEX *in, FFTW_COMPLEX *out, int sign, unsigned flags) {
    int n[2];
    n[0] = n0;
    n[1] = n1;
    return fftw_plan_dft(2, n, in, out, sign, flags);
}
fftw_plan fftw_plan_dft_3d(int n0, int n1, int n2, FFTW_COMPLEX *in, FFTW_COMPLEX *out, int sign,
                           unsigned flags) {
    int n[3];
    n[0] = n0;
    n[1] = n1;
    n[2] = n2;
    return fftw_plan_dft(3, n, in, out, sign, flags);
}

/* ------------------------------------------------------------------ real DFT */
/* default embedding of the last dimension, the rule of fftw_rdft2_pad (fftw_api.c:774-788):
 * complex side n/2+1; real side 2(n/2+1) when in place, n otherwise */
static void pad_embed(int rank, const int *n, const int *embed, int inplace, int cmplx, int *tmp,
                      const int **res) {
    if (!embed && rank > 0) {
        memcpy(tmp, n, sizeof(int) * (size_t)rank);
        if (inplace || cmplx) tmp[rank - 1] = (n[rank - 1] / 2 + 1) * (cmplx ? 1 : 2);
        *res = tmp;
    } else
        *res = embed;
}

fftw_plan fftw_plan_many_dft_r2c(int rank, const int *n, int howmany, FFTW_REAL_TYPE *in, const int *inembed,
                                 int istride, int idist, FFTW_COMPLEX *out, const int *onembed, int ostride,
                                 int odist, unsigned flags) {
    b2fft_geom g;
    int ti[B2FFT_MAXRANK], to[B2FFT_MAXRANK];
    const int *ie, *oe;
    const int inplace = ((void *)in == (void *)out);
    if (!kosher(rank, n, howmany)) return 0;
    memset(&g, 0, sizeof g);
    g.kind = B2FFT_R2C;
    g.sign = -1;
    pad_embed(rank, n, inembed, inplace, 0, ti, &ie);
    pad_embed(rank, n, onembed, inplace, 1, to, &oe);
    rowmajor(&g, rank, n, ie, oe, istride, ostride);
    g.vrank = 1;
    g.vdims[0].n = howmany;
    g.vdims[0].is = idist;
    g.vdims[0].os = odist;
    return finish(&g, in, out, flags);
}

fftw_plan fftw_plan_many_dft_c2r(int rank, const int *n, int howmany, FFTW_COMPLEX *in, const int *inembed,
                                 int istride, int idist, FFTW_REAL_TYPE *out, const int *onembed, int ostride,
                                 int odist, unsigned flags) {
    b2fft_geom g;
    int ti[B2FFT_MAXRANK], to[B2FFT_MAXRANK];
    const int *ie, *oe;
    const int inplace = ((void *)in == (void *)out);
    if (!kosher(rank, n, howmany)) return 0;
    /* out-of-place c2r may overwrite its input, as in the reference (fftw_api.c:684-685) */
    if (!inplace) flags |= FFTW_DESTROY_INPUT;
    memset(&g, 0, sizeof g);
    g.kind = B2FFT_C2R;
    g.sign = +1;
    pad_embed(rank, n, inembed, inplace, 1, ti, &ie);
    pad_embed(rank, n, onembed, inplace, 0, to, &oe);
    rowmajor(&g, rank, n, ie, oe, istride, ostride);
    g.vrank = 1;
    g.vdims[0].n = howmany;
    g.vdims[0].is = idist;
    g.vdims[0].os = odist;
    return finish(&g, in, out, flags);
}

fftw_plan fftw_plan_dft_r2c(int rank, const int *n, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out, unsigned flags) {
    return fftw_plan_many_dft_r2c(rank, n, 1, in, 0, 1, 1, out, 0, 1, 1, flags);
}
fftw_plan fftw_plan_dft_r2c_1d(int n, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out, unsigned flags) {
    return fftw_plan_dft_r2c(1, &n, in, out, flags);
}
fftw_plan fftw_plan_dft_r2c_2d(int n0, int n1, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out, unsigned flags) {
    int n[2];
    n[0] = n0;
    n[1] = n1;
    return fftw_plan_dft_r2c(2, n, in, out, flags);
}
fftw_plan fftw_plan_dft_r2c_3d(int n0, int n1, int n2, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out, unsigned flags) {
    int n[3];
    n[0] = n0;
    n[1] = n1;
    n[2] = n2;
    return fftw_plan_dft_r2c(3, n, in, out, flags);
}
fftw_plan fftw_plan_dft_c2r(int rank, const int *n, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out, unsigned flags) {
    return fftw_plan_many_dft_c2r(rank, n, 1, in, 0, 1, 1, out, 0, 1, 1, flags);
}
fftw_plan fftw_plan_dft_c2r_1d(int n, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out, unsigned flags) {
    return fftw_plan_dft_c2r(1, &n, in, out, flags);
}
fftw_plan fftw_plan_dft_c2r_2d(int n0, int n1, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out, unsigned flags) {
    int n[2];
    n[0] = n0;
    n[1] = n1;
    return fftw_plan_dft_c2r(2, n, in, out, flags);
}
fftw_plan fftw_plan_dft_c2r_3d(int n0, int n1, int n2, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out, unsigned flags) {
    int n[3];
    n[0] = n0;
    n[1] = n1;
    n[2] = n2;
    return fftw_plan_dft_c2r(3, n, in, out, flags);
}

/* ------------------------------------------------------------------ guru (interleaved) */
/* The guru interface hands over the loop nest directly (reference fftw_api.c:1172-1517);
 * it maps one to one onto the geometry descriptor.  Split-array variants are not on the hot
 * path: they return NULL, which callers such as the reference's bench treat as "cannot do"
 * (fftw/tests/fftw-bench.c:180-198). */
#define GURU_BODY(KIND, SIGN)                                                         \
    b2fft_geom g;                                                                     \
    int i;                                                                            \
    if (rank < 0 || rank > B2FFT_MAXRANK || howmany_rank < 0 ||                       \
        howmany_rank > B2FFT_MAXRANK)                                                 \
        return 0;                                                                     \
    memset(&g, 0, sizeof g);                                                          \
    g.kind = (KIND);                                                                  \
    g.sign = (SIGN) < 0 ? -1 : +1;                                                    \
    g.rank = rank;                                                                    \
    for (i = 0; i < rank; ++i) {                                                      \
        if (dims[i].n <= 0) return 0;                                                 \
        g.dims[i].n = dims[i].n;                                                      \
        g.dims[i].is = dims[i].is;                                                    \
        g.dims[i].os = dims[i].os;                                                    \
    }                                                                                 \
    g.vrank = howmany_rank;                                                           \
    for (i = 0; i < howmany_rank; ++i) {                                              \
        if (howmany_dims[i].n < 0) return 0;                                          \
        g.vdims[i].n = howmany_dims[i].n;                                             \
        g.vdims[i].is = howmany_dims[i].is;                                           \
        g.vdims[i].os = howmany_dims[i].os;                                           \
    }

fftw_plan fftw_plan_guru_dft(int rank, const fftw_iodim *dims, int howmany_rank, const fftw_iodim *howmany_dims,
                             FFTW_COMPLEX *in, FFTW_COMPLEX *out, int sign, unsigned flags) {
    GURU_BODY(B2FFT_C2C, sign)
    return finish(&g, in, out, flags);
}
fftw_plan fftw_plan_guru64_dft(int rank, const fftw_iodim64 *dims, int howmany_rank,
                               const fftw_iodim64 *howmany_dims, FFTW_COMPLEX *in, FFTW_COMPLEX *out, int sign,
                               unsigned flags) {
    GURU_BODY(B2FFT_C2C, sign)
    return finish(&g, in, out, flags);
}
fftw_plan fftw_plan_guru_dft_r2c(int rank, const fftw_iodim *dims, int howmany_rank,
                                 const fftw_iodim *howmany_dims, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out,
                                 unsigned flags) {
    GURU_BODY(B2FFT_R2C, -1)
    if (rank < 1) return 0;
    return finish(&g, in, out, flags);
}
fftw_plan fftw_plan_guru64_dft_r2c(int rank, const fftw_iodim64 *dims, int howmany_rank,
                                   const fftw_iodim64 *howmany_dims, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out,
                                   unsigned flags) {
    GURU_BODY(B2FFT_R2C, -1)
    if (rank < 1) return 0;
    return finish(&g, in, out, flags);
}
fftw_plan fftw_plan_guru_dft_c2r(int rank, const fftw_iodim *dims, int howmany_rank,
                                 const fftw_iodim *howmany_dims, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out,
                                 unsigned flags) {
    GURU_BODY(B2FFT_C2R, +1)
    if (rank < 1) return 0;
    return finish(&g, in, out, flags);
}
fftw_plan fftw_plan_guru64_dft_c2r(int rank, const fftw_iodim64 *dims, int howmany_rank,
                                   const fftw_iodim64 *howmany_dims, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out,
                                   unsigned flags) {
    GURU_BODY(B2FFT_C2R, +1)
    if (rank < 1) return 0;
    return finish(&g, in, out, flags);
}

/* split arrays: gather / transform / scatter on the device (csrc/split.cu); always the forward
 * transform, like the reference (fftw_api.c:1420-1450: "sign" is fixed, callers swap ri/ii) */
fftw_plan fftw_plan_guru_split_dft(int rank, const fftw_iodim *dims, int howmany_rank,
                                   const fftw_iodim *howmany_dims, FFTW_REAL_TYPE *ri, FFTW_REAL_TYPE *ii,
                                   FFTW_REAL_TYPE *ro, FFTW_REAL_TYPE *io, unsigned flags) {
    GURU_BODY(B2FFT_C2C, -1)
    return finish_split(&g, ri, ii, ro, io, flags);
}
fftw_plan fftw_plan_guru64_split_dft(int rank, const fftw_iodim64 *dims, int howmany_rank,
                                     const fftw_iodim64 *howmany_dims, FFTW_REAL_TYPE *ri, FFTW_REAL_TYPE *ii,
                                     FFTW_REAL_TYPE *ro, FFTW_REAL_TYPE *io, unsigned flags) {
    GURU_BODY(B2FFT_C2C, -1)
    return finish_split(&g, ri, ii, ro, io, flags);
}
fftw_plan fftw_plan_guru_split_dft_r2c(int rank, const fftw_iodim *dims, int howmany_rank,
                                       const fftw_iodim *howmany_dims, FFTW_REAL_TYPE *in, FFTW_REAL_TYPE *ro,
                                       FFTW_REAL_TYPE *io, unsigned flags) {
    GURU_BODY(B2FFT_R2C, -1)
    if (rank < 1) return 0;
    return finish_split(&g, in, 0, ro, io, flags);
}
fftw_plan fftw_plan_guru_split_dft_c2r(int rank, const fftw_iodim *dims, int howmany_rank,
                                       const fftw_iodim *howmany_dims, FFTW_REAL_TYPE *ri, FFTW_REAL_TYPE *ii,
                                       FFTW_REAL_TYPE *out, unsigned flags) {
    GURU_BODY(B2FFT_C2R, +1)
    if (rank < 1) return 0;
    return finish_split(&g, ri, ii, out, 0, flags);
}
fftw_plan fftw_plan_guru64_split_dft_r2c(int rank, const fftw_iodim64 *dims, int howmany_rank,
                                         const fftw_iodim64 *howmany_dims, FFTW_REAL_TYPE *in,
                                         FFTW_REAL_TYPE *ro, FFTW_REAL_TYPE *io, unsigned flags) {
    GURU_BODY(B2FFT_R2C, -1)
    if (rank < 1) return 0;
    return finish_split(&g, in, 0, ro, io, flags);
}
fftw_plan fftw_plan_guru64_split_dft_c2r(int rank, const fftw_iodim64 *dims, int howmany_rank,
                                         const fftw_iodim64 *howmany_dims, FFTW_REAL_TYPE *ri,
                                         FFTW_REAL_TYPE *ii, FFTW_REAL_TYPE *out, unsigned flags) {
    GURU_BODY(B2FFT_C2R, +1)
    if (rank < 1) return 0;
    return finish_split(&g, ri, ii, out, 0, flags);
}

/* ------------------------------------------------------------------ r2r (csrc/r2r.cu) */
static fftw_plan finish_r2r(int rank, const b2fft_dim *dims, int vrank, const b2fft_dim *vdims,
                            const fftw_r2r_kind *kind, void *in, void *out, unsigned flags) {
    fftw_plan p;
    b2fft_r2r *rr;
    int k[B2FFT_MAXRANK], i;
    if (flags & FFTW_WISDOM_ONLY) return 0;
    if (rank < 0 || rank > B2FFT_MAXRANK || (rank > 0 && !kind)) return 0;
    for (i = 0; i < rank; ++i) k[i] = (int)kind[i];
    PLANNER_LOCK();
    rr = b2fft_r2r_plan_create(rank, dims, vrank, vdims, k);
    PLANNER_UNLOCK();
    if (!rr) return 0;
    p = (fftw_plan)malloc(sizeof *p);
    if (!p) {
        b2fft_r2r_plan_destroy(rr);
        return 0;
    }
    memset(p, 0, sizeof *p);
    p->rr = rr;
    p->in = in;
    p->out = out;
    p->kind = -1;
    return p;
}
fftw_plan fftw_plan_many_r2r(int rank, const int *n, int howmany, FFTW_REAL_TYPE *in, const int *inembed,
                             int istride, int idist, FFTW_REAL_TYPE *out, const int *onembed, int ostride,
                             int odist, const fftw_r2r_kind *kind, unsigned flags) {
    b2fft_geom g;
    b2fft_dim v;
    if (!kosher(rank, n, howmany)) return 0;
    memset(&g, 0, sizeof g);
    rowmajor(&g, rank, n, inembed ? inembed : n, onembed ? onembed : n, istride, ostride);
    v.n = howmany;
    v.is = idist;
    v.os = odist;
    return finish_r2r(rank, g.dims, 1, &v, kind, in, out, flags);
}
fftw_plan fftw_plan_r2r(int rank, const int *n, FFTW_REAL_TYPE *in, FFTW_REAL_TYPE *out,
                        const fftw_r2r_kind *kind, unsigned flags) {
    return fftw_plan_many_r2r(rank, n, 1, in, 0, 1, 1, out, 0, 1, 1, kind, flags);
}
fftw_plan fftw_plan_r2r_1d(int n, FFTW_REAL_TYPE *in, FFTW_REAL_TYPE *out, fftw_r2r_kind kind, unsigned flags) {
    return fftw_plan_r2r(1, &n, in, out, &kind, flags);
}
fftw_plan fftw_plan_r2r_2d(int n0, int n1, FFTW_REAL_TYPE *in, FFTW_REAL_TYPE *out, fftw_r2r_kind kind0,
                           fftw_r2r_kind kind1, unsigned flags) {
    int n[2];
    fftw_r2r_kind k[2];
    n[0] = n0;
    n[1] = n1;
    k[0] = kind0;
    k[1] = kind1;
    return fftw_plan_r2r(2, n, in, out, k, flags);
}
fftw_plan fftw_plan_r2r_3d(int n0, int n1, int n2, FFTW_REAL_TYPE *in, FFTW_REAL_TYPE *out,
                           fftw_r2r_kind kind0, fftw_r2r_kind kind1, fftw_r2r_kind kind2, unsigned flags) {
    int n[3];
    fftw_r2r_kind k[3];
    n[0] = n0;
    n[1] = n1;
    n[2] = n2;
    k[0] = kind0;
    k[1] = kind1;
    k[2] = kind2;
    return fftw_plan_r2r(3, n, in, out, k, flags);
}
fftw_plan fftw_plan_guru_r2r(int rank, const fftw_iodim *dims, int howmany_rank, const fftw_iodim *howmany_dims,
                             FFTW_REAL_TYPE *in, FFTW_REAL_TYPE *out, const fftw_r2r_kind *kind,
                             unsigned flags) {
    GURU_BODY(B2FFT_C2C, -1)
    return finish_r2r(rank, g.dims, howmany_rank, g.vdims, kind, in, out, flags);
}
fftw_plan fftw_plan_guru64_r2r(int rank, const fftw_iodim64 *dims, int howmany_rank,
                               const fftw_iodim64 *howmany_dims, FFTW_REAL_TYPE *in, FFTW_REAL_TYPE *out,
                               const fftw_r2r_kind *kind, unsigned flags) {
    GURU_BODY(B2FFT_C2C, -1)
    return finish_r2r(rank, g.dims, howmany_rank, g.vdims, kind, in, out, flags);
}

/* ------------------------------------------------------------------ execution */
static void run(const fftw_plan p, void *in, void *out) {
    int rc;
    if (!p || !p->bp) return;
    rc = b2fft_execute(p->bp, in, out);
    if (rc != 0) {
        /* fftw_execute has no error channel (returns void, reference fftw_api.c:428-431);
         * say so loudly rather than leave stale output unnoticed */
        fprintf(stderr, "b200fft: fftw_execute failed (%d): %s\n", rc, b2fft_last_error());
    }
}
static void run_split(const fftw_plan p, void *ri, void *ii, void *ro, void *io) {
    int rc;
    if (!p || !p->sp) return;
    rc = b2fft_split_execute(p->sp, ri, ii, ro, io);
    if (rc != 0) fprintf(stderr, "b200fft: fftw_execute (split arrays) failed (%d): %s\n", rc, b2fft_last_error());
}
static void run_r2r(const fftw_plan p, void *in, void *out) {
    int rc;
    if (!p || !p->rr) return;
    rc = b2fft_r2r_execute(p->rr, in, out);
    if (rc != 0) fprintf(stderr, "b200fft: fftw_execute (r2r) failed (%d): %s\n", rc, b2fft_last_error());
}
void fftw_execute(const fftw_plan p) {
    if (p && p->rr)
        run_r2r(p, p->in, p->out);
    else if (p && p->sp)
        run_split(p, p->ri, p->ii, p->ro, p->io);
    else if (p)
        run(p, p->in, p->out);
}
void fftw_execute_dft(const fftw_plan p, FFTW_COMPLEX *in, FFTW_COMPLEX *out) { run(p, in, out); }
void fftw_execute_dft_r2c(const fftw_plan p, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out) { run(p, in, out); }
void fftw_execute_dft_c2r(const fftw_plan p, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out) { run(p, in, out); }
void fftw_execute_split_dft(const fftw_plan p, FFTW_REAL_TYPE *ri, FFTW_REAL_TYPE *ii, FFTW_REAL_TYPE *ro,
                            FFTW_REAL_TYPE *io) {
    if (p && p->sp && p->kind == B2FFT_C2C) run_split(p, ri, ii, ro, io);
}
void fftw_execute_split_dft_r2c(const fftw_plan p, FFTW_REAL_TYPE *in, FFTW_REAL_TYPE *ro, FFTW_REAL_TYPE *io) {
    if (p && p->sp && p->kind == B2FFT_R2C) run_split(p, in, 0, ro, io);
}
void fftw_execute_split_dft_c2r(const fftw_plan p, FFTW_REAL_TYPE *ri, FFTW_REAL_TYPE *ii, FFTW_REAL_TYPE *out) {
    if (p && p->sp && p->kind == B2FFT_C2R) run_split(p, ri, ii, out, 0);
}
void fftw_execute_r2r(const fftw_plan p, FFTW_REAL_TYPE *in, FFTW_REAL_TYPE *out) { run_r2r(p, in, out); }

/* ------------------------------------------------------------------ lifecycle */
void fftw_destroy_plan(fftw_plan p) {
    if (!p) return; /* NULL is legal (reference fftw_api.c:410) */
    PLANNER_LOCK();
    b2fft_plan_destroy(p->bp);
    b2fft_split_plan_destroy(p->sp);
    b2fft_r2r_plan_destroy(p->rr);
    PLANNER_UNLOCK();
    free(p);
}
void fftw_cleanup(void) {
    PLANNER_LOCK();
    b2fft_cleanup();
    PLANNER_UNLOCK();
}
void fftw_forget_wisdom(void) {
    PLANNER_LOCK();
    wis_n = 0;
    PLANNER_UNLOCK();
}
void fftw_set_timelimit(double t) { (void)t; }

/* The reference keeps these four in its threads library (fftw/threads/api.c:39-86).  A GPU
 * plan has no host worker threads; they are accepted so that threaded callers still link. */
int fftw_init_threads(void) { return 1; }
void fftw_plan_with_nthreads(int nthreads) { (void)nthreads; }
void fftw_cleanup_threads(void) { fftw_cleanup(); }
void fftw_make_planner_thread_safe(void) {}

/* wisdom I/O: "(b200fft_wisdom-1 <hex signature> ...)" */
static const char wis_head[] = "(b200fft_wisdom-1";
char *fftw_export_wisdom_to_string(void) {
    size_t cap, len;
    char *s;
    int i;
    PLANNER_LOCK(); /* (before wis_n is read: a concurrent planner must not truncate the export) */
    cap = sizeof wis_head + (size_t)wis_n * 28 + 32;
    s = (char *)malloc(cap);
    if (!s) {
        PLANNER_UNLOCK();
        return 0;
    }
    len = (size_t)sprintf(s, "%s", wis_head);
    for (i = 0; i < wis_n && len + 32 < cap; ++i) {
        len += (size_t)sprintf(s + len, " %016llx", (unsigned long long)wis_key[i]);
        if (wis_cfg[i] != WIS_UNMEASURED) len += (size_t)sprintf(s + len, ":%x", wis_cfg[i]); /* the measured winner */
    }
    PLANNER_UNLOCK();
    sprintf(s + len, ")\n");
    return s;
}
void fftw_export_wisdom_to_file(FILE *f) {
    char *s = fftw_export_wisdom_to_string();
    if (f && s) fputs(s, f);
    free(s);
}
int fftw_export_wisdom_to_filename(const char *filename) {
    FILE *f = fopen(filename, "w");
    if (!f) return 0;
    fftw_export_wisdom_to_file(f);
    return fclose(f) == 0;
}
void fftw_export_wisdom(fftw_write_char_func write_char, void *data) {
    char *s = fftw_export_wisdom_to_string(), *c;
    if (write_char && s)
        for (c = s; *c; ++c) write_char(*c, data);
    free(s);
}
int fftw_import_system_wisdom(void) { return 0; }
/* 1 on success, 0 on a malformed list (reference: fftw_import_wisdom_from_string, fftw_api.c:1139) */
int fftw_import_wisdom_from_string(const char *s) {
    const char *c;
    if (!s) return 0;
    while (*s == ' ' || *s == '\n' || *s == '\t') ++s;
    if (strncmp(s, wis_head, sizeof wis_head - 1) != 0) return 0;
    c = s + sizeof wis_head - 1;
    for (;;) {
        char *end;
        unsigned long long k;
        unsigned cfg = WIS_UNMEASURED;
        while (*c == ' ' || *c == '\n' || *c == '\t') ++c;
        if (*c == ')') return 1;
        if (!*c) return 0;
        k = strtoull(c, &end, 16);
        if (end == c) return 0;
        if (*end == ':') {
            char *e2;
            cfg = (unsigned)strtoul(end + 1, &e2, 16);
            if (e2 == end + 1) return 0;
            end = e2;
        }
        PLANNER_LOCK();
        wis_add((uint64_t)k, cfg);
        PLANNER_UNLOCK();
        c = end;
    }
}
/* reads characters until get() says EOF into a growing buffer and imports them */
static int import_stream(int (*get)(void *), void *data) {
    size_t cap = 4096, n = 0;
    char *buf = (char *)malloc(cap), *nb;
    int c, ok;
    if (!buf) return 0;
    while ((c = get(data)) != EOF) {
        if (n + 2 > cap) {
            nb = (char *)realloc(buf, cap *= 2);
            if (!nb) {
                free(buf);
                return 0;
            }
            buf = nb;
        }
        buf[n++] = (char)c;
    }
    buf[n] = 0;
    ok = fftw_import_wisdom_from_string(buf);
    free(buf);
    return ok;
}
static int get_from_file(void *f) { return fgetc((FILE *)f); }
int fftw_import_wisdom_from_file(FILE *f) {
    if (!f) return 0;
    return import_stream(get_from_file, f);
}
int fftw_import_wisdom_from_filename(const char *filename) {
    FILE *f = fopen(filename, "r");
    int ok;
    if (!f) return 0;
    ok = fftw_import_wisdom_from_file(f);
    fclose(f);
    return ok;
}
int fftw_import_wisdom(fftw_read_char_func read_char, void *data) {
    if (!read_char) return 0;
    return import_stream(read_char, data);
}

/* plan printing: the selected kernel configurations (reference prints its solver tree,
 * fftw_api.c:1046-1078) */
char *fftw_sprint_plan(const fftw_plan p) {
    int n;
    char *s;
    if (!p) return 0;
    if (p->rr) {
        n = b2fft_r2r_describe(p->rr, 0, 0);
        s = (char *)malloc((size_t)n + 1);
        if (s) b2fft_r2r_describe(p->rr, s, (size_t)n + 1);
        return s;
    }
    if (p->sp) {
        n = b2fft_split_describe(p->sp, 0, 0);
        s = (char *)malloc((size_t)n + 1);
        if (s) b2fft_split_describe(p->sp, s, (size_t)n + 1);
        return s;
    }
    n = b2fft_plan_describe(p->bp, 0, 0);
    s = (char *)malloc((size_t)n + 1);
    if (s) b2fft_plan_describe(p->bp, s, (size_t)n + 1);
    return s;
}
void fftw_fprint_plan(const fftw_plan p, FILE *f) {
    char *s = fftw_sprint_plan(p);
    if (s && f) fputs(s, f);
    free(s);
}
void fftw_print_plan(const fftw_plan p) { fftw_fprint_plan(p, stdout); }

/* memory: pinned host memory, so that fftw_execute can DMA it at full PCIe speed */
void *fftw_malloc(size_t n) { return b2fft_malloc(n); }
FFTW_REAL_TYPE *fftw_alloc_real(size_t n) { return (FFTW_REAL_TYPE *)b2fft_malloc(n * sizeof(FFTW_REAL_TYPE)); }
FFTW_COMPLEX *fftw_alloc_complex(size_t n) { return (FFTW_COMPLEX *)b2fft_malloc(n * sizeof(FFTW_COMPLEX)); }
void fftw_free(void *p) { b2fft_free(p); }

/* 5 N log2 N convention (fftw/libbench2/mflops.c:6-31), split evenly between adds and muls */
void fftw_flops(const fftw_plan p, double *add, double *mul, double *fmas) {
    double f = !p ? 0.0 : p->rr ? b2fft_r2r_flops(p->rr) : p->sp ? b2fft_split_flops(p->sp) : b2fft_plan_flops(p->bp);
    if (add) *add = 0.6 * f;
    if (mul) *mul = 0.4 * f;
    if (fmas) *fmas = 0.0;
}
/* cost = the plan's flop count, whatever kind of plan it is (r2r and split-array plans keep theirs in their own objects) */
static double plan_cost(const fftw_plan p) {
    return !p ? 0.0 : p->rr ? b2fft_r2r_flops(p->rr) : p->sp ? b2fft_split_flops(p->sp) : b2fft_plan_flops(p->bp);
}
double fftw_estimate_cost(const fftw_plan p) { return plan_cost(p); }
double fftw_cost(const fftw_plan p) { return plan_cost(p); }
int fftw_alignment_of(FFTW_REAL_TYPE *p) { return (int)(((uintptr_t)p) % 16); } /* fftw_api.c:14530 */
