/* oracle/fft_oracle.c -- CPU restatement of the reference DFT hot path (double precision).
 *
 * TEST INFRASTRUCTURE ONLY (see fft_oracle.h).  Parity status: PINNED against the
 * unmodified reference build oracle/_ref via tests/golden (tests/test_oracle.py).
 *
 * This is a restatement of WHAT the reference computes, following its plan structure
 * (Cooley-Tukey DIT split, O(n^2) leaf for primes, rank>=2 axis order, vrank loop,
 * rdft2 via a half-size complex DFT + hc2c pass), not a copy of its code.
 */
#include "fft_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------------------------
 * Twiddles.  w(m,n) = exp(+2*pi*i*m/n).  Octant reduction so that sin/cos are only
 * evaluated on [0, pi/4], computed in long double and rounded once -- the accuracy
 * recipe of real_cexp (fftw/fftw_api.c:18850-18892).
 * ------------------------------------------------------------------------------- */
static void oracle_cexp(long m, long n, double *c, double *s)
{
    long double cr, sr, t;
    long q;
    int neg_s = 0, swap = 0, neg_c = 0;
    static const long double TWOPI = 6.2831853071795864769252867665590057683943388L;

    m %= n;
    if (m < 0) m += n;
    /* now 0 <= m < n ; fold to 0 <= m <= n/2 using conjugate symmetry */
    if (2 * m > n) { m = n - m; neg_s = 1; }
    /* fold to 0 <= m <= n/4 using cos(pi - x) = -cos x */
    if (4 * m > n) { m = n - 2 * m; q = 2 * n; neg_c = 1; /* angle = pi*m'/n with m' = n-2m over 2n */
        /* angle a = 2pi*m_old/n ; pi - a = 2pi*(n-2m_old)/(2n) */
    } else { q = n; }
    /* angle = 2*pi*m/q, with 0 <= m/q <= 1/4 ; fold to <= 1/8 by swapping sin/cos */
    if (8 * m > q) { m = q - 4 * m; q = 4 * q; swap = 1; } /* pi/2 - a = 2pi*(q-4m)/(4q) */
    t = TWOPI * (long double) m / (long double) q;
    cr = cosl(t);
    sr = sinl(t);
    if (swap) { long double tmp = cr; cr = sr; sr = tmp; }
    if (neg_c) cr = -cr;
    if (neg_s) sr = -sr;
    *c = (double) cr;
    *s = (double) sr;
}

/* full table w[k] = exp(sign*2*pi*i*k/n), k = 0..n-1 */
static double *twiddle_table(long n, int sign)
{
    double *w = (double *) malloc(sizeof(double) * 2 * (size_t) n);
    long k;
    if (!w) return 0;
    for (k = 0; k < n; ++k) {
        double c, s;
        oracle_cexp(k, n, &c, &s);
        w[2 * k] = c;
        w[2 * k + 1] = sign < 0 ? -s : s;
    }
    return w;
}

static long smallest_factor(long n)
{
    long p;
    if (n % 4 == 0) return 4;
    if (n % 2 == 0) return 2;
    for (p = 3; p * p <= n; p += 2)
        if (n % p == 0) return p;
    return n;
}

/* ---------------------------------------------------------------------------------
 * One contiguous 1-D DFT of length n: in -> out (both contiguous, distinct).
 * W is the table for the ROOT size N (W[k] = w_N^k), ws = N/n so w_n^k = W[k*ws].
 *
 *  - n prime (or tiny): direct O(n^2) sum, the role of the `generic` solver
 *    (generic_apply, fftw/fftw_api.c:3429-3448) and of the n1_* leaf codelets
 *    (direct_apply, fftw/fftw_api.c:3183-3187).
 *  - otherwise one decimation-in-time split n = r*m, the role of ct_apply_dit
 *    (fftw/fftw_api.c:2079-2090): r child DFTs of size m on the decimated inputs
 *    (cld), then radix-r twiddle butterflies over the m columns, in place on the
 *    output (cldw = dftw_direct_apply, fftw/fftw_api.c:2315-2324).
 * ------------------------------------------------------------------------------- */
static void dft_rec(long n, const double *in, long is, double *out,
                    const double *W, long ws, double *tmp /* 2*r_max */)
{
    long r, m, j, k, q;

    if (n == 1) {
        out[0] = in[0];
        out[1] = in[1];
        return;
    }
    r = smallest_factor(n);
    if (r == n) {
        for (k = 0; k < n; ++k) {
            double sr = 0.0, si = 0.0;
            for (j = 0; j < n; ++j) {
                long e = ((j * k) % n) * ws;
                double wr = W[2 * e], wi = W[2 * e + 1];
                double xr = in[2 * j * is], xi = in[2 * j * is + 1];
                sr += xr * wr - xi * wi;
                si += xr * wi + xi * wr;
            }
            out[2 * k] = sr;
            out[2 * k + 1] = si;
        }
        return;
    }
    m = n / r;
    /* cld: Y_j = DFT_m(x[j + r*i]) stored at out[j*m .. j*m+m) */
    for (j = 0; j < r; ++j)
        dft_rec(m, in + 2 * j * is, is * r, out + 2 * j * m, W, ws * r, tmp);
    /* cldw: X[k + m*q] = sum_j  w_n^(j*k) * w_r^(j*q) * Y_j[k] */
    for (k = 0; k < m; ++k) {
        for (j = 0; j < r; ++j) {
            long e = (j * k) * ws; /* j*k < n */
            double wr = W[2 * e], wi = W[2 * e + 1];
            double yr = out[2 * (j * m + k)], yi = out[2 * (j * m + k) + 1];
            tmp[2 * j] = yr * wr - yi * wi;
            tmp[2 * j + 1] = yr * wi + yi * wr;
        }
        for (q = 0; q < r; ++q) {
            double sr = 0.0, si = 0.0;
            for (j = 0; j < r; ++j) {
                long e = ((j * q) % r) * m * ws; /* w_r^(jq) = w_n^(jq*m) */
                double wr = W[2 * e], wi = W[2 * e + 1];
                sr += tmp[2 * j] * wr - tmp[2 * j + 1] * wi;
                si += tmp[2 * j] * wi + tmp[2 * j + 1] * wr;
            }
            out[2 * (k + m * q)] = sr;
            out[2 * (k + m * q) + 1] = si;
        }
    }
}

/* context for repeated line transforms of one length */
typedef struct {
    long n;
    double *W, *a, *b, *tmp;
} line_ctx;

static int line_ctx_init(line_ctx *c, long n, int sign)
{
    c->n = n;
    c->W = twiddle_table(n, sign);
    c->a = (double *) malloc(sizeof(double) * 2 * (size_t) n);
    c->b = (double *) malloc(sizeof(double) * 2 * (size_t) n);
    c->tmp = (double *) malloc(sizeof(double) * 2 * (size_t) (n > 4 ? n : 4));
    return (c->W && c->a && c->b && c->tmp) ? 0 : -1;
}

static void line_ctx_free(line_ctx *c)
{
    free(c->W); free(c->a); free(c->b); free(c->tmp);
}

/* strided in-place-capable line transform: gather, transform, scatter */
static void line_apply(line_ctx *c, const double *in, long is, double *out, long os)
{
    long j, n = c->n;
    for (j = 0; j < n; ++j) {
        c->a[2 * j] = in[2 * j * is];
        c->a[2 * j + 1] = in[2 * j * is + 1];
    }
    dft_rec(n, c->a, 1, c->b, c->W, 1, c->tmp);
    for (j = 0; j < n; ++j) {
        out[2 * j * os] = c->b[2 * j];
        out[2 * j * os + 1] = c->b[2 * j + 1];
    }
}

/* row-major strides (in elements) of a rank-r array embedded in nembed,
 * as fftw_mktensor_rowmajor builds them (fftw/fftw_api.c:842-861) */
static void rowmajor_strides(int rank, const int *nembed, long stride, long *s)
{
    int i;
    if (rank <= 0) return;
    s[rank - 1] = stride;
    for (i = rank - 1; i > 0; --i) s[i - 1] = s[i] * nembed[i];
}

static int kosher(int rank, const int *n, int howmany)
{
    int i;
    if (rank < 0 || howmany < 0) return 0; /* fftw_many_kosherp, fftw/fftw_api.c:863-877 */
    for (i = 0; i < rank; ++i)
        if (n[i] <= 0) return 0;
    return 1;
}

/* all index tuples of dims except `axis`: calls line_apply along `axis` in place on `a` */
static void axis_pass_inplace(line_ctx *c, int rank, const int *n, const long *s,
                              int axis, double *a)
{
    long idx[16];
    int d;
    memset(idx, 0, sizeof(idx));
    for (;;) {
        long off = 0;
        for (d = 0; d < rank; ++d)
            if (d != axis) off += idx[d] * s[d];
        line_apply(c, a + 2 * off, s[axis], a + 2 * off, s[axis]);
        for (d = rank - 1; d >= 0; --d) {
            if (d == axis) continue;
            if (++idx[d] < n[d]) break;
            idx[d] = 0;
        }
        if (d < 0) break;
    }
}

/* ---------------------------------------------------------------------------------
 * oracle_dft_many: the vrank>=1 loop over `howmany` (vrank_geq1_apply,
 * fftw/fftw_api.c:4627-4638) around a rank>=2 decomposition that transforms the
 * trailing dimensions first and the leading dimension last, in place on the output
 * (rank_geq2_apply, fftw/fftw_api.c:4435-4445).
 * ------------------------------------------------------------------------------- */
int oracle_dft_many(int rank, const int *n, int howmany,
                    const double *in, const int *inembed, int istride, int idist,
                    double *out, const int *onembed, int ostride, int odist, int sign)
{
    long is[16], os[16], cs[16];
    long total = 1, t, v;
    int d;
    double *buf;

    if (!kosher(rank, n, howmany) || rank > 16) return -1;
    if (!inembed) inembed = n;
    if (!onembed) onembed = n;
    rowmajor_strides(rank, inembed, istride, is);
    rowmajor_strides(rank, onembed, ostride, os);
    rowmajor_strides(rank, n, 1, cs);
    for (d = 0; d < rank; ++d) total *= n[d];

    buf = (double *) malloc(sizeof(double) * 2 * (size_t) total);
    if (!buf) return -1;

    for (v = 0; v < howmany; ++v) {
        const double *pi = in + 2 * (long) v * idist;
        double *po = out + 2 * (long) v * odist;
        /* gather the transform into a compact row-major buffer */
        for (t = 0; t < total; ++t) {
            long rem = t, off = 0;
            for (d = rank - 1; d >= 0; --d) { off += (rem % n[d]) * is[d]; rem /= n[d]; }
            buf[2 * t] = pi[2 * off];
            buf[2 * t + 1] = pi[2 * off + 1];
        }
        for (d = rank - 1; d >= 0; --d) {
            line_ctx c;
            if (line_ctx_init(&c, n[d], sign)) { free(buf); return -1; }
            axis_pass_inplace(&c, rank, n, cs, d, buf);
            line_ctx_free(&c);
        }
        for (t = 0; t < total; ++t) {
            long rem = t, off = 0;
            for (d = rank - 1; d >= 0; --d) { off += (rem % n[d]) * os[d]; rem /= n[d]; }
            po[2 * off] = buf[2 * t];
            po[2 * off + 1] = buf[2 * t + 1];
        }
    }
    free(buf);
    return 0;
}

/* ---------------------------------------------------------------------------------
 * 1-D real <-> half-complex lines.
 *
 * even n: the HC2C_VIA_DFT route of rdft2-ct (ct_hc2c_mkplan fftw/fftw_api.c:5661-5788,
 * branch :5711-5724; ct_hc2c_apply_dit_dft :5579-5590): view the reals as m = n/2
 * complex numbers z[j] = x[2j] + i x[2j+1] (the r0/r1 even/odd pointers of
 * fftw_mkproblem_rdft2_d_3pointers, fftw/fftw_api.c:9358-9379), take DFT_m, then one
 * hc2c pass pairing column k with column m-k (ct_hc2c_direct_apply :5831-5845):
 *      X[k] = 1/2 [ (Z[k] + conj Z[m-k])  -  i w_n^k (Z[k] - conj Z[m-k]) ],  k = 0..m
 * odd n: a complex DFT of the real data (the rdft2-rdft / generic fallback,
 * rdft2_rdft_hc2c fftw/fftw_api.c:6482-6497), keeping outputs 0..n/2.
 * ------------------------------------------------------------------------------- */
typedef struct { long n; int sign; line_ctx c; double *wn, *z; } rline_ctx;

static int rline_init(rline_ctx *r, long n, int sign)
{
    r->n = n; r->sign = sign;
    r->wn = twiddle_table(n, sign);
    r->z = (double *) malloc(sizeof(double) * 2 * (size_t) n);
    if (!r->wn || !r->z) return -1;
    return line_ctx_init(&r->c, n % 2 == 0 ? n / 2 : n, sign);
}

static void rline_free(rline_ctx *r) { free(r->wn); free(r->z); line_ctx_free(&r->c); }

static int r2c_line(rline_ctx *rc, const double *x, long xs, double *X, long Xs)
{
    long j, k, n = rc->n;
    double *wn = rc->wn, *z = rc->z;
    if (n % 2 == 0) {
        long m = n / 2;
        line_ctx c = rc->c;
        for (j = 0; j < m; ++j) { z[2 * j] = x[(2 * j) * xs]; z[2 * j + 1] = x[(2 * j + 1) * xs]; }
        line_apply(&c, z, 1, z, 1);
        for (k = 0; k <= m; ++k) {
            long k0 = k % m, k1 = (m - k) % m;
            double ar = z[2 * k0], ai = z[2 * k0 + 1];
            double br = z[2 * k1], bi = -z[2 * k1 + 1];      /* conj Z[m-k] */
            double er = 0.5 * (ar + br), ei = 0.5 * (ai + bi); /* E[k] */
            double dr = 0.5 * (ar - br), di = 0.5 * (ai - bi); /* i O[k] */
            /* -i * (dr + i di) = di - i dr */
            double orr = di, oi = -dr;
            double wr, wi;
            if (k == m) { wr = -1.0; wi = 0.0; } else { wr = wn[2 * k]; wi = wn[2 * k + 1]; }
            X[2 * k * Xs] = er + (orr * wr - oi * wi);
            X[2 * k * Xs + 1] = ei + (orr * wi + oi * wr);
        }
    } else {
        line_ctx c = rc->c;
        for (j = 0; j < n; ++j) { z[2 * j] = x[j * xs]; z[2 * j + 1] = 0.0; }
        line_apply(&c, z, 1, z, 1);
        for (k = 0; k <= n / 2; ++k) { X[2 * k * Xs] = z[2 * k]; X[2 * k * Xs + 1] = z[2 * k + 1]; }
    }
    return 0;
}

/* c2r mirror (rdft2-ct-dif, ct_hc2c_apply_dif_dft fftw/fftw_api.c:5592-5603):
 *      Z[k] = (X[k] + conj X[m-k]) + i w_n^{-k} (X[k] - conj X[m-k]),  z = DFT^{+}_m(Z),
 *      x[2j] = Re z[j], x[2j+1] = Im z[j].   Unnormalised: c2r(r2c(x)) = n x.
 * The imaginary parts of X[0] (and X[n/2] for even n) are ignored, as in FFTW. */
static int c2r_line(rline_ctx *rc, const double *X, long Xs, double *x, long xs)
{
    long j, k, n = rc->n;
    double *wn = rc->wn, *z = rc->z;
    if (n % 2 == 0) {
        long m = n / 2;
        line_ctx c = rc->c;
        for (k = 0; k < m; ++k) {
            double ar = X[2 * k * Xs], ai = X[2 * k * Xs + 1];
            double br = X[2 * (m - k) * Xs], bi = -X[2 * (m - k) * Xs + 1];
            double sr, si, dr, di, tr, ti;
            if (k == 0) { ai = 0.0; bi = 0.0; }
            sr = ar + br; si = ai + bi;
            dr = ar - br; di = ai - bi;
            /* i * w * d */
            tr = dr * wn[2 * k] - di * wn[2 * k + 1];
            ti = dr * wn[2 * k + 1] + di * wn[2 * k];
            z[2 * k] = sr - ti;
            z[2 * k + 1] = si + tr;
        }
        line_apply(&c, z, 1, z, 1);
        for (j = 0; j < m; ++j) { x[(2 * j) * xs] = z[2 * j]; x[(2 * j + 1) * xs] = z[2 * j + 1]; }
    } else {
        line_ctx c = rc->c;
        z[0] = X[0]; z[1] = 0.0;
        for (k = 1; k <= n / 2; ++k) {
            z[2 * k] = X[2 * k * Xs]; z[2 * k + 1] = X[2 * k * Xs + 1];
            z[2 * (n - k)] = X[2 * k * Xs]; z[2 * (n - k) + 1] = -X[2 * k * Xs + 1];
        }
        line_apply(&c, z, 1, z, 1);
        for (j = 0; j < n; ++j) x[j * xs] = z[2 * j];
    }
    return 0;
}

/* ---------------------------------------------------------------------------------
 * N-D real transforms: rank>=2 rdft2 (rank_geq2_rdft2_apply_r2hc / _hc2r,
 * fftw/fftw_api.c:10111-10139): r2c = real DFT along the LAST dimension, then complex
 * DFTs along the remaining dimensions over the n_last/2+1 columns; c2r = the reverse
 * order.  Last-dimension sizes follow fftw_rdft2_complex_n (fftw/fftw_api.c:9252-9256).
 * ------------------------------------------------------------------------------- */
int oracle_r2c_many(int rank, const int *n, int howmany,
                    const double *in, const int *inembed, int istride, int idist,
                    double *out, const int *onembed, int ostride, int odist)
{
    long is[16], os[16], cs[16];
    int nc[16];
    long rows = 1, ctotal = 1, t, v;
    int d, nl;
    double *buf;
    rline_ctx rc;

    if (!kosher(rank, n, howmany) || rank < 1 || rank > 16) return -1;
    nl = n[rank - 1];
    for (d = 0; d < rank; ++d) nc[d] = n[d];
    nc[rank - 1] = nl / 2 + 1;
    if (!inembed) inembed = n;
    if (!onembed) onembed = nc;
    rowmajor_strides(rank, inembed, istride, is);
    rowmajor_strides(rank, onembed, ostride, os);
    rowmajor_strides(rank, nc, 1, cs);
    for (d = 0; d < rank - 1; ++d) rows *= n[d];
    ctotal = rows * nc[rank - 1];

    buf = (double *) malloc(sizeof(double) * 2 * (size_t) ctotal);
    if (!buf || rline_init(&rc, nl, -1)) return -1;

    for (v = 0; v < howmany; ++v) {
        const double *pi = in + (long) v * idist;
        double *po = out + 2 * (long) v * odist;
        for (t = 0; t < rows; ++t) {
            long rem = t, off = 0;
            for (d = rank - 2; d >= 0; --d) { off += (rem % n[d]) * is[d]; rem /= n[d]; }
            if (r2c_line(&rc, pi + off, is[rank - 1], buf + 2 * t * nc[rank - 1], 1)) return -1;
        }
        for (d = rank - 2; d >= 0; --d) {
            line_ctx c;
            if (line_ctx_init(&c, n[d], -1)) return -1;
            axis_pass_inplace(&c, rank, nc, cs, d, buf);
            line_ctx_free(&c);
        }
        for (t = 0; t < ctotal; ++t) {
            long rem = t, off = 0;
            for (d = rank - 1; d >= 0; --d) { off += (rem % nc[d]) * os[d]; rem /= nc[d]; }
            po[2 * off] = buf[2 * t];
            po[2 * off + 1] = buf[2 * t + 1];
        }
    }
    rline_free(&rc);
    free(buf);
    return 0;
}

int oracle_c2r_many(int rank, const int *n, int howmany,
                    const double *in, const int *inembed, int istride, int idist,
                    double *out, const int *onembed, int ostride, int odist)
{
    long is[16], os[16], cs[16];
    int nc[16];
    long rows = 1, ctotal = 1, t, v;
    int d, nl;
    double *buf;
    rline_ctx rc;

    if (!kosher(rank, n, howmany) || rank < 1 || rank > 16) return -1;
    nl = n[rank - 1];
    for (d = 0; d < rank; ++d) nc[d] = n[d];
    nc[rank - 1] = nl / 2 + 1;
    if (!inembed) inembed = nc;
    if (!onembed) onembed = n;
    rowmajor_strides(rank, inembed, istride, is);
    rowmajor_strides(rank, onembed, ostride, os);
    rowmajor_strides(rank, nc, 1, cs);
    for (d = 0; d < rank - 1; ++d) rows *= n[d];
    ctotal = rows * nc[rank - 1];

    buf = (double *) malloc(sizeof(double) * 2 * (size_t) ctotal);
    if (!buf || rline_init(&rc, nl, +1)) return -1;

    for (v = 0; v < howmany; ++v) {
        const double *pi = in + 2 * (long) v * idist;
        double *po = out + (long) v * odist;
        for (t = 0; t < ctotal; ++t) {
            long rem = t, off = 0;
            for (d = rank - 1; d >= 0; --d) { off += (rem % nc[d]) * is[d]; rem /= nc[d]; }
            buf[2 * t] = pi[2 * off];
            buf[2 * t + 1] = pi[2 * off + 1];
        }
        for (d = 0; d <= rank - 2; ++d) {
            line_ctx c;
            if (line_ctx_init(&c, n[d], +1)) return -1;
            axis_pass_inplace(&c, rank, nc, cs, d, buf);
            line_ctx_free(&c);
        }
        for (t = 0; t < rows; ++t) {
            long rem = t, off = 0;
            for (d = rank - 2; d >= 0; --d) { off += (rem % n[d]) * os[d]; rem /= n[d]; }
            if (c2r_line(&rc, buf + 2 * t * nc[rank - 1], 1, po + off, os[rank - 1])) return -1;
        }
    }
    rline_free(&rc);
    free(buf);
    return 0;
}

/* ---------------------------------------------------------------------------------
 * Slab-distributed 3-D form.  Block distribution: block = ceil(n/P)
 * (default_block, fftw/mpi/block.c:39-42); rank p owns indices
 * [p*block, min(n,(p+1)*block)) (block(), fftw/mpi/block.c:46-50); ranks past the
 * last block own nothing (local_size, fftw/mpi/api.c:218-232).
 * ------------------------------------------------------------------------------- */
static ptrdiff_t blk_default(ptrdiff_t n, int n_pes) { return (n + n_pes - 1) / n_pes; }
static ptrdiff_t blk_len(ptrdiff_t n, ptrdiff_t block, int which)
{
    ptrdiff_t d = n - which * block;
    return d <= 0 ? 0 : (d > block ? block : d);
}
static ptrdiff_t imax(ptrdiff_t a, ptrdiff_t b) { return a > b ? a : b; }

/* local_size_guru (fftw/mpi/api.c:248-352) specialised to rnk = 3, howmany = 1, with
 * input distributed over dim 0 and output over dim `odim` (0, or 1 = TRANSPOSED_OUT). */
static ptrdiff_t local_size_3d_common(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2,
                                      int n_pes, int my_pe, int odim,
                                      ptrdiff_t *ln_in, ptrdiff_t *ls_in,
                                      ptrdiff_t *ln_out, ptrdiff_t *ls_out)
{
    ptrdiff_t N = 1, b0 = blk_default(n0, n_pes), b1 = blk_default(n1, n_pes);
    ptrdiff_t in0, out_n, Nafter, bo = odim == 0 ? b0 : b1, no = odim == 0 ? n0 : n1;

    in0 = blk_len(n0, b0, my_pe);
    out_n = blk_len(no, bo, my_pe);
    *ln_in = in0;  *ls_in = in0 ? my_pe * b0 : 0;
    *ln_out = out_n; *ls_out = out_n ? my_pe * bo : 0;

    /* "extra space for transposed intermediate data" (fftw/mpi/api.c:278-293): the first
       of dims {0,1} that is undistributed on BOTH input and output gets distributed */
    if (odim == 0) /* dim 1 is undistributed on both sides */
        N = imax(N, n0 * blk_len(n1, b1, my_pe) * n2);
    /* dft-rank-geq2 (:295-300) and dft-rank-geq2 with dims swapped (:302-307) */
    Nafter = n1 * n2;
    N = imax(N, n0 * blk_len(Nafter, blk_default(Nafter, n_pes), my_pe));
    Nafter = n0 * n2;
    N = imax(N, n1 * blk_len(Nafter, blk_default(Nafter, n_pes), my_pe));
    /* input and output themselves (:347-348) */
    N = imax(N, imax(in0 * n1 * n2, odim == 0 ? out_n * n1 * n2 : out_n * n0 * n2));
    return N;
}

ptrdiff_t oracle_mpi_local_size_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2,
                                   int n_pes, int my_pe,
                                   ptrdiff_t *local_n0, ptrdiff_t *local_0_start)
{
    ptrdiff_t a, b;
    return local_size_3d_common(n0, n1, n2, n_pes, my_pe, 0, local_n0, local_0_start, &a, &b);
}

ptrdiff_t oracle_mpi_local_size_3d_transposed(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2,
                                              int n_pes, int my_pe,
                                              ptrdiff_t *local_n0, ptrdiff_t *local_0_start,
                                              ptrdiff_t *local_n1, ptrdiff_t *local_1_start)
{
    return local_size_3d_common(n0, n1, n2, n_pes, my_pe, 1,
                                local_n0, local_0_start, local_n1, local_1_start);
}

int oracle_mpi_dft_3d_local(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2,
                            int n_pes, int my_pe, int transposed_out, int sign,
                            const double *global_in, double *local_out)
{
    int n[3];
    ptrdiff_t ln0, ls0, ln1, ls1, i, j, k;
    double *full;
    n[0] = (int) n0; n[1] = (int) n1; n[2] = (int) n2;
    full = (double *) malloc(sizeof(double) * 2 * (size_t) (n0 * n1 * n2));
    if (!full) return -1;
    /* the distributed result equals the serial DFT, redistributed: the check
       mpi-bench itself performs (fftw/mpi/mpi-bench.c:236-257) */
    if (oracle_dft_many(3, n, 1, global_in, 0, 1, 1, full, 0, 1, 1, sign)) { free(full); return -1; }
    oracle_mpi_local_size_3d_transposed(n0, n1, n2, n_pes, my_pe, &ln0, &ls0, &ln1, &ls1);
    if (transposed_out) {
        /* local [local_n1][n0][n2]  =  global Y[i][ls1 + j][k] at local (j, i, k) */
        for (j = 0; j < ln1; ++j)
            for (i = 0; i < n0; ++i)
                for (k = 0; k < n2; ++k) {
                    ptrdiff_t g = (i * n1 + (ls1 + j)) * n2 + k, l = (j * n0 + i) * n2 + k;
                    local_out[2 * l] = full[2 * g];
                    local_out[2 * l + 1] = full[2 * g + 1];
                }
    } else {
        memcpy(local_out, full + 2 * ls0 * n1 * n2, sizeof(double) * 2 * (size_t) (ln0 * n1 * n2));
    }
    free(full);
    return 0;
}
