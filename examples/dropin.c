/* examples/dropin.c -- a plain C program written against the reference's public header only
 * (fftw/fftw3.h: plan, execute, destroy; precision by -DFFTW_SINGLE).  It builds unchanged against the
 * reference and against this engine; tests/test_abi.py compiles it with -std=c99 -pedantic -Werror and
 * runs it (emulation library on the CPU, product library on the B200).  Exit code 0 = all checks passed. */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "fftw3.h"

#ifdef FFTW_SINGLE
#define TOL 2e-5
#else
#define TOL 1e-11
#endif
#define PI 3.14159265358979323846

static double frand(unsigned *s) {
    *s = *s * 1664525u + 1013904223u;
    return (double)(*s >> 8) / (double)(1u << 24) - 0.5;
}

int main(void) {
    const int n = 48, hm = 3;
    unsigned seed = 1;
    int j, k, b, bad = 0;
    double err, nrm;
    FFTW_COMPLEX *x = (FFTW_COMPLEX *)fftw_malloc(sizeof(FFTW_COMPLEX) * (size_t)(n * hm));
    FFTW_COMPLEX *y = (FFTW_COMPLEX *)fftw_malloc(sizeof(FFTW_COMPLEX) * (size_t)(n * hm));
    FFTW_REAL_TYPE *r = fftw_alloc_real((size_t)n), *c = fftw_alloc_real((size_t)n);
    FFTW_COMPLEX *h = fftw_alloc_complex((size_t)(n / 2 + 1));
    fftw_plan p;
    if (!x || !y || !r || !c || !h) return 10;

    /* batched complex forward transform through the advanced interface */
    p = fftw_plan_many_dft(1, &n, hm, x, NULL, 1, n, y, NULL, 1, n, FFTW_FORWARD, FFTW_ESTIMATE);
    if (!p) return 11;
    for (j = 0; j < n * hm; ++j) {
        x[j][0] = (FFTW_REAL_TYPE)frand(&seed);
        x[j][1] = (FFTW_REAL_TYPE)frand(&seed);
    }
    fftw_execute(p);
    err = nrm = 0;
    for (b = 0; b < hm; ++b)
        for (k = 0; k < n; ++k) {
            double re = 0, im = 0;
            for (j = 0; j < n; ++j) {
                const double a = -2 * PI * (double)((j * k) % n) / n, cr = cos(a), sr = sin(a);
                re += x[b * n + j][0] * cr - x[b * n + j][1] * sr;
                im += x[b * n + j][0] * sr + x[b * n + j][1] * cr;
            }
            err += (y[b * n + k][0] - re) * (y[b * n + k][0] - re) + (y[b * n + k][1] - im) * (y[b * n + k][1] - im);
            nrm += re * re + im * im;
        }
    if (sqrt(err / nrm) > TOL) bad |= 1;
    fftw_destroy_plan(p);

    /* real -> half-complex and back (unnormalised: c2r(r2c(x)) = n x) */
    p = fftw_plan_dft_r2c_1d(n, r, h, FFTW_ESTIMATE);
    if (!p) return 12;
    for (j = 0; j < n; ++j) r[j] = (FFTW_REAL_TYPE)frand(&seed);
    fftw_execute(p);
    fftw_destroy_plan(p);
    p = fftw_plan_dft_c2r_1d(n, h, c, FFTW_ESTIMATE);
    if (!p) return 13;
    fftw_execute(p);
    fftw_destroy_plan(p);
    err = nrm = 0;
    for (j = 0; j < n; ++j) {
        err += (c[j] / n - r[j]) * (c[j] / n - r[j]);
        nrm += r[j] * r[j];
    }
    if (sqrt(err / nrm) > TOL) bad |= 2;

    /* DCT-II (FFTW_REDFT10) against its definition */
    p = fftw_plan_r2r_1d(n, r, c, FFTW_REDFT10, FFTW_ESTIMATE);
    if (!p) return 14;
    fftw_execute(p);
    fftw_destroy_plan(p);
    err = nrm = 0;
    for (k = 0; k < n; ++k) {
        double s = 0;
        for (j = 0; j < n; ++j) s += 2 * r[j] * cos(PI * (j + 0.5) * k / n);
        err += (c[k] - s) * (c[k] - s);
        nrm += s * s;
    }
    if (sqrt(err / nrm) > 10 * TOL) bad |= 4;

    /* error behaviour of the reference: invalid size -> NULL, destroy(NULL) legal */
    k = 0;
    if (fftw_plan_dft_1d(k, x, y, FFTW_FORWARD, FFTW_ESTIMATE) != NULL) bad |= 8;
    fftw_destroy_plan(NULL);

    fftw_free(x);
    fftw_free(y);
    fftw_free(r);
    fftw_free(c);
    fftw_free(h);
    fftw_cleanup();
    printf("dropin: %s (mask %d)\n", bad ? "FAILED" : "ok", bad);
    return bad;
}
