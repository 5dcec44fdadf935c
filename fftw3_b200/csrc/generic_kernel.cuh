// csrc/generic_kernel.cuh -- the general axis-pass kernel: any length m = prod(radix) with
// radices in {2,3,4,5,7,8,11,13,16}, any strides, complex or real (r2c / c2r) lines, optional
// four-step twiddle epilogue.  One CTA transforms a tile of C lines held in shared memory
// (Stockham auto-sort, ping-pong between two buffers); global memory is touched exactly
// once on the way in and once on the way out.
//
// Replaces, for one solver step of the reference: the Cooley-Tukey twiddle loop
// (dftw_direct_apply, fftw/fftw_api.c:2315-2324), the leaf codelet loop (direct_apply,
// fftw_api.c:3183-3187), the batch loop (vrank_geq1_apply, fftw_api.c:4627-4638), and for
// real data the hc2c post/pre-processing step (ct_hc2c_direct_apply, fftw_api.c:5831-5845;
// codelets hc2cfdftv_* / hc2cbdftv_*, r2cf_* / r2cb_*).
// Hot geometries are served by the specialised kernels in tile_kernel.cuh instead.
#pragma once
#include "butterfly.cuh"
#include "pass_params.h"

B2_D cpx_t b2_load_pt(const real_t *B2_RESTRICT src, long long off, long long pair, int vec) {
    if (vec) return *reinterpret_cast<const cpx_t *>(src + off);
    return cpx_t{src[off], src[off + pair]};
}
B2_D void b2_store_pt(real_t *B2_RESTRICT dst, long long off, long long pair, int vec, cpx_t v) {
    if (vec) {
        *reinterpret_cast<cpx_t *>(dst + off) = v;
    } else {
        dst[off] = v.x;
        dst[off + pair] = v.y;
    }
}

// one Stockham radix-R step over a tile: in/out are smem buffers of C lines, stride ls
template <int R>
B2_D void b2_radix_step(const cpx_t *in, cpx_t *out, const cpx_t *B2_RESTRICT tw, int m, int Ns,
                        int nlines, int ls, int tid, int nthreads, unsigned nbm, unsigned nbs, unsigned nsm,
                        unsigned nss) {
    const int nb = m / R;
    const int step = m / (Ns * R);
    const int total = nb * nlines;
    for (int b = tid; b < total; b += nthreads) {
        // exact multiply-shift divisions by nb and Ns (b2_pp_prepare_generic)
        const int c = (int)b2_fastdiv((unsigned)b, nbm, nbs), j = b - c * nb;
        const int jd = (int)b2_fastdiv((unsigned)j, nsm, nss), jm = j - jd * Ns;
        const cpx_t *x = in + (size_t)c * ls + j;
        cpx_t v[R];
#pragma unroll
        for (int k = 0; k < R; ++k) v[k] = x[k * nb];
        if (Ns > 1) {
            const int ts = jm * step;
#pragma unroll
            for (int k = 1; k < R; ++k) v[k] = cmul(v[k], tw[ts * k]);
        }
        Bf<R, real_t>::run(v);
        cpx_t *y = out + (size_t)c * ls + jd * Ns * R + jm;
#pragma unroll
        for (int k = 0; k < R; ++k) y[k * Ns] = v[k];
    }
}

struct GenericSmem {
    long long *soff, *doff;
    int *i2;
    cpx_t *a, *b;
};
B2_HD size_t b2_generic_smem_bytes(int C, int lstride) {
    return (size_t)C * (2 * sizeof(long long) + 2 * sizeof(int)) + 2 * (size_t)C * lstride * sizeof(cpx_t);
}

B2_D void b2_generic_pass_body(const PassParams &p, const real_t *B2_RESTRICT src,
                               real_t *B2_RESTRICT dst, long long bid, char *smem,
                               int b2_nthreads) {
    (void)b2_nthreads;
    GenericSmem S;
    S.a = reinterpret_cast<cpx_t *>(smem);
    S.b = S.a + (size_t)p.C * p.lstride;
    S.soff = reinterpret_cast<long long *>(S.b + (size_t)p.C * p.lstride);
    S.doff = S.soff + p.C;
    S.i2 = reinterpret_cast<int *>(S.doff + p.C);
    const int ls = p.lstride;

    // ---- tile coordinates ----
    const long long t0 = bid % p.tiles0;
    long long rest = bid / p.tiles0;
    const long long l0 = t0 * p.C;
    const int nlines = (int)((p.ln[0] - l0 < p.C) ? (p.ln[0] - l0) : p.C);
    long long sbase = 0, dbase = 0;
    int i2base = 0;
    for (int d = 1; d < p.nld; ++d) {
        const long long c = rest % p.ln[d];
        rest /= p.ln[d];
        sbase += c * p.ls[d];
        dbase += c * p.ld[d];
        if (d == p.epi_dim) i2base = (int)c;
    }
    B2_FOR_THREADS {
        for (int c = tid; c < nlines; c += B2_NTHREADS) {
            S.soff[c] = sbase + (l0 + c) * p.ls[0];
            S.doff[c] = dbase + (l0 + c) * p.ld[0];
            S.i2[c] = (p.epi_dim == 0) ? (int)(l0 + c) : i2base;
        }
    }
    B2_SYNC();

    // ---- load ----
    B2_FOR_THREADS {
        const int total = nlines * p.nload;
        for (int idx = tid; idx < total; idx += B2_NTHREADS) {
            int c, j;
            if (p.ld_col) {
                j = (nlines == p.C) ? (int)b2_fastdiv((unsigned)idx, p.gd_m[2], p.gd_s[2]) : idx / nlines;
                c = idx - j * nlines;
            } else {
                c = (int)b2_fastdiv((unsigned)idx, p.gd_m[0], p.gd_s[0]);
                j = idx - c * p.nload;
            }
            const long long off = S.soff[c] + (long long)j * p.s_axis;
            cpx_t v;
            if (p.rmode == RM_R2C_FULL)
                v = cpx_t{src[off], (real_t)0};
            else
                v = b2_load_pt(src, off, p.s_pair, p.vec_ld);
            if (p.swap_in) v = cswap(v);
            S.a[(size_t)c * ls + j] = v;
        }
    }
    B2_SYNC();

    // ---- c2r pre-processing: half-spectrum -> packed complex spectrum (swapped) ----
    if (p.rmode == RM_C2R_HALF) {
        B2_FOR_THREADS {
            const int m = p.m, hk = m / 2 + 1;
            const int total = nlines * hk;
            for (int idx = tid; idx < total; idx += B2_NTHREADS) {
                const int c = (int)b2_fastdiv((unsigned)idx, p.gd_m[3], p.gd_s[3]), k = idx - c * hk;
                cpx_t *X = S.a + (size_t)c * ls;
                const cpx_t A = X[k], B = cconj(X[m - k]);
                const cpx_t sum = cadd(A, B), dif = csub(A, B);
                const cpx_t D = cmul_pi(cmul(cconj(p.twr[k]), dif));
                const cpx_t zk = cadd(sum, D), zmk = cconj(csub(sum, D));
                if (k != 0 && k != m - k) X[m - k] = cswap(zmk);
                X[k] = cswap(zk);
            }
        }
        B2_SYNC();
    } else if (p.rmode == RM_C2R_FULL) {
        B2_FOR_THREADS {
            const int n = p.n, h = n / 2 + 1;
            const int total = nlines * n;
            // Hermitian extension into buffer b, swapped
            for (int idx = tid; idx < total; idx += B2_NTHREADS) {
                const int c = idx / n, k = idx - c * n;
                const cpx_t *X = S.a + (size_t)c * ls;
                cpx_t v = (k < h) ? X[k] : cconj(X[n - k]);
                if (k == 0 || 2 * k == n) v.y = 0;
                S.b[(size_t)c * ls + k] = cswap(v);
            }
        }
        B2_SYNC();
        cpx_t *t = S.a;
        S.a = S.b;
        S.b = t;
    }

    // ---- radix passes (Stockham, ping-pong a -> b) ----
    {
        int Ns = 1;
        for (int ps = 0; ps < p.npass; ++ps) {
            const int R = p.radix[ps];
            B2_FOR_THREADS {
                switch (R) {
                case 2: b2_radix_step<2>(S.a, S.b, p.tw, p.m, Ns, nlines, ls, tid, B2_NTHREADS, p.gnb_m[ps], p.gnb_s[ps], p.gns_m[ps], p.gns_s[ps]); break;
                case 3: b2_radix_step<3>(S.a, S.b, p.tw, p.m, Ns, nlines, ls, tid, B2_NTHREADS, p.gnb_m[ps], p.gnb_s[ps], p.gns_m[ps], p.gns_s[ps]); break;
                case 4: b2_radix_step<4>(S.a, S.b, p.tw, p.m, Ns, nlines, ls, tid, B2_NTHREADS, p.gnb_m[ps], p.gnb_s[ps], p.gns_m[ps], p.gns_s[ps]); break;
                case 5: b2_radix_step<5>(S.a, S.b, p.tw, p.m, Ns, nlines, ls, tid, B2_NTHREADS, p.gnb_m[ps], p.gnb_s[ps], p.gns_m[ps], p.gns_s[ps]); break;
                case 7: b2_radix_step<7>(S.a, S.b, p.tw, p.m, Ns, nlines, ls, tid, B2_NTHREADS, p.gnb_m[ps], p.gnb_s[ps], p.gns_m[ps], p.gns_s[ps]); break;
                case 8: b2_radix_step<8>(S.a, S.b, p.tw, p.m, Ns, nlines, ls, tid, B2_NTHREADS, p.gnb_m[ps], p.gnb_s[ps], p.gns_m[ps], p.gns_s[ps]); break;
                case 11: b2_radix_step<11>(S.a, S.b, p.tw, p.m, Ns, nlines, ls, tid, B2_NTHREADS, p.gnb_m[ps], p.gnb_s[ps], p.gns_m[ps], p.gns_s[ps]); break;
                case 13: b2_radix_step<13>(S.a, S.b, p.tw, p.m, Ns, nlines, ls, tid, B2_NTHREADS, p.gnb_m[ps], p.gnb_s[ps], p.gns_m[ps], p.gns_s[ps]); break;
                case 16: b2_radix_step<16>(S.a, S.b, p.tw, p.m, Ns, nlines, ls, tid, B2_NTHREADS, p.gnb_m[ps], p.gnb_s[ps], p.gns_m[ps], p.gns_s[ps]); break;
                default: break;
                }
            }
            B2_SYNC();
            cpx_t *t = S.a;
            S.a = S.b;
            S.b = t;
            Ns *= R;
        }
    }

    // ---- r2c post-processing: packed complex spectrum -> Hermitian half-spectrum ----
    if (p.rmode == RM_R2C_HALF) {
        B2_FOR_THREADS {
            const int m = p.m, hk = m / 2 + 1;
            const int total = nlines * hk;
            for (int idx = tid; idx < total; idx += B2_NTHREADS) {
                const int c = (int)b2_fastdiv((unsigned)idx, p.gd_m[3], p.gd_s[3]), k = idx - c * hk;
                cpx_t *Z = S.a + (size_t)c * ls;
                if (k == 0) {
                    const cpx_t z0 = Z[0];
                    Z[0] = cpx_t{z0.x + z0.y, (real_t)0};
                    Z[m] = cpx_t{z0.x - z0.y, (real_t)0};
                } else {
                    const cpx_t zk = Z[k], zc = cconj(Z[m - k]);
                    const cpx_t E = cscale(cadd(zk, zc), (real_t)0.5);
                    const cpx_t O = cmul_mi(cscale(csub(zk, zc), (real_t)0.5));
                    const cpx_t wO = cmul(p.twr[k], O);
                    Z[k] = cadd(E, wO);
                    if (k != m - k) Z[m - k] = cconj(csub(E, wO));
                }
            }
        }
        B2_SYNC();
    }

    // ---- store ----
    B2_FOR_THREADS {
        const int total = nlines * p.nstore;
        for (int idx = tid; idx < total; idx += B2_NTHREADS) {
            int c, k;
            if (p.st_col) {
                k = (nlines == p.C) ? (int)b2_fastdiv((unsigned)idx, p.gd_m[2], p.gd_s[2]) : idx / nlines;
                c = idx - k * nlines;
            } else {
                c = (int)b2_fastdiv((unsigned)idx, p.gd_m[1], p.gd_s[1]);
                k = idx - c * p.nstore;
            }
            cpx_t v = S.a[(size_t)c * ls + k];
            if (p.epi_dim >= 0) {
                const unsigned long long e = (unsigned long long)S.i2[c] * (unsigned long long)k;
                const cpx_t w = cmul(p.epi_hi[e >> p.epi_shift],
                                     p.epi_lo[e & ((1ull << p.epi_shift) - 1)]);
                v = cmul(v, w);
            }
            if (p.swap_out) v = cswap(v);
            const long long off = S.doff[c] + (long long)k * p.d_axis;
            if (p.rmode == RM_C2R_FULL)
                dst[off] = v.x;
            else
                b2_store_pt(dst, off, p.d_pair, p.vec_st, v);
        }
    }
}

#if B2_CUDA
#if defined(FFTW_SINGLE)
#define B2_GENERIC_BOUNDS __maxnreg__(56)  // 384-thread CTAs (lengths 3*2^k) fit three to an SM, 512-thread CTAs two
#else
#define B2_GENERIC_BOUNDS __launch_bounds__(512, 1)  // a radix-16 butterfly in double is 64 registers of data alone
#endif
__global__ void B2_GENERIC_BOUNDS
    b2_generic_pass_kernel(const __grid_constant__ PassParams p,
                                       const real_t *B2_RESTRICT src, real_t *B2_RESTRICT dst,
                                       long long bid0) {
    extern __shared__ __align__(16) char b2_smem[];
    b2_generic_pass_body(p, src, dst, bid0 + blockIdx.x, b2_smem, blockDim.x);
}
#endif
