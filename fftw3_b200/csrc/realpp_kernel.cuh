// csrc/realpp_kernel.cuh -- element-wise Hermitian split (r2c) / merge (c2r) used when the
// half-length complex transform of a real line is too long for one CTA and goes through
// the four-step path.  Same arithmetic as the post/pre-processing inside the generic kernel;
// the reference does this step with its hc2c twiddle codelets (hc2cfdftv_* / hc2cbdftv_*,
// fftw/rdft_simd/common/hc2cfdftv_16.c:235; driver ct_hc2c_direct_apply,
// fftw/fftw_api.c:5831-5845).
#pragma once
#include "selector.h"
#include "tile_kernel.cuh"

// grid: x = ceil((m/2+1)/threads) chunks, then lines.  src/dst may alias (in-place split).
B2_D void b2_realpp_body(const b2::RealPPParams &p, int post, const real_t *src, real_t *dst,
                         long long bid, int b2_nthreads) {
    const int hk = p.m / 2 + 1;
    const int chunks = (hk + p.pts_per_block - 1) / p.pts_per_block;
    const int chunk = (int)(bid % chunks);
    long long rest = bid / chunks;
    long long so = 0, dof = 0;
    for (int d = 0; d < p.nld; ++d) {
        const long long c = rest % p.ln[d];
        rest /= p.ln[d];
        so += c * p.ls[d];
        dof += c * p.ld[d];
    }
    B2_FOR_THREADS {
        const int m = p.m;
        const int kend = (chunk + 1) * p.pts_per_block < hk ? (chunk + 1) * p.pts_per_block : hk;
        const cpx_t *S = reinterpret_cast<const cpx_t *>(src + so);
        cpx_t *D = reinterpret_cast<cpx_t *>(dst + dof);
        // strides are in reals and even for complex arrays
        const long long ss = p.s_axis / 2, ds = p.d_axis / 2;
        constexpr int U = 4;  // pairs in flight per thread: all loads are issued before the first use
        for (int k0 = chunk * p.pts_per_block + tid; k0 < kend; k0 += U * B2_NTHREADS) {
            cpx_t zk[U], zc[U], w[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int k = k0 + u * B2_NTHREADS;
                if (k < kend) {
                    zk[u] = b2_ldcg(S + (long long)k * ss);
                    zc[u] = b2_ldcg(S + (long long)(m - k) * ss);
                    w[u] = p.twr[k];
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int k = k0 + u * B2_NTHREADS;
                if (k >= kend) continue;
                if (post) {
                    if (k == 0) {
                        const cpx_t z0 = zk[u];
                        D[0] = cpx_t{z0.x + z0.y, (real_t)0};
                        D[(long long)m * ds] = cpx_t{z0.x - z0.y, (real_t)0};
                    } else {
                        const cpx_t a = zk[u], b = cconj(zc[u]);
                        const cpx_t E = cscale(cadd(a, b), (real_t)0.5);
                        const cpx_t O = cmul_mi(cscale(csub(a, b), (real_t)0.5));
                        const cpx_t wO = cmul(w[u], O);
                        D[(long long)k * ds] = cadd(E, wO);
                        if (k != m - k) D[(long long)(m - k) * ds] = cconj(csub(E, wO));
                    }
                } else {
                    const cpx_t A = zk[u], B = cconj(zc[u]);
                    const cpx_t sum = cadd(A, B), dif = csub(A, B);
                    const cpx_t Dd = cmul_pi(cmul(cconj(w[u]), dif));
                    const cpx_t z1 = cadd(sum, Dd), z2 = cconj(csub(sum, Dd));
                    D[(long long)k * ds] = cswap(z1);
                    if (k != 0 && k != m - k) D[(long long)(m - k) * ds] = cswap(z2);
                }
            }
        }
    }
}

#if B2_CUDA
__global__ void b2_realpp_kernel(const __grid_constant__ b2::RealPPParams p, int post,
                                 const real_t *src, real_t *dst, long long bid0) {
    b2_realpp_body(p, post, src, dst, bid0 + blockIdx.x, blockDim.x);
}
#endif
