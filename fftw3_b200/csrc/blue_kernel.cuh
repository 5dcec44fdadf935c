// csrc/blue_kernel.cuh -- the element-wise steps of Bluestein's algorithm, which carries any
// length n (in particular primes > 13 and lengths with such a factor) on the power-of-two
// kernels:  X[k] = c[k] * sum_j (x[j] c[j]) b[k-j],  c[j] = exp(-i pi j^2 / n),  b = conj(c),
// the sum being a cyclic convolution of length M >= 2n-1 done with two length-M FFTs.
//
//   mode 0  chirp-in : work[l][j] = x[l][j] * c[j] for j < n, 0 for n <= j < M  (gathers the
//                      strided / real / Hermitian-half input into dense rows of M)
//   mode 1  multiply : work[l][j] *= bhat[j],  bhat = FFT_M(b wrapped) / M
//   mode 2  chirp-out: X[l][k] = work[l][k] * c[k] for k < nout (scatters to the output layout)
//
// The reference has the same algorithm as its `dft-bluestein` solver (dft_bluestein_apply,
// fftw/fftw_api.c:1643-1688; dft_bluestein_sequence :1598-1611 builds the chirp with j^2 reduced
// mod 2n, as BlueTables::chirp does; convolution size = next size with small factors,
// dft_bluestein_choose_transform_size :1739-1743).  The reference applies it to prime sizes and
// sends composite sizes with a large factor through Rader / generic sub-plans; here every axis
// whose length has a prime factor > 13 takes this path, with M the next power of two.
#pragma once
#include "selector.h"
#include "tile_kernel.cuh"

#define B2_BLUE_PTS 1024  // points of one line handled by one CTA

// grid: chunks of B2_BLUE_PTS points along the line (fastest), then lines.  Work rows are dense:
// line l (row-major over the line dims, dim 0 fastest) starts at complex offset l * M.
B2_D void b2_blue_body(const b2::BlueParams &p, const real_t *src, real_t *dst, long long bid, int b2_nthreads) {
    const int len = (p.mode == 2) ? p.nout : p.M;
    const int chunks = (len + B2_BLUE_PTS - 1) / B2_BLUE_PTS;
    const int chunk = (int)(bid % chunks);
    const long long line = bid / chunks;
    long long rest = line, so = 0, dof = 0;
    for (int d = 0; d < p.nld; ++d) {
        const long long c = rest % p.ln[d];
        rest /= p.ln[d];
        so += c * p.ls[d];
        dof += c * p.ld[d];
    }
    B2_FOR_THREADS {
        const int jend = (chunk + 1) * B2_BLUE_PTS < len ? (chunk + 1) * B2_BLUE_PTS : len;
        for (int j = chunk * B2_BLUE_PTS + tid; j < jend; j += B2_NTHREADS) {
            if (p.mode == 0) {
                cpx_t v = cpx_t{0, 0};
                if (j < p.n) {
                    if (p.src_real) {
                        v.x = src[so + (long long)j * p.s_axis];
                    } else {
                        const bool mirror = p.herm_in && j >= p.nin;  // x[j] = conj(x[n-j]) beyond the stored half
                        const long long jj = mirror ? p.n - j : j;
                        v.x = src[so + jj * p.s_axis];
                        v.y = src[so + jj * p.s_axis + 1];
                        if (mirror) v.y = -v.y;
                        if (p.herm_in && (j == 0 || 2 * j == p.n)) v.y = 0;  // the reference ignores these imaginary parts too
                    }
                    if (p.swap_in) v = cswap(v);
                    v = cmul(v, p.chirp[j]);
                }
                reinterpret_cast<cpx_t *>(dst)[line * p.M + j] = v;
            } else if (p.mode == 1) {
                const cpx_t *S = reinterpret_cast<const cpx_t *>(src);
                reinterpret_cast<cpx_t *>(dst)[line * p.M + j] = cmul(S[line * p.M + j], p.bhat[j]);
            } else {
                cpx_t v = cmul(reinterpret_cast<const cpx_t *>(src)[line * p.M + j], p.chirp[j]);
                if (p.swap_out) v = cswap(v);
                if (p.dst_real) {
                    dst[dof + (long long)j * p.d_axis] = v.x;
                } else {
                    dst[dof + (long long)j * p.d_axis] = v.x;
                    dst[dof + (long long)j * p.d_axis + 1] = v.y;
                }
            }
        }
    }
}

#if B2_CUDA
__global__ void b2_blue_kernel(const __grid_constant__ b2::BlueParams p, const real_t *src, real_t *dst,
                               long long bid0) {
    b2_blue_body(p, src, dst, bid0 + blockIdx.x, blockDim.x);
}
#endif
