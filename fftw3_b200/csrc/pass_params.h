// csrc/pass_params.h -- the descriptor of one "axis pass": batched length-n DFTs along one
// axis of a strided array, a tile of C lines per CTA.  It plays the role of the reference's
// problem_dft {sz, vecsz} after canonicalisation (fftw/fftw_api.h:1871-1875, tensor/iodim
// fftw_api.h:472-487) for ONE solver step (ct / direct / rank>=2 / vrank>=1 glue,
// fftw/fftw_api.c:2079-2102, 3160-3187, 4435-4445, 4627-4638 all collapse into it).
#pragma once
#include "b2_common.h"

#define B2_MAXPASS 14
#define B2_MAXLD 4
#define B2_MAXSCAT 16

enum { RM_C2C = 0, RM_R2C_HALF = 1, RM_C2R_HALF = 2, RM_R2C_FULL = 3, RM_C2R_FULL = 4 };

struct PassParams {
    // ---- transform ----
    int n;       // logical length along the axis
    int m;       // complex FFT length computed on chip: n, or n/2 in the half-length real modes
    int npass;   // radix passes, m = prod radix[]
    int radix[B2_MAXPASS];
    int rmode;   // RM_*
    int C;       // lines per CTA tile
    int lstride; // smem line stride, complex elements
    int nload, nstore;  // points per line read from / written to global memory
    int swap_in, swap_out;  // backward transform = swap re/im going in and coming out
    // ---- addressing, all in units of REALS ----
    long long s_axis, s_pair;  // src: point-to-point stride along the axis; re->im distance
    long long d_axis, d_pair;  // dst: same
    int nld;                   // line dims; dim 0 is the tile dim
    long long ln[B2_MAXLD], ls[B2_MAXLD], ld[B2_MAXLD];
    long long tiles0;          // ceil(ln[0] / C)
    int ld_col, st_col;        // 1: consecutive threads walk the tile dim first (column tile)
    int vec_ld, vec_st;        // 1: (re,im) adjacent and aligned -> one vector access
    // ---- twiddles ----
    const cpx_t *tw;   // exp(-2 pi i j / m), j < m
    const cpx_t *twr;  // exp(-2 pi i k / n), k <= n/2 (half-length real modes)
    // ---- four-step epilogue: dst[k][i2] *= w_N^(i2*k), i2 = coordinate along line dim epi_dim
    int epi_dim;           // -1: none
    int epi_shift;         // exponent e = (hi << epi_shift) | lo
    const cpx_t *epi_lo;   // w_N^lo
    const cpx_t *epi_hi;   // w_N^(hi << shift)
    // ---- two-pass real transform (plan_build_real.h): column pass that splits a pair of real
    // columns (rsplit), row pass that folds the Hermitian half on the way out (fold_n > 0)
    int rsplit;
    // ---- fused Bluestein row passes (EPI 6: x*chirp in, *bhat out; EPI 7: *chirp out, first blue_n points)
    int blue_mode;         // 0 none, 6, 7
    int blue_n;            // transform length n (the line holds M = m points)
    int blue_swap;         // user-level backward transform: swap re/im of x on the way in / of X on the way out
    const cpx_t *blue_chirp, *blue_bhat;
    int pmerge;            // row pass whose line is built from two adjacent Hermitian half rows (two-pass c2r)
    long long fold_n;      // real length n = n1 * n2; outputs k > n/2 are stored as conj at n - k
    int fold_n1;
    // three-pass real transform (n = n1*n2*n3): the row pass's output index is
    //   k1 + fold_cmul * (coordinate along line dim fold_dim) + fold_pmul * (point index);
    // fold_dim == 0: none (two-pass form, index k1 + fold_n1 * point).  That dim's ld is 0.
    int fold_dim, fold_cmul;
    long long fold_pmul;
    // ---- tile-kernel decode helpers, filled by b2_pp_prepare() just before a launch ----
    unsigned dv_m[B2_MAXLD], dv_s[B2_MAXLD];  // exact division by tiles0 (slot 0) and ln[1..] for indices < 2^31
    int lgC;                                  // log2(C), or -1 when C is not a power of two
    unsigned pf_ahead;                        // L2 prefetch distance in tiles (0: off)
    // ---- generic-kernel index helpers (b2_pp_prepare_generic): multiply-shift divisors ----
    unsigned gd_m[4], gd_s[4];                // by nload, nstore, C, m/2+1
    unsigned gnb_m[B2_MAXPASS], gnb_s[B2_MAXPASS];  // by m / radix[pass]
    unsigned gns_m[B2_MAXPASS], gns_s[B2_MAXPASS];  // by Ns(pass) = prod of earlier radices
    long long nblocks;                        // tiles in the whole pass
    // ---- scattered store of a column pass (slab exchange fused into the FFT, slab.cu): output row k
    // of the plane with coordinate cc along line dim scat_dim belongs to owner o = k / scat_blk and is
    // stored at scat_base[o] + (scat_coord0 + cc) * scat_pstride[o] + (k - o*scat_blk) * d_axis + line
    // offset (reals).  The bases are the owners' receive buffers (peer memory over NVLink, or this
    // rank's own).  scat_n == 0: off.
    int scat_n, scat_blk, scat_dim;
    unsigned scat_m, scat_s;                  // exact division by scat_blk
    long long scat_coord0;
    // hybrid form (scat_self1 = own rank + 1; 0: every owner is scattered): only the rows this rank owns
    // go to scat_base[self]; the others stay in place at dst + cc * scat_lplane + k * d_axis for the DMA
    int scat_self1;
    long long scat_lplane;
    real_t *scat_base[B2_MAXSCAT];
    long long scat_pstride[B2_MAXSCAT];
    // distance (reals) between consecutive rows of owner o's block (d_axis unless the destination is laid out
    // differently: the round-2 slab schedule stores chunk-major, rows of w columns); scat_local: every base is
    // memory of THIS GPU (staging or its own receive buffer), so narrow row segments cost nothing
    long long scat_rstride[B2_MAXSCAT];
    int scat_local;
};

// Division by an invariant d for 0 <= n < 2^31 (Granlund-Montgomery): with l = ceil(log2 d) and
// m = floor(2^(31+l) / d) + 1 (fits 32 bits), floor(n / d) = (n * m) >> (31 + l).
inline void b2_fastdiv_make(unsigned long long d, unsigned &m, unsigned &s) {
    unsigned l = 0;
    while ((1ull << l) < d) ++l;
    m = (unsigned)((1ull << (31 + l)) / d + 1);  // d < 2^31, so 31 + l <= 62
    s = 31 + l;
}
B2_HD unsigned b2_fastdiv(unsigned n, unsigned m, unsigned s) {
    return (unsigned)(((unsigned long long)n * m) >> s);
}
inline void b2_pp_prepare_generic(PassParams &p) {
    const unsigned long long d[4] = {(unsigned long long)(p.nload > 0 ? p.nload : 1), (unsigned long long)(p.nstore > 0 ? p.nstore : 1),
                                     (unsigned long long)(p.C > 0 ? p.C : 1), (unsigned long long)(p.m / 2 + 1)};
    for (int i = 0; i < 4; ++i) b2_fastdiv_make(d[i], p.gd_m[i], p.gd_s[i]);
    unsigned long long Ns = 1;
    for (int ps = 0; ps < p.npass && ps < B2_MAXPASS; ++ps) {
        const unsigned long long R = (unsigned long long)(p.radix[ps] > 0 ? p.radix[ps] : 1);
        b2_fastdiv_make((unsigned long long)p.m / R > 0 ? (unsigned long long)p.m / R : 1, p.gnb_m[ps], p.gnb_s[ps]);
        b2_fastdiv_make(Ns, p.gns_m[ps], p.gns_s[ps]);
        Ns *= R;
    }
}
inline void b2_pp_prepare(PassParams &p) {
    b2_fastdiv_make((unsigned long long)(p.tiles0 > 0 ? p.tiles0 : 1), p.dv_m[0], p.dv_s[0]);
    for (int d = 1; d < B2_MAXLD; ++d)
        b2_fastdiv_make((unsigned long long)(d < p.nld && p.ln[d] > 0 ? p.ln[d] : 1), p.dv_m[d], p.dv_s[d]);
    p.lgC = -1;
    p.nblocks = (p.tiles0 > 0 ? p.tiles0 : 1);
    for (int d = 1; d < p.nld; ++d) p.nblocks *= p.ln[d];
    for (int l = 0; l < 16; ++l)
        if ((1 << l) == p.C) p.lgC = l;
}
