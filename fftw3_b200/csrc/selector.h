// csrc/selector.h -- the kernel-configuration selector that replaces the reference's planner
// (solver registry + recursive search ranked by cost, fftw/fftw_api.c:14745-15710,
// ifftw_mkplan :15300-15426).  Given one axis of a canonical geometry it decides, without
// search: the radix factorisation, how many radix passes are fused per HBM round trip (all of
// them when the line fits on chip, otherwise a four-step split into two kernels), the batch
// tiling (lines per CTA, row or column tiles) and the launch shape.
#pragma once
#include <algorithm>
#include <string>
#include <vector>

#include "pass_params.h"
#include "twiddle.h"

namespace b2 {

enum Buf { B_IN = 0, B_OUT = 1, B_SCR = 2, B_WRK = 3, B_NBUF = 4 };  // SCR, WRK: plan-owned work arrays
enum StepKind { SK_PASS = 0, SK_RPOST = 1, SK_RPRE = 2, SK_BLUE = 3 };

// one line (batch) dimension of an axis pass; strides in reals
struct LDim {
    long long n, s, d;
    bool epi;
};

// element-wise Hermitian split / merge around a four-step real transform
struct RealPPParams {
    int m;                 // complex length (n/2)
    long long s_axis, d_axis;  // strides between consecutive complex points, in reals
    int nld;
    long long ln[B2_MAXLD], ls[B2_MAXLD], ld[B2_MAXLD];
    const cpx_t *twr;      // w_n^k, k <= m/2
    int pts_per_block;     // pairs (k, m-k) handled by one CTA
};

// element-wise steps of Bluestein's algorithm (blue_kernel.cuh)
#define B2_BLUE_MAXLD 8
struct BlueParams {
    int mode;          // 0 chirp-in, 1 multiply by bhat, 2 chirp-out
    int n, M;          // transform length, convolution length
    int nin;           // points stored in the source line: n, or n/2+1 for a Hermitian half
    int nout;          // points written per line by chirp-out: n, or n/2+1 (r2c)
    int src_real, dst_real, herm_in;
    int swap_in, swap_out;     // backward transform = swap re/im going in and coming out
    long long s_axis, d_axis;  // stride between consecutive points, in reals, of the user arrays
    int nld;
    long long ln[B2_BLUE_MAXLD], ls[B2_BLUE_MAXLD], ld[B2_BLUE_MAXLD];  // line dims, strides in reals
    const cpx_t *chirp;  // c[j] = exp(-i pi j^2 / n), j < n
    const cpx_t *bhat;   // FFT_M(conj(c) wrapped) / M
};

struct Step {
    int kind = SK_PASS;
    PassParams pp;
    RealPPParams rp;
    BlueParams bp;
    int src = B_IN, dst = B_OUT;
    bool inplace_ok = false;  // same locations read and written per CTA
    bool reads_input_layout = false;
    bool needs_tile_kernel = false;  // no generic-kernel form (two-pass real transform)
    long long nblocks = 0;
    int nthreads = 0;
    int gen_threads = 0;        // launch shape of the generic kernel for this step (fallback when
    PassParams gen_pp;          // the arrays given to execute are not vector-aligned)
    long long gen_nblocks = 0;
    size_t smem = 0;
    int fast_id = -1;           // >= 0: specialised tile kernel
    int chain_id = -1, chain_pos = 0;  // member of a chained launch (chain_build.h)
    // Grouped execution (engine.cu run_steps): the step's tiles, in launch order, fall into grp_L equal slabs, and
    // slab l of the step reads and writes only slab l of its arrays (planes of an N-D transform, transforms of a
    // batch).  Two consecutive steps with the same grp_L > 1 are run group by group so that the array between
    // them stays in the 126 MB L2.  grp_slab > 0: that array is plan scratch with this slab stride (reals): every
    // group is placed at the start of the scratch, which then never has to be written back to HBM.
    long long grp_L = 0, grp_slab = 0;
    std::vector<LDim> outer;    // extra line dims looped on the host
    std::string desc;
};

// ---- radix factorisation --------------------------------------------------------------
// m = prod radix[]; power-of-two part split into ceil(a/4) balanced radices <= 16, odd primes
// <= 13 as they are.  Returns the number of passes, or -1 when m has a larger prime factor.
inline int factorize(long long m, int *radix) {
    int np = 0;
    int a = 0;
    while (m % 2 == 0) {
        m /= 2;
        ++a;
    }
    const int odd[] = {3, 5, 7, 11, 13};
    for (int p : odd)
        while (m % p == 0) {
            if (np >= B2_MAXPASS) return -1;
            radix[np++] = p;
            m /= p;
        }
    if (m != 1) return -1;
    if (a > 0) {
        int k = (a + 3) / 4;
        for (int i = 0; i < k; ++i) {
            int bits = (a + (k - 1 - i)) / k;  // distribute evenly, larger first
            if (np >= B2_MAXPASS) return -1;
            radix[np++] = 1 << bits;
        }
    }
    return np;
}
inline bool factorable(long long m) {
    int r[B2_MAXPASS];
    return factorize(m, r) >= 0;
}

struct Caps {
    int row;  // largest m handled by one kernel when lines are contiguous
    int col;  // ... when the pass walks a strided axis (column tiles)
};
inline Caps caps() {
    Caps c;
    c.row = sizeof(real_t) == 4 ? 8192 : 4096;
    c.col = sizeof(real_t) == 4 ? 2048 : 1024;
    // experiments / tests: lower limits make the multi-pass splits reachable at small sizes
    if (const char *e = getenv("FFTW_B200_COL_CAP")) c.col = std::max(4, std::min(c.col, atoi(e)));
    if (const char *e = getenv("FFTW_B200_ROW_CAP")) c.row = std::max(4, std::min(c.row, atoi(e)));
    return c;
}

struct AxisSpec {
    int n = 1;
    int rmode = RM_C2C;
    long long s_axis = 2, s_pair = 1, d_axis = 2, d_pair = 1;
    bool swap_in = false, swap_out = false;
    bool src_vec = true, dst_vec = true;  // (re,im) adjacent and every offset even
    long long epi_N = 0;                  // > 0: multiply by w_N^(i2*k) on the way out
    std::vector<LDim> lines;
};

inline int ilog2_ceil(unsigned long long v) {
    int b = 0;
    while ((1ull << b) < v) ++b;
    return b;
}

// Builds one generic axis-pass step.  Returns false (with the error set) when impossible.
inline bool make_pass_step(Step &st, const AxisSpec &a) {
    PassParams &p = st.pp;
    memset(&p, 0, sizeof p);
    st.kind = SK_PASS;
    p.n = a.n;
    p.rmode = a.rmode;
    const bool half = (a.rmode == RM_R2C_HALF || a.rmode == RM_C2R_HALF);
    p.m = half ? a.n / 2 : a.n;
    p.npass = factorize(p.m, p.radix);
    if (p.npass < 0) {
        set_error("length %d has a prime factor > 13 (Bluestein path not built yet)", p.m);
        return false;
    }
    switch (a.rmode) {
    case RM_C2C: p.nload = p.nstore = p.n; p.lstride = p.n; break;
    case RM_R2C_HALF: p.nload = p.m; p.nstore = p.m + 1; p.lstride = p.m + 1; break;
    case RM_C2R_HALF: p.nload = p.m + 1; p.nstore = p.m; p.lstride = p.m + 1; break;
    case RM_R2C_FULL: p.nload = p.n; p.nstore = p.n / 2 + 1; p.lstride = p.n; break;
    case RM_C2R_FULL: p.nload = p.n / 2 + 1; p.nstore = p.n; p.lstride = p.n; break;
    }
    if ((p.lstride & 1) == 0) p.lstride += 1;  // odd line stride: column tiles hit distinct banks
    p.swap_in = a.swap_in;
    p.swap_out = a.swap_out;
    p.s_axis = a.s_axis;
    p.s_pair = a.s_pair;
    p.d_axis = a.d_axis;
    p.d_pair = a.d_pair;
    p.vec_ld = a.src_vec && a.s_pair == 1 && a.rmode != RM_R2C_FULL;
    p.vec_st = a.dst_vec && a.d_pair == 1 && a.rmode != RM_C2R_FULL;

    // ---- line dims: drop n == 1, pick the tile dim, merge contiguous neighbours ----
    std::vector<LDim> L;
    for (const LDim &d : a.lines)
        if (d.n > 1) L.push_back(d);
    const bool src_row = a.s_axis > 0 && a.s_axis <= 2, dst_row = a.d_axis > 0 && a.d_axis <= 2;
    int tile = -1;
    if (!dst_row) {
        for (size_t i = 0; i < L.size(); ++i)
            if (L[i].d > 0 && L[i].d <= 2 && (tile < 0 || L[i].d < L[tile].d)) tile = (int)i;
        if (tile >= 0) p.st_col = 1;
    }
    if (tile < 0 && !src_row) {
        for (size_t i = 0; i < L.size(); ++i)
            if (L[i].s > 0 && L[i].s <= 2 && (tile < 0 || L[i].s < L[tile].s)) tile = (int)i;
        if (tile >= 0) p.ld_col = 1;
    }
    if (tile >= 0 && p.st_col && !src_row && L[tile].s > 0 && L[tile].s <= 2) p.ld_col = 1;
    if (tile < 0)
        for (size_t i = 0; i < L.size(); ++i)
            if (tile < 0 || L[i].d < L[tile].d) tile = (int)i;
    std::vector<LDim> rest;
    LDim t0 = {1, 0, 0, false};
    if (tile >= 0) t0 = L[tile];
    for (size_t i = 0; i < L.size(); ++i)
        if ((int)i != tile) rest.push_back(L[i]);
    std::sort(rest.begin(), rest.end(), [](const LDim &x, const LDim &y) { return x.d < y.d; });
    // merge rest[i+1] into rest[i] when contiguous in both layouts
    for (size_t i = 0; i + 1 < rest.size();) {
        if (!rest[i].epi && !rest[i + 1].epi && rest[i + 1].s == rest[i].n * rest[i].s &&
            rest[i + 1].d == rest[i].n * rest[i].d) {
            rest[i].n *= rest[i + 1].n;
            rest.erase(rest.begin() + i + 1);
        } else
            ++i;
    }
    // and into the tile dim (never when it carries the four-step coordinate)
    if (!t0.epi && tile >= 0)
        for (size_t i = 0; i < rest.size();) {
            if (!rest[i].epi && rest[i].s == t0.n * t0.s && rest[i].d == t0.n * t0.d) {
                t0.n *= rest[i].n;
                rest.erase(rest.begin() + i);
                i = 0;
            } else
                ++i;
        }
    p.nld = 1;
    p.ln[0] = t0.n;
    p.ls[0] = t0.s;
    p.ld[0] = t0.d;
    p.epi_dim = -1;
    if (t0.epi) p.epi_dim = 0;
    st.outer.clear();
    for (const LDim &d : rest) {
        if (p.nld < B2_MAXLD) {
            if (d.epi) p.epi_dim = p.nld;
            p.ln[p.nld] = d.n;
            p.ls[p.nld] = d.s;
            p.ld[p.nld] = d.d;
            ++p.nld;
        } else {
            if (d.epi) {
                set_error("too many batch dimensions around a four-step axis");
                return false;
            }
            st.outer.push_back(d);
        }
    }

    // ---- batch tiling ----
    const size_t csz = sizeof(cpx_t);
    const size_t budget = (size_t)max_smem_optin() - 2048;
    long long Cwant;
    if (p.ld_col || p.st_col)
        Cwant = 128 / (long long)csz;  // 128-byte column segments
    else
        Cwant = std::max<long long>(1, (sizeof(real_t) == 4 ? 4096 : 2048) / p.lstride);
    Cwant = std::min<long long>(Cwant, p.ln[0]);
    long long Cmax = (long long)(budget / (2 * (size_t)p.lstride * csz + 32));
    if (Cmax < 1) {
        set_error("length %d does not fit in shared memory", p.m);
        return false;
    }
    long long Chalf = (long long)((budget / 2) / (2 * (size_t)p.lstride * csz + 32));
    long long C = std::min(Cwant, Cmax);
    if (Chalf >= 1 && Chalf * 2 >= C) C = std::min(C, Chalf);  // keep two CTAs per SM if cheap
    p.C = (int)std::max<long long>(1, C);
    p.tiles0 = (p.ln[0] + p.C - 1) / p.C;
    long long nb = p.tiles0;
    for (int d = 1; d < p.nld; ++d) nb *= p.ln[d];
    st.nblocks = nb;
    long long work = (long long)p.C * p.m / 8;
    int nt = (int)std::min<long long>(512, std::max<long long>(64, (work + 31) / 32 * 32));
    st.nthreads = nt;
    st.smem = (size_t)p.C * (2 * sizeof(long long) + 2 * sizeof(int)) + 2 * (size_t)p.C * p.lstride * csz;

    // ---- twiddles ----
    p.tw = twiddles().get((unsigned long long)p.m, (unsigned long long)p.m, 1);
    if (!p.tw) return false;
    if (half) {
        p.twr = twiddles().get((unsigned long long)p.n, (unsigned long long)p.m / 2 + 1, 1);
        if (!p.twr) return false;
    }
    if (a.epi_N > 0 && p.epi_dim >= 0) {
        const int bits = ilog2_ceil((unsigned long long)a.epi_N);
        p.epi_shift = (bits + 1) / 2;
        const unsigned long long lo = 1ull << p.epi_shift;
        const unsigned long long hi = ((unsigned long long)a.epi_N >> p.epi_shift) + 1;
        p.epi_lo = twiddles().get((unsigned long long)a.epi_N, lo, 1);
        p.epi_hi = twiddles().get((unsigned long long)a.epi_N, hi, lo);
        if (!p.epi_lo || !p.epi_hi) return false;
    } else
        p.epi_dim = -1;

    char buf[256];
    std::string rs;
    for (int i = 0; i < p.npass; ++i) rs += (i ? "x" : "") + std::to_string(p.radix[i]);
    static const char *rmn[] = {"c2c", "r2c", "c2r", "r2c-odd", "c2r-odd"};
    snprintf(buf, sizeof buf, "%s n=%d radix=%s tile=%dx%d(%s->%s) ctas=%lld threads=%d smem=%zuB%s",
             rmn[p.rmode], p.n, rs.c_str(), p.C, p.m, p.ld_col ? "col" : "row", p.st_col ? "col" : "row",
             st.nblocks, st.nthreads, st.smem, p.epi_dim >= 0 ? " +twiddle" : "");
    st.desc = buf;
    b2_pp_prepare_generic(p);
    return true;
}

}  // namespace b2
