// csrc/plan_build.h -- geometry -> list of kernel steps.  This is where the reference's
// solver tree (rank>=2 splitting fftw/fftw_api.c:4508-4558, vrank>=1 :4712-4755, Cooley-Tukey
// ct_mkplan :2158-2270, rdft2-ct ct_hc2c_mkplan :5661-5788, rank>=2 rdft2 :10219-10282) is
// re-expressed as a flat list: one step per axis (two for a four-step axis), each step one
// kernel launch that sweeps the data once.
#pragma once
#include <math.h>

#include "b200fft.h"
#include "blue_tables.h"
#include "selector.h"

struct b2fft_comm_s;

inline bool b2_have_tile_kernel(int n, int mode, int epi);  // tile_dispatch.cuh
inline bool b2_have_any_tile_kernel(int n, int mode, int epi);  // the same, mixed-radix kernels included

namespace b2 {

// set while a plan is rebuilt without the steps that exist only as tile kernels (two-pass real
// transform, fused Bluestein passes) -- engine.cu: unaligned arrays, or a tile kernel that the
// selector turned down after all
inline bool &no_two_pass_real() {
    static thread_local bool v = false;
    return v;
}

struct StepPlan : Step {
    int src_fixed = -1;            // -1: wherever the data currently is
    std::vector<int> dst_allowed;  // candidate destination buffers
};

struct PlanCore {
    b2fft_geom g;
    std::vector<StepPlan> steps;
    bool in_place = false;
    bool empty = false;         // howmany == 0: a legal no-op
    long long in_reals = 0;     // extent of the user's arrays, in reals
    long long out_reals = 0;
    long long in_dense = 0, out_dense = 0;  // reals actually touched
    long long scr_reals = 0;
    long long wrk_reals = 0;    // second plan-owned work array (Bluestein rows)
    // Negative strides (the reference's tensors carry signed strides, fftw/fftw_api.h:472-487): the caller's pointer is
    // element 0, which is then not the lowest address.  in_off / out_off = distance (reals) from the lowest address
    // touched to the pointer; in_reals / out_reals stay the extent of the whole span.  Such geometries run on the
    // generic kernel (signed 64-bit offsets throughout); work arrays in the output layout get the same offset.
    bool neg = false;
    long long in_off = 0, out_off = 0;
    double flops = 0;
    std::string desc;
};

struct GDim {
    long long n, is, os;
};

// four-step split n = n1 * n2 (n1: strided/column pass with twiddle epilogue, n2: row pass)
inline bool split_four_step(long long n, long long *n1o, long long *n2o) {
    const Caps c = caps();
    long long best = -1;
    double bestd = 1e300;
    for (long long a = 2; a * a <= n * (long long)c.col; ++a) {
        if (a > c.col) break;
        if (n % a) continue;
        long long b = n / a;
        if (b > c.col || !factorable(a) || !factorable(b)) continue;
        double d = fabs(log((double)a) - log((double)b));
        // a factor pair that runs on tile kernels beats a better balanced one that does not (for a power of two
        // every pair in range has them, so nothing changes there): larger factor = column pass + twiddle, smaller
        // = row -> column pass
        const long long hi = a > b ? a : b, lo = a > b ? b : a;
        d += 100.0 * ((b2_have_any_tile_kernel((int)hi, 1, 1) ? 0 : 1) + (b2_have_any_tile_kernel((int)lo, 2, 0) ? 0 : 1));
        if (d < bestd) {
            bestd = d;
            best = a;
        }
    }
    if (best < 0) return false;
    long long a = best, b = n / best;
    if (a < b) std::swap(a, b);  // larger factor to the column pass, shorter rows for pass B tiles
    if (const char *e = getenv("FFTW_B200_SPLIT_N1")) {  // experiments: force the column-pass length
        const long long f = atoll(e);
        if (f > 1 && n % f == 0 && f <= c.col && n / f <= c.col && factorable(f) && factorable(n / f)) {
            a = f;
            b = n / f;
        }
    }
    *n1o = a;
    *n2o = b;
    return true;
}

// six-step split n = n1 * n2 * n3 for lengths beyond the square of the column limit (2^22 fp32,
// 2^20 fp64): all three factors within the column limit, as balanced as possible, n1 >= n2 >= n3
inline bool split_three(long long n, long long *n1o, long long *n2o, long long *n3o) {
    const Caps c = caps();
    double bestd = 1e300;
    long long ba = -1, bb = -1;
    for (long long a = 2; a <= c.col; ++a) {
        if (n % a || !factorable(a)) continue;
        const long long r = n / a;
        for (long long b = 2; b <= a; ++b) {
            if (r % b) continue;
            const long long d = r / b;
            if (d > b || d < 2 || !factorable(b) || !factorable(d)) continue;
            const double spread = log((double)a) - log((double)d);
            if (spread < bestd) {
                bestd = spread;
                ba = a;
                bb = b;
            }
        }
    }
    if (ba < 0) return false;
    *n1o = ba;
    *n2o = bb;
    *n3o = n / (ba * bb);
    return true;
}

inline long long dst_extent(const StepPlan &s) {
    if (s.kind == SK_BLUE) {
        const BlueParams &b = s.bp;
        long long lines = 1, e = 0;
        for (int d = 0; d < b.nld; ++d) {
            lines *= b.ln[d];
            e += (b.ln[d] - 1) * b.ld[d];
        }
        if (b.mode != 2) return lines * b.M * 2;
        return e + (long long)(b.nout - 1) * b.d_axis + (b.dst_real ? 1 : 2);
    }
    if (s.kind != SK_PASS) {
        long long e = (long long)s.rp.m * s.rp.d_axis + 2;
        for (int d = 0; d < s.rp.nld; ++d) e += (s.rp.ln[d] - 1) * s.rp.ld[d];
        return e;
    }
    const PassParams &p = s.pp;
    long long e = (long long)(p.nstore - 1) * p.d_axis + p.d_pair + 1;
    if (p.rsplit) e = (long long)(p.m / 2) * p.d_axis + 4;  // rows 0..m/2, two complex per column pair
    for (int d = 0; d < p.nld; ++d) e += (p.ln[d] - 1) * p.ld[d];
    for (const LDim &o : s.outer) e += (o.n - 1) * o.d;
    return e;
}

inline bool lines_same(const std::vector<LDim> &L) {
    for (const LDim &d : L)
        if (d.s != d.d) return false;
    return true;
}

inline bool add_c2c_axis(PlanCore &pc, long long n, long long s_axis, long long d_axis,
                         const std::vector<LDim> &lines, bool swap_in, bool swap_out,
                         const std::vector<int> &dst_allowed);

inline bool axis_direct_ok(long long n, bool row) {
    const Caps c = caps();
    if (n <= (row ? c.row : c.col)) return factorable(n);
    long long a, b, d;
    return split_four_step(n, &a, &b) || split_three(n, &a, &b, &d);
}

struct BlueSpec {
    long long n = 0;
    long long s_axis = 2, d_axis = 2;  // reals
    bool src_real = false, dst_real = false, herm_in = false;
    long long nout = 0;                // 0: n
    bool swap_in = false, swap_out = false;
    std::vector<LDim> lines;           // (src, dst) strides in reals
};

// One axis by Bluestein's algorithm: chirp-in to dense rows of M = 2^k >= 2n-1 in a plan-owned
// work array, FFT_M, multiply by bhat, inverse FFT_M, chirp-out to the destination layout.
inline bool add_bluestein_axis(PlanCore &pc, const BlueSpec &b, const std::vector<int> &dst_allowed) {
    const long long n = b.n;
    long long M = 1;
    while (M < 2 * n - 1) M *= 2;
    if (!axis_direct_ok(M, true)) {
        set_error("length %lld: the Bluestein convolution length %lld is beyond the engine's limit", n, M);
        return false;
    }
    BlueParams bp;
    memset(&bp, 0, sizeof bp);
    bp.n = (int)n;
    bp.M = (int)M;
    bp.nin = b.herm_in ? (int)(n / 2 + 1) : (int)n;
    bp.nout = (int)(b.nout ? b.nout : n);
    bp.src_real = b.src_real;
    bp.dst_real = b.dst_real;
    bp.herm_in = b.herm_in;
    bp.swap_in = b.swap_in;
    bp.swap_out = b.swap_out;
    bp.s_axis = b.s_axis;
    bp.d_axis = b.d_axis;
    long long L = 1;
    for (const LDim &l : b.lines) {
        if (l.n <= 1) continue;
        if (bp.nld >= B2_BLUE_MAXLD) {
            set_error("Bluestein axis with more than %d batch dimensions", B2_BLUE_MAXLD);
            return false;
        }
        bp.ln[bp.nld] = l.n;
        bp.ls[bp.nld] = l.s;
        bp.ld[bp.nld] = l.d;
        ++bp.nld;
        L *= l.n;
    }
    bp.chirp = blue_tables().chirp((unsigned long long)n);
    bp.bhat = blue_tables().bhat((unsigned long long)n, (unsigned long long)M);
    if (!bp.chirp || !bp.bhat) return false;
    auto mk = [&](int mode, const char *what) {
        StepPlan st;
        st.kind = SK_BLUE;
        st.bp = bp;
        st.bp.mode = mode;
        const long long len = mode == 2 ? bp.nout : M;
        st.nblocks = L * ((len + 1023) / 1024);
        st.nthreads = 256;
        st.smem = 0;
        char buf[160];
        snprintf(buf, sizeof buf, "bluestein %s n=%lld M=%lld lines=%lld ctas=%lld", what, n, M, L, st.nblocks);
        st.desc = buf;
        return st;
    };
    const std::vector<int> work = {B_WRK, B_SCR};
    // Contiguous complex lines and M within one CTA's reach: two fused row passes
    //   [x * chirp, zero-pad, FFT_M, * bhat] -> work,  [FFT_M^-1, * chirp, first n points] -> destination
    // instead of five sweeps over the padded rows.
    bool line_even = true;
    for (int d = 0; d < bp.nld; ++d)
        if ((bp.ls[d] | bp.ld[d]) & 1) line_even = false;
    if (!b.src_real && !b.dst_real && !b.herm_in && bp.nout == bp.n && b.s_axis == 2 && b.d_axis == 2 && line_even &&
        !no_two_pass_real() && !getenv("FFTW_B200_NO_FUSED_BLUESTEIN") && b2_have_tile_kernel((int)M, 0, 6) &&
        b2_have_tile_kernel((int)M, 0, 7)) {
        for (int pass = 0; pass < 2; ++pass) {
            AxisSpec a;
            a.n = (int)M;
            a.s_axis = a.d_axis = 2;
            long long wstride = 2 * M;
            for (int d = 0; d < bp.nld; ++d) {
                a.lines.push_back(LDim{bp.ln[d], pass == 0 ? bp.ls[d] : wstride, pass == 0 ? wstride : bp.ld[d], false});
                wstride *= bp.ln[d];
            }
            a.swap_in = a.swap_out = (pass == 1);
            StepPlan st;
            if (!make_pass_step(st, a)) return false;
            st.pp.blue_mode = pass == 0 ? 6 : 7;
            st.pp.blue_n = (int)n;
            st.pp.blue_swap = pass == 0 ? b.swap_in : b.swap_out;
            st.pp.blue_chirp = bp.chirp;
            st.pp.blue_bhat = bp.bhat;
            st.needs_tile_kernel = true;
            st.inplace_ok = false;
            st.dst_allowed = pass == 0 ? work : dst_allowed;
            pc.steps.push_back(st);
        }
        return true;
    }
    {
        StepPlan st = mk(0, "chirp-in");
        st.dst_allowed = work;
        pc.steps.push_back(st);
    }
    const std::vector<LDim> rows = {LDim{L, 2 * M, 2 * M, false}};
    if (!add_c2c_axis(pc, M, 2, 2, rows, false, false, work)) return false;
    {
        StepPlan st = mk(1, "multiply");
        st.inplace_ok = true;
        st.dst_allowed = work;
        pc.steps.push_back(st);
    }
    if (!add_c2c_axis(pc, M, 2, 2, rows, true, true, work)) return false;
    {
        StepPlan st = mk(2, "chirp-out");
        st.dst_allowed = dst_allowed;
        pc.steps.push_back(st);
    }
    return true;
}

// Append the step(s) for one complex axis.  `lines` carry (src, dst) strides in reals.
// first/last: whether the backward swap is applied on the way in / out.
inline bool add_c2c_axis(PlanCore &pc, long long n, long long s_axis, long long d_axis,
                         const std::vector<LDim> &lines, bool swap_in, bool swap_out,
                         const std::vector<int> &dst_allowed) {
    const Caps c = caps();
    const bool row = (s_axis <= 2 && d_axis <= 2);
    const long long cap = row ? c.row : c.col;
    const bool both = swap_in && swap_out;
    if (!axis_direct_ok(n, row)) {
        BlueSpec b;
        b.n = n;
        b.s_axis = s_axis;
        b.d_axis = d_axis;
        b.swap_in = swap_in;
        b.swap_out = swap_out;
        b.lines = lines;
        return add_bluestein_axis(pc, b, dst_allowed);
    }
    if (n <= cap) {
        AxisSpec a;
        a.n = (int)n;
        a.s_axis = s_axis;
        a.d_axis = d_axis;
        a.lines = lines;
        a.swap_in = swap_in;
        a.swap_out = swap_out;
        StepPlan st;
        if (!make_pass_step(st, a)) return false;
        st.inplace_ok = (s_axis == d_axis) && lines_same(lines);
        st.dst_allowed = dst_allowed;
        pc.steps.push_back(st);
        return true;
    }
    long long n1, n2;
    bool two_pass = split_four_step(n, &n1, &n2);
    if (two_pass && !opt("FFTW_B200_NO_PREFER3", B2_CFG_NO_PREFER3)) {
        // Measured on B200 (profiles/r1_long1d_probe.md): three passes through the 128..256-point kernels
        // (4 CTAs per SM, ~0.35 ms per GiB pass) beat two passes through the 2048-point (fp32) or
        // 1024-point (fp64) kernels, which run one CTA per SM: fp32 2^22 1.31 ms in two passes against
        // 2^23 / 2^24 at 1.17 / 1.14 ms in three; fp64 2^20 1.34 ms against 2^21 at 1.07 ms.
        const long long big = sizeof(real_t) == 4 ? 2048 : 1024;
        long long a3, b3, c3;
        if (std::max(n1, n2) >= big && split_three(n, &a3, &b3, &c3) && a3 < big) two_pass = false;
    }
    if (!two_pass) {
        // ---- three passes, n = n1 * n2 * n3 (j = j1*n2*n3 + j2*n3 + j3, k = k1 + n1*k2 + n1*n2*k3):
        //   1. n1-point FFTs over j1 for every (j2, j3), times w_n^(k1*(j2*n3 + j3))      [k1][j2][j3]
        //   2. inside every row k1: n2-point FFTs over j2, times w_(n2*n3)^(k2*j3)        [k1][k2][j3]
        //   3. n3-point FFTs over the contiguous j3, stored transposed at k1 + n1*k2 + n1*n2*k3
        // i.e. the four-step split applied to n and again to each of its rows; passes 1 and 2 are
        // column passes with a twiddle epilogue, pass 3 is the row -> column kernel with two line dims.
        // (The reference recurses the same way: ct_mkplan fftw/fftw_api.c:2158-2270 plans the child
        // of size m = n/r with another Cooley-Tukey step.)
        long long n3;
        if (!split_three(n, &n1, &n2, &n3)) {
            set_error("length %lld: no split into two or three factors <= %d", n, c.col);
            return false;
        }
        const long long np = n2 * n3;
        {
            AxisSpec a;
            a.n = (int)n1;
            a.s_axis = s_axis * np;
            a.d_axis = d_axis * np;
            a.lines = lines;
            a.lines.push_back(LDim{np, s_axis, d_axis, true});
            a.swap_in = swap_in;
            a.swap_out = both;  // swap . swap = identity between the passes of one axis: every pass of a
                                // backward transform swaps on both sides, so that a kernel can take the
                                // direction as ONE compile-time flag (coltma_kernel.cuh)
            a.epi_N = n;
            StepPlan st;
            if (!make_pass_step(st, a)) return false;
            st.inplace_ok = (s_axis == d_axis) && lines_same(lines);
            st.dst_allowed = dst_allowed;
            pc.steps.push_back(st);
        }
        {
            AxisSpec a;
            a.n = (int)n2;
            a.swap_in = a.swap_out = both;
            a.s_axis = a.d_axis = d_axis * n3;
            for (const LDim &l : lines) a.lines.push_back(LDim{l.n, l.d, l.d, false});
            a.lines.push_back(LDim{n1, d_axis * np, d_axis * np, false});
            a.lines.push_back(LDim{n3, d_axis, d_axis, true});
            a.epi_N = np;
            StepPlan st;
            if (!make_pass_step(st, a)) return false;
            st.inplace_ok = true;
            st.dst_allowed = dst_allowed;
            pc.steps.push_back(st);
        }
        {
            AxisSpec a;
            a.n = (int)n3;
            a.s_axis = d_axis;
            a.d_axis = d_axis * n1 * n2;
            for (const LDim &l : lines) a.lines.push_back(LDim{l.n, l.d, l.d, false});
            a.lines.push_back(LDim{n1, d_axis * np, d_axis, false});
            a.lines.push_back(LDim{n2, d_axis * n3, d_axis * n1, false});
            a.swap_in = both;
            a.swap_out = swap_out;
            StepPlan st;
            if (!make_pass_step(st, a)) return false;
            st.inplace_ok = false;
            st.dst_allowed = dst_allowed;
            pc.steps.push_back(st);
        }
        return true;
    }
    {  // pass A: n2 interleaved sub-sequences of length n1, twiddle w_n^(i2*k1) on the way out
        AxisSpec a;
        a.n = (int)n1;
        a.s_axis = s_axis * n2;
        a.d_axis = d_axis * n2;
        a.lines = lines;
        a.lines.push_back(LDim{n2, s_axis, d_axis, true});
        a.swap_in = swap_in;
        a.swap_out = both;
        a.epi_N = n;
        StepPlan st;
        if (!make_pass_step(st, a)) return false;
        st.inplace_ok = (s_axis == d_axis) && lines_same(lines);
        st.dst_allowed = dst_allowed;
        pc.steps.push_back(st);
    }
    {  // pass B: length-n2 rows for each k1, written transposed: X[k1 + n1*k2]
        AxisSpec a;
        a.n = (int)n2;
        a.s_axis = d_axis;
        a.d_axis = d_axis * n1;
        for (const LDim &l : lines) a.lines.push_back(LDim{l.n, l.d, l.d, false});
        a.lines.push_back(LDim{n1, d_axis * n2, d_axis, false});
        a.swap_in = both;
        a.swap_out = swap_out;
        StepPlan st;
        if (!make_pass_step(st, a)) return false;
        st.inplace_ok = false;
        st.dst_allowed = dst_allowed;
        pc.steps.push_back(st);
    }
    return true;
}

// Decide which buffer every step writes.  Exhaustive over <= 2^steps choices.
inline bool assign_buffers(PlanCore &pc) {
    const int ns = (int)pc.steps.size();
    if (ns == 0) return true;
    auto norm = [&](int b) { return (pc.in_place && b == B_IN) ? B_OUT : b; };
    std::vector<int> best, cur(ns);
    double bestc = 1e300;
    // iterative DFS
    std::vector<int> choice(ns, 0);
    std::vector<std::vector<int>> allowed(ns);
    for (int i = 0; i < ns; ++i) {
        for (int b : pc.steps[i].dst_allowed) {
            int nb = norm(b);
            if (std::find(allowed[i].begin(), allowed[i].end(), nb) == allowed[i].end())
                allowed[i].push_back(nb);
        }
        if (allowed[i].empty()) return false;
    }
    long long combos = 1;
    for (int i = 0; i < ns; ++i) combos *= (long long)allowed[i].size();
    for (long long code = 0; code < combos; ++code) {
        long long c = code;
        int state = norm(B_IN);
        double cost = 0;
        bool ok = true;
        for (int i = 0; i < ns && ok; ++i) {
            int dst = allowed[i][c % allowed[i].size()];
            c /= allowed[i].size();
            const StepPlan &s = pc.steps[i];
            if (s.src_fixed >= 0 && norm(s.src_fixed) != state) ok = false;
            if (dst == state && !s.inplace_ok) ok = false;
            if (dst == B_IN && !pc.in_place && state != B_IN &&
                std::find(s.dst_allowed.begin(), s.dst_allowed.end(), (int)B_IN) == s.dst_allowed.end())
                ok = false;
            if (dst != state && s.inplace_ok) cost += 1.0;
            if (dst == B_SCR) cost += 0.25;
            cur[i] = dst;
            state = dst;
        }
        if (!ok || state != B_OUT) continue;
        if (cost < bestc) {
            bestc = cost;
            best = cur;
        }
    }
    if (best.empty()) {
        set_error("no valid buffer assignment for this plan");
        return false;
    }
    int state = norm(B_IN);
    pc.scr_reals = 0;
    pc.wrk_reals = 0;
    for (int i = 0; i < ns; ++i) {
        pc.steps[i].src = state;
        pc.steps[i].dst = best[i];
        state = best[i];
        if (best[i] == B_SCR) pc.scr_reals = std::max(pc.scr_reals, dst_extent(pc.steps[i]));
        if (best[i] == B_WRK) pc.wrk_reals = std::max(pc.wrk_reals, dst_extent(pc.steps[i]));
    }
    return true;
}

// extent of the span the dims address, in units; *below = how far the span reaches below element 0 (negative strides)
inline long long extent_of(const std::vector<GDim> &dims, bool in, long long unit, long long *below = nullptr) {
    long long hi = 0, lo = 0;
    for (const GDim &d : dims) {
        if (d.n == 0) return 0;
        const long long reach = (d.n - 1) * (in ? d.is : d.os);
        (reach >= 0 ? hi : lo) += reach;
    }
    if (below) *below = -lo * unit;
    return (hi - lo + 1) * unit;
}

inline bool build_c2c(PlanCore &pc) {
    const b2fft_geom &g = pc.g;
    std::vector<GDim> T, V;
    double N = 1, H = 1;
    for (int i = 0; i < g.rank; ++i) {
        if (g.dims[i].n <= 0) {
            set_error("non-positive transform size");
            return false;
        }
        N *= (double)g.dims[i].n;
        T.push_back(GDim{g.dims[i].n, g.dims[i].is, g.dims[i].os});
    }
    for (int i = 0; i < g.vrank; ++i) {
        if (g.vdims[i].n < 0) {
            set_error("negative howmany");
            return false;
        }
        H *= (double)g.vdims[i].n;
        V.push_back(GDim{g.vdims[i].n, g.vdims[i].is, g.vdims[i].os});
    }
    std::vector<GDim> all = T;
    all.insert(all.end(), V.begin(), V.end());
    pc.in_reals = extent_of(all, true, 2, &pc.in_off);
    pc.out_reals = extent_of(all, false, 2, &pc.out_off);
    pc.neg = pc.in_off > 0 || pc.out_off > 0;
    pc.in_dense = pc.out_dense = (long long)(N * H * 2);
    pc.flops = 5.0 * N * (N > 1 ? log2(N) : 0) * H;
    if (H == 0) {
        pc.empty = true;
        return true;
    }
    if (pc.neg && pc.in_place)
        for (const GDim &d : all)
            if (d.n > 1 && d.is != d.os) {  // (the location-set comparison below assumes positive strides)
                set_error("in-place transform with negative strides needs identical input and output strides");
                return false;
            }
    // In place with different input and output strides (e.g. a batched transform written back transposed):
    // solvable when both stride sets address the same locations (reference: fftw_tensor_inplace_locations
    // fftw_api.c:17298, checked in fftw_mkproblem_dft :4090-4094).  The transform then runs through the
    // plan's work arrays in the OUTPUT layout and one last sweep copies it over the input.
    bool relayout = false;
    if (pc.in_place) {
        std::vector<std::pair<long long, long long>> li, lo;
        for (const GDim &d : all)
            if (d.n > 1) {
                if (d.is != d.os) relayout = true;
                li.push_back({d.is, d.n});
                lo.push_back({d.os, d.n});
            }
        // same location sets <=> equal after sorting by stride and merging contiguous dims
        // (fftw_tensor_compress_contiguous on both sides, then fftw_tensor_equal)
        auto compress = [](std::vector<std::pair<long long, long long>> &v) {
            std::sort(v.begin(), v.end());
            for (size_t i = 0; i + 1 < v.size();) {
                if (v[i + 1].first == v[i].first * v[i].second) {
                    v[i].second *= v[i + 1].second;
                    v.erase(v.begin() + i + 1);
                } else
                    ++i;
            }
        };
        compress(li);
        compress(lo);
        if (relayout && li != lo) {
            set_error("in-place transform whose input and output strides address different locations");
            return false;
        }
    }
    const std::vector<int> fft_dst = relayout ? std::vector<int>{B_SCR, B_WRK} : std::vector<int>{B_OUT, B_SCR};
    // axes: smallest output stride first
    std::vector<int> order;
    for (int i = 0; i < (int)T.size(); ++i)
        if (T[i].n > 1) order.push_back(i);
    std::sort(order.begin(), order.end(), [&](int a, int b) { return T[a].os < T[b].os; });
    const bool bwd = g.sign > 0;
    if (order.empty()) {
        // all transform dims have length 1: a (strided) copy, done by a length-1 "pass"
        AxisSpec a;
        a.n = 1;
        for (const GDim &d : all) a.lines.push_back(LDim{d.n, 2 * d.is, 2 * d.os, false});
        StepPlan st;
        if (!make_pass_step(st, a)) return false;
        st.inplace_ok = lines_same(a.lines);
        st.dst_allowed = fft_dst;
        pc.steps.push_back(st);
    }
    std::vector<size_t> axis_first_step;
    for (size_t oi = 0; oi < order.size(); ++oi) {
        const int ax = order[oi];
        const bool first = (oi == 0), last = (oi + 1 == order.size());
        axis_first_step.push_back(pc.steps.size());
        std::vector<LDim> lines;
        for (int i = 0; i < (int)T.size(); ++i)
            if (i != ax) lines.push_back(LDim{T[i].n, 2 * (first ? T[i].is : T[i].os), 2 * T[i].os, false});
        for (const GDim &d : V) lines.push_back(LDim{d.n, 2 * (first ? d.is : d.os), 2 * d.os, false});
        // backward = swap re/im, forward transform, swap: applied per axis (two swaps between axes cancel)
        (void)last;
        if (!add_c2c_axis(pc, T[ax].n, 2 * (first ? T[ax].is : T[ax].os), 2 * T[ax].os, lines, bwd, bwd, fft_dst))
            return false;
    }
    // Grouped execution of the first two axes (the reference walks the same slabs one after the other: rank>=2
    // applies its child plan per leading index, rank_geq2_apply fftw/fftw_api.c:4435-4445): when every other
    // dimension is outside the plane of the two axes in the output layout, the two passes can run plane group by
    // plane group.  Single-kernel axes only.
    axis_first_step.push_back(pc.steps.size());
    if (order.size() >= 2 && !relayout && axis_first_step[1] - axis_first_step[0] == 1 && axis_first_step[2] - axis_first_step[1] == 1) {
        std::vector<GDim> ds;
        for (const GDim &d : all)
            if (d.n > 1) ds.push_back(d);
        std::sort(ds.begin(), ds.end(), [](const GDim &a, const GDim &b) { return a.os < b.os; });
        bool nested = ds.size() >= 3 && ds[0].os == T[order[0]].os && ds[0].n == T[order[0]].n && ds[1].os == T[order[1]].os &&
                      ds[1].n == T[order[1]].n;
        long long L = 1;
        for (size_t k = 0; nested && k + 1 < ds.size(); ++k)
            if (ds[k + 1].os < ds[k].os * ds[k].n) nested = false;
        for (size_t k = 2; nested && k < ds.size(); ++k) L *= ds[k].n;
        if (nested && L > 1) {
            pc.steps[axis_first_step[0]].grp_L = L;
            pc.steps[axis_first_step[1]].grp_L = L;
        }
    }
    if (relayout) {  // work array (output layout) -> the user's array
        AxisSpec a;
        a.n = 1;
        for (const GDim &d : all) a.lines.push_back(LDim{d.n, 2 * d.os, 2 * d.os, false});
        StepPlan st;
        if (!make_pass_step(st, a)) return false;
        st.inplace_ok = false;
        st.dst_allowed = {B_OUT};
        pc.steps.push_back(st);
    }
    if (!assign_buffers(pc)) return false;
    if (pc.neg)  // work arrays share the output layout: the same span, the same offset to element 0
        for (const StepPlan &st : pc.steps) {
            if (st.dst == B_SCR) pc.scr_reals = std::max(pc.scr_reals, pc.out_reals);
            if (st.dst == B_WRK) pc.wrk_reals = std::max(pc.wrk_reals, pc.out_reals);
        }
    return true;
}

}  // namespace b2
