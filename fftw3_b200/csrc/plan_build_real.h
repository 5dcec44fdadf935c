// csrc/plan_build_real.h -- r2c / c2r geometries -> kernel steps.  The reference reaches
// these through problem_rdft2 (fftw/fftw_api.c:9298-9379) and the solvers rdft2-ct
// (ct_hc2c_mkplan :5661-5788: half-length complex FFT + hc2c twiddle pass), rdft2 direct
// leaves (:7139-7157) and rank>=2 rdft2 (:10111-10139, 10219-10282: real transform on the last
// dimension, complex transforms on the others).  Here: the last-dimension real transform is a
// length-n/2 complex FFT with the Hermitian split / merge fused into the same kernel.
#pragma once
#include "plan_build.h"

namespace b2 {

// n = n1 * n2 for the two-pass real transform: column pass of n1 over n2/2 column pairs, row pass of
// n2 over n1/2+1 rows.  Both need their tile kernel; prefer balanced factors, then the shorter column.
inline bool split_two_pass_real(long long n, long long *n1o, long long *n2o, bool r2c = true) {
    long long best = -1, bestmax = 1ll << 62;
    for (long long a = 4; a < n; a *= 2) {
        if (n % a) break;
        const long long b = n / a;
        if (b < 4 || (b & (b - 1))) continue;
        if (a > (1 << 20) || b > (1 << 20)) continue;
        // r2c: column pass n1 with pair split, row->col pass n2; c2r: column pass n2 with twiddle,
        // row->col pass n1 with pair merge
        if (r2c ? (!b2_have_tile_kernel((int)a, 1, 2) || !b2_have_tile_kernel((int)b, 2, 0))
                : (!b2_have_tile_kernel((int)b, 1, 1) || !b2_have_tile_kernel((int)a, 2, 5)))
            continue;
        const long long mx = std::max(a, b);
        if (mx < bestmax) {
            bestmax = mx;
            best = a;
        }
    }
    if (const char *e = getenv("FFTW_B200_SPLIT_N1")) {
        const long long f = atoll(e);
        if (f >= 4 && n % f == 0 &&
            (r2c ? (b2_have_tile_kernel((int)f, 1, 2) && b2_have_tile_kernel((int)(n / f), 2, 0))
                 : (b2_have_tile_kernel((int)(n / f), 1, 1) && b2_have_tile_kernel((int)f, 2, 5))))
            best = f;
    }
    if (best < 0) return false;
    *n1o = best;
    *n2o = n / best;
    return true;
}

// three-pass real split n = q1 * q2 * q3 through the short (<= 256-point) kernels: pair-split column
// pass q1, twiddled column pass q2 inside every row, row -> column pass q3 with the Hermitian fold
inline bool split_three_pass_real(long long n, long long *q1o, long long *q2o, long long *q3o) {
    if (n & (n - 1)) return false;
    long long cap = 256;
    if (const char *e = getenv("FFTW_B200_REAL3_CAP")) cap = std::max(32ll, atoll(e));
    for (long long a = cap; a >= 32; a /= 2) {
        if (n % a || !b2_have_tile_kernel((int)a, 1, 2)) continue;
        const long long Q = n / a;
        // q2 >= q3, as balanced as possible
        for (long long c = 32; c * c <= Q; c *= 2) {
            const long long b = Q / c;
            if (b * c != Q || b > cap || c > b) continue;
            if (!b2_have_tile_kernel((int)b, 1, 1) || !b2_have_tile_kernel((int)c, 2, 0)) continue;
            *q1o = a;
            *q2o = b;
            *q3o = c;  // keeps the last (largest c) found: the most balanced pair
        }
        if (*q1o == a) return true;
    }
    return false;
}

inline void fill_realpp(StepPlan &st, int post, int m, long long n_real, long long s_axis,
                        long long d_axis, const std::vector<LDim> &lines, bool use_src_for_both) {
    st.kind = post ? SK_RPOST : SK_RPRE;
    RealPPParams &r = st.rp;
    memset(&r, 0, sizeof r);
    r.m = m;
    r.s_axis = s_axis;
    r.d_axis = d_axis;
    r.nld = 0;
    st.outer.clear();
    for (const LDim &l : lines) {
        if (l.n <= 1) continue;
        LDim x = l;
        if (use_src_for_both) x.d = x.s;
        if (r.nld < B2_MAXLD) {
            r.ln[r.nld] = x.n;
            r.ls[r.nld] = x.s;
            r.ld[r.nld] = x.d;
            ++r.nld;
        } else
            st.outer.push_back(x);
    }
    r.twr = twiddles().get((unsigned long long)n_real, (unsigned long long)m / 2 + 1, 1);
    st.nthreads = 256;
    r.pts_per_block = 1024;
    long long nb = (m / 2 + 1 + r.pts_per_block - 1) / r.pts_per_block;
    for (int d = 0; d < r.nld; ++d) nb *= r.ln[d];
    st.nblocks = nb;
    st.smem = 0;
    char buf[128];
    snprintf(buf, sizeof buf, "%s n=%lld ctas=%lld", post ? "hermitian-split" : "hermitian-merge",
             n_real, nb);
    st.desc = buf;
}

// compact scratch layout for the lines of a big real transform: [lines...][m] complex
inline std::vector<long long> compact_strides(const std::vector<LDim> &lines, long long m) {
    std::vector<long long> cs(lines.size());
    long long s = 2 * m;
    for (size_t i = 0; i < lines.size(); ++i) {
        cs[i] = s;
        s *= lines[i].n;
    }
    return cs;
}

inline bool build_real(PlanCore &pc) {
    const b2fft_geom &g = pc.g;
    const bool r2c = (g.kind == B2FFT_R2C);
    if (g.rank < 1) {
        set_error("real transforms need rank >= 1");
        return false;
    }
    // is: strides of the INPUT array in its own elements, os: of the OUTPUT array.
    // r2c: input real, output complex; c2r: input complex, output real.
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
    const int L = g.rank - 1;
    const long long nl = T[L].n, hc = nl / 2 + 1;
    const long long ru = 1, cu = 2;                // reals per element of a real / complex array
    const long long iu = r2c ? ru : cu, ou = r2c ? cu : ru;
    {   // extents: the last dim holds nl reals on the real side and hc complex on the other.  Strides are signed (the
        // pointer is element 0): hi / lo = how far the span reaches above / below it, in elements
        long long hi_i = 0, lo_i = 0, hi_o = 0, lo_o = 0, di = 1, dn = 1;
        auto reach = [](long long n, long long st, long long &hi, long long &lo) {
            const long long r = (n - 1) * st;
            (r >= 0 ? hi : lo) += r;
        };
        for (int i = 0; i < g.rank; ++i) {
            long long ni = (i == L) ? (r2c ? nl : hc) : T[i].n;
            long long no = (i == L) ? (r2c ? hc : nl) : T[i].n;
            reach(ni, T[i].is, hi_i, lo_i);
            reach(no, T[i].os, hi_o, lo_o);
            di *= ni;
            dn *= no;
        }
        bool zero = false;
        for (const GDim &d : V) {
            if (d.n == 0) {
                zero = true;
                continue;
            }
            reach(d.n, d.is, hi_i, lo_i);
            reach(d.n, d.os, hi_o, lo_o);
            di *= d.n;
            dn *= d.n;
        }
        pc.in_reals = zero ? 0 : (hi_i - lo_i + 1) * iu;
        pc.out_reals = zero ? 0 : (hi_o - lo_o + 1) * ou;
        pc.in_off = zero ? 0 : -lo_i * iu;
        pc.out_off = zero ? 0 : -lo_o * ou;
        pc.neg = pc.in_off > 0 || pc.out_off > 0;
        pc.in_dense = (zero ? 0 : di) * iu;
        pc.out_dense = (zero ? 0 : dn) * ou;
    }
    pc.flops = 2.5 * N * (N > 1 ? log2(N) : 0) * H;
    if (H == 0) {
        pc.empty = true;
        return true;
    }
    // Negative strides (guru interfaces; the reference's tensors are signed, fftw/fftw_api.h:472-487): served where the real
    // transform is ONE real step plus complex passes over the leading dimensions, all on the generic kernel (signed offsets
    // throughout) and in place in the caller's arrays; the long multi-pass and Bluestein real paths keep positive strides.
    bool any_neg = false;
    for (const GDim &d : T)
        if (d.n > 1 && (d.is < 0 || d.os < 0)) any_neg = true;
    for (const GDim &d : V)
        if (d.n > 1 && (d.is < 0 || d.os < 0)) any_neg = true;
    if (pc.in_place) {
        // same bytes: stride in reals on the real side == 2 * stride in complex on the other
        bool ok = true;
        for (int i = 0; i < g.rank; ++i) {
            if (i == L || T[i].n == 1) continue;
            if (T[i].is * iu != T[i].os * ou) ok = false;
        }
        for (const GDim &d : V)
            if (d.n > 1 && d.is * iu != d.os * ou) ok = false;
        if (!ok) {
            set_error("in-place real transform: input and output rows do not coincide");
            return false;
        }
    }
    const Caps c = caps();
    const bool even = (nl % 2 == 0);
    const long long m = even ? nl / 2 : nl;
    const bool unit_rows = r2c ? (T[L].is == 1 && T[L].os == 1) : (T[L].is == 1 && T[L].os == 1);
    const long long cap = unit_rows ? c.row : c.col;

    // lines of the last-dimension real step, (src, dst) strides in reals
    std::vector<LDim> rl;
    for (int i = 0; i < L; ++i) rl.push_back(LDim{T[i].n, T[i].is * iu, T[i].os * ou, false});
    for (const GDim &d : V) rl.push_back(LDim{d.n, d.is * iu, d.os * ou, false});
    bool evens_src = true, evens_dst = true;  // all complex-side offsets even -> vector access
    for (const LDim &l : rl) {
        if (l.n <= 1) continue;
        if (l.s & 1) evens_src = false;
        if (l.d & 1) evens_dst = false;
    }

    auto leading_axes = [&](bool on_input) -> bool {
        // complex transforms over dims 0..L-1 with hc columns; r2c: forward in place on OUT,
        // c2r: backward on IN (the reference destroys multi-dim c2r input too)
        std::vector<int> order;
        for (int i = 0; i < L; ++i)
            if (T[i].n > 1) order.push_back(i);
        auto st_of = [&](int i) { return 2 * (on_input ? T[i].is : T[i].os); };
        std::sort(order.begin(), order.end(), [&](int a, int b) { return st_of(a) < st_of(b); });
        for (size_t oi = 0; oi < order.size(); ++oi) {
            const int ax = order[oi];
            std::vector<LDim> lines;
            for (int i = 0; i < g.rank; ++i) {
                if (i == ax) continue;
                long long n = (i == L) ? hc : T[i].n;
                lines.push_back(LDim{n, st_of(i), st_of(i), false});
            }
            for (const GDim &d : V) {
                long long s = 2 * (on_input ? d.is : d.os);
                lines.push_back(LDim{d.n, s, s, false});
            }
            const bool bwd = !r2c;
            std::vector<int> allowed = on_input ? std::vector<int>{B_IN, B_SCR} : std::vector<int>{B_OUT, B_SCR};
            if (any_neg) allowed.pop_back();  // (a work array would need this side's span and offset: stay in the caller's array)
            if (!add_c2c_axis(pc, T[ax].n, st_of(ax), st_of(ax), lines, bwd, bwd, allowed))
                return false;
        }
        return true;
    };

    if (!r2c && !leading_axes(true)) return false;
    const size_t first_real_step = pc.steps.size();

    // lengths the radix kernels cannot do (a prime factor > 13, or odd and longer than one CTA's
    // line): Bluestein on the full length, reading the real / Hermitian-half line directly
    const bool direct_ok = even ? axis_direct_ok(m, unit_rows) : (nl <= cap && factorable(nl));
    if (any_neg && !(direct_ok && (even ? m <= cap : true))) {
        set_error("negative strides: real transforms of this length (more than one pass, or Bluestein) need positive strides");
        return false;
    }
    if (!direct_ok) {
        BlueSpec b;
        b.n = nl;
        b.lines = rl;
        if (r2c) {
            b.src_real = true;
            b.s_axis = T[L].is;
            b.d_axis = 2 * T[L].os;
            b.nout = hc;
        } else {
            b.herm_in = true;
            b.dst_real = true;
            b.s_axis = 2 * T[L].is;
            b.d_axis = T[L].os;
            b.swap_in = b.swap_out = true;
        }
        if (!add_bluestein_axis(pc, b, {B_OUT})) return false;
    } else if (even && m <= cap) {
        AxisSpec a;
        a.n = (int)nl;
        a.rmode = r2c ? RM_R2C_HALF : RM_C2R_HALF;
        if (r2c) {
            a.s_axis = 2 * T[L].is;
            a.s_pair = T[L].is;
            a.d_axis = 2 * T[L].os;
            a.d_pair = 1;
            a.src_vec = (T[L].is == 1) && evens_src;
            a.dst_vec = evens_dst;
        } else {
            a.s_axis = 2 * T[L].is;
            a.s_pair = 1;
            a.d_axis = 2 * T[L].os;
            a.d_pair = T[L].os;
            a.src_vec = evens_src;
            a.dst_vec = (T[L].os == 1) && evens_dst;
            a.swap_out = true;
        }
        a.lines = rl;
        StepPlan st;
        if (!make_pass_step(st, a)) return false;
        st.inplace_ok = pc.in_place;  // row-for-row overlap was verified above
        if (pc.in_place && (st.pp.ld_col || st.pp.st_col || (T[L].is != 1 && T[L].n > 1))) {
            // a CTA must own whole rows for the in-place real step to be race-free
            if (T[L].is != 1 || T[L].os != 1) {
                set_error("in-place real transform with a strided last dimension");
                return false;
            }
        }
        st.dst_allowed = {B_OUT};
        pc.steps.push_back(st);
    } else if (!even) {
        if (nl > cap) {
            set_error("odd real length %lld > %lld is not supported", nl, cap);
            return false;
        }
        AxisSpec a;
        a.n = (int)nl;
        a.rmode = r2c ? RM_R2C_FULL : RM_C2R_FULL;
        if (r2c) {
            a.s_axis = T[L].is;
            a.d_axis = 2 * T[L].os;
            a.dst_vec = evens_dst;
        } else {
            a.s_axis = 2 * T[L].is;
            a.d_axis = T[L].os;
            a.src_vec = evens_src;
            a.swap_out = true;
        }
        a.lines = rl;
        StepPlan st;
        if (!make_pass_step(st, a)) return false;
        st.inplace_ok = pc.in_place;
        if (pc.in_place && (T[L].is != 1 || T[L].os != 1)) {
            set_error("in-place real transform with a strided last dimension");
            return false;
        }
        st.dst_allowed = {B_OUT};
        pc.steps.push_back(st);
    } else {
        // ---- four-step half-length complex transform through a compact scratch ----
        long long n1, n2;
        std::vector<LDim> nz;
        for (const LDim &l : rl)
            if (l.n > 1) nz.push_back(l);
        if (!split_four_step(m, &n1, &n2)) {
            // ---- half-length transform beyond the four-step limit (n > 2^23 fp32 / 2^21 fp64): the
            // three-pass complex split of plan_build.h on the packed pairs, Hermitian split / merge
            // as a separate in-place sweep.  Unit-stride lines only.
            if (T[L].is != 1 || T[L].os != 1 || !evens_src || !evens_dst) {
                set_error("real length %lld with a strided last dimension is beyond the four-step limit", nl);
                return false;
            }
            if (r2c) {
                std::vector<LDim> cl;
                for (const LDim &l : nz) cl.push_back(LDim{l.n, l.s, l.d, false});
                if (!add_c2c_axis(pc, m, 2, 2, cl, false, false, {B_OUT, B_SCR})) return false;
                StepPlan st;
                std::vector<LDim> ol;
                for (const LDim &l : nz) ol.push_back(LDim{l.n, l.d, l.d, false});
                fill_realpp(st, 1, (int)m, nl, 2, 2, ol, false);
                if (!st.rp.twr) return false;
                st.inplace_ok = true;
                st.src_fixed = B_OUT;
                st.dst_allowed = {B_OUT};
                pc.steps.push_back(st);
                if (!leading_axes(false)) return false;
                return assign_buffers(pc);
            }
            const std::vector<long long> cs3 = compact_strides(nz, m);
            {
                StepPlan st;
                std::vector<LDim> ol;
                for (size_t i = 0; i < nz.size(); ++i) ol.push_back(LDim{nz[i].n, nz[i].s, cs3[i], false});
                fill_realpp(st, 0, (int)m, nl, 2, 2, ol, false);
                if (!st.rp.twr) return false;
                st.dst_allowed = {B_SCR};
                pc.steps.push_back(st);
            }
            std::vector<LDim> cl;
            for (size_t i = 0; i < nz.size(); ++i) cl.push_back(LDim{nz[i].n, cs3[i], nz[i].d, false});
            if (!add_c2c_axis(pc, m, 2, 2, cl, false, true, {B_OUT, B_SCR})) return false;
            return assign_buffers(pc);
        }
        long long q1 = 0, q2 = 0;
        long long t1 = 0, t2 = 0, t3 = 0;
        // ---- three passes through the short kernels (four CTAs per SM) instead of two through the
        // 1024-point ones (one CTA per SM): see profiles/r1_real3_probe.md
        // Measured on B200 (profiles/r1_real3_probe.md, 2^20 x 256 fp32): 1.27 ms against 1.21 ms for the two-pass
        // plan -- at 2^20 the factors are 256 x 64 x 64 and the 64-point kernels' 8 KB tiles do not reach the
        // 256-point kernels' bandwidth.  So: default only where no two-pass split exists (n > 2^22 fp32 /
        // 2^20 fp64, where it replaces four passes), FFTW_B200_THREE_PASS_REAL=1 forces it from 2^19 on.
        long long u1 = 0, u2 = 0;
        const bool want3 = opt("FFTW_B200_THREE_PASS_REAL", B2_CFG_THREE_PASS_REAL) ? nl >= (1ll << 19) : !split_two_pass_real(nl, &u1, &u2);
        if (r2c && !no_two_pass_real() && !opt("FFTW_B200_NO_TWO_PASS_REAL", B2_CFG_NO_TWO_PASS_REAL) && !getenv("FFTW_B200_NO_THREE_PASS_REAL") &&
            want3 && nl < (1ll << 31) && (nl & (nl - 1)) == 0 && T[L].is == 1 && T[L].os == 1 && evens_src && evens_dst &&
            split_three_pass_real(nl, &t1, &t2, &t3)) {
            const long long rows = t1 / 2 + 1, Q = t2 * t3;
            const std::vector<long long> cs3 = compact_strides(nz, rows * Q);
            {   // pass 1: IN viewed as complex [t1][Q/2] -> SCR [t1/2+1][Q], pair split + w_n^(k1*j)
                AxisSpec a;
                a.n = (int)t1;
                a.s_axis = Q;
                a.d_axis = 2 * Q;
                a.lines.push_back(LDim{Q / 2, 2, 4, true});
                for (size_t i = 0; i < nz.size(); ++i) a.lines.push_back(LDim{nz[i].n, nz[i].s, cs3[i], false});
                a.epi_N = nl;
                StepPlan st;
                if (!make_pass_step(st, a)) return false;
                st.pp.rsplit = 1;
                st.needs_tile_kernel = true;
                st.dst_allowed = {B_SCR};
                pc.steps.push_back(st);
            }
            {   // pass 2: inside every row k1, t2-point FFTs over j2 (stride t3), times w_Q^(k2*j3); in place
                AxisSpec a;
                a.n = (int)t2;
                a.s_axis = a.d_axis = 2 * t3;
                a.lines.push_back(LDim{t3, 2, 2, true});
                a.lines.push_back(LDim{rows, 2 * Q, 2 * Q, false});
                for (size_t i = 0; i < nz.size(); ++i) a.lines.push_back(LDim{nz[i].n, cs3[i], cs3[i], false});
                a.epi_N = Q;
                StepPlan st;
                if (!make_pass_step(st, a)) return false;
                st.inplace_ok = true;
                st.needs_tile_kernel = true;
                st.src_fixed = B_SCR;
                st.dst_allowed = {B_SCR};
                pc.steps.push_back(st);
            }
            {   // pass 3: rows of t3 for every (k1, k2), output index k1 + t1*k2 + t1*t2*k3, folded
                AxisSpec a;
                a.n = (int)t3;
                a.s_axis = 2;
                a.d_axis = 2 * t1 * t2;
                a.lines.push_back(LDim{rows, 2 * Q, 2, false});
                a.lines.push_back(LDim{t2, 2 * t3, 2 * t1, false});
                for (size_t i = 0; i < nz.size(); ++i) a.lines.push_back(LDim{nz[i].n, cs3[i], nz[i].d, false});
                StepPlan st;
                if (!make_pass_step(st, a)) return false;
                PassParams &q = st.pp;
                q.fold_n = nl;
                q.fold_n1 = (int)t1;
                q.fold_dim = 0;
                for (int d = 1; d < q.nld; ++d)
                    if (q.ln[d] == t2 && q.ls[d] == 2 * t3 && q.ld[d] == 2 * t1 && q.fold_dim == 0) q.fold_dim = d;
                if (q.fold_dim == 0) {
                    set_error("three-pass real transform: fold dimension lost in canonicalisation");
                    return false;
                }
                q.ld[q.fold_dim] = 0;  // the k2 offset enters through the folded index
                q.fold_cmul = (int)t1;
                q.fold_pmul = t1 * t2;
                st.needs_tile_kernel = true;
                st.src_fixed = B_SCR;
                st.dst_allowed = {B_OUT};
                pc.steps.push_back(st);
            }
            if (!leading_axes(false)) return false;
            return assign_buffers(pc);
        }
        if (r2c && !no_two_pass_real() && !opt("FFTW_B200_NO_TWO_PASS_REAL", B2_CFG_NO_TWO_PASS_REAL) && (nl & (nl - 1)) == 0 &&
            T[L].is == 1 && T[L].os == 1 && evens_src && evens_dst && split_two_pass_real(nl, &q1, &q2)) {
            // ---- two passes: real column pairs (FFT + split + twiddle), then rows with a Hermitian fold
            const long long rows = q1 / 2 + 1;
            const std::vector<long long> cs2 = compact_strides(nz, rows * q2);
            {   // pass 1: IN viewed as complex [q1][q2/2] -> SCR [q1/2+1][q2] complex
                AxisSpec a;
                a.n = (int)q1;
                a.s_axis = q2;      // reals between consecutive j1 at fixed j2
                a.d_axis = 2 * q2;  // reals between consecutive k1 rows of the intermediate
                a.lines.push_back(LDim{q2 / 2, 2, 4, true});
                for (size_t i = 0; i < nz.size(); ++i) a.lines.push_back(LDim{nz[i].n, nz[i].s, cs2[i], false});
                a.epi_N = nl;
                StepPlan st;
                if (!make_pass_step(st, a)) return false;
                st.pp.rsplit = 1;
                st.needs_tile_kernel = true;
                st.dst_allowed = {B_SCR};
                pc.steps.push_back(st);
            }
            {   // pass 2: rows of q2 for k1 = 0..q1/2, stored transposed at k1 + q1*k2 or folded to n - that
                AxisSpec a;
                a.n = (int)q2;
                a.s_axis = 2;
                a.d_axis = 2 * T[L].os * q1;
                a.lines.push_back(LDim{rows, 2 * q2, 2 * T[L].os, false});
                for (size_t i = 0; i < nz.size(); ++i) a.lines.push_back(LDim{nz[i].n, cs2[i], nz[i].d, false});
                StepPlan st;
                if (!make_pass_step(st, a)) return false;
                st.pp.fold_n = nl;
                st.pp.fold_n1 = (int)q1;
                st.needs_tile_kernel = true;
                st.src_fixed = B_SCR;
                st.dst_allowed = {B_OUT};
                pc.steps.push_back(st);
            }
            if (nz.size() == 1 && nz[0].n > 1) {  // one batch dim: the two passes can run transform group by transform group
                StepPlan &a = pc.steps[pc.steps.size() - 2], &b = pc.steps.back();
                a.grp_L = b.grp_L = nz[0].n;
                a.grp_slab = b.grp_slab = cs2[0];  // the intermediate's stride per transform, in reals
            }
            (void)first_real_step;
            if (!leading_axes(false)) return false;
            return assign_buffers(pc);
        }
        // (c2r: opt-in.  Measured on B200, 2^20 x 256: 1.56 ms against 1.33 ms for merge + two passes --
        // the unfold + twiddle column pass runs 1744 instructions per thread at one 1024-thread CTA
        // per SM; 2^16 x 4096: 1.01 against 1.07 ms.)
        if (!r2c && opt("FFTW_B200_TWO_PASS_C2R", B2_CFG_TWO_PASS_C2R) && !no_two_pass_real() && (nl & (nl - 1)) == 0 &&
            T[L].is == 1 && T[L].os == 1 && evens_src && evens_dst && split_two_pass_real(nl, &q1, &q2, false)) {
            // ---- two passes, the mirror image of the r2c pair: columns k1 <= q1/2 of the Hermitian-
            // unfolded spectrum (backward FFT over k2 + twiddle), then per pair of output columns a
            // backward FFT over k1 whose input is merged from two Hermitian half rows
            const long long rows = q1 / 2 + 1;
            const std::vector<long long> cs2 = compact_strides(nz, rows * q2);
            {   // pass 1: IN (n/2+1 complex) -> SCR [q2][q1/2+1]
                AxisSpec a;
                a.n = (int)q2;
                a.s_axis = 2 * q1;
                a.d_axis = 2 * rows;
                a.lines.push_back(LDim{rows, 2, 2, true});
                for (size_t i = 0; i < nz.size(); ++i) a.lines.push_back(LDim{nz[i].n, nz[i].s, cs2[i], false});
                a.swap_in = a.swap_out = true;
                a.epi_N = nl;
                StepPlan st;
                if (!make_pass_step(st, a)) return false;
                st.pp.fold_n = nl;
                st.pp.fold_n1 = (int)q1;
                st.needs_tile_kernel = true;
                st.dst_allowed = {B_SCR};
                pc.steps.push_back(st);
            }
            {   // pass 2: rows 2c, 2c+1 of SCR -> OUT viewed as complex [q1][q2/2], transposed store
                AxisSpec a;
                a.n = (int)q1;
                a.s_axis = 2;
                a.d_axis = q2;
                a.lines.push_back(LDim{q2 / 2, 4 * rows, 2, false});
                for (size_t i = 0; i < nz.size(); ++i) a.lines.push_back(LDim{nz[i].n, cs2[i], nz[i].d, false});
                a.swap_out = true;
                StepPlan st;
                if (!make_pass_step(st, a)) return false;
                st.pp.pmerge = 1;
                st.needs_tile_kernel = true;
                st.src_fixed = B_SCR;
                st.dst_allowed = {B_OUT};
                pc.steps.push_back(st);
            }
            return assign_buffers(pc);
        }
        const std::vector<long long> cs = compact_strides(nz, m);
        if (r2c) {
            {   // A: IN (pairs of reals) -> SCR compact, column pass + twiddle
                AxisSpec a;
                a.n = (int)n1;
                a.s_axis = 2 * T[L].is * n2;
                a.s_pair = T[L].is;
                a.d_axis = 2 * n2;
                a.src_vec = (T[L].is == 1) && evens_src;
                for (size_t i = 0; i < nz.size(); ++i) a.lines.push_back(LDim{nz[i].n, nz[i].s, cs[i], false});
                a.lines.push_back(LDim{n2, 2 * T[L].is, 2, true});
                a.epi_N = m;
                StepPlan st;
                if (!make_pass_step(st, a)) return false;
                st.dst_allowed = {B_SCR};
                pc.steps.push_back(st);
            }
            {   // B: SCR -> OUT rows (Z[k] at complex position k), transposed store
                AxisSpec a;
                a.n = (int)n2;
                a.s_axis = 2;
                a.d_axis = 2 * T[L].os * n1;
                a.dst_vec = evens_dst;
                for (size_t i = 0; i < nz.size(); ++i) a.lines.push_back(LDim{nz[i].n, cs[i], nz[i].d, false});
                a.lines.push_back(LDim{n1, 2 * n2, 2 * T[L].os, false});
                StepPlan st;
                if (!make_pass_step(st, a)) return false;
                st.src_fixed = B_SCR;
                st.dst_allowed = {B_OUT};
                pc.steps.push_back(st);
            }
            {   // Hermitian split, in place on OUT
                StepPlan st;
                std::vector<LDim> ol;
                for (const LDim &l : nz) ol.push_back(LDim{l.n, l.d, l.d, false});
                fill_realpp(st, 1, (int)m, nl, 2 * T[L].os, 2 * T[L].os, ol, false);
                if (!st.rp.twr) return false;
                st.inplace_ok = true;
                st.src_fixed = B_OUT;
                st.dst_allowed = {B_OUT};
                pc.steps.push_back(st);
            }
        } else {
            {   // Hermitian merge: IN -> SCR compact (swapped for the backward transform)
                StepPlan st;
                std::vector<LDim> ol;
                for (size_t i = 0; i < nz.size(); ++i) ol.push_back(LDim{nz[i].n, nz[i].s, cs[i], false});
                fill_realpp(st, 0, (int)m, nl, 2 * T[L].is, 2, ol, false);
                if (!st.rp.twr) return false;
                st.dst_allowed = {B_SCR};
                pc.steps.push_back(st);
            }
            {   // A: in place on SCR
                AxisSpec a;
                a.n = (int)n1;
                a.s_axis = 2 * n2;
                a.d_axis = 2 * n2;
                for (size_t i = 0; i < nz.size(); ++i) a.lines.push_back(LDim{nz[i].n, cs[i], cs[i], false});
                a.lines.push_back(LDim{n2, 2, 2, true});
                a.epi_N = m;
                StepPlan st;
                if (!make_pass_step(st, a)) return false;
                st.inplace_ok = true;
                st.src_fixed = B_SCR;
                st.dst_allowed = {B_SCR};
                pc.steps.push_back(st);
            }
            {   // B: SCR -> OUT as pairs of reals, transposed store, un-swap
                AxisSpec a;
                a.n = (int)n2;
                a.s_axis = 2;
                a.d_axis = 2 * T[L].os * n1;
                a.d_pair = T[L].os;
                a.dst_vec = (T[L].os == 1) && evens_dst;
                a.swap_out = true;
                for (size_t i = 0; i < nz.size(); ++i) a.lines.push_back(LDim{nz[i].n, cs[i], nz[i].d, false});
                a.lines.push_back(LDim{n1, 2 * n2, 2 * T[L].os, false});
                StepPlan st;
                if (!make_pass_step(st, a)) return false;
                st.src_fixed = B_SCR;
                st.dst_allowed = {B_OUT};
                pc.steps.push_back(st);
            }
        }
    }
    (void)first_real_step;
    if (r2c && !leading_axes(false)) return false;
    return assign_buffers(pc);
}

}  // namespace b2
