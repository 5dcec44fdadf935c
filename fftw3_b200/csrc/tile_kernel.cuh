// csrc/tile_kernel.cuh -- specialised tile kernels: the hot path.  Instantiated for power-of-two lengths with TileSched
// (tile_dispatch.cuh) and for 95 other smooth lengths with mixed-radix schedules (MixSched; mixtile*.cu).
//
// One CTA transforms a tile of C lines of N points.  Each thread keeps E points of one line
// in registers across ALL radix passes (compile-time schedule, radices <= E); between passes
// the points are exchanged through padded shared memory (Stockham auto-sort), the first pass
// reads global memory straight into registers and the last pass writes registers straight
// back, so a line costs exactly one HBM read and one HBM write however many radix passes it
// needs.  Twiddles: one accurate table value per butterfly (fp64-generated, rounded once),
// the other R-2 by a short product tree.
//
//   MODE_ROW     lines contiguous in src and dst (batched 1-D, last axis of N-D)
//   MODE_COL     lines strided, neighbouring lines contiguous (other axes of N-D, four-step
//                pass A): a warp walks C neighbouring lines -> C*sizeof(complex) segments
//   MODE_ROW2COL contiguous lines in, transposed out (four-step pass B)
//
// These replace the reference's twiddle codelets t1*/t2*/t3* and q1* (fftw/dft_simd/common/
// t3fv_32.c:482, q1fv_8.c:542; drivers dftw_direct_apply fftw/fftw_api.c:2315-2324,
// dftw_directsq_apply :2613) and leaf codelets n1*/n2* (n1fv_128.c:1851; direct_apply :3183).
// Anything these kernels do not cover runs in the generic kernel (generic_kernel.cuh).
#pragma once
#include "butterfly.cuh"
#include "plan_build.h"

enum { MODE_ROW = 0, MODE_COL = 1, MODE_ROW2COL = 2 };

constexpr int b2_ilog2(int v) { return v <= 1 ? 0 : 1 + b2_ilog2(v / 2); }

template <int N, int E> struct TileSched {
    static constexpr int LOGN = b2_ilog2(N), LOGE = b2_ilog2(E);
    static constexpr int LOGR = LOGE < 4 ? LOGE : 4;  // butterflies up to radix 16 (butterfly.cuh)
    static constexpr int NP = (LOGN + LOGR - 1) / LOGR;
    static constexpr int bits(int i) { return (LOGN + (NP - 1 - i)) / NP; }
    static constexpr int radix(int i) { return 1 << bits(i); }
    static constexpr int ns(int i) { return i == 0 ? 1 : ns(i - 1) * radix(i - 1); }
    static constexpr int PER = -1;  // shared-memory padding period: the precision's default
    static constexpr bool MIXED = false;
    static constexpr int xp(int) { return 0; }  // (per-exchange layouts are a MixSched matter)
    static constexpr int xa(int) { return 0; }
    static constexpr int xls(int) { return -1; }
};

// Per-exchange shared-memory layouts of a mixed-radix schedule.  The exchange between radix passes X and X+1 is written
// by one access pattern (R points NS apart per butterfly) and read by another (E points T apart per thread), both known
// at compile time, and the buffer is rewritten in every exchange -- so every exchange may have its own layout
//   f(o) = o + (o / P) * A   with line stride LS,
// chosen to minimise bank conflicts.  The table (mix_layouts.h) is produced by scripts/analysis/smem_conflicts.py, an
// exact enumeration of the kernel's shared-memory accesses calibrated against ncu's conflict counters.
template <int N, int E> struct MixLayout {
    static constexpr int NX = 0;
    static constexpr int get(int, int) { return 0; }
};
#define B2_MIX_LAYOUT(N, E, NXV, ...)                \
    template <> struct MixLayout<N, E> {             \
        static constexpr int NX = NXV;               \
        static constexpr int get(int x, int f) {     \
            constexpr int t[NXV][3] = {__VA_ARGS__}; \
            return t[x][f];                          \
        }                                            \
    };
#include "mix_layouts.h"

// Mixed-radix schedule for a length that is not a power of two: N = R0 * R1 * ..., every radix divides E
// (a thread holds E points and runs E / R butterflies of radix R in a pass).  Radices: whatever Bf<R>
// exists for (butterfly.cuh: 2 3 4 5 7 8 16 and the composites 6 9 10 12 20 24 25 30).
template <int... RS> struct MixRadices {
    static constexpr int NP = sizeof...(RS);
    static constexpr int get(int i) {
        constexpr int r[] = {RS...};
        return r[i];
    }
    static constexpr int prod(int i) { return i == 0 ? 1 : prod(i - 1) * get(i - 1); }  // of the first i radices
    static constexpr bool all_divide(int e, int i) { return i == NP ? true : (e % get(i) == 0 && all_divide(e, i + 1)); }
};
template <int N, int E, int... RS> struct MixSched {
    typedef MixRadices<RS...> RR;
    static constexpr int NP = RR::NP;
    static constexpr int radix(int i) { return RR::get(i); }
    static constexpr int ns(int i) { return RR::prod(i); }
    static_assert(RR::prod(NP) == N, "the radices must multiply to N");
    static_assert(N % E == 0 && RR::all_divide(E, 0), "every radix must divide E, and E must divide N");
    // the first pass writes R0 neighbouring points per butterfly, butterflies of neighbouring threads R0 apart:
    // one pad element per R0 makes that stride odd (an odd R0 needs none)
    static constexpr int PER = (RR::get(0) % 2 == 0) ? RR::get(0) : 0;
    static constexpr bool MIXED = true;
    typedef MixLayout<N, E> ML;
    static constexpr bool TABLE = ML::NX == NP - 1 && NP > 1;
    static constexpr int xp(int X) { return TABLE ? ML::get(X, 0) : PER; }
    static constexpr int xa(int X) { return TABLE ? ML::get(X, 1) : (PER ? 1 : 0); }
    static constexpr int xls(int X) { return TABLE ? ML::get(X, 2) : -1; }  // -1: the kernel's default line stride
};

#if B2_CUDA
#define B2_TREGS(ctx) (ctx.vbase)
#else
#define B2_TREGS(ctx) (ctx.vbase + (size_t)tid * E)
#endif

struct TileCtx {
    const PassParams *p;
    const real_t *src;
    real_t *dst;
    cpx_t *sm;
    cpx_t *vbase;
    long long l0;         // first line of the tile along the tile dim
    long long sbase, dbase;
    int nlines;
    int i2base;
    int scoord;           // coordinate along the scatter dim (EPI 8)
    int fcoord;           // coordinate along the fold dim (three-pass real transform)
    int nthreads;
};

// index -> padded index, one pad element per PER (0: no padding)
template <int PER> B2_HD int b2_pad(int o) {
    if (PER == 0) return o;
    if ((PER & (PER - 1)) == 0) return o + (o >> b2_ilog2(PER > 0 ? PER : 1));
    return o + (int)((unsigned)o / (unsigned)(PER > 0 ? PER : 1));
}

// Streaming data is read through L2 only (ld.global.cg): it is never reused from L1, and when
// passes are chained inside one launch the producer may be another SM, whose writes a stale L1
// line would hide.
B2_D cpx_t b2_ldcg(const cpx_t *p) {
#if B2_CUDA
    if (sizeof(real_t) == 4) {
        const float2 v = __ldcg(reinterpret_cast<const float2 *>(p));
        return cpx_t{(real_t)v.x, (real_t)v.y};
    } else {
        const double2 v = __ldcg(reinterpret_cast<const double2 *>(p));
        return cpx_t{(real_t)v.x, (real_t)v.y};
    }
#else
    return *p;
#endif
}

// L2 prefetch hint for the 128-byte line holding *p
B2_D void b2_prefetch_l2(const void *p) {
#if B2_CUDA
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#else
    (void)p;
#endif
}

// w[k] = w1^k, k = 1..R-1, by a product tree of depth <= log2(R)
template <int R> B2_D void b2_twiddle_powers(cpx_t w1, cpx_t *w) {
    w[1] = w1;
#pragma unroll
    for (int k = 2; k < R; ++k) {
        if ((k & 1) == 0)
            w[k] = cmul(w[k / 2], w[k / 2]);
        else
            w[k] = cmul(w[k - 1], w1);
    }
}

// EPI: 0 none, 1 four-step twiddle on the way out, 2 split of a pair of real columns + twiddle
template <int N, int E, int MODE, int EPI, class SCHED = TileSched<N, E>> struct TileFFT {
    typedef SCHED S;
    static constexpr int T = N / E;  // threads per line
    static constexpr int PER = S::PER < 0 ? (sizeof(real_t) == 4 ? 16 : 8) : S::PER;  // padding period (0: none)
    static constexpr int PERD = PER > 0 ? PER : 1;
    static constexpr bool COLMAP = (MODE == MODE_COL);
    // smem line stride (complex): row-mapped passes pad inside the line; any kernel that walks
    // neighbouring lines with neighbouring lanes needs an odd stride
    static constexpr int LSRAW = (COLMAP || PER == 0) ? N : (N + N / PERD);
    static constexpr int LS = (MODE == MODE_ROW) ? LSRAW : (LSRAW | 1);

    static B2_HD int phys(int o) { return COLMAP ? o : b2_pad<PER>(o); }
    // phys(o + k*STEP) for k = 0..: when STEP is a multiple of the padding period (or the whole
    // run stays inside one period) the padding separates and k*STEP' becomes an immediate offset
    template <int STEP, int SPAN> static B2_HD int phys_k(int o, int po, int k) {
        if (COLMAP || PER == 0) return o + k * STEP;
        if (STEP % PERD == 0) return po + k * (STEP + STEP / PERD);
        if (SPAN > 0 && STEP * SPAN <= PER && (PERD % (STEP * (SPAN > 0 ? SPAN : 1))) == 0) return po + k * STEP;  // o is a multiple of STEP*SPAN
        return b2_pad<PER>(o + k * STEP);
    }
    // ---- exchange X (between radix passes X and X+1): its own layout for mixed-radix schedules, the one above otherwise
    static constexpr int xper(int X) { return S::MIXED ? S::xp(X) : PER; }
    static constexpr int xadd(int X) { return S::MIXED ? S::xa(X) : (PER ? 1 : 0); }
    static constexpr int xls(int X) { return (COLMAP || !S::MIXED || S::xls(X) < 0) ? LS : S::xls(X); }
    static constexpr int xls_max(int X) { return X < 0 ? LS : (xls(X) > xls_max(X - 1) ? xls(X) : xls_max(X - 1)); }
    template <int X> static B2_HD int physx(int o) {
        constexpr int P = xper(X), A = xadd(X), PD = P > 0 ? P : 1;
        if (COLMAP || P == 0) return o;
        if (A == 1) return b2_pad<P>(o);
        if ((P & (P - 1)) == 0) return o + (o >> b2_ilog2(PD)) * A;
        return o + (int)((unsigned)o / (unsigned)PD) * A;
    }
    template <int X, int STEP, int SPAN> static B2_HD int physx_k(int o, int po, int k) {
        constexpr int P = xper(X), A = xadd(X), PD = P > 0 ? P : 1;
        if (COLMAP || P == 0) return o + k * STEP;
        if (STEP % PD == 0) return po + k * (STEP + STEP / PD * A);
        if (SPAN > 0 && STEP * SPAN <= P && (PD % (STEP * (SPAN > 0 ? SPAN : 1))) == 0) return po + k * STEP;  // the run stays inside one period
        return physx<X>(o + k * STEP);
    }
    static B2_HD size_t smem_bytes(int C) { return (size_t)C * xls_max(S::NP - 2) * sizeof(cpx_t); }

    // thread -> (line c, slot t).  Every thread of the CTA owns a slot: a chained launch, whose
    // CTAs have the thread count of the widest step, retires its surplus warps before it gets here
    // (a per-thread "live" guard costs the 64-register kernels 20-80 bytes of spills).
    static B2_HD bool map(int tid, int C, int lgC, int &c, int &t) {
        if (COLMAP) {
            if (lgC >= 0) {
                t = tid >> lgC;
                c = tid & (C - 1);
            } else {
                t = tid / C;
                c = tid - t * C;
            }
        } else {
            c = tid / T;
            t = tid - c * T;
        }
        return true;
    }

    // tile index -> coordinates: exact multiply-shift divisions (b2_pp_prepare), 31-bit index
    static B2_D void decode(const PassParams &p, unsigned bid, long long &l0, int &nlines, long long &sbase,
                            long long &dbase, int &i2base, int *scoord = nullptr, int *fcoord = nullptr) {
        const int C = p.C;
        unsigned rest = b2_fastdiv(bid, p.dv_m[0], p.dv_s[0]);
        const unsigned t0 = bid - rest * (unsigned)p.tiles0;
        l0 = (long long)t0 * C;
        nlines = (int)((p.ln[0] - l0 < C) ? (p.ln[0] - l0) : C);
        sbase = 0;
        dbase = 0;
        i2base = 0;
        for (int d = 1; d < p.nld; ++d) {
            const unsigned qd = b2_fastdiv(rest, p.dv_m[d], p.dv_s[d]);
            const unsigned cc = rest - qd * (unsigned)p.ln[d];
            rest = qd;
            sbase += (long long)cc * p.ls[d];
            dbase += (long long)cc * p.ld[d];
            if (d == p.epi_dim) i2base = (int)cc;
            if (EPI == 8 && scoord && d == p.scat_dim) *scoord = (int)cc;
            if (MODE == MODE_ROW2COL && fcoord && d == p.fold_dim) *fcoord = (int)cc;
        }
    }

    template <int P> static B2_D void butterflies(const TileCtx &x, cpx_t *v, int c, int t, bool active) {
        constexpr int R = S::radix(P), NS = S::ns(P), NB = E / R;
        constexpr bool LAST = (P == S::NP - 1);
        constexpr int X = LAST ? 0 : P;  // the exchange this pass writes (none after the last pass)
        const cpx_t *tw = x.p->tw;
        cpx_t *line = x.sm + (size_t)c * xls(X);
#pragma unroll
        for (int m = 0; m < NB; ++m) {
            const int j = t + m * T;
            cpx_t a[R];
#pragma unroll
            for (int k = 0; k < R; ++k) a[k] = v[m + k * NB];
            const int jm = (NS & (NS - 1)) == 0 ? (j & (NS - 1)) : (int)((unsigned)j % (unsigned)NS);
            if (NS > 1) {
                cpx_t w[R];
                b2_twiddle_powers<R>(tw[jm * (N / (NS * R))], w);
#pragma unroll
                for (int k = 1; k < R; ++k) a[k] = cmul(a[k], w[k]);
            }
            Bf<R, real_t>::run(a);
            if (LAST) {
#pragma unroll
                for (int k = 0; k < R; ++k) v[m + k * NB] = a[k];
            } else if (active) {
                const int o = (j - jm) * R + jm;  // (j / NS) * NS * R + jm
                const int po = physx<X>(o);
                // (mixed radix: the R points of a butterfly lie inside one block of NS*R, so a period that is a multiple
                // of NS*R leaves their offsets immediate)
#pragma unroll
                for (int k = 0; k < R; ++k)
                    line[NS == 1 ? physx_k<X, 1, R>(o, po, k) : physx_k<X, NS, (S::MIXED ? R : 0)>(o, po, k)] = a[k];
            }
        }
    }

    template <int P> static B2_D void passes_from(TileCtx &x) {
        // exchange: everybody has written pass P-1's outputs; read this pass's inputs
        const int b2_nthreads = x.nthreads;
        (void)b2_nthreads;
        B2_SYNC();
        B2_FOR_THREADS {
            cpx_t *v = B2_TREGS(x);
            int c, t;
            if (map(tid, x.p->C, x.p->lgC, c, t)) {
                const cpx_t *line = x.sm + (size_t)c * xls(P - 1);
                const int pt = physx<P - 1>(t);
#pragma unroll
                for (int q = 0; q < E; ++q) v[q] = line[physx_k<P - 1, T, 0>(t, pt, q)];
            }
        }
        if (P < S::NP - 1) B2_SYNC();  // the next write must not overtake a neighbour's read
        B2_FOR_THREADS {
            cpx_t *v = B2_TREGS(x);
            int c, t;
            const bool live = map(tid, x.p->C, x.p->lgC, c, t);
            butterflies<P>(x, v, c, t, live);
        }
        if constexpr (P + 1 < S::NP) passes_from<P + 1>(x);
    }

    // w_N^e from the two-level table of the four-step epilogue
    static B2_D cpx_t epi_root(const PassParams &p, unsigned long long e) {
        return cmul(p.epi_hi[e >> p.epi_shift], p.epi_lo[e & ((1ull << p.epi_shift) - 1)]);
    }

    // EPI == 2, first pass of the two-pass real transform.  The line is the complex FFT Z of a pair
    // of real columns (x_a + i x_b); per k1 <= N/2 form the two real spectra
    //   Xa = (Z[k1] + conj Z[N-k1]) / 2,   Xb = -i (Z[k1] - conj Z[N-k1]) / 2,
    // multiply by the four-step twiddles w_n^(k1*j2) of their columns j2 = 2*pair, 2*pair+1 and store
    // both as one vector at row k1 of the [N/2+1][n2] intermediate.  The pairing stays inside the
    // line, so it goes through this CTA's shared memory only.
    // Reference: the same split is the hc2c step of rdft2-ct (ct_hc2c_direct_apply
    // fftw/fftw_api.c:5831-5845) applied to column pairs instead of the whole half-length transform.
    static B2_D void rsplit_exit(TileCtx &x, const PassParams &p, cpx_t *cdst, long long bid) {
        (void)bid;
        const int b2_nthreads = x.nthreads;
        (void)b2_nthreads;
        const int C = p.C;
        B2_SYNC();  // the last exchange has been read
        B2_FOR_THREADS {
            cpx_t *v = B2_TREGS(x);
            int c, t;
            if (map(tid, C, p.lgC, c, t)) {
                cpx_t *line = x.sm + (size_t)c * LS;
                const int pt = phys(t);
#pragma unroll
                for (int q = 0; q < E; ++q) line[phys_k<T, 0>(t, pt, q)] = v[q];
            }
        }
        B2_SYNC();
        B2_FOR_THREADS {
            int c, t;
            map(tid, C, p.lgC, c, t);
            if (c < x.nlines) {
                const cpx_t *line = x.sm + (size_t)c * LS;
                const unsigned long long j2 = 2ull * (unsigned long long)(x.l0 + c);  // even column of the pair
                const long long ds = p.d_axis / 2;
                cpx_t *row0 = cdst + (x.dbase + (x.l0 + c) * p.ld[0]) / 2;
                struct alignas(16) Pair {  // one 16-byte vector store in single precision, two in double
                    cpx_t a, b;
                };
                constexpr int HQ = E / 2;  // k1 = t + q*T, q < E/2 covers 0 .. N/2-1
                const cpx_t wa0 = epi_root(p, j2 * (unsigned long long)t), was = epi_root(p, j2 * (unsigned long long)T);
                const cpx_t w10 = epi_root(p, (unsigned long long)t), w1s = epi_root(p, (unsigned long long)T);
                cpx_t pa[HQ], p1[HQ];
                pa[0] = cpx_t{1, 0};
                p1[0] = cpx_t{1, 0};
                if (HQ > 1) {
                    b2_twiddle_powers<HQ>(was, pa);
                    b2_twiddle_powers<HQ>(w1s, p1);
                }
#pragma unroll
                for (int q = 0; q < HQ; ++q) {
                    const int k1 = t + q * T;
                    const cpx_t za = line[phys(k1)], zb = cconj(line[phys((N - k1) & (N - 1))]);
                    const cpx_t xa = cscale(cadd(za, zb), (real_t)0.5);
                    const cpx_t xb = cmul_mi(cscale(csub(za, zb), (real_t)0.5));
                    const cpx_t wa = q ? cmul(wa0, pa[q]) : wa0;
                    const cpx_t w1 = q ? cmul(w10, p1[q]) : w10;
                    Pair o;
                    o.a = cmul(xa, wa);
                    o.b = cmul(xb, cmul(wa, w1));
                    *reinterpret_cast<Pair *>(row0 + (long long)k1 * ds) = o;
                }
                if (t == 0) {  // k1 = N/2 pairs with itself
                    const int k1 = N / 2;
                    const cpx_t za = line[phys(k1)];
                    const cpx_t wa = epi_root(p, j2 * (unsigned long long)k1);
                    Pair o;
                    o.a = cmul(cpx_t{za.x, 0}, wa);
                    o.b = cmul(cpx_t{za.y, 0}, cmul(wa, epi_root(p, (unsigned long long)k1)));
                    *reinterpret_cast<Pair *>(row0 + (long long)k1 * ds) = o;
                }
            }
        }
    }

    // EPI == 3: r2c of n = 2N reals per contiguous line.  The line holds Z = FFT_N(x[2j] + i x[2j+1]);
    //   X[k] = E + w_n^k O,  X[N-k] = conj(E - w_n^k O),  E = (Z[k] + conj Z[N-k]) / 2,
    //   O = -i (Z[k] - conj Z[N-k]) / 2,  and X[0] = Re Z0 + Im Z0, X[N] = Re Z0 - Im Z0.
    // Same arithmetic as the generic kernel's r2c post-processing and as the reference's hc2c step
    // (ct_hc2c_direct_apply fftw/fftw_api.c:5831-5845, codelets hc2cfdftv_*).
    static B2_D void r2c_row_exit(TileCtx &x, const PassParams &p, cpx_t *cdst) {
        const int b2_nthreads = x.nthreads;
        (void)b2_nthreads;
        const int C = p.C;
        B2_SYNC();
        B2_FOR_THREADS {
            cpx_t *v = B2_TREGS(x);
            int c, t;
            if (map(tid, C, p.lgC, c, t)) {
                cpx_t *line = x.sm + (size_t)c * LS;
                const int pt = phys(t);
#pragma unroll
                for (int q = 0; q < E; ++q) line[phys_k<T, 0>(t, pt, q)] = v[q];
            }
        }
        B2_SYNC();
        B2_FOR_THREADS {
            int c, t;
            map(tid, C, p.lgC, c, t);
            if (c < x.nlines) {
                const cpx_t *line = x.sm + (size_t)c * LS;
                cpx_t *D = cdst + (x.dbase + (x.l0 + c) * p.ld[0]) / 2;
                constexpr int HQ = E / 2;  // k = t + q*T, q < E/2 covers 0 .. N/2-1
                cpx_t zk[HQ], zc[HQ], w[HQ];
#pragma unroll
                for (int q = 0; q < HQ; ++q) {
                    const int k = t + q * T;
                    zk[q] = line[phys(k)];
                    zc[q] = cconj(line[phys((N & (N - 1)) == 0 ? ((N - k) & (N - 1)) : (k ? N - k : 0))]);
                    w[q] = p.twr[k];
                }
#pragma unroll
                for (int q = 0; q < HQ; ++q) {
                    const int k = t + q * T;
                    if (k == 0) {
                        D[0] = cpx_t{zk[q].x + zk[q].y, (real_t)0};
                        D[N] = cpx_t{zk[q].x - zk[q].y, (real_t)0};
                    } else {
                        const cpx_t Ev = cscale(cadd(zk[q], zc[q]), (real_t)0.5);
                        const cpx_t Ov = cmul_mi(cscale(csub(zk[q], zc[q]), (real_t)0.5));
                        const cpx_t wO = cmul(w[q], Ov);
                        D[k] = cadd(Ev, wO);
                        D[N - k] = cconj(csub(Ev, wO));
                    }
                }
                if (t == 0) D[N / 2] = cconj(line[phys(N / 2)]);  // k = N/2: w = -i, X = conj(Z)
            }
        }
    }

    // EPI == 4: c2r producing n = 2N reals per contiguous line from N+1 complex.  Merge
    //   Z[k] = (A + B) + D,  Z[N-k] = conj((A + B) - D),  A = X[k], B = conj X[N-k],
    //   D = i conj(w_n^k) (A - B),
    // written re/im-swapped into the line (the backward transform = swap, forward FFT, swap).
    // Reference: hc2cb* codelets driven by ct_hc2c_apply_dif_dft fftw/fftw_api.c:5591-5603.
    static B2_D void c2r_row_enter(TileCtx &x, const PassParams &p, const cpx_t *csrc) {
        const int b2_nthreads = x.nthreads;
        (void)b2_nthreads;
        const int C = p.C;
        B2_FOR_THREADS {
            int c, t;
            map(tid, C, p.lgC, c, t);
            if (c < x.nlines) {
                cpx_t *line = x.sm + (size_t)c * LS;
                const cpx_t *X = csrc + (x.sbase + (x.l0 + c) * p.ls[0]) / 2;
                constexpr int HQ = E / 2;
                cpx_t A[HQ], B[HQ], w[HQ];
#pragma unroll
                for (int q = 0; q < HQ; ++q) {
                    const int k = t + q * T;
                    A[q] = b2_ldcg(X + k);
                    B[q] = cconj(b2_ldcg(X + (N - k)));
                    w[q] = p.twr[k];
                }
#pragma unroll
                for (int q = 0; q < HQ; ++q) {
                    const int k = t + q * T;
                    const cpx_t sum = cadd(A[q], B[q]), dif = csub(A[q], B[q]);
                    const cpx_t Dd = cmul_pi(cmul(cconj(w[q]), dif));
                    line[phys(k)] = cswap(cadd(sum, Dd));
                    if (k != 0) line[phys(N - k)] = cswap(cconj(csub(sum, Dd)));
                }
                if (t == 0) {  // k = N/2 pairs with itself: w = -i, D = -(A - conj A)
                    const cpx_t a = b2_ldcg(X + N / 2);
                    line[phys(N / 2)] = cswap(cpx_t{2 * a.x, -2 * a.y});
                }
            }
        }
        B2_SYNC();
    }

    // EPI == 5, second pass of the two-pass c2r.  The line's input is two adjacent rows of N/2+1
    // complex, Ua = U[.][2c] and Ub = U[.][2c+1] (Hermitian in k1); the length-N complex input whose
    // backward transform is the real pair (x[..2c], x[..2c+1]) is
    //   V[k1] = Ua[k1] + i Ub[k1],   V[N-k1] = conj Ua[k1] + i conj Ub[k1],
    // written re/im-swapped into the line (backward = swap, forward FFT, swap).
    static B2_D void pair_merge_enter(TileCtx &x, const PassParams &p, const cpx_t *csrc) {
        const int b2_nthreads = x.nthreads;
        (void)b2_nthreads;
        const int C = p.C;
        B2_FOR_THREADS {
            int c, t;
            map(tid, C, p.lgC, c, t);
            if (c < x.nlines) {
                cpx_t *line = x.sm + (size_t)c * LS;
                const cpx_t *Ua = csrc + (x.sbase + (x.l0 + c) * p.ls[0]) / 2;
                const cpx_t *Ub = Ua + (N / 2 + 1);
                constexpr int HQ = E / 2;
                cpx_t a[HQ], b[HQ];
#pragma unroll
                for (int q = 0; q < HQ; ++q) {
                    a[q] = b2_ldcg(Ua + t + q * T);
                    b[q] = b2_ldcg(Ub + t + q * T);
                }
#pragma unroll
                for (int q = 0; q < HQ; ++q) {
                    const int k = t + q * T;
                    const cpx_t ib = cmul_pi(b[q]);                 // i Ub
                    const cpx_t icb = cmul_pi(cconj(b[q]));         // i conj Ub
                    line[phys(k)] = cswap(cadd(a[q], ib));
                    if (k != 0) line[phys(N - k)] = cswap(cadd(cconj(a[q]), icb));
                }
                if (t == 0) {
                    const cpx_t an = b2_ldcg(Ua + N / 2), bn = b2_ldcg(Ub + N / 2);
                    line[phys(N / 2)] = cswap(cadd(an, cmul_pi(bn)));
                }
            }
        }
        B2_SYNC();
    }

    // second pass of the two-pass real transform: output index kk of the length-n spectrum goes to
    // kk when kk <= n/2, else its conjugate goes to n - kk; rows k1 = 0 and n1/2 only keep the former
    static B2_D bool fold_target(const PassParams &p, long long k1, long long k2, int fc, long long &idx, bool &conj) {
        const long long kk = p.fold_dim > 0 ? k1 + (long long)p.fold_cmul * fc + p.fold_pmul * k2
                                            : k1 + (long long)p.fold_n1 * k2;
        if (2 * kk <= p.fold_n) {
            idx = kk;
            conj = false;
            return true;
        }
        if (k1 == 0 || 2 * k1 == p.fold_n1) return false;
        idx = p.fold_n - kk;
        conj = true;
        return true;
    }

    // registers -> global memory, v[q] holds output index t + q*T of line c (EPI 0, 1, 6, 7, 8)
    static B2_D void store_regs(const TileCtx &x, const PassParams &p, cpx_t *cdst, cpx_t *v, int c, int t) {
        const long long base = (x.dbase + (x.l0 + c) * p.ld[0]) / 2;
        const long long ds = p.d_axis / 2;
        if (EPI == 1) {
            // w_Nbig^(i2*k), k = t + q*T: base * step^q, both from the two-level table
            const unsigned long long i2 = (p.epi_dim == 0) ? (unsigned long long)(x.l0 + c)
                                                           : (unsigned long long)x.i2base;
            const unsigned long long mask = (1ull << p.epi_shift) - 1;
            const unsigned long long e0 = i2 * (unsigned long long)t, e1 = i2 * (unsigned long long)T;
            const cpx_t wb = cmul(p.epi_hi[e0 >> p.epi_shift], p.epi_lo[e0 & mask]);
            const cpx_t ws = cmul(p.epi_hi[e1 >> p.epi_shift], p.epi_lo[e1 & mask]);
            cpx_t wq[E];
            wq[0] = cpx_t{1, 0};
            b2_twiddle_powers<E>(ws, wq);
#pragma unroll
            for (int q = 0; q < E; ++q) v[q] = cmul(v[q], q ? cmul(wb, wq[q]) : wb);
        }
        if (EPI == 6) {  // Bluestein: spectrum of the padded line times bhat = FFT_M(conj chirp) / M
#pragma unroll
            for (int q = 0; q < E; ++q) v[q] = cmul(v[q], p.blue_bhat[t + q * T]);
        }
        if (p.swap_out) {
#pragma unroll
            for (int q = 0; q < E; ++q) v[q] = cswap(v[q]);
        }
        cpx_t *ps = cdst + (base + (long long)t * ds);
        if (EPI == 8) {
            // slab exchange fused into the store: row k goes to its owner's receive buffer
            // (peer memory over NVLink, 128-byte segments per warp row), dbase excludes the
            // scatter dim (its stride is per owner)
            const long long plane = p.scat_coord0 + x.scoord;
            const long long loff = x.dbase + (x.l0 + c) * p.ld[0];
#pragma unroll
            for (int q = 0; q < E; ++q) {
                const unsigned k = (unsigned)(t + q * T);
                const unsigned o = b2_fastdiv(k, p.scat_m, p.scat_s);
                const long long r = (long long)(k - o * (unsigned)p.scat_blk);
                cpx_t *po = reinterpret_cast<cpx_t *>(p.scat_base[o]) +
                            (plane * p.scat_pstride[o] + r * p.scat_rstride[o] + loff) / 2;
                if (p.scat_self1 > 0 && (int)o != p.scat_self1 - 1)
                    po = cdst + ((long long)x.scoord * p.scat_lplane + (long long)k * p.d_axis + loff) / 2;
                *po = v[q];
            }
        } else if (EPI == 7) {  // Bluestein, inverse FFT done: X[k] = conv[k] * chirp[k], k < n only
#pragma unroll
            for (int q = 0; q < E; ++q) {
                const int k = t + q * T;
                if (k < p.blue_n) {
                    cpx_t a = cmul(v[q], p.blue_chirp[k]);
                    if (p.blue_swap) a = cswap(a);
                    ps[q * T] = a;
                }
            }
        } else if (ds == 1) {
#pragma unroll
            for (int q = 0; q < E; ++q) ps[q * T] = v[q];
        } else {
            const long long step = (long long)T * ds;
#pragma unroll
            for (int q = 0; q < E; ++q) {
                *ps = v[q];
                ps += step;
            }
        }
    }

    static B2_D void run(const PassParams &p, const real_t *B2_RESTRICT src, real_t *B2_RESTRICT dst,
                         long long bid, char *smem, int b2_nthreads) {
        (void)b2_nthreads;
        TileCtx x;
        x.p = &p;
        x.src = src;
        x.dst = dst;
        x.sm = reinterpret_cast<cpx_t *>(smem);
        x.nthreads = B2_NTHREADS;
        B2_REGS(cpx_t, vreg, E);
#if B2_CUDA
        x.vbase = vreg;
#else
        x.vbase = vreg_store.data();
#endif
        const int C = p.C;
        x.scoord = 0;
        x.fcoord = 0;
        decode(p, (unsigned)bid, x.l0, x.nlines, x.sbase, x.dbase, x.i2base, &x.scoord, &x.fcoord);
        const cpx_t *csrc = reinterpret_cast<const cpx_t *>(src);
        cpx_t *cdst = reinterpret_cast<cpx_t *>(dst);

        // ---- load (global -> registers) + first radix pass ----
        if (EPI == 4) c2r_row_enter(x, p, csrc);
        if (EPI == 5) pair_merge_enter(x, p, csrc);
        B2_FOR_THREADS {
            cpx_t *v = B2_TREGS(x);
            int c, t;
            const bool live = map(tid, C, p.lgC, c, t);
            const bool active = live && c < x.nlines;
            // While a tile is being transformed nothing of its CTA is in flight to HBM: ask L2 now
            // for the tile that the CTA taking over this slot will load (pf_ahead tiles on), one
            // hint per 128-byte line of that tile.
            if (p.pf_ahead > 0 && (unsigned long long)bid + p.pf_ahead < (unsigned long long)p.nblocks) {
                long long nl0, nsb, ndb;
                int nnl, ni2;
                decode(p, (unsigned)bid + p.pf_ahead, nl0, nnl, nsb, ndb, ni2);
                constexpr int PER = 128 / (int)sizeof(cpx_t);  // points per 128-byte line
                const cpx_t *nb = csrc + (nsb + nl0 * p.ls[0]) / 2;
                if (COLMAP) {
                    const int spr = (nnl + PER - 1) / PER;  // 128-byte lines per point row, usually 1
                    const long long as = p.s_axis / 2;
                    for (int s = tid; s < N * spr; s += B2_NTHREADS) {
                        const int k = spr == 1 ? s : s / spr, j = s - k * spr;
                        b2_prefetch_l2(nb + (long long)k * as + j * PER);
                    }
                } else {
                    constexpr int SPL = N / PER;  // 128-byte lines per (contiguous) line of the tile
                    const long long lstep = p.ls[0] / 2;
                    for (int s = tid; s < nnl * SPL; s += B2_NTHREADS) {
                        const int c2 = s / SPL, j = s - c2 * SPL;
                        b2_prefetch_l2(nb + (long long)c2 * lstep + j * PER);
                    }
                }
            }
            if (EPI == 4 || EPI == 5) {
                // c2r: the merged input comes from shared memory (c2r_row_enter / pair_merge_enter)
                if (active) {
                    const cpx_t *line = x.sm + (size_t)c * LS;
                    const int pt = phys(t);
#pragma unroll
                    for (int q = 0; q < E; ++q) v[q] = line[phys_k<T, 0>(t, pt, q)];
                } else {
#pragma unroll
                    for (int q = 0; q < E; ++q) v[q] = cpx_t{0, 0};
                }
            } else if (active) {
                // all strides are in reals and even here (vector access was checked by the selector)
                const long long as = p.s_axis / 2;
                const cpx_t *pl = csrc + ((x.sbase + (x.l0 + c) * p.ls[0]) / 2 + (long long)t * as);
                if (MODE == MODE_COL && p.fold_n > 0) {
                    // two-pass c2r, first pass: element (k2, k1) of the full spectrum is X[kk] when
                    // kk = k1 + n1*k2 <= n/2, else conj X[n - kk] (Hermitian unfold of the stored half)
                    // (32-bit index arithmetic: the selector only takes this path for n < 2^30, unit stride)
                    const cpx_t *fb = csrc + x.sbase / 2;
                    const int fn = (int)p.fold_n, dk = p.fold_n1 * T;
                    int kk = (int)(x.l0 + c) + p.fold_n1 * t;
#pragma unroll
                    for (int q = 0; q < E; ++q) {
                        const bool far = 2 * kk > fn;
                        cpx_t a = b2_ldcg(fb + (far ? fn - kk : kk));
                        if (far) a.y = -a.y;
                        v[q] = a;
                        kk += dk;
                    }
                } else if (EPI == 6) {
                    // Bluestein, first FFT: x[j] * chirp[j] for j < n, zero padding up to M = N
#pragma unroll
                    for (int q = 0; q < E; ++q) {
                        const int j = t + q * T;
                        cpx_t a = cpx_t{0, 0};
                        if (j < p.blue_n) {
                            a = b2_ldcg(pl + q * T);
                            if (p.blue_swap) a = cswap(a);
                            a = cmul(a, p.blue_chirp[j]);
                        }
                        v[q] = a;
                    }
                } else if (as == 1) {  // contiguous lines: the offsets fold into the load instructions
#pragma unroll
                    for (int q = 0; q < E; ++q) v[q] = b2_ldcg(pl + q * T);
                } else {
                    const long long step = (long long)T * as;
#pragma unroll
                    for (int q = 0; q < E; ++q) {
                        v[q] = b2_ldcg(pl);
                        pl += step;
                    }
                }
                if (p.swap_in) {
#pragma unroll
                    for (int q = 0; q < E; ++q) v[q] = cswap(v[q]);
                }
            } else {
#pragma unroll
                for (int q = 0; q < E; ++q) v[q] = cpx_t{0, 0};
            }
            if (EPI != 4 && EPI != 5) butterflies<0>(x, v, c, t, live);
        }
        if (EPI == 4 || EPI == 5) {
            B2_SYNC();  // every thread holds its inputs before pass 0 overwrites the line
            B2_FOR_THREADS {
                cpx_t *v = B2_TREGS(x);
                int c, t;
                const bool live = map(tid, C, p.lgC, c, t);
                butterflies<0>(x, v, c, t, live);
            }
        }
        if constexpr (S::NP > 1) passes_from<1>(x);

        // ---- store (registers -> global), v[q] holds output index t + q*T ----
        if (EPI == 2) {
            rsplit_exit(x, p, cdst, bid);
        } else if (EPI == 3) {
            r2c_row_exit(x, p, cdst);
        } else if (MODE != MODE_ROW2COL) {
            B2_FOR_THREADS {
                cpx_t *v = B2_TREGS(x);
                int c, t;
                const bool live = map(tid, C, p.lgC, c, t);
                if (live && c < x.nlines) store_regs(x, p, cdst, v, c, t);
            }
        } else {
            // transposed exit: through shared memory, then neighbouring lanes walk the lines
            B2_SYNC();
            B2_FOR_THREADS {
                cpx_t *v = B2_TREGS(x);
                int c, t;
                if (map(tid, C, p.lgC, c, t)) {
                    cpx_t *line = x.sm + (size_t)c * LS;
                    const int pt = phys(t);
#pragma unroll
                    for (int q = 0; q < E; ++q) line[phys_k<T, 0>(t, pt, q)] = v[q];
                }
            }
            B2_SYNC();
            B2_FOR_THREADS {
                const int nl = x.nlines;
                const long long ds = p.d_axis / 2;
                const long long base = (x.dbase + x.l0 * p.ld[0]) / 2;
                const long long lstep = p.ld[0] / 2;
                if (nl == C && p.lgC >= 0 && B2_NTHREADS == C * T) {
                    // full tile: thread -> (line c, points k0 + q*T), no per-element division
                    const int c = tid & (C - 1), k0 = tid >> p.lgC;
                    const cpx_t *sp = x.sm + (size_t)c * LS;
                    const int pk = phys(k0);
                    cpx_t *ps = cdst + (base + (long long)c * lstep + (long long)k0 * ds);
                    const long long step = (long long)T * ds;
                    constexpr int H = (E + 1) / 2;  // two half batches: 16 values in flight would spill
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        cpx_t a[H];
#pragma unroll
                        for (int q = 0; q < H; ++q) a[q] = (h * H + q < E) ? sp[phys_k<T, 0>(k0, pk, h * H + q)] : cpx_t{0, 0};
                        if (p.swap_out) {
#pragma unroll
                            for (int q = 0; q < H; ++q) a[q] = cswap(a[q]);
                        }
                        if (p.fold_n > 0) {
                            cpx_t *fb = cdst + x.dbase / 2;
#pragma unroll
                            for (int q = 0; q < H; ++q) {
                                long long idx;
                                bool cj;
                                if (h * H + q < E && fold_target(p, x.l0 + c, k0 + (h * H + q) * T, x.fcoord, idx, cj))
                                    fb[idx * lstep] = cj ? cconj(a[q]) : a[q];
                            }
                        } else {
#pragma unroll
                            for (int q = 0; q < H; ++q) {
                                if (h * H + q < E) *ps = a[q];
                                ps += step;
                            }
                        }
                    }
                } else {
                    for (int idx = tid; idx < nl * N; idx += B2_NTHREADS) {
                        const int k = idx / nl, c = idx - k * nl;
                        cpx_t a = x.sm[(size_t)c * LS + phys(k)];
                        if (p.swap_out) a = cswap(a);
                        if (p.fold_n > 0) {
                            long long fi;
                            bool cj;
                            if (fold_target(p, x.l0 + c, k, x.fcoord, fi, cj)) cdst[x.dbase / 2 + fi * lstep] = cj ? cconj(a) : a;
                        } else
                            cdst[base + (long long)c * lstep + (long long)k * ds] = a;
                    }
                }
            }
        }
    }
};

#if B2_CUDA
template <int N, int E, int MODE, int EPI, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB)
    b2_tile_kernel(const __grid_constant__ PassParams p, const real_t *B2_RESTRICT src,
                   real_t *B2_RESTRICT dst, long long bid0) {
    extern __shared__ __align__(16) char b2_smem[];
    TileFFT<N, E, MODE, EPI>::run(p, src, dst, bid0 + blockIdx.x, b2_smem, blockDim.x);
}
#endif
