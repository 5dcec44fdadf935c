// csrc/coltma_kernel.cuh -- strided-axis ("column") passes of 512 .. 2048 points, round-2 design.
//
// Why: the round-1 column kernels (tile_kernel.cuh MODE_COL, colpipe_kernel.cuh) were issue-bound, not
// memory-bound: 72 thread-instructions per point, only 34 of them floating point -- the rest address
// arithmetic, per-tile index decoding done by every warp, 64-bit shared-memory exchanges and twiddle
// product trees -- and one CTA per SM whose load / butterfly / store phases could not overlap
// (profiles/r1_ncu_full_colpipe1024.txt).  The data-movement skeleton of this kernel (TMA load ->
// registers -> STG, nothing else) runs at 6.1 TB/s (profiles/r2_tma_colcopy.md), so the structure is:
//
//   * persistent CTA walking tiles of C lines x N points (a row of the tile = C*8 bytes);
//   * the tile is brought on chip by TMA (cp.async.bulk.tensor, 256-row boxes, hardware swizzle so
//     that the register fill is bank-conflict free) into a landing buffer L.  The moment all threads
//     hold their 16 points in registers the NEXT tile's load is issued into L: it overlaps the whole
//     butterfly work and the stores of the current tile;
//   * the Stockham exchanges between radix passes go through a separate buffer X that holds HALF the
//     lines; the CTA is split into two line-groups that take turns on X (named barriers), so one
//     group's exchange overlaps the other group's butterflies;
//   * exchanges are 128-bit (two adjacent points per LDS / STS) wherever the radix allows, twiddles
//     come from per-pass shared-memory tables [j mod Ns][k] (one accurate table value per factor, no
//     product tree), every shared-memory address is base + immediate;
//   * one thread decodes the next tile's coordinates; everybody else reads them from shared memory.
//
// Arithmetic: the same radix schedule and butterflies as TileFFT (tile_kernel.cuh).  These kernels
// replace the reference's twiddle codelets driven over strided columns (dftw_direct_apply
// fftw/fftw_api.c:2315-2324, q1fv_8.c:542; rank>=2 column sweeps rank_geq2_apply :4435-4445).
// CUDA fp32 build only: fp64 and the emulation build keep the round-1 kernels for these geometries.
#pragma once
#include "tile_kernel.cuh"

#if B2_CUDA && defined(FFTW_SINGLE)
#define B2_HAVE_COLTMA 1
#include <cuda.h>

namespace b2tma {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int n) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n));
}
__device__ __forceinline__ void mbar_expect(uint64_t *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nB2W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra B2D;\nbra B2W;\nB2D:\n}" ::"r"(
            smem_u32(b)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load5(uint32_t dst, const CUtensorMap *m, uint64_t *bar, int c0, int c1, int c2, int c3,
                                          int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(dst),
        "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}
// named barriers: sync = wait for COUNT threads, arrive = count me in and go on
template <int ID, int COUNT> __device__ __forceinline__ void bar_sync() {
    asm volatile("bar.sync %0, %1;" ::"n"(ID), "n"(COUNT) : "memory");
}
template <int ID, int COUNT> __device__ __forceinline__ void bar_arrive() {
    asm volatile("bar.arrive %0, %1;" ::"n"(ID), "n"(COUNT) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// the driver entry point, through the runtime (no link against libcuda)
inline EncodeTiledFn encode_tiled() {
    static EncodeTiledFn f = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        cudaDriverEntryPointQueryResult q;
        void *fp = nullptr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            f = (EncodeTiledFn)fp;
        else
            cudaGetLastError();
    }
    return f;
}

}  // namespace b2tma

struct alignas(16) ColTmaMeta {  // what a thread needs to know about a tile, written by the decoding thread
    long long dbase;             // dst offset (reals) of the tile's line 0, point 0, without the axis term
    int l0, nlines, i2base, scoord;
};

struct alignas(16) ColTmaPair {  // two adjacent fp32 complex points: one LDS.128 / STS.128
    cpx_t a, b;
};

template <int N> struct ColTmaShape {
    static constexpr int E = 16;
    static constexpr int C = (int)(128 / sizeof(cpx_t)) / (N >= 2048 ? 2 : 1);  // lines per tile: 128-byte rows, 64 at 2048 points
    static constexpr int T = N / E;                     // threads per line
    static constexpr int THREADS = C * T;
    static constexpr int CG = C / 2, GT = THREADS / 2;  // lines / threads per group
    static constexpr int ROWB = C * (int)sizeof(cpx_t);
    static constexpr int LBYTES = N * ROWB;
    // X line stride (complex): the 8 lanes x 16 bytes of one quarter-warp must tile 128 bytes of banks
    static constexpr int XLS = N + (CG == 8 ? 2 : 4);
    static constexpr int XBYTES = CG * XLS * (int)sizeof(cpx_t);
    typedef TileSched<N, E> S;
    static constexpr int twsize(int P) { return P <= 0 ? 0 : S::ns(P) * (S::radix(P) + 2); }  // complex entries of pass P's table
    static constexpr int twoff(int P) { return P <= 1 ? 0 : twoff(P - 1) + twsize(P - 1); }
    static constexpr int TWBYTES = (twoff(S::NP - 1) + twsize(S::NP - 1)) * (int)sizeof(cpx_t);
    static constexpr int OFF_X = LBYTES;
    static constexpr int OFF_TW = OFF_X + ((XBYTES + 127) / 128) * 128;
    static constexpr int OFF_META = OFF_TW + ((TWBYTES + 127) / 128) * 128;
    static constexpr int SMEM = OFF_META + 2 * (int)sizeof(ColTmaMeta) + 64 + 1024;  // + slack to align L to 1024 bytes
    static constexpr int NBOX = (N + 255) / 256, BOXROWS = N < 256 ? N : 256;
};

// ---- PART 2: the kernel body ----
// EPI: 0 plain, 1 four-step twiddle on the way out, 2 split of a pair of real columns + twiddle (first pass of the
//      two-pass r2c), 8 rows scattered to their owners (slab exchange)
// BWD: backward transform = swap re/im on the way in and out (compile time: the swap is pure register renaming)
template <int N, int EPI, int BWD> struct ColTma {
    typedef ColTmaShape<N> H;
    typedef typename H::S S;
    typedef ColTmaPair Pair;
    static constexpr int E = H::E, C = H::C, T = H::T, THREADS = H::THREADS, CG = H::CG, GT = H::GT;
    static_assert(sizeof(real_t) == 4, "fp32 build only");
    static_assert(S::NP == 3, "three radix passes");
    static_assert(S::radix(S::NP - 1) == 8, "the last pass is a radix-8 pass with pair ownership");

    // ---- tile index -> coordinates (one thread) ----
    static __device__ void decode(const PassParams &p, unsigned bid, ColTmaMeta &m, int *coord) {
        unsigned rest = b2_fastdiv(bid, p.dv_m[0], p.dv_s[0]);
        const unsigned t0 = bid - rest * (unsigned)p.tiles0;
        m.l0 = (int)(t0 * (unsigned)C);
        const long long left = p.ln[0] - (long long)m.l0;
        m.nlines = (int)(left < C ? left : C);
        long long db = 0;
        m.i2base = 0;
        m.scoord = 0;
        coord[0] = m.l0;
        coord[1] = coord[2] = coord[3] = 0;
        for (int d = 1; d < p.nld; ++d) {
            const unsigned qd = b2_fastdiv(rest, p.dv_m[d], p.dv_s[d]);
            const unsigned cc = rest - qd * (unsigned)p.ln[d];
            rest = qd;
            db += (long long)cc * p.ld[d];
            coord[d] = (int)cc;
            if (d == p.epi_dim) m.i2base = (int)cc;
            if (EPI == 8 && d == p.scat_dim) m.scoord = (int)cc;
        }
        m.dbase = db;
    }

    static __device__ __forceinline__ void issue_tile(const CUtensorMap *tm, const PassParams &p, unsigned bid, ColTmaMeta *meta,
                                                      uint32_t Ladr, uint64_t *full) {
        int coord[4];
        ColTmaMeta m;
        decode(p, bid, m, coord);
        *meta = m;
        b2tma::mbar_expect(full, (uint32_t)H::LBYTES);  // (release: orders the meta store before the waiters' acquire)
#pragma unroll
        for (int r = 0; r < H::NBOX; ++r)
            b2tma::tma_load5(Ladr + (uint32_t)(r * H::BOXROWS * H::ROWB), tm, full, coord[0], r * H::BOXROWS, coord[1], coord[2],
                             coord[3]);
    }

    // w_N^e from the two-level table of the four-step epilogue
    static __device__ __forceinline__ cpx_t epi_root(const PassParams &p, unsigned long long e) {
        return cmul(p.epi_hi[e >> p.epi_shift], p.epi_lo[e & ((1ull << p.epi_shift) - 1)]);
    }

    // One exchange through X by this thread's group: `wr` stores a pass's outputs, `rd` loads the next pass's
    // inputs.  The groups alternate on X, A (group 0) first.  Barrier ids: 1 / 2 inside group A / B,
    // 3 = "A has read, X is B's", 4 = "B has read, X is A's".
    template <class WR, class RD> static __device__ __forceinline__ void exchange(int g, bool first_use, WR wr, RD rd) {
        if (g == 0) {
            if (!first_use) b2tma::bar_sync<4, THREADS>();
        } else
            b2tma::bar_sync<3, THREADS>();
        wr();
        if (g == 0)
            b2tma::bar_sync<1, GT>();
        else
            b2tma::bar_sync<2, GT>();
        rd();
        if (g == 0)
            b2tma::bar_arrive<3, THREADS>();
        else
            b2tma::bar_arrive<4, THREADS>();
    }

    static __device__ void run(const CUtensorMap *tm, const PassParams &p, real_t *B2_RESTRICT dst, long long first, long long end,
                               long long stride, unsigned char *smem_raw) {
        // L must sit on a 1024-byte boundary (swizzle atom)
        const uint32_t raw = b2tma::smem_u32(smem_raw);
        const uint32_t pad = (1024u - (raw & 1023u)) & 1023u;
        unsigned char *sm = smem_raw + pad;
        const uint32_t Ladr = raw + pad;
        cpx_t *X = reinterpret_cast<cpx_t *>(sm + H::OFF_X);
        cpx_t *TW = reinterpret_cast<cpx_t *>(sm + H::OFF_TW);
        ColTmaMeta *meta = reinterpret_cast<ColTmaMeta *>(sm + H::OFF_META);
        uint64_t *full = reinterpret_cast<uint64_t *>(sm + H::OFF_META + 2 * sizeof(ColTmaMeta));

        const int tid = threadIdx.x;
        const int g = tid / GT, lt = tid - g * GT, lane = tid & 31, wg = lt >> 5;
        // thread -> (line cl of the group, slot t): a warp holds all CG lines of the group and 32/CG slots whose
        // rows fall into different swizzle phases; any bijection works for the FFT itself
        int cl, t;
        if (CG == 8) {
            cl = lane & 7;
            t = ((lane >> 4) & 1) | ((wg & 1) << 1) | (((lane >> 3) & 1) << 2) | ((wg >> 1) << 3);
        } else {
            cl = lane & 3;
            t = ((lane >> 2) & 1) | (((lane >> 4) & 1) << 1) | (((lane >> 3) & 1) << 2) | (wg << 3);
        }
        const int c = g * CG + cl;  // line of the tile

        // ---- one-time setup: barrier, twiddle tables TW[off(P) + jm*(R+2) + k] = w_(NS*R)^(jm*k) ----
        if (tid == 0) {
            b2tma::mbar_init(full, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        fill_table<1>(TW, p.tw, tid);
        fill_table<2>(TW, p.tw, tid);
        __syncthreads();
        if (first >= end) return;
        if (tid == 0) issue_tile(tm, p, (unsigned)first, &meta[0], Ladr, full);

        // per-thread constant addresses.  Landing buffer: element (k, c) sits in row k at the 16-byte chunk
        // (c/2) ^ swizzle(k); k = t + q*T with T a multiple of 8, so the swizzle term depends on t only
        const int chunk = (c >> 1) ^ (CG == 8 ? (t & 7) : ((t >> 1) & 3));
        const unsigned char *Lthr = sm + t * H::ROWB + chunk * 16 + (c & 1) * 8;
        cpx_t *Xl = X + cl * H::XLS;  // this thread's line in X
        cpx_t *cdst = reinterpret_cast<cpx_t *>(dst);
        const long long ds = p.d_axis / 2, lds0 = p.ld[0] / 2;

        bool first_use = true;
        unsigned it = 0;
        for (long long tile = first; tile < end; tile += stride, ++it) {
            b2tma::mbar_wait(full, it & 1);
            const ColTmaMeta m = meta[it & 1];
            cpx_t v[E];
#pragma unroll
            for (int q = 0; q < E; ++q) v[q] = *reinterpret_cast<const cpx_t *>(Lthr + q * T * H::ROWB);
            __syncthreads();  // L is free: the next tile's load overlaps everything below
            if (tid == 0 && tile + stride < end) issue_tile(tm, p, (unsigned)(tile + stride), &meta[(it + 1) & 1], Ladr, full);
            if (BWD) {
#pragma unroll
                for (int q = 0; q < E; ++q) v[q] = cswap(v[q]);
            }

            // ---- pass 0: strided ownership (element t + q*T), Ns = 1, outputs o = j*R + k ----
            {
                constexpr int R = S::radix(0), NB = E / R;
#pragma unroll
                for (int mb = 0; mb < NB; ++mb) {
                    cpx_t a[R];
#pragma unroll
                    for (int k = 0; k < R; ++k) a[k] = v[mb + k * NB];
                    Bf<R, real_t>::run(a);
#pragma unroll
                    for (int k = 0; k < R; ++k) v[mb + k * NB] = a[k];
                }
                exchange(
                    g, first_use,
                    [&] {
#pragma unroll
                        for (int mb = 0; mb < NB; ++mb) {
                            Pair *o = reinterpret_cast<Pair *>(Xl + (t + mb * T) * R);
#pragma unroll
                            for (int k = 0; k < R; k += 2) o[k / 2] = Pair{v[mb + k * NB], v[mb + (k + 1) * NB]};
                        }
                    },
                    [&] { load_next<1>(Xl, t, v); });
                first_use = false;
            }
            // ---- middle pass ----
            twiddle_butterflies<1>(TW, t, v);
            exchange(
                g, false, [&] { store_pass<1>(Xl, t, v); }, [&] { load_next<2>(Xl, t, v); });
            // ---- last pass: pair ownership, v[2k+e] = output 2t + e + k*2T ----
            twiddle_butterflies<2>(TW, t, v);
            if (EPI == 2) {
                // real-pair split: X_a, X_b of output k1 need Z[k1] AND Z[N-k1], which another thread of the line holds:
                // one more exchange through X (natural order), then every thread pairs its own k1 < N/2 with the mirror
                cpx_t zb[E / 2];
                exchange(
                    g, false,
                    [&] {
                        Pair *o = reinterpret_cast<Pair *>(Xl + 2 * t);
#pragma unroll
                        for (int k = 0; k < E / 2; ++k) o[k * T] = Pair{v[2 * k], v[2 * k + 1]};
                    },
                    [&] {
#pragma unroll
                        for (int q = 0; q < E / 2; ++q) {
                            const int k1 = 2 * t + (q & 1) + (q >> 1) * 2 * T;  // q = 2k + e, k < E/4: k1 < N/2
                            zb[q] = Xl[(N - k1) & (N - 1)];
                        }
                    });
                if (c < m.nlines) rsplit_store(p, m, cdst, ds, c, t, v, zb);
            } else if (c < m.nlines)
                store_tile(p, m, cdst, ds, lds0, c, t, v);
        }
        if (g == 0 && !first_use) b2tma::bar_sync<4, THREADS>();  // pairs group B's last arrive
    }

// ---- PART 3: pass helpers, store, kernel entry, tensor map ----
    // twiddle table of pass P (P = 1, 2): one accurate table value per (j mod Ns, k), rows padded to R + 2
    template <int P> static __device__ __forceinline__ void fill_table(cpx_t *TW, const cpx_t *tw, int tid) {
        constexpr int R = S::radix(P), NS = S::ns(P), OFF = H::twoff(P);
        for (int i = tid; i < NS * R; i += THREADS) {
            const int jm = i / R, k = i - jm * R;
            TW[OFF + jm * (R + 2) + k] = tw[(jm * k) * (N / (NS * R))];
        }
    }

    // inputs of pass P from X.  Radix 8 (two butterflies per thread): pair ownership, elements
    // 2t + e + k*2T as v[2k+e], 128-bit loads.  Radix 16: strided ownership t + q*T, 64-bit loads.
    template <int P> static __device__ __forceinline__ void load_next(const cpx_t *Xl, int t, cpx_t *v) {
        constexpr int R = S::radix(P);
        if (R == 8) {
            const Pair *s = reinterpret_cast<const Pair *>(Xl + 2 * t);
#pragma unroll
            for (int q = 0; q < E / 2; ++q) {
                const Pair x = s[q * T];
                v[2 * q] = x.a;
                v[2 * q + 1] = x.b;
            }
        } else {
#pragma unroll
            for (int q = 0; q < E; ++q) v[q] = Xl[t + q * T];
        }
    }

    // twiddles from the pass's table, then the butterflies, all in registers
    template <int P> static __device__ __forceinline__ void twiddle_butterflies(const cpx_t *TW, int t, cpx_t *v) {
        constexpr int R = S::radix(P), NS = S::ns(P);
        static_assert(R == 8 || R == E, "radix 8 (two butterflies per thread) or 16 (one)");
        const cpx_t *tab = TW + H::twoff(P);
        if (R == 8) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int jm = (2 * t + e) & (NS - 1);
                const Pair *w = reinterpret_cast<const Pair *>(tab + jm * (R + 2));
                cpx_t a[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) a[k] = v[2 * k + e];
#pragma unroll
                for (int k2 = 0; k2 < 4; ++k2) {
                    const Pair ww = w[k2];
                    if (k2) a[2 * k2] = cmul(a[2 * k2], ww.a);
                    a[2 * k2 + 1] = cmul(a[2 * k2 + 1], ww.b);
                }
                Bf<8, real_t>::run(a);
#pragma unroll
                for (int k = 0; k < 8; ++k) v[2 * k + e] = a[k];
            }
        } else {
            const int jm = t & (NS - 1);
            const Pair *w = reinterpret_cast<const Pair *>(tab + jm * (R + 2));
#pragma unroll
            for (int k2 = 0; k2 < E / 2; ++k2) {
                const Pair ww = w[k2];
                if (k2) v[2 * k2] = cmul(v[2 * k2], ww.a);
                v[2 * k2 + 1] = cmul(v[2 * k2 + 1], ww.b);
            }
            Bf<E, real_t>::run(v);
        }
    }

    // outputs of the middle pass P into X: o = (j - jm)*R + jm + k*Ns
    template <int P> static __device__ __forceinline__ void store_pass(cpx_t *Xl, int t, const cpx_t *v) {
        constexpr int R = S::radix(P), NS = S::ns(P);
        if (R == 8) {  // butterflies j = 2t, 2t+1 are neighbours in the output too: one 128-bit store per k
            const int j = 2 * t, jm = j & (NS - 1);
            Pair *o = reinterpret_cast<Pair *>(Xl + (j - jm) * R + jm);
#pragma unroll
            for (int k = 0; k < R; ++k) o[k * (NS / 2)] = Pair{v[2 * k], v[2 * k + 1]};
        } else {
            const int jm = t & (NS - 1);
            cpx_t *o = Xl + (t - jm) * R + jm;
#pragma unroll
            for (int k = 0; k < R; ++k) o[k * NS] = v[k];
        }
    }

    // EPI 2, first pass of the two-pass real transform (same arithmetic as TileFFT::rsplit_exit, tile_kernel.cuh).  The line
    // is the complex FFT Z of a pair of real columns x_a + i x_b; per k1 <= N/2
    //   X_a = (Z[k1] + conj Z[N-k1]) / 2,   X_b = -i (Z[k1] - conj Z[N-k1]) / 2,
    // times the four-step twiddles w_n^(k1*j2), w_n^(k1*(j2+1)) of their columns j2 = 2*line, j2 + 1, stored as one 16-byte
    // vector at row k1 of the [N/2+1][n2] intermediate.  v[2k+e] = Z[k1], zb[2k+e] = Z[N-k1] for k1 = 2t + e + k*2T, k < E/4;
    // thread t = 0 also stores row N/2 (it holds Z[N/2] = v[E/2]).  Reference: the hc2c step of rdft2-ct
    // (ct_hc2c_direct_apply fftw/fftw_api.c:5831-5845) applied to column pairs.
    static __device__ __forceinline__ void rsplit_store(const PassParams &p, const ColTmaMeta &m, cpx_t *cdst, long long ds, int c,
                                                        int t, const cpx_t *v, const cpx_t *zb) {
        const unsigned long long j2 = 2ull * (unsigned long long)(m.l0 + c);  // even column of the pair
        Pair *row0 = reinterpret_cast<Pair *>(cdst + m.dbase / 2) + (m.l0 + c);  // (two complex per line in the intermediate)
        const long long dsp = ds / 2;                                            // row pitch in Pairs
        // wa(k1) = w^(j2*k1) = A_e * S^k,  w1(k1) = w^k1 = B_e * U^k   with k1 = 2t + e + k*2T
        const cpx_t S = epi_root(p, j2 * (unsigned long long)(2 * T)), U = epi_root(p, (unsigned long long)(2 * T));
        cpx_t Sk[E / 4], Uk[E / 4];
        Sk[0] = Uk[0] = cpx_t{1, 0};
        b2_twiddle_powers<E / 4>(S, Sk);
        b2_twiddle_powers<E / 4>(U, Uk);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const cpx_t A = epi_root(p, j2 * (unsigned long long)(2 * t + e)), B = epi_root(p, (unsigned long long)(2 * t + e));
#pragma unroll
            for (int k = 0; k < E / 4; ++k) {
                const int k1 = 2 * t + e + k * 2 * T;
                const cpx_t za = v[2 * k + e], zc = cconj(zb[2 * k + e]);
                const cpx_t xa = cscale(cadd(za, zc), (real_t)0.5);
                const cpx_t xb = cmul_mi(cscale(csub(za, zc), (real_t)0.5));
                const cpx_t wa = k ? cmul(A, Sk[k]) : A;
                const cpx_t w1 = k ? cmul(B, Uk[k]) : B;
                row0[(long long)k1 * dsp] = Pair{cmul(xa, wa), cmul(xb, cmul(wa, w1))};
            }
        }
        if (t == 0) {  // k1 = N/2 pairs with itself
            const int k1 = N / 2;
            const cpx_t za = v[E / 2];
            const cpx_t wa = epi_root(p, j2 * (unsigned long long)k1);
            row0[(long long)k1 * dsp] = Pair{cmul(cpx_t{za.x, 0}, wa), cmul(cpx_t{za.y, 0}, cmul(wa, epi_root(p, (unsigned long long)k1)))};
        }
    }

    // epilogue + global stores; v[2q+e] holds output index 2t + e + q*2T of line c
    static __device__ __forceinline__ void store_tile(const PassParams &p, const ColTmaMeta &m, cpx_t *cdst, long long ds,
                                                      long long lds0, int c, int t, cpx_t *v) {
        if (EPI == 1) {
            const unsigned long long i2 = (p.epi_dim == 0) ? (unsigned long long)(m.l0 + c) : (unsigned long long)m.i2base;
            const cpx_t ws = epi_root(p, i2 * (unsigned long long)(2 * T));
            cpx_t wq[E / 2];
            wq[0] = cpx_t{1, 0};
            b2_twiddle_powers<E / 2>(ws, wq);
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const cpx_t wb = epi_root(p, i2 * (unsigned long long)(2 * t + e));
#pragma unroll
                for (int q = 0; q < E / 2; ++q) v[2 * q + e] = cmul(v[2 * q + e], q ? cmul(wb, wq[q]) : wb);
            }
        }
        if (BWD) {
#pragma unroll
            for (int q = 0; q < E; ++q) v[q] = cswap(v[q]);
        }
        const long long loff = m.dbase / 2 + (long long)(m.l0 + c) * lds0;
        if (EPI == 8) {
            const long long plane = p.scat_coord0 + m.scoord;
#pragma unroll
            for (int q = 0; q < E; ++q) {
                const unsigned k = (unsigned)(2 * t + (q & 1) + (q >> 1) * 2 * T);
                const unsigned o = b2_fastdiv(k, p.scat_m, p.scat_s);
                const long long r = (long long)(k - o * (unsigned)p.scat_blk);
                cpx_t *po = reinterpret_cast<cpx_t *>(p.scat_base[o]) + (plane * p.scat_pstride[o] + r * p.scat_rstride[o]) / 2 + loff;
                if (p.scat_self1 > 0 && (int)o != p.scat_self1 - 1)
                    po = cdst + ((long long)m.scoord * p.scat_lplane + (long long)k * p.d_axis) / 2 + loff;
                *po = v[q];
            }
        } else {
            cpx_t *ps = cdst + loff + (long long)(2 * t) * ds;
            const long long step = (long long)(2 * T) * ds;
#pragma unroll
            for (int q = 0; q < E / 2; ++q) {
                ps[0] = v[2 * q];
                ps[ds] = v[2 * q + 1];
                ps += step;
            }
        }
    }
};

template <int N, int EPI, int BWD>
__global__ void __launch_bounds__(ColTmaShape<N>::THREADS, ColTmaShape<N>::THREADS <= 512 ? 2 : 1)
    b2_coltma_kernel(const __grid_constant__ CUtensorMap tm, const __grid_constant__ PassParams p, real_t *B2_RESTRICT dst,
                     long long bid0, long long end) {
    extern __shared__ __align__(1024) unsigned char b2_tma_smem[];
    ColTma<N, EPI, BWD>::run(&tm, p, dst, bid0 + blockIdx.x, end, gridDim.x, b2_tma_smem);
}

// host side: the tensor map of a pass's source array.  dims: [tile dim][axis][line dims 1..3], 8-byte elements
// (one fp32 complex point), so the tile dim must have unit stride (the selector's MODE_COL condition).
inline bool b2_coltma_make_map(CUtensorMap *tm, const PassParams &p, const real_t *src, int C, int N) {
    b2tma::EncodeTiledFn enc = b2tma::encode_tiled();
    if (!enc || p.ls[0] != 2 || p.nld > 4) return false;
    cuuint64_t dims[5] = {(cuuint64_t)p.ln[0], (cuuint64_t)N, 1, 1, 1};
    cuuint64_t strides[4] = {(cuuint64_t)p.s_axis * sizeof(real_t), 0, 0, 0};
    for (int d = 1; d < 4; ++d) {
        if (d < p.nld) {
            dims[1 + d] = (cuuint64_t)p.ln[d];
            strides[d] = (cuuint64_t)p.ls[d] * sizeof(real_t);
        } else
            strides[d] = strides[d - 1] * dims[d];  // a dim of extent 1: any legal stride
    }
    for (int d = 0; d < 4; ++d)
        if (strides[d] == 0 || (strides[d] & 15) || strides[d] >= (1ull << 40)) return false;
    cuuint32_t box[5] = {(cuuint32_t)C, (cuuint32_t)(N < 256 ? N : 256), 1, 1, 1};
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    const CUtensorMapSwizzle sw = C * sizeof(cpx_t) == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 5, const_cast<real_t *>(src), dims, strides, box, es,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}
#endif  // B2_CUDA && FFTW_SINGLE
