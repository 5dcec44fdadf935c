// csrc/tile_dispatch.cuh -- table of instantiated tile kernels and the rule that picks one
// for a step (the "selector over kernel configurations" for the hot geometries).
#pragma once
#include "colpipe_kernel.cuh"
#include "coltma_kernel.cuh"
#include "tile_entry.h"

#ifndef B2_MINB256
#define B2_MINB256 4
#endif
template <int N, int E, int MODE, int EPI, int C>
struct TileInst {
    static constexpr int THREADS = C * (N / E);
    static constexpr int MINB = E >= 32 ? (THREADS <= 256 ? 2 : 1)
                                        : THREADS <= 128 ? 6 : THREADS <= 256 ? B2_MINB256 : THREADS <= 512 ? 2 : 1;
    static int launch(const PassParams &p, const real_t *src, real_t *dst, long long b0, long long nb,
                      int threads, size_t smem, b2::stream_t st) {
#if B2_CUDA
        // L2 prefetch distance: the tile whose CTA will take over this CTA's slot (one wave on)
        static int wave = -1;
        if (wave < 0) {
            int occ = 1;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, b2_tile_kernel<N, E, MODE, EPI, THREADS, MINB>,
                                                              threads, smem) != cudaSuccess || occ < 1) {
                cudaGetLastError();
                occ = 1;
            }
            wave = occ * b2::sm_count();
            if (const char *e = getenv("FFTW_B200_PF")) wave = atoi(e) < 0 ? 0 : atoi(e);
        }
        // Measured on B200 (profiles/r1_axis_probe.md): the hint pays for short rows and for column
        // tiles with a moderate axis stride (+5..9 %), and costs 2..13 % for long rows and for
        // strides of MiBs (a TLB miss per hint) -- so it is only switched on where it pays.
        const bool pays = (MODE == MODE_COL) ? (p.s_axis * (long long)sizeof(real_t) < (1ll << 20)) : (N <= 128);
        PassParams q = p;
        static const bool pf_forced = getenv("FFTW_B200_PF") != nullptr;  // read once: getenv per launch shows at 10 us steps
        q.pf_ahead = (pays || pf_forced) ? (unsigned)wave : 0u;
        b2_tile_kernel<N, E, MODE, EPI, THREADS, MINB><<<(unsigned)nb, threads, smem, st>>>(q, src, dst, b0);
        return 0;
#else
        (void)st;
        std::vector<char> sm(smem + 64);
        for (long long b = 0; b < nb; ++b) TileFFT<N, E, MODE, EPI>::run(p, src, dst, b0 + b, sm.data(), threads);
        return 0;
#endif
    }
    static size_t smem_for(int c) { return TileFFT<N, E, MODE, EPI>::smem_bytes(c); }
    static void init() {
#if B2_CUDA
        cudaFuncSetAttribute(b2_tile_kernel<N, E, MODE, EPI, THREADS, MINB>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)TileFFT<N, E, MODE, EPI>::smem_bytes(C));
#endif
    }
    static TileEntry entry() {
        return TileEntry{N, E, MODE, EPI, C, THREADS, TileFFT<N, E, MODE, EPI>::smem_bytes(C), &smem_for, &launch, &init, 0};
    }
};

// persistent pipelined column kernel: grid = resident CTAs, each walks a stripe of tiles
template <int N, int E, int EPI, int C>
struct TilePipeInst {
    static constexpr int THREADS = C * (N / E);
    static constexpr int MINB = THREADS <= 256 ? 4 : THREADS <= 512 ? 2 : 1;
    static int launch(const PassParams &p, const real_t *src, real_t *dst, long long b0, long long nb,
                      int threads, size_t smem, b2::stream_t st) {
#if B2_CUDA
        if (((uintptr_t)src & 15) != 0)  // 16-byte async copies need a 16-byte aligned array
            return TileInst<N, E, MODE_COL, EPI, C>::launch(p, src, dst, b0, nb, threads,
                                                             TileFFT<N, E, MODE_COL, EPI>::smem_bytes(C), st);
        static int wave = -1;
        if (wave < 0) {
            int occ = 1;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, b2_colpipe_kernel<N, E, EPI, THREADS, MINB>, threads,
                                                              smem) != cudaSuccess || occ < 1) {
                cudaGetLastError();
                occ = 1;
            }
            wave = occ * b2::sm_count();
        }
        PassParams q = p;
        q.pf_ahead = 0;
        const long long grid = nb < wave ? nb : wave;
        b2_colpipe_kernel<N, E, EPI, THREADS, MINB><<<(unsigned)grid, threads, smem, st>>>(q, src, dst, b0, b0 + nb);
        return 0;
#else
        (void)st;
        std::vector<char> sm(smem + 64);
        const long long grid = nb < 3 ? nb : 3;  // few "CTAs": every one walks a stripe of several tiles
        for (long long b = 0; b < grid; ++b) TilePipe<N, E, EPI>::run(p, src, dst, b0 + b, b0 + nb, grid, sm.data(), threads);
        return 0;
#endif
    }
    static size_t smem_for(int c) { return TilePipe<N, E, EPI>::smem_bytes(c); }
    static void init() {
#if B2_CUDA
        cudaFuncSetAttribute(b2_colpipe_kernel<N, E, EPI, THREADS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)TilePipe<N, E, EPI>::smem_bytes(C));
#endif
    }
    static TileEntry entry() {
        return TileEntry{N, E, MODE_COL, EPI, C, THREADS, TilePipe<N, E, EPI>::smem_bytes(C), &smem_for, &launch, &init, 1};
    }
};

#if defined(B2_HAVE_COLTMA)
// persistent TMA column kernel (coltma_kernel.cuh): grid = resident CTAs, each walks a stripe of tiles
template <int N, int EPI>
struct ColTmaInst {
    typedef ColTmaShape<N> H;
    static int launch(const PassParams &p, const real_t *src, real_t *dst, long long b0, long long nb,
                      int threads, size_t smem, b2::stream_t st) {
        (void)threads, (void)smem;
        CUtensorMap tm;
        if (((uintptr_t)src & 15) != 0 || !b2_coltma_make_map(&tm, p, src, H::C, N))  // TMA needs 16-byte aligned arrays / strides
            return TileInst<N, 16, MODE_COL, EPI, H::C>::launch(p, src, dst, b0, nb, H::THREADS,
                                                                 TileFFT<N, 16, MODE_COL, EPI>::smem_bytes(H::C), st);
        static int wave = -1;
        if (wave < 0) {
            int occ = 1;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, b2_coltma_kernel<N, EPI, 0>, H::THREADS, H::SMEM) != cudaSuccess || occ < 1) {
                cudaGetLastError();
                occ = 1;
            }
            wave = occ * b2::sm_count();
        }
        PassParams q = p;
        q.pf_ahead = 0;
        const long long grid = nb < wave ? nb : wave;
        if (p.swap_in)  // the selector takes this kernel only with swap_in == swap_out
            b2_coltma_kernel<N, EPI, 1><<<(unsigned)grid, H::THREADS, H::SMEM, st>>>(tm, q, dst, b0, b0 + nb);
        else
            b2_coltma_kernel<N, EPI, 0><<<(unsigned)grid, H::THREADS, H::SMEM, st>>>(tm, q, dst, b0, b0 + nb);
        return 0;
    }
    static size_t smem_for(int) { return H::SMEM; }
    static void init() {
        cudaFuncSetAttribute(b2_coltma_kernel<N, EPI, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, H::SMEM);
        cudaFuncSetAttribute(b2_coltma_kernel<N, EPI, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, H::SMEM);
    }
    static TileEntry entry() { return TileEntry{N, 16, MODE_COL, EPI, H::C, H::THREADS, (size_t)H::SMEM, &smem_for, &launch, &init, 2}; }
};
#endif

// X(N, E, MODE, EPI, C): every instantiated tile kernel, in table order.
// B2_TILE_LIST: the complex kernels (also the ones a chained launch can run);
// B2_TILE_LIST_REAL: the variants with a real-transform prologue / epilogue (EPI 2, 3, 4).
#define B2_TILE_ALLMODES(X, N, E, CROW, CCOL) \
    X(N, E, MODE_ROW, 0, CROW) X(N, E, MODE_COL, 0, CCOL) X(N, E, MODE_COL, 1, CCOL) X(N, E, MODE_ROW2COL, 0, CCOL)
#define B2_TILE_REALMODES(X, N, E, CROW, CCOL) \
    X(N, E, MODE_COL, 2, CCOL) X(N, E, MODE_ROW, 3, CROW) X(N, E, MODE_ROW, 4, CROW) X(N, E, MODE_ROW2COL, 5, CCOL) \
    X(N, E, MODE_ROW, 6, CROW) X(N, E, MODE_ROW, 7, CROW) X(N, E, MODE_COL, 8, CCOL)
// column-tile widths of the long column kernels (lines per CTA); overridable for experiments
#ifndef B2_CC512
#define B2_CC512 16
#endif
#ifndef B2_CC1024
#define B2_CC1024 16
#endif
#ifndef B2_CC2048
#define B2_CC2048 8
#endif
#ifndef B2_E512
#define B2_E512 16
#endif
#ifndef B2_E1024
#define B2_E1024 16
#endif
#if defined(FFTW_SINGLE)
#define B2_TILE_SIZES(A, X)              \
    A(X, 32, 8, 64, 16)                  \
    A(X, 64, 8, 32, 16)                  \
    A(X, 128, 8, 16, 16)                 \
    A(X, 256, 16, 16, 16)                \
    A(X, 512, B2_E512, 8, B2_CC512)      \
    A(X, 1024, B2_E1024, 4, B2_CC1024)   \
    A(X, 2048, 16, 2, B2_CC2048)
#define B2_TILE_LIST(X) B2_TILE_SIZES(B2_TILE_ALLMODES, X) X(4096, 16, MODE_ROW, 0, 1) X(8192, 16, MODE_ROW, 0, 1)
#define B2_TILE_LIST_REAL(X)               \
    B2_TILE_SIZES(B2_TILE_REALMODES, X)    \
    X(4096, 16, MODE_ROW, 3, 1) X(4096, 16, MODE_ROW, 4, 1) X(8192, 16, MODE_ROW, 3, 1) X(8192, 16, MODE_ROW, 4, 1) \
    X(4096, 16, MODE_ROW, 6, 1) X(4096, 16, MODE_ROW, 7, 1) X(8192, 16, MODE_ROW, 6, 1) X(8192, 16, MODE_ROW, 7, 1)
#else
#define B2_TILE_SIZES(A, X) \
    A(X, 32, 8, 64, 8)      \
    A(X, 64, 8, 32, 8)      \
    A(X, 128, 8, 16, 8)     \
    A(X, 256, 8, 8, 8)      \
    A(X, 512, 8, 4, 8)      \
    A(X, 1024, 8, 2, 8)
#define B2_TILE_LIST(X) B2_TILE_SIZES(B2_TILE_ALLMODES, X) X(2048, 8, MODE_ROW, 0, 1) X(4096, 8, MODE_ROW, 0, 1)
#define B2_TILE_LIST_REAL(X)               \
    B2_TILE_SIZES(B2_TILE_REALMODES, X)    \
    X(2048, 8, MODE_ROW, 3, 1) X(2048, 8, MODE_ROW, 4, 1) X(4096, 8, MODE_ROW, 3, 1) X(4096, 8, MODE_ROW, 4, 1) \
    X(2048, 8, MODE_ROW, 6, 1) X(2048, 8, MODE_ROW, 7, 1) X(4096, 8, MODE_ROW, 6, 1) X(4096, 8, MODE_ROW, 7, 1)
#endif

// P(N, E, EPI, C): the pipelined column kernels; listed after the plain ones, preferred by the selector
#define B2_PIPE_EPIS(P, N, E, C) P(N, E, 0, C) P(N, E, 1, C) P(N, E, 8, C)
#if defined(FFTW_SINGLE)
#define B2_PIPE_LIST(P) B2_PIPE_EPIS(P, 512, B2_E512, B2_CC512) B2_PIPE_EPIS(P, 1024, B2_E1024, B2_CC1024) B2_PIPE_EPIS(P, 2048, 16, B2_CC2048)
#else
#define B2_PIPE_LIST(P) B2_PIPE_EPIS(P, 512, 8, 8) B2_PIPE_EPIS(P, 1024, 8, 8)
#endif

// T(N, EPI): the TMA column kernels (fp32 CUDA build only); preferred over both older forms
#if defined(B2_HAVE_COLTMA)
#define B2_TMA_EPIS(T, N) T(N, 0) T(N, 1) T(N, 2) T(N, 8)
#define B2_TMA_LIST(T) B2_TMA_EPIS(T, 512) B2_TMA_EPIS(T, 1024) B2_TMA_EPIS(T, 2048)
#else
#define B2_TMA_LIST(T)
#endif

inline const std::vector<TileEntry> &b2_tile_table() {
#define B2_X_ENTRY(N, E, MODE, EPI, C) TileInst<N, E, MODE, EPI, C>::entry(),
#define B2_P_ENTRY(N, E, EPI, C) TilePipeInst<N, E, EPI, C>::entry(),
#define B2_T_ENTRY(N, EPI) ColTmaInst<N, EPI>::entry(),
    static const std::vector<TileEntry> tab = [] {
        std::vector<TileEntry> t = {B2_TILE_LIST(B2_X_ENTRY) B2_TILE_LIST_REAL(B2_X_ENTRY) B2_PIPE_LIST(B2_P_ENTRY)
                                        B2_TMA_LIST(B2_T_ENTRY)};
        int nmix = 0;
        const TileEntry *mix = b2::mix_entries(&nmix);  // lengths that are not powers of two (mixtile.cu)
        t.insert(t.end(), mix, mix + nmix);
        mix = b2::mix_entries_more(&nmix);
        t.insert(t.end(), mix, mix + nmix);
        mix = b2::mix_entries_cols(&nmix);
        t.insert(t.end(), mix, mix + nmix);
        return t;
    }();
#undef B2_X_ENTRY
#undef B2_P_ENTRY
#undef B2_T_ENTRY
    return tab;
}

// does a tile kernel exist for (length, mode, epilogue)?  (plan_build_real.h asks before it commits
// to the two-pass real path, which has no generic-kernel form)
inline bool b2_have_tile_kernel(int n, int mode, int epi) {
    if (getenv("FFTW_B200_GENERIC_ONLY")) return false;
    for (const TileEntry &e : b2_tile_table())
        if (e.n == n && e.mode == mode && e.epi == epi && e.pipe != 3) return true;
    return false;
}

inline bool b2_have_any_tile_kernel(int n, int mode, int epi) {
    if (getenv("FFTW_B200_GENERIC_ONLY")) return false;
    for (const TileEntry &e : b2_tile_table())
        if (e.n == n && e.mode == mode && e.epi == epi && (e.pipe != 3 || !b2::opt("FFTW_B200_NO_MIX", b2::B2_CFG_NO_MIX))) return true;
    return false;
}

inline void b2_init_tile_kernels() {
    for (const TileEntry &e : b2_tile_table()) e.init();
}

inline void b2_select_tile_kernel(b2::StepPlan &s) {
    s.fast_id = -1;
    if (getenv("FFTW_B200_GENERIC_ONLY")) return;
    PassParams &p = s.pp;
    if (s.kind != b2::SK_PASS || !p.vec_ld || !p.vec_st) return;
    const bool real_half = (p.rmode == RM_R2C_HALF || p.rmode == RM_C2R_HALF);
    if (p.rmode != RM_C2C && !real_half) return;
    if (real_half && (p.s_axis != 2 || p.d_axis != 2 || p.s_pair != 1 || p.d_pair != 1)) return;
    if ((p.s_axis | p.d_axis) & 1) return;
    if (p.s_axis < 0 || p.d_axis < 0) return;  // negative strides: generic kernel (signed offsets throughout)
    for (int d = 0; d < p.nld; ++d)
        if (((p.ls[d] | p.ld[d]) & 1) || p.ls[d] < 0 || p.ld[d] < 0) return;
    for (const b2::LDim &o : s.outer)
        if (o.s < 0 || o.d < 0) return;
    for (const b2::LDim &o : s.outer)
        if ((o.s | o.d) & 1) return;
    int mode;
    if (p.s_axis == 2 && p.d_axis == 2)
        mode = MODE_ROW;
    else if (p.s_axis != 2 && p.d_axis != 2 && p.ls[0] == 2 && p.ld[0] == (p.rsplit ? 4 : 2) && p.ln[0] > 1)
        mode = MODE_COL;
    else if (p.s_axis == 2 && p.d_axis != 2 && p.ld[0] == 2 && p.ln[0] > 1)
        mode = MODE_ROW2COL;
    else
        return;
    const int epi = p.rmode == RM_R2C_HALF ? 3 : p.rmode == RM_C2R_HALF ? 4 : p.blue_mode ? p.blue_mode : p.pmerge ? 5 : p.rsplit ? 2 : p.scat_n > 0 ? 8 : (p.epi_dim >= 0 ? 1 : 0);
    if (p.blue_mode && mode != MODE_ROW) return;
    if (real_half && mode != MODE_ROW) return;
    if ((epi == 1 || epi == 2 || epi == 8) && mode != MODE_COL) return;
    if (epi == 8 && p.epi_dim >= 0) return;
    if (p.rsplit && p.epi_dim != 0) return;
    if (p.fold_n > 0 && mode != (p.swap_in ? MODE_COL : MODE_ROW2COL)) return;  // fold on the way out / unfold on the way in
    if (p.pmerge && mode != MODE_ROW2COL) return;
    if (p.fold_n > 0 && p.swap_in && (p.ls[0] != 2 || p.fold_n >= (1ll << 30))) return;  // unfold load: unit stride, 32-bit indices
    const std::vector<TileEntry> &tab = b2_tile_table();
    // pipelined column kernels: 16-byte async copies -> every stride a multiple of 16 bytes
    bool pipe_ok = !b2::opt("FFTW_B200_NO_PIPE", b2::B2_CFG_NO_PIPE);
    if (sizeof(real_t) == 4) {
        if (p.s_axis & 3) pipe_ok = false;
        for (int d = 1; d < p.nld; ++d)
            if (p.ls[d] & 3) pipe_ok = false;
        for (const b2::LDim &o : s.outer)
            if (o.s & 3) pipe_ok = false;
    }
    if (p.fold_n > 0) pipe_ok = false;  // the Hermitian unfold load is not a plain tile copy
    // Measured on B200 (profiles/r1_colpipe_probe.md): the pipelined kernels win where one CTA fills the
    // SM (1024 / 2048 points: 0.519 -> 0.476 ms, 0.683 -> 0.645 ms per GiB pass) and lose a few per cent
    // at 512 points (two CTAs per SM already overlap) and for axis strides of hundreds of KiB
    if (!getenv("FFTW_B200_PIPE_ALL") && (p.m < 1024 || p.s_axis * (long long)sizeof(real_t) >= (256ll << 10))) pipe_ok = false;
    if (const char *e = getenv("FFTW_B200_PIPE_MIN_N"))
        if (p.m < atoi(e)) pipe_ok = false;
    // TMA column kernels (coltma_kernel.cuh): every stride a multiple of 16 bytes, plain tile loads, and the
    // direction as one flag (plan_build.h gives every pass of a backward axis swap_in == swap_out).  The
    // store scattered to REMOTE owners (fused slab exchange) stays on the older kernels unless FFTW_B200_TMA_SCATTER
    // is set: a warp of this kernel stores 64-byte row segments, which is what made remote NVLink writes slow in round 1.
    bool tma_ok = !b2::opt("FFTW_B200_NO_TMA", b2::B2_CFG_NO_TMA) && sizeof(real_t) == 4 && p.fold_n == 0 && p.swap_in == p.swap_out &&
                  p.nld <= 4 && s.outer.empty();
    if (epi == 8 && p.scat_self1 == 0 && !p.scat_local && !getenv("FFTW_B200_TMA_SCATTER")) tma_ok = false;  // (local stores are fine)
    if (p.s_axis & 3) tma_ok = false;
    for (int d = 1; d < p.nld; ++d)
        if (p.ls[d] & 3) tma_ok = false;
    // Measured on B200 (profiles/r2_col_probe.md): 512 points at a pitch of MiBs (the x axis of a 512^3 cube:
    // every row of a tile in another 2 MiB page) is the one case where the round-1 kernel is faster (0.461 against
    // 0.513 ms per GiB); 1024 / 2048 points at such pitches are equal or better with TMA
    if (p.m <= 512 && p.s_axis * (long long)sizeof(real_t) >= (1ll << 20) && !getenv("FFTW_B200_TMA_ALL")) tma_ok = false;
    if (const char *e = getenv("FFTW_B200_TMA_MIN_N"))
        if (p.m < atoi(e)) tma_ok = false;
    // mixed-radix kernels (mixtile.cu): plain complex passes only (their lengths never meet the real-transform epilogues)
    const bool mix_ok = !b2::opt("FFTW_B200_NO_MIX", b2::B2_CFG_NO_MIX) && p.fold_n == 0;
    const bool mix_alt = getenv("FFTW_B200_MIX_ALT") != nullptr;
    for (int pass = 3; pass >= 0 && s.fast_id < 0; --pass)
    for (size_t i = 0; i < tab.size(); ++i) {
        const TileEntry &e = tab[i];
        if ((pass == 3 && !mix_ok) || (pass == 2 && !tma_ok) || (pass == 1 && !pipe_ok)) break;
        if (e.n != p.m || e.mode != mode || e.epi != epi || e.pipe != pass) continue;
        if (pass == 3 && mix_alt) {  // experiments: the LAST shape listed for this length instead of the first
            size_t j = i;
            for (size_t k = i + 1; k < tab.size(); ++k)
                if (tab[k].n == e.n && tab[k].mode == e.mode && tab[k].epi == e.epi && tab[k].pipe == 3) j = k;
            if (j != i) {
                i = j - 1;
                continue;
            }
        }
        s.fast_id = (int)i;
        p.C = e.c;
        p.tiles0 = (p.ln[0] + p.C - 1) / p.C;
        long long nb = p.tiles0;
        for (int d = 1; d < p.nld; ++d) nb *= p.ln[d];
        if (nb >= (1ll << 31)) {  // the tile kernels decode a 31-bit tile index
            s.fast_id = -1;
            continue;
        }
        s.nblocks = nb;
        s.nthreads = e.threads;
        s.smem = e.smem;
        static const char *mn[] = {"row", "col", "row->col"};
        static const char *en[] = {"", "+twiddle", "+real-pair split+twiddle", "+hermitian split (r2c)", "+hermitian merge (c2r)", "+real-pair merge", "+chirp in, x bhat out (bluestein)", "+chirp out (bluestein)", "+scatter to the owners' receive buffers (slab exchange)"};
        char buf[256];
        std::string rs;
        for (int i2 = 0; i2 < p.npass; ++i2) rs += (i2 ? "x" : "") + std::to_string(p.radix[i2]);
        snprintf(buf, sizeof buf, "%s n=%d tile-kernel[%s%s%s] E=%d regs-resident, tile=%dx%d ctas=%lld threads=%d smem=%zuB",
                 real_half ? (epi == 3 ? "r2c" : "c2r") : "c2c", real_half ? p.n : p.m, mn[mode], e.pipe == 3 ? ", mixed radix" : e.pipe == 2 ? ", persistent, TMA-fed, two line groups" : e.pipe ? ", persistent, async-copy pipelined" : "", p.fold_n > 0 ? (p.swap_in ? "+hermitian unfold+twiddle" : "+hermitian fold") : en[epi], e.e, p.C, p.m, nb, e.threads, e.smem);
        s.desc = buf;
        if (e.pipe == 3 && e.sched) s.desc += std::string(" radices ") + e.sched;
        return;
    }
}

inline int b2_launch_tile_kernel(int id, const PassParams &p, const real_t *src, real_t *dst, long long b0,
                                 long long nb, int threads, size_t smem, b2::stream_t st) {
    PassParams q = p;
    b2_pp_prepare(q);
    q.pf_ahead = 0;
    return b2_tile_table()[id].launch(q, src, dst, b0, nb, threads, smem, st);
}
