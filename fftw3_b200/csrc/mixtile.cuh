// csrc/mixtile.cuh -- tile kernels for the common lengths that are NOT powers of two: kernel, instance wrapper and list macros
// (the lists themselves are in mixtile.cu, mixtile2.cu and mixtile3.cu, separate translation units so that they compile side by
// side with engine.cu instead of after it).
//
// Same kernel body as the power-of-two tile kernels (tile_kernel.cuh TileFFT: a thread keeps E points of a line in
// registers across all radix passes, one exchange through padded shared memory between passes, global memory touched
// once on the way in and once on the way out), instantiated with a compile-time MIXED-RADIX schedule (MixSched):
// n = 1000 runs as 10 x 10 x 10 with E = 10, n = 1536 as 24 x 8 x 8 with E = 24, n = 6561 as 9 x 9 x 9 x 9 with E = 27.
// The radix-6/10/12/14/15/20/21/24/28/30 butterflies are prime-factor compositions of the 2/3/4/5/7/8 ones, radix 9 is 3 x 3
// (butterfly.cuh).  Before this file such lengths ran on the generic kernel (generic_kernel.cuh: run-time radices,
// every pass through two shared-memory buffers; 16-23 % of HBM bandwidth, profiles/r1_nonpow2_probe.log).
//
// Reference: these kernels stand where the reference composes its generated codelets n1_{3,5,6,7,9,10,12,15,20,25}
// and t1_* for such lengths (fftw/dft_scalar/codelets/codlist.c:57-109; drivers direct_apply fftw/fftw_api.c:3183-3187,
// dftw_direct_apply :2315-2324).
//
// Row form for every listed length; r2c / c2r row forms where E is even; column / column+twiddle / row->column forms (what the
// other axes of an N-D transform and the two passes of a long 1-D transform need) for the lengths that have a column tile width.
#pragma once
#include "tile_entry.h"
#include "tile_kernel.cuh"

#if B2_CUDA
template <class S, int N, int E, int MODE, int EPI, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB)
    b2_mix_kernel(const __grid_constant__ PassParams p, const real_t *B2_RESTRICT src, real_t *B2_RESTRICT dst,
                  long long bid0) {
    extern __shared__ __align__(16) char b2_smem[];
    TileFFT<N, E, MODE, EPI, S>::run(p, src, dst, bid0 + blockIdx.x, b2_smem, blockDim.x);
}
#endif

template <class S, int N, int E, int MODE, int EPI, int C>
struct MixInst {
    typedef TileFFT<N, E, MODE, EPI, S> F;
    static constexpr int THREADS = C * (N / E);
    // registers a thread should get: its E points plus a butterfly's temporaries and addressing
    static constexpr int NEED = (2 * E * (int)(sizeof(real_t) / 4) + (E >= 40 ? 112 : sizeof(real_t) == 4 ? 40 : 56) + 7) / 8 * 8;
    static constexpr int TW = (THREADS + 31) / 32 * 32;
    static constexpr int MINB0 = 65536 / (TW * NEED);
    static constexpr int MINB = MINB0 < 1 ? 1 : MINB0 > 8 ? 8 : MINB0;
    static_assert(THREADS <= 1024, "tile too wide");
    static_assert(C * (size_t)F::xls_max(S::NP - 2) * sizeof(cpx_t) <= 200 * 1024, "tile does not fit in shared memory");

    static int launch(const PassParams &p, const real_t *src, real_t *dst, long long b0, long long nb, int threads,
                      size_t smem, b2::stream_t st) {
#if B2_CUDA
        // same L2 prefetch policy as the power-of-two kernels (tile_dispatch.cuh TileInst::launch)
        static int wave = -1;
        if (wave < 0) {
            int occ = 1;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, b2_mix_kernel<S, N, E, MODE, EPI, THREADS, MINB>, threads,
                                                              smem) != cudaSuccess || occ < 1) {
                cudaGetLastError();
                occ = 1;
            }
            wave = occ * b2::sm_count();
        }
        const bool pays = (MODE == MODE_COL) ? (p.s_axis * (long long)sizeof(real_t) < (1ll << 20)) : (N <= 128);
        PassParams q = p;
        q.pf_ahead = pays ? (unsigned)wave : 0u;
        b2_mix_kernel<S, N, E, MODE, EPI, THREADS, MINB><<<(unsigned)nb, threads, smem, st>>>(q, src, dst, b0);
        return 0;
#else
        (void)st;
        std::vector<char> sm(smem + 64);
        for (long long b = 0; b < nb; ++b) F::run(p, src, dst, b0 + b, sm.data(), threads);
        return 0;
#endif
    }
    static size_t smem_for(int c) { return F::smem_bytes(c); }
    static void init() {
#if B2_CUDA
        cudaFuncSetAttribute(b2_mix_kernel<S, N, E, MODE, EPI, THREADS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)F::smem_bytes(C));
#endif
    }
    static TileEntry entry(const char *sched) {
        return TileEntry{N, E, MODE, EPI, C, THREADS, F::smem_bytes(C), &smem_for, &launch, &init, 3, sched};
    }
};

// ROW(N, E, CROW, radices...): batched contiguous lines only.  ALL(N, E, CROW, CCOL, radices...): also the column forms.
// CROW is any count that brings the CTA near 256 threads; CCOL is a power of two (lines side by side: CCOL complex
// per row segment) with CCOL * N / E <= 1024 threads and the tile in shared memory.
#define B2_MIX_ROW(N, E, CROW, ...) MixInst<MixSched<N, E, __VA_ARGS__>, N, E, MODE_ROW, 0, CROW>::entry(#__VA_ARGS__),
#define B2_MIX_ALL(N, E, CROW, CCOL, ...)                                                  \
    MixInst<MixSched<N, E, __VA_ARGS__>, N, E, MODE_ROW, 0, CROW>::entry(#__VA_ARGS__),     \
    MixInst<MixSched<N, E, __VA_ARGS__>, N, E, MODE_COL, 0, CCOL>::entry(#__VA_ARGS__),     \
    MixInst<MixSched<N, E, __VA_ARGS__>, N, E, MODE_COL, 1, CCOL>::entry(#__VA_ARGS__),     \
    MixInst<MixSched<N, E, __VA_ARGS__>, N, E, MODE_ROW2COL, 0, CCOL>::entry(#__VA_ARGS__),

// r2c / c2r rows of 2N reals: the N-point row kernel with the Hermitian split on the way out (EPI 3) / merge on the way in
// (EPI 4) -- tile_kernel.cuh r2c_row_exit / c2r_row_enter pair k with N-k inside the line's shared memory; E must be even
#define B2_MIX_REAL(N, E, CROW, ...)                                                       \
    MixInst<MixSched<N, E, __VA_ARGS__>, N, E, MODE_ROW, 3, CROW>::entry(#__VA_ARGS__),     \
    MixInst<MixSched<N, E, __VA_ARGS__>, N, E, MODE_ROW, 4, CROW>::entry(#__VA_ARGS__),
#define B2_MIX_ONE(N, E, MODE, EPI, C, ...) MixInst<MixSched<N, E, __VA_ARGS__>, N, E, MODE, EPI, C>::entry(#__VA_ARGS__),
// the three column forms alone (a length may be listed with several shapes: the selector takes the FIRST row of the
// table that fits, FFTW_B200_MIX_ALT=1 the last)
#define B2_MIX_COLS(N, E, CCOL, ...)                                                       \
    MixInst<MixSched<N, E, __VA_ARGS__>, N, E, MODE_COL, 0, CCOL>::entry(#__VA_ARGS__),     \
    MixInst<MixSched<N, E, __VA_ARGS__>, N, E, MODE_COL, 1, CCOL>::entry(#__VA_ARGS__),     \
    MixInst<MixSched<N, E, __VA_ARGS__>, N, E, MODE_ROW2COL, 0, CCOL>::entry(#__VA_ARGS__),

#if defined(FFTW_SINGLE)
#define B2_CC(c32, c64) c32
#define B2_MIX_ALL32(N, E, CROW, CCOL, ...) B2_MIX_ALL(N, E, CROW, CCOL, __VA_ARGS__)
#else
#define B2_CC(c32, c64) c64
#define B2_MIX_ALL32(N, E, CROW, CCOL, ...) B2_MIX_ROW(N, E, CROW, __VA_ARGS__)  // beyond the fp64 column cap (selector.h caps())
#endif

