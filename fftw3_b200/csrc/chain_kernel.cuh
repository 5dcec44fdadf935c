// csrc/chain_kernel.cuh -- several dependent passes in ONE launch, with the intermediate data
// kept in L2.
//
// A chain is a run of consecutive plan steps (tile kernels and/or the Hermitian split/merge)
// whose work separates into G independent groups (a plane of a 3-D array, one transform of a
// batch): step i+1 of group g reads only what step i of group g wrote.  The grid is laid out
// as time slots; slot tau holds the tiles of step i for group tau - i*lag.  A CTA of step i > 0
// first waits (acquire-polls a counter) until every tile of step i-1 of its group has finished;
// those tiles have smaller block indices, so they were dispatched earlier and need nothing from
// later blocks: the wait cannot deadlock.  `lag` is chosen so that, when a consumer is
// dispatched, its producers left the SMs long ago -- the poll normally succeeds at once --
// while the group's data (a few MiB) is still resident in the 126 MB L2.
//
// Effect: the z and y passes of a 3-D transform, or the two four-step passes plus the
// Hermitian split of a long real transform, cost one HBM read and one HBM write instead of
// two or three of each.
//
// The reference does the same steps as separate sweeps over the array through the CPU caches
// (rank_geq2_apply fftw/fftw_api.c:4435-4445, ct_apply_dit :2079-2090,
// ct_hc2c_apply_dit_dft :5579-5590).
#pragma once
#include "realpp_kernel.cuh"
#include "tile_dispatch.cuh"

#define B2_CHAIN_MAX 3

struct ChainStep {
    int kind;         // SK_PASS (tile kernel) / SK_RPOST / SK_RPRE
    int mode, epi;    // tile kernel variant (with pp.m)
    long long tpg;    // tiles per group
    int nthreads;     // threads this step's tiles use (<= the launch's CTA size)
    const real_t *src;
    real_t *dst;
    PassParams pp;
    b2::RealPPParams rp;
};

struct ChainParams {
    int nsteps;
    int lag;                  // time slots between a producer and its consumer
    long long G;              // groups
    long long bpg;            // blocks per time slot = sum of tpg
    unsigned *counters;       // [nsteps][G], zeroed before the launch
    unsigned *errflag;        // set when a wait gave up (a bug, never expected)
    ChainStep st[B2_CHAIN_MAX];
};

B2_D void b2_chain_run_step(const ChainStep &s, long long bid, char *smem, int b2_nthreads) {
    if (s.kind == b2::SK_PASS) {
        const PassParams &pp = s.pp;
#define B2_X_RUN(N, E, MODE, EPI, C)                                                        \
    if (pp.m == N && s.mode == MODE && s.epi == (int)(EPI)) {                            \
        TileFFT<N, E, MODE, EPI>::run(pp, s.src, s.dst, bid, smem, b2_nthreads);            \
        return;                                                                             \
    }
        B2_TILE_LIST(B2_X_RUN)
#undef B2_X_RUN
    } else {
        b2_realpp_body(s.rp, s.kind == b2::SK_RPOST, s.src, s.dst, bid, b2_nthreads);
    }
}

#if B2_CUDA
__global__ void __launch_bounds__(1024, 1)
    b2_chain_kernel(const __grid_constant__ ChainParams cp, long long bid0) {
    extern __shared__ __align__(16) char b2_smem[];
    const long long bid = bid0 + blockIdx.x;
    const long long tau = bid / cp.bpg;
    long long r = bid - tau * cp.bpg;
    int i = 0;
    while (i + 1 < cp.nsteps && r >= cp.st[i].tpg) {
        r -= cp.st[i].tpg;
        ++i;
    }
    const long long g = tau - (long long)i * cp.lag;
    if (g < 0 || g >= cp.G) return;  // ramp-up / ramp-down slots
    if (i > 0) {
        if (threadIdx.x == 0) {
            const unsigned need = (unsigned)cp.st[i - 1].tpg;
            const unsigned *c = cp.counters + (size_t)(i - 1) * cp.G + g;
            unsigned v;
            unsigned long long spins = 0;
            for (;;) {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(c) : "memory");
                if (v >= need) break;
                __nanosleep(200);
                if (++spins > (1ull << 24)) {  // several seconds: give up loudly instead of hanging
                    atomicExch(cp.errflag, 1u);
                    break;
                }
            }
        }
        __syncthreads();
    }
    // surplus warps (the CTA is as wide as the widest step) retire here: barriers count only
    // threads that have not exited, and thread 0 -- the one that signals -- is always live
    if ((int)threadIdx.x >= cp.st[i].nthreads) return;
    b2_chain_run_step(cp.st[i], g * cp.st[i].tpg + r, b2_smem, cp.st[i].nthreads);
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(cp.counters + (size_t)i * cp.G + g, 1u);
    }
}
#endif

// host: run a chain (emulation: sequentially, which trivially respects the dependencies)
inline int b2_launch_chain(const ChainParams &cp, int threads, size_t smem, b2::stream_t st) {
    const long long slots = cp.G + (long long)cp.lag * (cp.nsteps - 1);
    const long long nblocks = slots * cp.bpg;
#if B2_CUDA
    static size_t smem_set[64] = {0};  // per device (function attributes are); the opt-in only ever grows
    const int dv = b2::current_device() & 63;
    if (smem > smem_set[dv]) {
        B2_CK(cudaFuncSetAttribute(b2_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set[dv] = smem;
    }
    B2_CK(cudaMemsetAsync(cp.counters, 0, sizeof(unsigned) * (size_t)cp.nsteps * cp.G, st));
    const long long maxgrid = 1ll << 30;
    for (long long b0 = 0; b0 < nblocks; b0 += maxgrid) {
        const long long nb = std::min(maxgrid, nblocks - b0);
        b2_chain_kernel<<<(unsigned)nb, threads, smem, st>>>(cp, b0);
        ++b2::launch_counter();
        B2_CK(cudaGetLastError());
    }
    return 0;
#else
    (void)st;
    std::vector<char> sm(smem + 64);
    std::vector<unsigned> done((size_t)cp.nsteps * cp.G, 0);
    for (long long bid = 0; bid < nblocks; ++bid) {
        const long long tau = bid / cp.bpg;
        long long r = bid - tau * cp.bpg;
        int i = 0;
        while (i + 1 < cp.nsteps && r >= cp.st[i].tpg) {
            r -= cp.st[i].tpg;
            ++i;
        }
        const long long g = tau - (long long)i * cp.lag;
        if (g < 0 || g >= cp.G) continue;
        if (i > 0 && done[(size_t)(i - 1) * cp.G + g] != (unsigned)cp.st[i - 1].tpg) {
            b2::set_error("chain: dependency not met in block order (step %d group %lld)", i, g);
            return -9;
        }
        b2_chain_run_step(cp.st[i], g * cp.st[i].tpg + r, sm.data(), cp.st[i].nthreads);
        ++done[(size_t)i * cp.G + g];
    }
    ++b2::launch_counter();
    return 0;
#endif
}
