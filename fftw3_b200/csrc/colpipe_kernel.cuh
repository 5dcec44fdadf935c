// csrc/colpipe_kernel.cuh -- persistent, software-pipelined column tile kernels for the long
// strided axes (512 .. 2048 points), where a tile of C lines fills the register file and only
// one or two CTAs fit an SM (profiles/r1_axis_probe.md: 76 / 64 / 48 % of the HBM copy peak
// because the load, butterfly and store phases of that one CTA cannot overlap).
//
// Same arithmetic as TileFFT<N, E, MODE_COL, EPI> (tile_kernel.cuh); what changes is how a tile
// gets on chip.  A CTA owns a stripe of tiles (tile, tile + grid, ...).  The global loads do not go
// through registers: the tile is copied asynchronously (cp.async, 16-byte granules, L2 only) into
// the CTA's shared-memory buffer in "landing" layout [point k][line c] (a row = one 128-byte
// segment).  The same buffer then serves as the Stockham exchange buffer of the radix passes, and
// the moment its last exchange has been read -- before the last butterflies, the epilogue twiddles
// and the global stores of this tile -- the copy of the CTA's NEXT tile is issued into it.  Loads
// of tile i+1 therefore overlap the last pass and the stores of tile i, with no second buffer
// (a 1024 x 16 fp32 tile is 128 KB: two do not fit an SM).
//
// Replaces, like tile_kernel.cuh, the reference's twiddle codelets driven over strided columns
// (dftw_direct_apply fftw/fftw_api.c:2315-2324; rank>=2 column sweeps rank_geq2_apply :4435-4445).
#pragma once
#include "tile_kernel.cuh"

B2_D void b2_cp_async16(void *sdst, const void *gsrc) {
#if B2_CUDA
    const unsigned sa = (unsigned)__cvta_generic_to_shared(sdst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gsrc) : "memory");
#else
    memcpy(sdst, gsrc, 16);
#endif
}
B2_D void b2_cp_async_commit() {
#if B2_CUDA
    asm volatile("cp.async.commit_group;" ::: "memory");
#endif
}
B2_D void b2_cp_async_wait_all() {
#if B2_CUDA
    asm volatile("cp.async.wait_group 0;" ::: "memory");
#endif
}

template <int N, int E, int EPI> struct TilePipe {
    typedef TileFFT<N, E, MODE_COL, EPI> F;
    typedef TileSched<N, E> S;
    static constexpr int T = N / E;
    static constexpr int VPC = 16 / (int)sizeof(cpx_t);  // complex elements per 16-byte granule
    static B2_HD size_t smem_bytes(int C) {
        const size_t a = F::smem_bytes(C), b = (size_t)N * C * sizeof(cpx_t);
        return a > b ? a : b;
    }

    // asynchronous copy of tile `bid` into the landing layout sm[k * C + c]
    static B2_D void prefetch(const PassParams &p, const cpx_t *csrc, unsigned bid, cpx_t *sm, int b2_nthreads) {
        (void)b2_nthreads;
        long long l0, sb, db;
        int nl, i2;
        F::decode(p, bid, l0, nl, sb, db, i2);
        const int C = p.C;
        const cpx_t *base = csrc + (sb + l0 * p.ls[0]) / 2;
        const long long as = p.s_axis / 2;
        B2_FOR_THREADS {
            if (nl == C) {
                // full tile: thread -> (row k0 + i*rows_per_sweep, granule j); C/VPC granules per row
                const int lgG = p.lgC - (VPC == 2 ? 1 : 0);
                const int j = tid & ((1 << lgG) - 1), k0 = tid >> lgG, dk = B2_NTHREADS >> lgG;
                const cpx_t *g = base + (long long)k0 * as + j * VPC;
                cpx_t *d = sm + k0 * C + j * VPC;
                const long long gstep = (long long)dk * as;
                const int dstep = dk * C;
                for (int k = k0; k < N; k += dk) {
                    b2_cp_async16(d, g);
                    g += gstep;
                    d += dstep;
                }
            } else {
                // ragged last tile of a row of tiles: through registers
                for (int idx = tid; idx < N * nl; idx += B2_NTHREADS) {
                    const int k = idx / nl, c = idx - k * nl;
                    sm[k * C + c] = b2_ldcg(base + (long long)k * as + c);
                }
            }
            b2_cp_async_commit();
        }
    }

    template <int P>
    static B2_D void passes(TileCtx &x, const cpx_t *csrc, long long next, bool has_next) {
        const int b2_nthreads = x.nthreads;
        (void)b2_nthreads;
        B2_SYNC();  // everybody has written pass P-1's outputs
        B2_FOR_THREADS {
            cpx_t *v = B2_TREGS(x);
            int c, t;
            F::map(tid, x.p->C, x.p->lgC, c, t);
            const cpx_t *line = x.sm + (size_t)c * F::LS;
#pragma unroll
            for (int q = 0; q < E; ++q) v[q] = line[t + q * T];
        }
        if (P < S::NP - 1) {
            B2_SYNC();  // the next write must not overtake a neighbour's read
        } else if (has_next) {
            B2_SYNC();  // the buffer is free: start the copy of this CTA's next tile
            prefetch(*x.p, csrc, (unsigned)next, x.sm, x.nthreads);
        }
        B2_FOR_THREADS {
            cpx_t *v = B2_TREGS(x);
            int c, t;
            F::map(tid, x.p->C, x.p->lgC, c, t);
            F::template butterflies<P>(x, v, c, t, true);
        }
        if constexpr (P + 1 < S::NP) passes<P + 1>(x, csrc, next, has_next);
    }

    static B2_D void run(const PassParams &p, const real_t *B2_RESTRICT src, real_t *B2_RESTRICT dst,
                         long long first, long long end, long long stride, char *smem, int b2_nthreads) {
        (void)b2_nthreads;
        static_assert(S::NP >= 2, "pipelined column kernels are for multi-pass lengths");
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
        const cpx_t *csrc = reinterpret_cast<const cpx_t *>(src);
        cpx_t *cdst = reinterpret_cast<cpx_t *>(dst);
        if (first >= end) return;
        prefetch(p, csrc, (unsigned)first, x.sm, x.nthreads);
        for (long long tile = first; tile < end; tile += stride) {
            x.scoord = 0;
            x.fcoord = 0;
            F::decode(p, (unsigned)tile, x.l0, x.nlines, x.sbase, x.dbase, x.i2base, &x.scoord);
            b2_cp_async_wait_all();
            B2_SYNC();  // the tile has landed (every thread waited for its own copies)
            B2_FOR_THREADS {
                cpx_t *v = B2_TREGS(x);
                int c, t;
                F::map(tid, C, p.lgC, c, t);
                if (c < x.nlines) {
                    const cpx_t *col = x.sm + t * C + c;
#pragma unroll
                    for (int q = 0; q < E; ++q) v[q] = col[q * T * C];
                    if (p.swap_in) {
#pragma unroll
                        for (int q = 0; q < E; ++q) v[q] = cswap(v[q]);
                    }
                } else {
#pragma unroll
                    for (int q = 0; q < E; ++q) v[q] = cpx_t{0, 0};
                }
            }
            B2_SYNC();  // landing layout read; pass 0 rewrites the buffer in exchange layout
            B2_FOR_THREADS {
                cpx_t *v = B2_TREGS(x);
                int c, t;
                F::map(tid, C, p.lgC, c, t);
                F::template butterflies<0>(x, v, c, t, true);
            }
            const long long next = tile + stride;
            passes<1>(x, csrc, next, next < end);
            B2_FOR_THREADS {
                cpx_t *v = B2_TREGS(x);
                int c, t;
                F::map(tid, C, p.lgC, c, t);
                if (c < x.nlines) F::store_regs(x, p, cdst, v, c, t);
            }
        }
    }
};

#if B2_CUDA
template <int N, int E, int EPI, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB)
    b2_colpipe_kernel(const __grid_constant__ PassParams p, const real_t *B2_RESTRICT src,
                      real_t *B2_RESTRICT dst, long long bid0, long long end) {
    extern __shared__ __align__(16) char b2_smem[];
    TilePipe<N, E, EPI>::run(p, src, dst, bid0 + blockIdx.x, end, gridDim.x, b2_smem, blockDim.x);
}
#endif
