// csrc/b2_common.h -- shared definitions for the B200 FFT engine kernels.
//
// The kernels are written once and compiled twice:
//   * by nvcc for sm_100a (the product: fftw3_b200/lib/{fp32,fp64}/libfftw3.so);
//   * by g++ with -DB2_EMULATE into a TEST-ONLY library (tests/_emu/) in which every CTA is
//     executed by a host loop over its threads, region by region between barriers.  The
//     emulation exists so that the selector and every index computation can be checked
//     against the oracle in this GPU-less container; it is never built into the product
//     library, which fails loudly when no CUDA device is present.
//
// Kernel bodies are therefore written as a sequence of B2_FOR_THREADS { ... } regions
// separated by B2_SYNC(); per-thread state that lives across regions is declared with
// B2_REGS / re-bound with B2_BIND.
#pragma once
#include <stddef.h>
#include <stdint.h>

#if defined(FFTW_SINGLE)
typedef float real_t;
#define B2_PREC_NAME "fp32"
#define b2 b2_f32  // both precision builds may be loaded into one process: keep their C++ symbols apart
#else
typedef double real_t;
#define B2_PREC_NAME "fp64"
#define b2 b2_f64
#endif

#if defined(__CUDACC__) && !defined(B2_EMULATE)
#define B2_CUDA 1
#define B2_HD __host__ __device__ __forceinline__
#define B2_D __device__ __forceinline__
#define B2_SYNC() __syncthreads()
// one "iteration": the hardware thread itself
#define B2_FOR_THREADS for (int tid = threadIdx.x, b2_once_ = 1; b2_once_; b2_once_ = 0)
#define B2_NTHREADS ((int)blockDim.x)
#define B2_REGS(type, name, count) type name[count]
#define B2_BIND(type, name, count)
#define B2_RESTRICT __restrict__
#else
#define B2_CUDA 0
#define B2_HD inline
#define B2_D inline
#define B2_SYNC() ((void)0)
#define B2_FOR_THREADS for (int tid = 0; tid < b2_nthreads; ++tid)
#define B2_NTHREADS (b2_nthreads)
#include <vector>
#define B2_REGS(type, name, count) std::vector<type> name##_store((size_t)b2_nthreads * (count))
#define B2_BIND(type, name, count) type *name = &name##_store[(size_t)tid * (count)]
#define B2_RESTRICT __restrict__
#endif

template <typename T> struct alignas(2 * sizeof(T)) cpx {
    T x, y;
};
typedef cpx<real_t> cpx_t;

template <typename T> B2_HD cpx<T> cadd(cpx<T> a, cpx<T> b) { return cpx<T>{a.x + b.x, a.y + b.y}; }
template <typename T> B2_HD cpx<T> csub(cpx<T> a, cpx<T> b) { return cpx<T>{a.x - b.x, a.y - b.y}; }
template <typename T> B2_HD cpx<T> cmul(cpx<T> a, cpx<T> b) {
    return cpx<T>{a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x};
}
// Packed fp32 arithmetic (Blackwell FADD2 / FFMA2: two fp32 operations per issue slot), opt-in with
// -DB2_F32X2=1.  Tried in round 1 because the tile kernels are issue-bound wherever only one or two
// CTAs fit an SM and half of their FP instructions are complex additions.  Result (static SASS of
// b2_tile_kernel<4096,16,row>): 348 FADD -> 96 FADD2 + 96 FFMA2, but +67 MOV and +18 FMUL to build
// aligned register pairs, i.e. 1744 -> 1664 instructions (-4.6 %), and 48..116 bytes of spills in every
// 64-register kernel.  Not worth it; left off.
#ifndef B2_F32X2
#define B2_F32X2 0
#endif
#if B2_CUDA && defined(FFTW_SINGLE) && B2_F32X2 && defined(__CUDA_ARCH__)
template <> B2_HD cpx<float> cadd<float>(cpx<float> a, cpx<float> b) {
    const float2 r = __fadd2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y));
    return cpx<float>{r.x, r.y};
}
template <> B2_HD cpx<float> csub<float>(cpx<float> a, cpx<float> b) {
    const float2 r = __ffma2_rn(make_float2(b.x, b.y), make_float2(-1.0f, -1.0f), make_float2(a.x, a.y));
    return cpx<float>{r.x, r.y};
}
#endif
template <typename T> B2_HD cpx<T> cconj(cpx<T> a) { return cpx<T>{a.x, -a.y}; }
template <typename T> B2_HD cpx<T> cswap(cpx<T> a) { return cpx<T>{a.y, a.x}; }
// a * (-i)
template <typename T> B2_HD cpx<T> cmul_mi(cpx<T> a) { return cpx<T>{a.y, -a.x}; }
// a * (+i)
template <typename T> B2_HD cpx<T> cmul_pi(cpx<T> a) { return cpx<T>{-a.y, a.x}; }
template <typename T> B2_HD cpx<T> cscale(cpx<T> a, T s) { return cpx<T>{a.x * s, a.y * s}; }
