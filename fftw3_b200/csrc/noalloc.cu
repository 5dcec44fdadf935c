// csrc/noalloc.cu -- b2fft_c2c_1d_static_scratch: a batched 1-D c2c entry point that allocates NOTHING.
//
// No counterpart in the reference; added for callers that cannot
// call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a
// variant of the engine's two-pass ("four-step", n = n1 * n2) long 1-D transform -- the plan path keeps its
// intermediate array in a cudaMalloc'ed scratch buffer and its twiddles in cudaMalloc'ed tables
// (plan_build.h add_c2c_axis, twiddle.h) -- with both in ONE file-scope static __device__ array:
//
//     b2_na_scratch = [ work area: B2FFT_NOALLOC_MAX_POINTS complex | w_n table: B2FFT_NOALLOC_MAX_N complex ]
//
// (8.5 MiB in the fp32 library, 17 MiB in the fp64 one; every GPU that loads the library holds it until
// exit).  The entry point calls no cudaMalloc / cudaFree / allocating library routine, builds no plan and
// touches no cache of the engine; sizes beyond the maxima are rejected with an error.
//
// Every call shares the array.  Nothing in it has to start at zero: a call first writes the w_n table
// (kernel 1), then the whole part of the work area it will read (kernel 2), then reads it (kernel 3), all on
// the caller's stream -- so the previous call's contents are never observed.  Calls that OVERLAP on the GPU
// (different streams, or different host threads without stream ordering) race on the array: serialise them.
//
// Host code must not hand the array's NAME to a kernel (in host code the name is a host-side stand-in, not a
// device address): the address is taken once per device with cudaGetSymbolAddress, with that device current,
// and passed to the kernels as an ordinary pointer argument.
#include <mutex>

#include "b200fft.h"
#include "device_rt.h"

using namespace b2;

#if B2_CUDA
static __device__ __align__(256) cpx_t b2_na_scratch[B2FFT_NOALLOC_MAX_POINTS + B2FFT_NOALLOC_MAX_N];

#define B2_NA_CB 4  // lines per CTA

// w[j] = exp(-2 pi i j / n), j < n: sincospi in double (n is a power of two, so 2j/n is exact), rounded once
__global__ void b2_na_twiddle_kernel(cpx_t *w, int n) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    double s, c;
    sincospi(-2.0 * (double)j / (double)n, &s, &c);
    w[j] = cpx_t{(real_t)c, (real_t)s};
}

// One pass of the four-step transform: CTA = B2_NA_CB neighbouring lines of m points, radix-2 Stockham in
// shared memory (ping-pong), twiddles w_(2Ns)^k = w_n^(k n / 2Ns) from the table.
//   point j of line l is read from  src[b*n + l*ls_s + j*ps_s]  and output k goes to  dst[b*n + l*ls_d + k*ps_d],
//   times w_n^(l*k) when `epi` (first pass).  swap_in / swap_out: backward transform = swap re/im.
__global__ void b2_na_pass_kernel(const cpx_t *__restrict__ src, cpx_t *__restrict__ dst, const cpx_t *__restrict__ w, int n,
                                  int m, long long ls_s, long long ps_s, long long ls_d, long long ps_d, int epi, int swap_in,
                                  int swap_out) {
    extern __shared__ __align__(16) unsigned char b2_na_smem[];
    cpx_t *bufa = reinterpret_cast<cpx_t *>(b2_na_smem), *bufb = bufa + B2_NA_CB * m;
    const int half = m / 2, tid = threadIdx.x;
    const int c = tid % B2_NA_CB, j = tid / B2_NA_CB;  // line of the tile, butterfly index < m/2
    const long long l = (long long)blockIdx.x * B2_NA_CB + c;
    const long long base = (long long)blockIdx.y * n;
    const cpx_t *s = src + base + l * ls_s;
    for (int q = 0; q < 2; ++q) {
        cpx_t a = s[(long long)(j + q * half) * ps_s];
        if (swap_in) a = cswap(a);
        bufa[c * m + j + q * half] = a;
    }
    __syncthreads();
    cpx_t *from = bufa, *to = bufb;
    for (int Ns = 1; Ns < m; Ns *= 2) {
        const int k = j & (Ns - 1);
        const cpx_t a = from[c * m + j];
        const cpx_t b = cmul(from[c * m + j + half], w[(long long)k * (n / (2 * Ns))]);
        const int o = ((j - k) << 1) + k;
        to[c * m + o] = cadd(a, b);
        to[c * m + o + Ns] = csub(a, b);
        __syncthreads();
        cpx_t *t = from;
        from = to;
        to = t;
    }
    cpx_t *d = dst + base + l * ls_d;
    for (int q = 0; q < 2; ++q) {
        const int k = j + q * half;
        cpx_t a = from[c * m + k];
        if (epi) a = cmul(a, w[l * k]);  // l * k < n2 * n1 = n
        if (swap_out) a = cswap(a);
        d[(long long)k * ps_d] = a;
    }
}
#endif

extern "C" int b2fft_c2c_1d_static_scratch(ptrdiff_t n, ptrdiff_t howmany, const void *in, void *out, int sign,
                                           void *cuda_stream) {
#if B2_CUDA
    if (n < 64 || n > B2FFT_NOALLOC_MAX_N || (n & (n - 1)) != 0) {
        set_error("b2fft_c2c_1d_static_scratch: n = %td must be a power of two in [64, %d]", n, (int)B2FFT_NOALLOC_MAX_N);
        return -1;
    }
    if (howmany < 1 || howmany > 65535 || n * howmany > (ptrdiff_t)B2FFT_NOALLOC_MAX_POINTS) {
        set_error("b2fft_c2c_1d_static_scratch: n * howmany = %td exceeds the static scratch array (%d complex points)",
                  n * howmany, (int)B2FFT_NOALLOC_MAX_POINTS);
        return -2;
    }
    if (!in || !out || sign == 0) {
        set_error("b2fft_c2c_1d_static_scratch: bad arguments");
        return -3;
    }
    // the array's device address: once per device, with that device current (the caller's current device)
    static std::mutex mu;
    static cpx_t *addr[64] = {nullptr};
    int dev = -1;
    B2_CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_error("b2fft_c2c_1d_static_scratch: device ordinal %d out of range", dev);
        return -4;
    }
    cpx_t *base;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!addr[dev]) {
            cpx_t *p = nullptr;
            B2_CK(cudaGetSymbolAddress((void **)&p, b2_na_scratch));
            addr[dev] = p;
        }
        base = addr[dev];
    }
    cpx_t *work = base, *w = base + B2FFT_NOALLOC_MAX_POINTS;
    int lg = 0;
    while (((ptrdiff_t)1 << lg) < n) ++lg;
    const int n1 = 1 << ((lg + 1) / 2), n2 = (int)(n / n1);  // 8 <= n2 <= n1 <= 256
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int bwd = sign > 0;
    b2_na_twiddle_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(w, (int)n);
    // pass A: n1-point FFTs down the n2 interleaved sub-sequences, times w_n^(i2*k1):  in[j1*n2 + i2] -> work[k1*n2 + i2]
    b2_na_pass_kernel<<<dim3((unsigned)(n2 / B2_NA_CB), (unsigned)howmany), B2_NA_CB * (n1 / 2),
                        2 * B2_NA_CB * n1 * sizeof(cpx_t), st>>>((const cpx_t *)in, work, w, (int)n, n1, 1, n2, 1, n2, 1, bwd, 0);
    // pass B: n2-point FFTs along the rows k1, stored transposed:  work[k1*n2 + j2] -> out[k1 + n1*k2]
    b2_na_pass_kernel<<<dim3((unsigned)(n1 / B2_NA_CB), (unsigned)howmany), B2_NA_CB * (n2 / 2),
                        2 * B2_NA_CB * n2 * sizeof(cpx_t), st>>>(work, (cpx_t *)out, w, (int)n, n2, n2, 1, 1, n1, 0, 0, bwd);
    launch_counter() += 3;
    B2_CK(cudaGetLastError());
    return 0;
#else
    (void)n, (void)howmany, (void)in, (void)out, (void)sign, (void)cuda_stream;
    set_error("b2fft_c2c_1d_static_scratch: not part of the emulation build (it exists for CUDA callers that cannot allocate)");
    return -100;
#endif
}
