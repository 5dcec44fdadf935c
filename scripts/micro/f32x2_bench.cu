// Microbenchmark: issue rate of scalar FADD/FFMA against the packed FADD2/FFMA2 forms on sm_100a.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o f32x2_bench f32x2_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
template <int MODE> __global__ void k(float2 *out, float2 seed) {
    float2 a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = make_float2(seed.x + i + threadIdx.x, seed.y - i);
    const float2 w = make_float2(seed.y, seed.x);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) { a[i].x += w.x; a[i].y += w.y; }
            if (MODE == 1) a[i] = __fadd2_rn(a[i], w);
            if (MODE == 2) { a[i].x = fmaf(a[i].x, w.x, w.y); a[i].y = fmaf(a[i].y, w.y, w.x); }
            if (MODE == 3) a[i] = __ffma2_rn(a[i], w, w);
        }
    }
    float2 s = make_float2(0, 0);
#pragma unroll
    for (int i = 0; i < 8; ++i) { s.x += a[i].x; s.y += a[i].y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE> void run(const char *name, float2 *out) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148 * 8, 256>>>(out, make_float2(1.0001f, 0.9999f));
    cudaEventRecord(e0);
    k<MODE><<<148 * 8, 256>>>(out, make_float2(1.0001f, 0.9999f));
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double ops = 148.0 * 8 * 256 * ITERS * 16;  // scalar fp32 operations
    printf("%-8s %.3f ms  %.2f T scalar-op/s\n", name, ms, ops / ms * 1e-9);
}
int main() {
    float2 *out; cudaMalloc(&out, 148 * 8 * 256 * sizeof(float2));
    run<0>("FADD", out); run<1>("FADD2", out); run<2>("FFMA", out); run<3>("FFMA2", out);
    return 0;
}
