// Microbenchmark: the slab exchange's copy pattern between two B200 over NVLink, one process, two devices.
// Source W[ln0][n1][n2] complex64 on GPU a; a block [ln0][b1 rows][w columns] goes to R[ln0][b1][n2] on GPU b
// (rows of w*8 bytes at a pitch of n2*8 bytes), both directions at once.  Compares the copy engines (one 3-D
// async memcpy per direction) with an SM copy kernel (16-byte loads / stores, G CTAs) for several row widths.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o peer_copy peer_copy.cu
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#define CK(x)                                                                               \
    do {                                                                                    \
        cudaError_t e_ = (x);                                                               \
        if (e_ != cudaSuccess) {                                                            \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(1);                                                                        \
        }                                                                                   \
    } while (0)

// rows of `wv` 16-byte vectors; row r of plane z: src + (z*sy + r)*spitch, dst + (z*dy + r)*dpitch  (pitches in vectors)
__global__ void copy_kernel(const int4 *__restrict__ src, int4 *__restrict__ dst, int wv, long long rows, int ry, long long sy,
                            long long dy, long long spitch, long long dpitch) {
    const long long total = rows * wv;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / wv;
        const int x = (int)(i - row * wv);
        const long long z = row / ry, r = row - z * ry;
        dst[(z * dy + r) * dpitch + x] = __ldcs(src + (z * sy + r) * spitch + x);
    }
}

int main(int argc, char **argv) {
    int nd = 0;
    CK(cudaGetDeviceCount(&nd));
    const bool local_only = nd < 2 || (argc > 1 && atoi(argv[1]) == 1);
    const long long n2 = 2048, n1 = 2048, ln0 = 128, b1 = 1024;  // W: 4 GiB per GPU, block: half the rows
    void *W[2], *R[2];
    cudaStream_t st[2];
    cudaEvent_t e0[2], e1[2];
    for (int d = 0; d < 2; ++d) {
        CK(cudaSetDevice(local_only ? 0 : d));
        if (!local_only) {
            cudaError_t e = cudaDeviceEnablePeerAccess(1 - d, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CK(e);
            cudaGetLastError();
        }
        CK(cudaMalloc(&W[d], ln0 * n1 * n2 * 8));
        CK(cudaMalloc(&R[d], ln0 * b1 * n2 * 8));
        CK(cudaMemset(W[d], 1, ln0 * n1 * n2 * 8));
        CK(cudaStreamCreateWithFlags(&st[d], cudaStreamNonBlocking));
        CK(cudaEventCreate(&e0[d]));
        CK(cudaEventCreate(&e1[d]));
    }
    printf("%s: W[%lld][%lld][%lld] c64, block of %lld rows per plane, both directions at once\n", local_only ? "LOCAL (same GPU)" : "PEER (NVLink)",
           ln0, n1, n2, b1);
    for (long long w : {128ll, 256ll, 512ll, 1024ll, 2048ll}) {
        const double bytes = (double)ln0 * b1 * w * 8;
        for (int mode = 0; mode < 5; ++mode) {  // 0: copy engine; 1..4: SM kernel with 8, 16, 32, 64 CTAs
            const int G = mode == 0 ? 0 : (4 << mode);
            const int reps = 4;
            for (int it = 0; it < reps + 1; ++it) {
                for (int d = 0; d < 2; ++d) {
                    const int src = d, dst = local_only ? d : 1 - d;
                    CK(cudaSetDevice(local_only ? 0 : d));
                    if (it == 1) CK(cudaEventRecord(e0[d], st[d]));
                    if (mode == 0) {
                        cudaMemcpy3DParms q = {};
                        q.srcPtr = make_cudaPitchedPtr(W[src], n2 * 8, n2 * 8, n1);
                        q.srcPos = make_cudaPos(0, 0, 0);
                        q.dstPtr = make_cudaPitchedPtr(R[dst], n2 * 8, n2 * 8, b1);
                        q.dstPos = make_cudaPos(0, 0, 0);
                        q.extent = make_cudaExtent(w * 8, b1, ln0);
                        q.kind = cudaMemcpyDeviceToDevice;
                        CK(cudaMemcpy3DAsync(&q, st[d]));
                    } else {
                        copy_kernel<<<G, 512, 0, st[d]>>>((const int4 *)W[src], (int4 *)R[dst], (int)(w / 2), ln0 * b1, (int)b1, n1, b1,
                                                         n2 / 2, n2 / 2);
                    }
                }
            }
            double gbs[2];
            for (int d = 0; d < 2; ++d) {
                CK(cudaSetDevice(local_only ? 0 : d));
                CK(cudaEventRecord(e1[d], st[d]));
                CK(cudaEventSynchronize(e1[d]));
                float ms;
                CK(cudaEventElapsedTime(&ms, e0[d], e1[d]));
                gbs[d] = bytes * reps / ms * 1e-6;
            }
            if (mode == 0)
                printf("rows of %5lld B: copy engine %6.0f / %6.0f GB/s |", w * 8, gbs[0], gbs[1]);
            else
                printf(" SM x%-2d %6.0f / %6.0f |", G, gbs[0], gbs[1]);
            CK(cudaGetLastError());
        }
        printf("\n");
        fflush(stdout);
    }
    return 0;
}
