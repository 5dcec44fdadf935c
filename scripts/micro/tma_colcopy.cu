// Microbenchmark: the data-movement skeleton of a strided-axis ("column") FFT pass on sm_100a.
// A tile = N points x C neighbouring lines (rows of C*8 bytes at a pitch of `inner`*8 bytes) of a
// [outer][N][inner] complex64 array.  It asks what the memory system gives, with no butterflies at all:
//   mode 0  TMA load -> smem ring -> TMA store                      (pure DMA ceiling, one thread works)
//   mode 1  TMA load -> smem -> LDS.64 to registers -> STG.64       (one buffer, next load issued as soon
//           as the tile sits in registers: the structure coltma_kernel.cuh uses)
//   mode 2  TMA load -> smem -> registers -> smem -> TMA store      (one buffer per CTA, overlap only
//           across the k CTAs of an SM)
//   mode 3  LDG.64 (column mapping) -> registers -> STG.64          (what the round-1 tile kernels do)
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tma_colcopy tma_colcopy.cu
// Results on B200: profiles/r2_tma_colcopy.md
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
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

typedef CUresult (*EncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiled encode_tiled() {
    static EncodeTiled f = nullptr;
    if (!f) {
        cudaDriverEntryPointQueryResult q;
        CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&f, cudaEnableDefault, &q));
        if (q != cudaDriverEntryPointSuccess || !f) {
            printf("cuTensorMapEncodeTiled not found\n");
            exit(1);
        }
    }
    return f;
}

// [outer][N][inner] of 8-byte elements; box = C x rows x 1
static CUtensorMap make_map(void *base, long long inner, int N, long long outer, int C, int rows) {
    CUtensorMap m;
    cuuint64_t dims[3] = {(cuuint64_t)inner, (cuuint64_t)N, (cuuint64_t)outer};
    cuuint64_t strides[2] = {(cuuint64_t)inner * 8, (cuuint64_t)inner * 8 * N};
    cuuint32_t box[3] = {(cuuint32_t)C, (cuuint32_t)rows, 1};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = encode_tiled()(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, base, dims, strides, box, es,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        printf("cuTensorMapEncodeTiled failed: %d\n", (int)r);
        exit(1);
    }
    return m;
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int n) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n));
}
__device__ __forceinline__ void mbar_expect(uint64_t *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(smem_u32(b)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load3(void *dst, const CUtensorMap *m, uint64_t *bar, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
                     smem_u32(dst)),
                 "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void tma_store3(const CUtensorMap *m, const void *src, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(m), "r"(smem_u32(src)),
                 "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int K> __device__ __forceinline__ void tma_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(K) : "memory");
}
__device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

constexpr int ROWS = 256;  // rows per TMA box (box dims are limited to 256)

// ---- KERNEL (part 2 follows) ----
template <int N, int C, int MODE, int NBUF>
__global__ void __launch_bounds__(N *C / 16, 1)
    colcopy(const __grid_constant__ CUtensorMap min, const __grid_constant__ CUtensorMap mout, const float2 *__restrict__ A,
            float2 *__restrict__ B, long long inner, long long outer) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int T = N / 16;        // threads per line
    constexpr int TILE = N * C * 8;  // bytes
    constexpr int NB = (N + ROWS - 1) / ROWS, RB = N < ROWS ? N : ROWS;
    float2 *buf = reinterpret_cast<float2 *>(smem_raw);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + (size_t)NBUF * TILE);
    const int tid = threadIdx.x;
    const long long tiles_i = inner / C, ntiles = tiles_i * outer;
    if (MODE != 3) {
        if (tid == 0) {
            for (int b = 0; b < NBUF; ++b) mbar_init(&full[b], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
    }
    auto issue_load = [&](long long tile, int b) {
        const int o = (int)(tile / tiles_i), i0 = (int)(tile % tiles_i) * C;
        mbar_expect(&full[b], TILE);
#pragma unroll
        for (int r = 0; r < NB; ++r)
            tma_load3(reinterpret_cast<unsigned char *>(buf) + (size_t)b * TILE + (size_t)r * RB * C * 8, &min, &full[b], i0, r * RB, o);
    };
    auto issue_store = [&](long long tile, int b) {
        const int o = (int)(tile / tiles_i), i0 = (int)(tile % tiles_i) * C;
#pragma unroll
        for (int r = 0; r < NB; ++r)
            tma_store3(&mout, reinterpret_cast<unsigned char *>(buf) + (size_t)b * TILE + (size_t)r * RB * C * 8, i0, r * RB, o);
        tma_commit();
    };
    const int c = tid % C, t = tid / C;
    if (MODE == 0) {
        if (tid != 0) return;
        long long it = 0, tl = blockIdx.x, tload = tl;
        int nl = 0;
        for (; nl < NBUF - 1 && tload < ntiles; ++nl, tload += gridDim.x) issue_load(tload, nl % NBUF);
        for (; tl < ntiles; tl += gridDim.x, ++it) {
            const int b = (int)(it % NBUF);
            if (NBUF == 1) {
                tma_wait_read<0>();
                issue_load(tl, 0);
            }
            mbar_wait(&full[b], (uint32_t)((it / NBUF) & 1));
            issue_store(tl, b);
            if (NBUF > 1 && tload < ntiles) {
                tma_wait_read<1>();  // the buffer of iteration it-1 is the next load's target
                issue_load(tload, nl % NBUF);
                ++nl;
                tload += gridDim.x;
            }
        }
        tma_wait_read<0>();
        return;
    }
    if (MODE == 1) {
        long long it = 0;
        if (tid == 0 && blockIdx.x < ntiles) issue_load(blockIdx.x, 0);
        for (long long tl = blockIdx.x; tl < ntiles; tl += gridDim.x, ++it) {
            mbar_wait(&full[0], (uint32_t)(it & 1));
            float2 v[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) v[q] = buf[(t + q * T) * C + c];
            __syncthreads();
            if (tid == 0 && tl + gridDim.x < ntiles) issue_load(tl + gridDim.x, 0);
            const int o = (int)(tl / tiles_i), i0 = (int)(tl % tiles_i) * C;
            float2 *d = B + ((long long)o * N + t) * inner + i0 + c;
#pragma unroll
            for (int q = 0; q < 16; ++q) d[(long long)q * T * inner] = v[q];
        }
        return;
    }
    if (MODE == 2) {
        long long it = 0;
        for (long long tl = blockIdx.x; tl < ntiles; tl += gridDim.x, ++it) {
            if (tid == 0) {
                tma_wait_read<0>();  // the previous store has read the buffer
                issue_load(tl, 0);
            }
            mbar_wait(&full[0], (uint32_t)(it & 1));
            float2 v[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) v[q] = buf[(t + q * T) * C + c];
            __syncthreads();
#pragma unroll
            for (int q = 0; q < 16; ++q) buf[(t + q * T) * C + c] = make_float2(v[q].y, v[q].x);
            fence_async();
            __syncthreads();
            if (tid == 0) issue_store(tl, 0);
        }
        if (tid == 0) tma_wait_read<0>();
        return;
    }
    if (MODE == 3) {
        for (long long tl = blockIdx.x; tl < ntiles; tl += gridDim.x) {
            const int o = (int)(tl / tiles_i), i0 = (int)(tl % tiles_i) * C;
            const float2 *s = A + ((long long)o * N + t) * inner + i0 + c;
            float2 *d = B + ((long long)o * N + t) * inner + i0 + c;
            float2 v[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) v[q] = __ldcg(s + (long long)q * T * inner);
#pragma unroll
            for (int q = 0; q < 16; ++q) d[(long long)q * T * inner] = v[q];
        }
    }
}

static int g_sms = 148;

template <int N, int C, int MODE, int NBUF> void run(long long inner, int per_sm, float2 *A, float2 *B, size_t elems) {
    const long long outer = (long long)(elems / ((size_t)N * inner));
    if (outer < 1 || inner % C) return;
    constexpr int TILE = N * C * 8;
    const size_t smem = MODE == 3 ? 0 : (size_t)NBUF * TILE + 64;
    // occupancy is steered by padding the dynamic shared memory: exactly per_sm CTAs fit
    size_t want = ((size_t)228 * 1024 / per_sm - 1536) / 128 * 128;
    if (want < smem) return;
    auto kern = colcopy<N, C, MODE, NBUF>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)want));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, N * C / 16, want));
    if (occ < 1) return;
    const int rows = N < ROWS ? N : ROWS;
    CUtensorMap mi = make_map(A, inner, N, outer, C, rows), mo = make_map(B, inner, N, outer, C, rows);
    const long long ntiles = inner / C * outer;
    const int grid = (int)std::min<long long>(ntiles, (long long)g_sms * occ);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int i = 0; i < 2; ++i) kern<<<grid, N * C / 16, want>>>(mi, mo, A, B, inner, outer);
    CK(cudaDeviceSynchronize());
    const int reps = 10;
    CK(cudaEventRecord(e0));
    for (int i = 0; i < reps; ++i) kern<<<grid, N * C / 16, want>>>(mi, mo, A, B, inner, outer);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    ms /= reps;
    const double bytes = 2.0 * N * inner * outer * 8;
    printf("N=%4d C=%2d (%3d B rows) mode=%d nbuf=%d pitch=%8lld B ctas/SM=%d threads=%4d : %.3f ms  %.2f TB/s\n", N, C, C * 8, MODE,
           NBUF, inner * 8, occ, N * C / 16, ms, bytes / ms * 1e-9);
    fflush(stdout);
}

int main() {
    cudaDeviceProp pr;
    CK(cudaGetDeviceProperties(&pr, 0));
    g_sms = pr.multiProcessorCount;
    const size_t elems = (size_t)1 << 27;  // 1 GiB of complex64
    float2 *A, *B;
    CK(cudaMalloc(&A, elems * 8));
    CK(cudaMalloc(&B, elems * 8));
    CK(cudaMemset(A, 0, elems * 8));
    CK(cudaMemset(B, 0, elems * 8));
    printf("%s, %d SMs; 1 GiB in + 1 GiB out per launch\n", pr.name, g_sms);
    for (long long inner : {1024ll, 2048ll, 262144ll}) {
        run<1024, 16, 3, 1>(inner, 1, A, B, elems);
        run<1024, 16, 1, 1>(inner, 1, A, B, elems);
        run<1024, 8, 1, 1>(inner, 2, A, B, elems);
        run<1024, 8, 2, 1>(inner, 2, A, B, elems);
        run<1024, 16, 2, 1>(inner, 1, A, B, elems);
        run<1024, 8, 0, 3>(inner, 1, A, B, elems);
        run<1024, 16, 0, 1>(inner, 1, A, B, elems);
        run<2048, 8, 3, 1>(inner, 1, A, B, elems);
        run<2048, 8, 1, 1>(inner, 1, A, B, elems);
        run<2048, 8, 2, 1>(inner, 1, A, B, elems);
        run<2048, 4, 1, 1>(inner, 2, A, B, elems);
        run<2048, 8, 0, 1>(inner, 1, A, B, elems);
        run<512, 16, 3, 1>(inner, 2, A, B, elems);
        run<512, 16, 1, 1>(inner, 2, A, B, elems);
        run<512, 16, 2, 1>(inner, 2, A, B, elems);
        run<512, 8, 2, 1>(inner, 5, A, B, elems);
        run<512, 16, 0, 3>(inner, 1, A, B, elems);
    }
    return 0;
}
