// csrc/device_rt.h -- the only place that talks to the CUDA runtime (or, in the TEST-ONLY
// -DB2_EMULATE build, to malloc/memcpy).  Everything above it (selector, plans, C-ABI) is
// identical in both builds, which is what lets the CPU tests exercise the selector and the
// kernels' index arithmetic without a GPU.  The product build has no CPU execution path:
// with no CUDA device every plan is NULL and b2fft_last_error() says why.
#pragma once
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#if defined(__linux__)
#include <sys/syscall.h>
#include <unistd.h>
#endif

#include <atomic>
#include <mutex>
#include <string>

#include "b2_common.h"

namespace b2 {

inline std::string &last_error() {
    static thread_local std::string e;
    return e;
}
inline void set_error(const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    last_error() = buf;
    if (getenv("FFTW_B200_VERBOSE")) fprintf(stderr, "[b200fft] %s\n", buf);
}

// Experiment switches that FFTW_MEASURE is allowed to toggle per candidate plan: opt(name) answers like getenv(name)
// unless the calling thread is building a candidate with that switch forced on or off (b2fft_plan_create_tuned).
enum { B2_CFG_NO_TMA = 1, B2_CFG_NO_PIPE = 2, B2_CFG_NO_PREFER3 = 4, B2_CFG_THREE_PASS_REAL = 8, B2_CFG_NO_TWO_PASS_REAL = 16,
       B2_CFG_TWO_PASS_C2R = 32, B2_CFG_NO_MIX = 64, B2_CFG_ALL = 127 };
inline int &cfg_override() {
    static thread_local int v = -1;  // -1: the environment decides; >= 0: bit mask of switches that are ON (all others OFF)
    return v;
}
inline const char *opt(const char *name, int bit) {
    const int c = cfg_override();
    if (c >= 0) return (c & bit) ? "1" : nullptr;
    return getenv(name);
}

inline unsigned long long &launch_counter() {
    static unsigned long long n = 0;
    return n;
}

#if B2_CUDA
typedef cudaStream_t stream_t;
#define B2_CK(expr)                                                                      \
    do {                                                                                 \
        cudaError_t e_ = (expr);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            b2::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, \
                          __LINE__);                                                     \
            return (int)e_;                                                              \
        }                                                                                \
    } while (0)

inline int &device_ordinal() {
    static int d = -1;
    return d;
}
// Every extern "C" entry point opens one of these first: whatever device the engine makes current inside the call,
// the CALLER's current device is what it was when the call returns (a host application on cuda:1 must not find
// itself on cuda:0 after planning an FFT).
struct DeviceScope {
    int prev = -1;
    DeviceScope() {
        if (cudaGetDevice(&prev) != cudaSuccess) {
            cudaGetLastError();
            prev = -1;
        }
    }
    ~DeviceScope() {
        int cur = -1;
        if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
    }
    DeviceScope(const DeviceScope &) = delete;
    DeviceScope &operator=(const DeviceScope &) = delete;
};
// 0 when a usable device is selected (and made current for the rest of the enclosing DeviceScope)
inline int ensure_device() {
    static std::atomic<int> state{0};  // 0 unknown, 1 ok, -1 none
    static std::mutex init_mu;
    if (state.load(std::memory_order_acquire) == 1) {
        cudaSetDevice(device_ordinal());
        return 0;
    }
    std::lock_guard<std::mutex> lk(init_mu);
    if (state.load() == 1) {
        cudaSetDevice(device_ordinal());
        return 0;
    }
    if (state.load() == -1) return -1;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        set_error("no CUDA device available (%s): the B200 engine has no CPU fallback",
                  e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        state.store(-1);
        return -1;
    }
    int d = device_ordinal();
    if (d < 0) {
        const char *s = getenv("FFTW_CUDA_DEVICE");
        if (!s) s = getenv("LOCAL_RANK");
        d = s ? atoi(s) : 0;
        if (d < 0 || d >= n) d = d % n;
    }
    device_ordinal() = d;
    if (cudaSetDevice(d) != cudaSuccess) {
        set_error("cudaSetDevice(%d) failed", d);
        state.store(-1);
        return -1;
    }
    state.store(1, std::memory_order_release);
    return 0;
}
// a plan runs on the device it was created on, whatever the engine's device is by now (b2fft_set_device)
inline int ensure_device_of(int dev) {
    if (ensure_device() != 0) return -1;
    if (dev >= 0 && dev != device_ordinal() && cudaSetDevice(dev) != cudaSuccess) {
        set_error("cudaSetDevice(%d) failed: %s", dev, cudaGetErrorString(cudaGetLastError()));
        return -1;
    }
    return 0;
}
inline int current_device() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess) {
        cudaGetLastError();
        d = 0;
    }
    return d;
}
inline void *dev_malloc(size_t bytes) {
    void *p = nullptr;
    if (bytes == 0) bytes = 16;
    if (cudaMalloc(&p, bytes) != cudaSuccess) {
        set_error("cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    return p;
}
inline void dev_free(void *p) {
    if (p && cudaFree(p) != cudaSuccess) cudaGetLastError();  // never leave a stale error behind
}
inline int dev_upload(void *dst, const void *src, size_t bytes) {
    B2_CK(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
    return 0;
}
// NUMA node of the current device (sysfs: /sys/bus/pci/devices/<domain:bus:dev.fn>/numa_node), or -1
inline int device_numa_node() {
    char id[32] = {0};
    if (cudaDeviceGetPCIBusId(id, sizeof id, current_device()) != cudaSuccess) {
        cudaGetLastError();
        return -1;
    }
    for (char *c = id; *c; ++c)
        if (*c >= 'A' && *c <= 'Z') *c = (char)(*c - 'A' + 'a');
    char path[96];
    snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/numa_node", id);
    FILE *f = fopen(path, "r");
    if (!f) return -1;
    int node = -1;
    if (fscanf(f, "%d", &node) != 1) node = -1;
    fclose(f);
    return node;
}
// Pinned host memory for fftw_malloc.  With one process per GPU on a two-socket box (8 ranks: GPUs 0-3 on node 0, 4-7 on
// node 1) nothing ties a rank's allocations to its GPU's socket, and host-array executes of 8 ranks then push 4 GiB each
// per step across the socket interconnect (round 1: 45.7 ms per rank at N = 1, 259.8 ms at N = 8).  The pages are
// requested on the GPU's own NUMA node (set_mempolicy(MPOL_PREFERRED) around the allocation, restored afterwards);
// where the policy call is not permitted the allocation is simply not steered.  FFTW_B200_NO_NUMA=1 switches it off.
inline void *host_malloc(size_t bytes) {
    void *p = nullptr;
    if (bytes == 0) bytes = 16;
    if (ensure_device() != 0) return nullptr;
    bool steered = false;
#if defined(__linux__) && defined(SYS_set_mempolicy)
    const int node = getenv("FFTW_B200_NO_NUMA") ? -1 : device_numa_node();
    if (node >= 0 && node < 1024) {
        unsigned long mask[16] = {0};
        mask[node / (8 * sizeof(unsigned long))] |= 1ul << (node % (8 * sizeof(unsigned long)));
        steered = syscall(SYS_set_mempolicy, 1 /* MPOL_PREFERRED */, mask, (unsigned long)(8 * sizeof mask)) == 0;
    }
#endif
    const cudaError_t e = cudaHostAlloc(&p, bytes, cudaHostAllocPortable);
    if (e == cudaSuccess && steered) memset(p, 0, bytes < (1u << 20) ? bytes : (size_t)(1u << 20));  // (pinned pages are resident already)
#if defined(__linux__) && defined(SYS_set_mempolicy)
    if (steered) syscall(SYS_set_mempolicy, 0 /* MPOL_DEFAULT */, nullptr, 0ul);
#endif
    if (e != cudaSuccess) {
        set_error("cudaHostAlloc(%zu) failed: %s", bytes, cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    return p;
}
inline void host_free(void *p) {
    if (p) cudaFreeHost(p);
}
// 1 if p is device-accessible memory (device or managed), 0 if plain/pinned host memory
inline int is_device_ptr(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}
inline int max_smem_optin() {
    static int v = 0;
    if (!v) {
        cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, device_ordinal());
        if (v <= 0) v = 227 * 1024;
    }
    return v;
}
inline int sm_count() {
    static int v = 0;
    if (!v) {
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device_ordinal());
        if (v <= 0) v = 148;
    }
    return v;
}
#else  // ---------------------------------------------------------------- emulation
typedef void *stream_t;
struct DeviceScope {};
inline int ensure_device_of(int) { return 0; }
inline int current_device() { return 0; }
#define B2_CK(expr)                 \
    do {                            \
        int e_ = (expr);            \
        if (e_ != 0) return e_;     \
    } while (0)
inline int ensure_device() { return 0; }
inline void *dev_malloc(size_t bytes) { return malloc(bytes ? bytes : 16); }
inline void dev_free(void *p) { free(p); }
inline int dev_upload(void *dst, const void *src, size_t bytes) {
    memcpy(dst, src, bytes);
    return 0;
}
inline void *host_malloc(size_t bytes) { return malloc(bytes ? bytes : 16); }
inline void host_free(void *p) { free(p); }
inline int is_device_ptr(const void *) { return 1; }  // emulation: "device" memory is the heap
inline int max_smem_optin() { return 227 * 1024; }
inline int sm_count() { return 148; }
#endif

}  // namespace b2
