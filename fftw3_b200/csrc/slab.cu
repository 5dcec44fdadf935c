// csrc/slab.cu -- slab-distributed 3-D complex DFT, one process per GPU: the
// fftw_mpi_plan_dft_3d form of the reference re-expressed over NCCL / NVLink.
//
// Reference (specification only -- fftw/mpi is not even compiled by the fork):
//   geometry   fftw/mpi/block.c:39-50 (block = ceil(n/P)), fftw/mpi/api.c:248-346 (local sizes)
//   algorithm  fftw/mpi/dft-rank-geq2-transposed.c:47-70,113-198: local (n1 x n2) FFTs on the
//              local_n0 planes, global transpose n0 <-> n1 of n2-long rows, then length-n0 FFTs;
//              without TRANSPOSED_OUT a second global transpose back (dft-rank-geq2.c:40-58)
//   exchange   fftw/mpi/transpose-alltoall.c:49-104 (local transposes around MPI_Alltoall)
//
// Here the local re-transposes around the all-to-all disappear: rank r sends, for each of its
// planes i0 and each peer p, the contiguous run out1[i0][p's n1-block][*] straight from the
// stage-1 output, and the receiver drops it at recv[i0][local j1][*]; the last FFT pass reads
// that layout with a strided gather and writes the FFTW transposed-output layout
// [local_n1][n0][n2] (fftw/doc/mpi.texi:470-509).  The exchange is one ncclGroup of
// ncclSend/ncclRecv per chunk of planes, issued on a second stream so that the exchange of
// chunk k overlaps the local FFTs of chunk k+1.
#include <dlfcn.h>

#include <algorithm>
#include <vector>

#include "b200fft.h"
#include "device_rt.h"
#if B2_CUDA
#include <cuda.h>
#endif

using namespace b2;

// ------------------------------------------------------------------ block distribution
static ptrdiff_t blk_default(ptrdiff_t n, int P) { return (n + P - 1) / P; }
static ptrdiff_t blk_len(ptrdiff_t n, ptrdiff_t block, int pe) {
    const ptrdiff_t left = n - (ptrdiff_t)pe * block;
    return left <= 0 ? 0 : std::min(left, block);
}

// block0 / block1: block sizes of the n0 (input) and n1 (transposed output) distributions, 0 = ceil(n/P);
// a block size that needs more than P blocks cannot be distributed (fftw/mpi/api.c:578-582): returns 0
extern "C" ptrdiff_t b2fft_slab_local_size_3d_blocks(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, ptrdiff_t block0,
                                                    ptrdiff_t block1, int P, int pe, ptrdiff_t *local_n0,
                                                    ptrdiff_t *local_0_start, ptrdiff_t *local_n1,
                                                    ptrdiff_t *local_1_start);
extern "C" ptrdiff_t b2fft_slab_local_size_3d_transposed(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, int P,
                                                         int pe, ptrdiff_t *local_n0, ptrdiff_t *local_0_start,
                                                         ptrdiff_t *local_n1, ptrdiff_t *local_1_start) {
    return b2fft_slab_local_size_3d_blocks(n0, n1, n2, 0, 0, P, pe, local_n0, local_0_start, local_n1, local_1_start);
}
extern "C" ptrdiff_t b2fft_slab_local_size_3d_blocks(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, ptrdiff_t block0,
                                                    ptrdiff_t block1, int P, int pe, ptrdiff_t *local_n0,
                                                    ptrdiff_t *local_0_start, ptrdiff_t *local_n1,
                                                    ptrdiff_t *local_1_start) {
    const ptrdiff_t b0 = block0 > 0 ? block0 : blk_default(n0, P), b1 = block1 > 0 ? block1 : blk_default(n1, P);
    if ((n0 + b0 - 1) / b0 > P || (n1 + b1 - 1) / b1 > P) return 0;
    const ptrdiff_t l0 = blk_len(n0, b0, pe), l1 = blk_len(n1, b1, pe);
    if (local_n0) *local_n0 = l0;
    if (local_0_start) *local_0_start = l0 ? pe * b0 : 0;
    if (local_n1) *local_n1 = l1;
    if (local_1_start) *local_1_start = l1 ? pe * b1 : 0;
    // upper bound on what any intermediate layout of the reference's algorithms needs
    // (fftw/mpi/api.c:269-306): keep the same contract so callers allocate identically
    ptrdiff_t N = 1;
    N = std::max(N, n0 * l1 * n2);  // transposed intermediate
    const ptrdiff_t a1 = n1 * n2, a0 = n0 * n2;
    N = std::max(N, n0 * blk_len(a1, blk_default(a1, P), pe));
    N = std::max(N, n1 * blk_len(a0, blk_default(a0, P), pe));
    N = std::max(N, std::max(l0 * n1 * n2, l1 * n0 * n2));
    return N;
}

// ------------------------------------------------------------------ NCCL (loaded at run time)
typedef struct { char internal[128]; } nccl_uid_t;
typedef void *nccl_comm_t;
struct NcclApi {
    void *h = nullptr;
    int (*GetUniqueId)(nccl_uid_t *) = nullptr;
    int (*CommInitRank)(nccl_comm_t *, int, nccl_uid_t, int) = nullptr;
    int (*CommDestroy)(nccl_comm_t) = nullptr;
    int (*Send)(const void *, size_t, int, int, nccl_comm_t, void *) = nullptr;
    int (*Recv)(void *, size_t, int, int, nccl_comm_t, void *) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, nccl_comm_t, void *) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, nccl_comm_t, void *) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
static NcclApi *nccl() {
    static NcclApi api;
    static int state = 0;
    if (state == 1) return &api;
    if (state == -1) return nullptr;
    const char *names[] = {getenv("B2FFT_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) {
        if (!nm) continue;
        api.h = dlopen(nm, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        if (!api.h) api.h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (api.h) break;
    }
    if (!api.h) {
        set_error("NCCL not found: import torch first or set B2FFT_NCCL_LIB (%s)", dlerror());
        state = -1;
        return nullptr;
    }
#define B2_NSYM(field, sym)                                       \
    *(void **)(&api.field) = dlsym(api.h, sym);                   \
    if (!api.field) {                                             \
        set_error("NCCL symbol %s missing", sym);                 \
        state = -1;                                               \
        return nullptr;                                           \
    }
    B2_NSYM(GetUniqueId, "ncclGetUniqueId")
    B2_NSYM(CommInitRank, "ncclCommInitRank")
    B2_NSYM(CommDestroy, "ncclCommDestroy")
    B2_NSYM(Send, "ncclSend")
    B2_NSYM(Recv, "ncclRecv")
    B2_NSYM(AllGather, "ncclAllGather")
    B2_NSYM(AllReduce, "ncclAllReduce")
    B2_NSYM(GroupStart, "ncclGroupStart")
    B2_NSYM(GroupEnd, "ncclGroupEnd")
    B2_NSYM(GetErrorString, "ncclGetErrorString")
#undef B2_NSYM
    state = 1;
    return &api;
}
#define B2_NCCL(call)                                                                  \
    do {                                                                               \
        int r_ = (call);                                                               \
        if (r_ != 0) {                                                                 \
            set_error("%s failed: %s", #call, nccl()->GetErrorString(r_));             \
            return 1000 + r_;                                                          \
        }                                                                              \
    } while (0)

struct b2fft_comm_s {
    int rank = 0, size = 1;
    nccl_comm_t nc = nullptr;
};

extern "C" int b2fft_nccl_unique_id(void *id128) {
    NcclApi *a = nccl();
    if (!a) return -1;
    DeviceScope caller;
    if (ensure_device() != 0) return -1;
    nccl_uid_t id;
    B2_NCCL(a->GetUniqueId(&id));
    memcpy(id128, &id, sizeof id);
    return 0;
}
extern "C" b2fft_comm *b2fft_comm_create(const void *id128, int n_pes, int my_pe) {
    if (n_pes < 1 || my_pe < 0 || my_pe >= n_pes) {
        set_error("bad communicator shape");
        return nullptr;
    }
    DeviceScope caller;
    if (ensure_device() != 0) return nullptr;
    b2fft_comm *c = new b2fft_comm_s();
    c->rank = my_pe;
    c->size = n_pes;
    if (n_pes > 1) {
        NcclApi *a = nccl();
        if (!a) {
            delete c;
            return nullptr;
        }
        nccl_uid_t id;
        memcpy(&id, id128, sizeof id);
        int r = a->CommInitRank(&c->nc, n_pes, id, my_pe);
        if (r != 0) {
            set_error("ncclCommInitRank failed: %s", a->GetErrorString(r));
            delete c;
            return nullptr;
        }
    }
    return c;
}
extern "C" b2fft_comm *b2fft_comm_self(void) {
    b2fft_comm *c = new b2fft_comm_s();
    return c;
}
// test hook: a communicator that only carries (rank, size); its plans can run the local stages,
// the exchange is then done by the caller (tests/test_slab_gloo.py does it over gloo)
extern "C" b2fft_comm *b2fft_comm_detached(int n_pes, int my_pe) {
    b2fft_comm *c = new b2fft_comm_s();
    c->rank = my_pe;
    c->size = n_pes;
    return c;
}
extern "C" void b2fft_comm_destroy(b2fft_comm *c) {
    if (!c) return;
    if (c->nc && nccl()) nccl()->CommDestroy(c->nc);
    delete c;
}
extern "C" int b2fft_comm_rank(const b2fft_comm *c) { return c ? c->rank : 0; }
extern "C" int b2fft_comm_size(const b2fft_comm *c) { return c ? c->size : 1; }

// ------------------------------------------------------------------ the slab plan
#include "plan_struct.h"

struct SlabMsg {
    int peer;
    long long soff, roff, cnt;  // in complex elements
};

struct SlabPlan {
    // n2 is the length of a contiguous row in complex elements = (last transform dim n2f) x (howmany
    // interleaved transforms hm, fftw_mpi_plan_many_dft's tuple layout): everything but the FFT along
    // that dim treats the row as opaque
    ptrdiff_t n0, n1, n2, n2f = 0, hm = 1;
    bool transposed_in = false;
    // kind: B2FFT_C2C, or the real forms (fftw_mpi_plan_dft_r2c_3d / _c2r_3d): the contiguous row then
    // holds n2f/2+1 complex, the real side is padded to 2(n2f/2+1) reals per row
    // (fftw/mpi/rdft2-rank-geq2-transposed.c:47-70, fftw/doc/mpi.texi "Multi-dimensional MPI DFTs of Real Data").
    // c2r runs the pipeline backwards: n0 FFTs (pA), reverse exchange, local 2-D c2r (pB).
    int kind = B2FFT_C2C;
    bool c2r_tin = false;            // c2r with FFTW_MPI_TRANSPOSED_IN: the input is [local_n1][n0][n2c]
    // 2-D real data (fftw_mpi_plan_dft_r2c_2d / _c2r_2d): the half-complex dimension itself is the one that
    // gets distributed after the exchange -- n1 = n1f/2+1 complex columns, rows of one element (n2 = 1),
    // and the local real step is a batched 1-D r2c / c2r over the n1f-long rows
    bool real2d = false;
    ptrdiff_t n1f = 0;
    b2fft_plan *pA = nullptr, *pB = nullptr;
    int P, pe, sign;
    bool transposed_out;
    ptrdiff_t b0, b1, ln0, ln1, s0, s1;
    b2fft_comm *comm;
    // stage 1: (n1 x n2) FFTs on chunks of planes, in -> S;  stage 2: length-n0 FFTs, R -> out
    int nchunks = 1;
    ptrdiff_t chunk = 0;                 // planes per chunk (the last may be shorter)
    b2fft_plan *p1 = nullptr, *p1_tail = nullptr, *p2 = nullptr;
    cpx_t *S = nullptr, *R = nullptr;
    size_t s_elems = 0, r_elems = 0;
    cpx_t *dev_in = nullptr, *dev_out = nullptr;  // staging when the caller passes host arrays
    size_t alloc_elems = 0;
#if B2_CUDA
    cudaStream_t sc = nullptr, sn = nullptr;
    std::vector<cudaEvent_t> ev;
    cudaEvent_t t0 = nullptr, t1 = nullptr, t2 = nullptr, t3 = nullptr;
#endif
    float ms[3] = {0, 0, 0};
    // direct peer-memory exchange (CUDA IPC): every rank's S and R mapped into this process
    bool p2p = false;
    std::vector<cpx_t *> peerS, peerR;
    float *flag = nullptr;  // tiny device buffer for the cross-rank barrier
    // fused exchange: the n1-axis pass of stage 1 stores every output row straight into its owner's
    // receive buffer (tile kernel EPI 8) -- no send-side DMA, no second trip through HBM
    bool fused = false;
    // hybrid: the n1 pass stores the rows this rank owns straight into its own receive buffer and leaves
    // the others in place for the DMA, which then skips the local 1/P of the exchange
    bool hybrid = false;
    // ---- round-2 schedule ("v2") for transposed-output c2c: chunks run along the CONTIGUOUS dimension ----
    // The planes of the slab are split into v2_ng groups, the rows into v2_nch chunks of v2_w columns.  Per group g:
    //   row pass over the group's planes, in -> W (W = the caller's output array), then per chunk k:
    //     n1 pass: reads W[i0][i1][chunk] and stores row k1 of plane i0 chunk-major (scatter epilogue, local stores):
    //        rows this rank owns      -> its own receive buffer  R_k[s0 + i0][j1][w]
    //        rows of the other ranks  -> staging                 S_k[i0][k1][w]
    //     exchange: ONE 2-D copy per peer, runs of ln1 * w elements (512 KiB at 2048^3 on 8 GPUs) from S_k straight
    //     into the peer's R_k at its final position (copy engines, peers in the shifted order pe + d)
    //   after the last group's copies of chunk k: cross-rank barrier, then the n0 pass R_k[i0][j1][w] -> out[j1][k0][chunk].
    // The exchange of a chunk overlaps the n1 passes that follow it, the NEXT group's row pass, and the n0 passes of the
    // chunks before it; only the first group's row pass and the last n0 pass are exposed.  (Round 1 chunked along n0:
    // its last stage could only start after the whole exchange.  Message shape of the reference:
    // fftw/mpi/transpose-alltoall.c:224-227.)  One rank: every row is "owned", no staging, no copies.
    bool v2 = false;
    int v2_nch = 1, v2_lag = 2, v2_ng = 1;
    ptrdiff_t v2_w = 0, v2_gp = 0;               // chunk width (columns), planes per group
    b2fft_plan *v2_row[2] = {nullptr, nullptr};  // one group's row pass: [0] out of place, [1] in place; built on first use
    b2fft_plan *v2_n1 = nullptr, *v2_n0 = nullptr;
#if B2_CUDA
    // DMA exchange: one stream per peer so that the copies to different peers run on different copy engines
    std::vector<cudaStream_t> sx;
    std::vector<cudaEvent_t> evx;
    cudaEvent_t evb = nullptr;
    cudaStream_t sb = nullptr;               // v2: barriers run beside the copy stream
    std::vector<cudaEvent_t> v2_e1, v2_ec, v2_e2;  // per chunk: n1 pass done, copies done, every rank's copies done
    cudaEvent_t v2_eb = nullptr, v2_ta = nullptr, v2_tb = nullptr;
    cudaEvent_t ev_in = nullptr;  // orders an execute after the caller's earlier work on the default stream
    // v2 cross-rank signalling without kernels: word F[k*P + src] of MY flag array is set to the execute's epoch by
    // rank `src` (a 4-byte copy behind its data copies of chunk k, on its copy stream); my n0 pass of chunk k waits
    // for the P-1 words with cuStreamWaitValue32.  Row k = v2_nch holds "rank src has finished reading its receive
    // buffer for this epoch".  (Round 1 and the fallback use a 1-float NCCL all-reduce as the barrier: a kernel that
    // needs a free SM slot on every GPU at the same time, behind persistent FFT kernels that fill all of them.)
    unsigned *F = nullptr, *epoch_dev = nullptr;
    std::vector<unsigned *> peerF;
    unsigned epoch = 0;
    bool v2_flags = false;
    cudaStream_t s2 = nullptr;  // v2: the n0 passes, so that a chunk still in flight never blocks the n1 passes
    cudaEvent_t v2_e3 = nullptr;
#endif
};

// the send buffer of the round-1 schedule and of the staged API: allocated when first needed (the v2 schedule
// works in the caller's output array instead, which is what lets 2048^3 run on one 180 GB GPU)
static bool slab_ensure_S(SlabPlan &s) {
    if (!s.S) s.S = (cpx_t *)dev_malloc(s.s_elems * sizeof(cpx_t));
    return s.S != nullptr;
}

// switch the stage-1 plans to the scattered store; all-or-nothing
static void slab_enable_fused(SlabPlan &s) {
    s.fused = false;
    s.hybrid = false;
    if (getenv("B2FFT_SLAB_NO_FUSED") || (int)s.peerR.size() != s.P || s.P > B2_MAXSCAT) return;
    // Measured on 2 B200 (profiles/r1_slab_fused_2gpu.md): the scattered store wins when a tile row is a
    // full 128-byte segment (n1 <= 1024: 1024^3 7.09 ms against 8.12 ms with the DMA exchange) and loses
    // when the 2048-point kernel's 8-line tiles turn the NVLink writes into 64-byte packets (2048^3:
    // 72.6 against 63.8 ms).  B2FFT_SLAB_FUSED=1 forces it.
    const long long col_limit = (long long)(sizeof(real_t) == 4 ? 1024 : 1024);
    // Above that limit (and with B2FFT_SLAB_HYBRID=1 anywhere): hybrid -- only the rows this rank owns are
    // scattered (local stores), the copy engines move the P-1 remote blocks and no longer the local one.
    bool hybrid = s.P > 1 && ((s.n1 > col_limit && !getenv("B2FFT_SLAB_FUSED")) || getenv("B2FFT_SLAB_HYBRID"));
    if (hybrid && getenv("B2FFT_SLAB_NO_HYBRID")) return;
    std::vector<long long> ps(s.P);
    for (int o = 0; o < s.P; ++o) ps[o] = (long long)blk_len(s.n1, s.b1, o) * s.n2 * 2;
    bool ok = true;
    for (b2fft_plan *p : {s.p1, s.p1_tail})
        if (p) ok = ok && b2_plan_enable_scatter(p, s.P, (long long)s.b1, (long long)s.n1 * s.n2 * 2, ps.data(),
                                                 hybrid ? s.pe : -1) == 0;
    ok = ok && (s.p1 || s.p1_tail || s.ln0 == 0);
    s.fused = ok && !hybrid;
    s.hybrid = ok && hybrid;
}

#if B2_CUDA
static int slab_setup_p2p(SlabPlan &s);
static bool slab_setup_v2(SlabPlan &s);
#define B2_SLAB_FLAG_BYTES (65 * B2_MAXSCAT * sizeof(unsigned))  // <= 64 chunks + the "done" row, <= B2_MAXSCAT ranks
// cuStreamWaitValue32 through the runtime (no link against libcuda); nullptr when the driver does not offer it
typedef CUresult (*StreamWaitValueFn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static StreamWaitValueFn stream_wait_value() {
    static StreamWaitValueFn f = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        cudaDriverEntryPointQueryResult q;
        void *fp = nullptr;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fp, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            f = (StreamWaitValueFn)fp;
        else
            cudaGetLastError();
    }
    return f;
}
#endif

static b2fft_plan *make_stage1(const SlabPlan &s, ptrdiff_t planes) {
    if (planes <= 0) return nullptr;
    b2fft_geom g;
    memset(&g, 0, sizeof g);
    if (s.kind == B2FFT_R2C && s.real2d) {
        // real [rows][2*n1c] -> complex [rows][n1c]
        g.kind = B2FFT_R2C;
        g.sign = -1;
        g.rank = 1;
        g.dims[0] = b2fft_dim{s.n1f, 1, 1};
        g.vrank = 1;
        g.vdims[0] = b2fft_dim{planes, 2 * s.n1, s.n1};
        g.in_place = 0;
        return b2fft_plan_create(&g);
    }
    if (s.kind == B2FFT_R2C) {
        // real [planes][n1][2*n2c] -> complex [planes][n1][n2c]
        g.kind = B2FFT_R2C;
        g.sign = -1;
        g.rank = 2;
        // (s.n2 = n2c * howmany: with interleaved tuples the n2 transform has stride howmany on both sides)
        g.dims[0] = b2fft_dim{s.n1, 2 * s.n2, s.n2};
        g.dims[1] = b2fft_dim{s.n2f, s.hm, s.hm};
        g.vrank = s.hm > 1 ? 2 : 1;
        g.vdims[0] = b2fft_dim{planes, s.n1 * 2 * s.n2, s.n1 * s.n2};
        g.vdims[1] = b2fft_dim{s.hm, 1, 1};
        g.in_place = 0;
        return b2fft_plan_create(&g);
    }
    g.kind = B2FFT_C2C;
    g.sign = s.sign;
    g.rank = 2;
    g.dims[0] = b2fft_dim{s.n1, s.n2, s.n2};
    g.dims[1] = b2fft_dim{s.n2f, s.hm, s.hm};
    g.vrank = s.hm > 1 ? 2 : 1;
    g.vdims[0] = b2fft_dim{planes, s.n1 * s.n2, s.n1 * s.n2};
    g.vdims[1] = b2fft_dim{s.hm, 1, 1};
    g.in_place = 0;
    return b2fft_plan_create(&g);
}
static b2fft_plan *make_stage2(const SlabPlan &s) {
    if (s.ln1 <= 0) return nullptr;
    b2fft_geom g;
    memset(&g, 0, sizeof g);
    g.kind = B2FFT_C2C;
    g.sign = s.sign;
    g.rank = 1;
    g.vrank = 2;
    if (s.transposed_out) {
        // R[i0][j1][i2] -> out[j1][k0][i2]
        g.dims[0] = b2fft_dim{s.n0, s.ln1 * s.n2, s.n2};
        g.vdims[0] = b2fft_dim{s.ln1, s.n2, s.n0 * s.n2};
        g.vdims[1] = b2fft_dim{s.n2, 1, 1};
        g.in_place = 0;
    } else {
        g.dims[0] = b2fft_dim{s.n0, s.ln1 * s.n2, s.ln1 * s.n2};
        g.vdims[0] = b2fft_dim{s.ln1, s.n2, s.n2};
        g.vdims[1] = b2fft_dim{s.n2, 1, 1};
        g.in_place = 1;
    }
    return b2fft_plan_create(&g);
}

// c2r, first step: backward length-n0 FFTs.  Transposed input [ln1][n0][rl] -> receive layout
// [n0][ln1][rl]; natural input: in place on the receive layout (filled by the forward exchange).
static b2fft_plan *make_c2r_a(const SlabPlan &s) {
    if (s.ln1 <= 0) return nullptr;
    b2fft_geom g;
    memset(&g, 0, sizeof g);
    g.kind = B2FFT_C2C;
    g.sign = +1;
    g.rank = 1;
    g.vrank = 2;
    if (s.c2r_tin) {
        g.dims[0] = b2fft_dim{s.n0, s.n2, s.ln1 * s.n2};
        g.vdims[0] = b2fft_dim{s.ln1, s.n0 * s.n2, s.n2};
        g.in_place = 0;
    } else {
        g.dims[0] = b2fft_dim{s.n0, s.ln1 * s.n2, s.ln1 * s.n2};
        g.vdims[0] = b2fft_dim{s.ln1, s.n2, s.n2};
        g.in_place = 1;
    }
    g.vdims[1] = b2fft_dim{s.n2, 1, 1};
    return b2fft_plan_create(&g);
}
// c2r, last step: local 2-D c2r, complex [ln0][n1][n2c] -> real [ln0][n1][2*n2c]
static b2fft_plan *make_c2r_b(const SlabPlan &s) {
    if (s.ln0 <= 0) return nullptr;
    b2fft_geom g;
    memset(&g, 0, sizeof g);
    g.kind = B2FFT_C2R;
    g.sign = +1;
    if (s.real2d) {
        g.rank = 1;
        g.dims[0] = b2fft_dim{s.n1f, 1, 1};
        g.vrank = 1;
        g.vdims[0] = b2fft_dim{s.ln0, s.n1, 2 * s.n1};
        g.in_place = 0;
        return b2fft_plan_create(&g);
    }
    g.rank = 2;
    g.dims[0] = b2fft_dim{s.n1, s.n2, 2 * s.n2};
    g.dims[1] = b2fft_dim{s.n2f, s.hm, s.hm};
    g.vrank = s.hm > 1 ? 2 : 1;
    g.vdims[0] = b2fft_dim{s.ln0, s.n1 * s.n2, s.n1 * 2 * s.n2};
    g.vdims[1] = b2fft_dim{s.hm, 1, 1};
    g.in_place = 0;
    return b2fft_plan_create(&g);
}

// messages of the n0 <-> n1 exchange for the planes [i0a, i0b) of every rank's local range
// (same chunk index on every rank).  forward = S -> R; the reverse exchange swaps the roles.
static void slab_messages(const SlabPlan &s, int chunk, std::vector<SlabMsg> &sends, std::vector<SlabMsg> &recvs) {
    sends.clear();
    recvs.clear();
    for (int p = 0; p < s.P; ++p) {
        const ptrdiff_t ln1p = blk_len(s.n1, s.b1, p), ln0p = blk_len(s.n0, s.b0, p);
        // my planes of this chunk -> peer p's n1 block
        const ptrdiff_t a = std::min<ptrdiff_t>(s.ln0, chunk * s.chunk), b = std::min<ptrdiff_t>(s.ln0, a + s.chunk);
        if (ln1p > 0)
            for (ptrdiff_t i0 = a; i0 < b; ++i0)
                sends.push_back(SlabMsg{p, (long long)((i0 * s.n1 + p * s.b1) * s.n2), 0, (long long)(ln1p * s.n2)});
        // peer p's planes of this chunk -> my n1 block
        const ptrdiff_t pa = std::min<ptrdiff_t>(ln0p, chunk * s.chunk), pb = std::min<ptrdiff_t>(ln0p, pa + s.chunk);
        if (s.ln1 > 0)
            for (ptrdiff_t i0 = pa; i0 < pb; ++i0)
                recvs.push_back(SlabMsg{p, 0, (long long)(((p * s.b0 + i0) * s.ln1) * s.n2), (long long)(s.ln1 * s.n2)});
    }
}

void b2_slab_destroy(b2fft_plan_s *p) {
    SlabPlan *s = p->slab;
    if (!s) return;
    b2fft_plan_destroy(s->p1);
    b2fft_plan_destroy(s->p1_tail);
    b2fft_plan_destroy(s->p2);
    b2fft_plan_destroy(s->pA);
    b2fft_plan_destroy(s->pB);
    b2fft_plan_destroy(s->v2_row[0]);
    b2fft_plan_destroy(s->v2_row[1]);
    b2fft_plan_destroy(s->v2_n1);
    b2fft_plan_destroy(s->v2_n0);
#if B2_CUDA
    for (std::vector<cudaEvent_t> *v : {&s->v2_e1, &s->v2_ec, &s->v2_e2})
        for (cudaEvent_t e : *v)
            if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : {s->v2_eb, s->v2_ta, s->v2_tb, s->ev_in})
        if (e) cudaEventDestroy(e);
    if (s->sb) cudaStreamDestroy(s->sb);
    if (s->s2) cudaStreamDestroy(s->s2);
    if (s->v2_e3) cudaEventDestroy(s->v2_e3);
    for (int q = 0; q < (int)s->peerF.size(); ++q)
        if (q != s->pe && s->peerF[q]) cudaIpcCloseMemHandle(s->peerF[q]);
    if (s->F) cudaFree(s->F);
    if (s->epoch_dev) cudaFree(s->epoch_dev);
    for (int q = 0; q < (int)s->peerS.size(); ++q)
        if (q != s->pe) {
            if (s->peerS[q]) cudaIpcCloseMemHandle(s->peerS[q]);
            if (s->peerR[q]) cudaIpcCloseMemHandle(s->peerR[q]);
        }
    if (s->flag) cudaFree(s->flag);
    for (cudaStream_t x : s->sx)
        if (x) cudaStreamDestroy(x);
    for (cudaEvent_t e : s->evx)
        if (e) cudaEventDestroy(e);
    if (s->evb) cudaEventDestroy(s->evb);
#endif
    dev_free(s->S);
    dev_free(s->R);
    dev_free(s->dev_in);
    dev_free(s->dev_out);
#if B2_CUDA
    for (cudaEvent_t e : s->ev) cudaEventDestroy(e);
    for (cudaEvent_t e : {s->t0, s->t1, s->t2, s->t3})
        if (e) cudaEventDestroy(e);
    if (s->sc) cudaStreamDestroy(s->sc);
    if (s->sn) cudaStreamDestroy(s->sn);
#endif
    delete s;
    p->slab = nullptr;
}

static void slab_describe(SlabPlan &sr, b2fft_plan *bp) {
    SlabPlan *s = &sr;
    const ptrdiff_t n0 = s->n0, n1 = s->n1, n2 = s->n2f;
    char buf[640];
    snprintf(buf, sizeof buf,
             "  slab 3-D %s %td x %td x %td%s, %td interleaved transform(s), over %d ranks (rank %d: planes %td+%td, out rows %td+%td), %s output\n"
             "  stage 1: %d chunk(s) of %td plane(s), (n1 x n2) FFTs in->send; exchange: %s\n"
             "  stage 2: length-%td FFTs recv->out\n",
             s->kind == B2FFT_C2C ? "c2c" : s->kind == B2FFT_R2C ? "r2c" : "c2r (n0 FFTs, reverse exchange, local 2-D c2r)", n0, n1, n2,
             s->transposed_in ? (s->kind == B2FFT_C2C ? " (first two dims exchanged: FFTW_MPI_TRANSPOSED_IN)" : " (transposed input)") : "", s->hm, s->P, s->pe,
             s->s0, s->ln0, s->s1, s->ln1, s->transposed_out ? "transposed" : "natural",
             s->nchunks, s->chunk,
             s->v2 ? "v2: plane groups x chunks along the contiguous dimension; the n1 pass stores chunk-major (own rows -> receive buffer, others -> staging), one 2-D DMA per peer, group and chunk into peer memory; the next group's row pass and the n0 passes overlap it"
             : s->fused ? "fused into the n1 pass: rows stored straight into their owner's receive buffer (peer memory over NVLink)"
             : s->hybrid ? "local block stored by the n1 pass, remote blocks by strided DMA into peer memory (CUDA IPC over NVLink), NCCL barrier"
             : s->P == 1 ? "local copy" : s->p2p ? "strided DMA into peer memory (CUDA IPC over NVLink), NCCL barrier"
                                                : "ncclSend/ncclRecv per plane and peer",
             n0);
    bp->pc.desc = buf;
}

static b2fft_plan *slab_plan_create(int kind, ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, ptrdiff_t howmany,
                                    b2fft_comm *c, int sign, int transposed_in, int transposed_out,
                                    ptrdiff_t block0 = 0, ptrdiff_t block1 = 0, bool real2d = false);
// block0: block size of the distributed INPUT dimension (n0, or n1 with transposed_in), block1: of the
// distributed OUTPUT dimension; 0 = default (fftw_mpi_plan_many_dft's block / tblock, fftw/mpi/api.c:592-636)
extern "C" b2fft_plan *b2fft_slab_plan_dft_3d_blocks(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, ptrdiff_t howmany,
                                                     ptrdiff_t block0, ptrdiff_t block1, b2fft_comm *c, int sign,
                                                     int transposed_in, int transposed_out, unsigned flags) {
    (void)flags;
    return slab_plan_create(B2FFT_C2C, n0, n1, n2, howmany, c, sign, transposed_in, transposed_out, block0, block1);
}

extern "C" b2fft_plan *b2fft_slab_plan_dft_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, ptrdiff_t howmany,
                                              b2fft_comm *c, int sign, int transposed_in, int transposed_out,
                                              unsigned flags) {
    (void)flags;
    return slab_plan_create(B2FFT_C2C, n0, n1, n2, howmany, c, sign, transposed_in, transposed_out);
}
// kind = B2FFT_R2C (real [ln0][n1][2(n2/2+1)] -> complex, natural or transposed output) or B2FFT_C2R
// (complex, natural or transposed input -> real [ln0][n1][2(n2/2+1)]); n2 is the logical real length
extern "C" b2fft_plan *b2fft_slab_plan_many_rdft_3d(int kind, ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2,
                                                    ptrdiff_t howmany, b2fft_comm *c, int transposed_in,
                                                    int transposed_out, unsigned flags) {
    (void)flags;
    if (kind != B2FFT_R2C && kind != B2FFT_C2R) return nullptr;
    if ((kind == B2FFT_R2C && transposed_in) || (kind == B2FFT_C2R && transposed_out)) {
        set_error("slab real transform: a transposed REAL array is not supported");
        return nullptr;
    }
    return slab_plan_create(kind, n0, n1, n2, howmany, c, kind == B2FFT_R2C ? -1 : +1, transposed_in, transposed_out);
}
// the same with explicit block sizes: block0 over n0 (the dimension the REAL array is distributed over), block1 over n1 (the
// dimension the transposed complex array is distributed over); 0 = default (fftw_mpi_plan_many_dft_r2c / _c2r iblock / oblock,
// reference fftw/mpi/api.c:790-858)
extern "C" b2fft_plan *b2fft_slab_plan_many_rdft_3d_blocks(int kind, ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2,
                                                           ptrdiff_t howmany, ptrdiff_t block0, ptrdiff_t block1,
                                                           b2fft_comm *c, int transposed_in, int transposed_out,
                                                           unsigned flags) {
    (void)flags;
    if (kind != B2FFT_R2C && kind != B2FFT_C2R) return nullptr;
    if ((kind == B2FFT_R2C && transposed_in) || (kind == B2FFT_C2R && transposed_out)) {
        set_error("slab real transform: a transposed REAL array is not supported");
        return nullptr;
    }
    if (block0 < 0 || block1 < 0) {
        set_error("slab real transform: negative block size");
        return nullptr;
    }
    return slab_plan_create(kind, n0, n1, n2, howmany, c, kind == B2FFT_R2C ? -1 : +1, transposed_in, transposed_out, block0,
                            block1);
}
extern "C" b2fft_plan *b2fft_slab_plan_rdft_3d(int kind, ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, b2fft_comm *c,
                                               int transposed_in, int transposed_out, unsigned flags) {
    (void)flags;
    if (kind != B2FFT_R2C && kind != B2FFT_C2R) return nullptr;
    if ((kind == B2FFT_R2C && transposed_in) || (kind == B2FFT_C2R && transposed_out)) {
        set_error("slab real transform: a transposed REAL array is not supported");
        return nullptr;
    }
    return slab_plan_create(kind, n0, n1, n2, 1, c, kind == B2FFT_R2C ? -1 : +1, transposed_in, transposed_out);
}

// 2-D real data: n1 is the logical real length of the rows
extern "C" b2fft_plan *b2fft_slab_plan_rdft_2d(int kind, ptrdiff_t n0, ptrdiff_t n1, b2fft_comm *c, int transposed_in,
                                               int transposed_out, unsigned flags) {
    (void)flags;
    if (kind != B2FFT_R2C && kind != B2FFT_C2R) return nullptr;
    if ((kind == B2FFT_R2C && transposed_in) || (kind == B2FFT_C2R && transposed_out)) {
        set_error("slab real transform: a transposed REAL array is not supported");
        return nullptr;
    }
    return slab_plan_create(kind, n0, n1, 1, 1, c, kind == B2FFT_R2C ? -1 : +1, transposed_in, transposed_out, 0, 0, true);
}
// ... with explicit block sizes: block0 over n0, block1 over the n1/2+1 complex columns; 0 = default
extern "C" b2fft_plan *b2fft_slab_plan_rdft_2d_blocks(int kind, ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t block0, ptrdiff_t block1,
                                                      b2fft_comm *c, int transposed_in, int transposed_out, unsigned flags) {
    (void)flags;
    if (kind != B2FFT_R2C && kind != B2FFT_C2R) return nullptr;
    if ((kind == B2FFT_R2C && transposed_in) || (kind == B2FFT_C2R && transposed_out)) {
        set_error("slab real transform: a transposed REAL array is not supported");
        return nullptr;
    }
    if (block0 < 0 || block1 < 0) {
        set_error("slab real transform: negative block size");
        return nullptr;
    }
    return slab_plan_create(kind, n0, n1, 1, 1, c, kind == B2FFT_R2C ? -1 : +1, transposed_in, transposed_out, block0, block1, true);
}

static b2fft_plan *slab_plan_create(int kind, ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, ptrdiff_t howmany,
                                    b2fft_comm *c, int sign, int transposed_in, int transposed_out,
                                    ptrdiff_t block0, ptrdiff_t block1, bool real2d) {
    if (!c || n0 <= 0 || n1 <= 0 || n2 <= 0) {
        set_error("slab plan: bad arguments");
        return nullptr;
    }
    if (howmany < 1) {
        set_error("slab plan: howmany must be positive");
        return nullptr;
    }
    DeviceScope caller;
    if (ensure_device() != 0) return nullptr;
    SlabPlan *s = new SlabPlan();
    // FFTW_MPI_TRANSPOSED_IN: the input is [local_n1][n0][n2], block-distributed over n1 -- that is the
    // ordinary slab input of the problem with its first two dimensions exchanged, and (the DFT being
    // separable) that problem's transposed output is this problem's natural output and vice versa
    // (fftw/mpi/dft-rank-geq2-transposed.c:113-198 plans the same way: XM(dft_rank_geq2_transposed)
    // serves both flags by swapping which half of the algorithm runs first).
    s->kind = kind;
    s->transposed_in = transposed_in != 0;
    if (kind == B2FFT_C2C && transposed_in) {
        std::swap(n0, n1);
        transposed_out = !transposed_out;
    }
    s->c2r_tin = (kind == B2FFT_C2R && transposed_in);
    s->real2d = real2d;
    if (real2d) {
        s->n1f = n1;
        n1 = n1 / 2 + 1;  // complex columns: the dimension the exchange distributes
    }
    s->n0 = n0;
    s->n1 = n1;
    s->n2f = n2;
    s->hm = howmany;
    if (kind != B2FFT_C2C && !real2d) n2 = n2 / 2 + 1;  // complex elements per row
    n2 *= howmany;
    s->n2 = n2;
    s->P = c->size;
    s->pe = c->rank;
    s->sign = sign < 0 ? -1 : +1;
    s->transposed_out = transposed_out != 0;
    s->comm = c;
    // inside the plan dimension "0" is the one distributed on input, "1" the one distributed after the
    // exchange: with the c2c TRANSPOSED_IN swap above the caller's input / output block sizes already line up
    s->b0 = block0 > 0 ? block0 : blk_default(n0, s->P);
    s->b1 = block1 > 0 ? block1 : blk_default(n1, s->P);
    if ((n0 + s->b0 - 1) / s->b0 > s->P || (n1 + s->b1 - 1) / s->b1 > s->P) {
        set_error("slab plan: more blocks than ranks");
        delete s;
        return nullptr;
    }
    s->alloc_elems = (size_t)b2fft_slab_local_size_3d_blocks(n0, n1, n2, s->b0, s->b1, s->P, s->pe, &s->ln0, &s->s0, &s->ln1, &s->s1);
    // chunks of planes: enough to overlap, few enough to keep the messages large
    const ptrdiff_t maxl0 = s->b0;
    s->nchunks = (int)std::max<ptrdiff_t>(1, std::min<ptrdiff_t>(8, maxl0 / 8));
    if (const char *e = getenv("B2FFT_SLAB_CHUNKS")) s->nchunks = (int)std::max<ptrdiff_t>(1, std::min<ptrdiff_t>(atoi(e), maxl0));
    if (s->P == 1) s->nchunks = 1;
    s->chunk = (maxl0 + s->nchunks - 1) / s->nchunks;
    s->nchunks = (int)((maxl0 + s->chunk - 1) / s->chunk);
    b2fft_plan *bp = new b2fft_plan_s();
    bp->device = current_device();
    bp->slab = s;
    bool ok = true;
    if (kind == B2FFT_C2R) {
        if (s->ln1 > 0) {
            s->pA = make_c2r_a(*s);
            ok = ok && s->pA;
        }
        if (ok && s->ln0 > 0) {
            s->pB = make_c2r_b(*s);
            ok = ok && s->pB;
        }
    } else if (s->ln0 > 0) {
        const ptrdiff_t full = std::min(s->chunk, s->ln0);
        s->p1 = make_stage1(*s, full);
        ok = ok && s->p1;
        const ptrdiff_t tail = s->ln0 % s->chunk;
        if (ok && tail && s->ln0 > s->chunk) {
            s->p1_tail = make_stage1(*s, tail);
            ok = ok && s->p1_tail;
        }
    }
    if (ok && s->ln1 > 0 && kind != B2FFT_C2R) {
        s->p2 = make_stage2(*s);
        ok = ok && s->p2;
    }
    s->s_elems = (size_t)std::max<ptrdiff_t>(1, s->ln0 * n1 * n2);
    s->r_elems = (size_t)std::max<ptrdiff_t>(1, n0 * s->ln1 * n2);
    // v2 schedule: transposed-output c2c (after the TRANSPOSED_IN swap above).  Every rank decides alike.
    // (CUDA build only: the emulation build drives the stages one by one through b2fft_slab_stage)
    s->v2 = B2_CUDA && kind == B2FFT_C2C && s->transposed_out && !getenv("B2FFT_SLAB_NO_V2");
    if (ok) {
        if (!s->v2) ok = slab_ensure_S(*s);  // v2 works in the caller's output array
        s->R = (cpx_t *)dev_malloc(s->r_elems * sizeof(cpx_t));
        ok = ok && s->R;
    }
#if B2_CUDA
    if (ok) {
        ok = cudaStreamCreateWithFlags(&s->sc, cudaStreamNonBlocking) == cudaSuccess &&
             cudaStreamCreateWithFlags(&s->sn, cudaStreamNonBlocking) == cudaSuccess;
        s->ev.resize(s->nchunks + 1);
        for (cudaEvent_t &e : s->ev) ok = ok && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess;
        for (cudaEvent_t *e : {&s->t0, &s->t1, &s->t2, &s->t3}) ok = ok && cudaEventCreate(e) == cudaSuccess;
        if (ok) ok = slab_setup_p2p(*s) == 0;
        if (s->v2 && s->P > 1 && !s->p2p) s->v2 = false;  // no peer mapping: grouped send / receive (round-1 schedule)
        if (ok && s->p2p && kind != B2FFT_C2R && !s->v2) slab_enable_fused(*s);
    }
#endif
    if (ok && s->P == 1 && kind != B2FFT_C2R) {  // one rank: the "exchange" is the scatter into this rank's own receive buffer
        s->peerR.assign(1, s->R);
        if (!s->v2) slab_enable_fused(*s);
    }
#if B2_CUDA
    if (ok && s->v2 && !slab_setup_v2(*s)) {
        // not available for this geometry (e.g. n1 without a scattered-store tile kernel): round-1 schedule
        s->v2 = false;
        b2fft_plan_destroy(s->v2_n1);
        b2fft_plan_destroy(s->v2_n0);
        s->v2_n1 = s->v2_n0 = nullptr;
        ok = slab_ensure_S(*s);
        if (ok && (s->p2p || s->P == 1)) slab_enable_fused(*s);
    }
#endif
    if (!ok) {
        b2_slab_destroy(bp);
        delete bp;
        return nullptr;
    }
    double N = (double)n0 * (real2d ? s->n1f : n1) * s->n2f;
    bp->pc.flops = (kind == B2FFT_C2C ? 5.0 : 2.5) * N * log2(N) * (double)howmany / s->P;
    slab_describe(*s, bp);
    return bp;
}

#if B2_CUDA
static bool slab_setup_v2(SlabPlan &s) {
    // chunk width: whole 128-byte segments and rows of >= 1 KiB; else fewer chunks
    // measured on 2048^3 (profiles/r2_slab_8gpu.md): on 8 GPUs 4 chunks 16.6-16.9 ms, 8 chunks 18.7-19.8, 16 chunks 19.6-20.9
    // (exchange-bound: every extra chunk adds a barrier and copy start-up on the critical path); on 2 GPUs the
    // pipeline is compute-bound and 8 chunks are as good as 4
    int nch = s.P > 2 ? 4 : 8;
    if (const char *e = getenv("B2FFT_SLAB_V2_CHUNKS")) nch = std::max(1, atoi(e));
    while (nch > 1 && (s.n2 % nch != 0 || (s.n2 / nch) % 16 != 0 || s.n2 / nch < 128)) --nch;
    s.v2_nch = nch;
    s.v2_w = s.n2 / nch;
    if (const char *e = getenv("B2FFT_SLAB_V2_LAG")) s.v2_lag = std::max(0, atoi(e));
    // plane groups: the same split on every rank (even blocks only), groups of >= 8 planes
    int ng = s.P > 1 ? 2 : 1;
    if (const char *e = getenv("B2FFT_SLAB_V2_GROUPS")) ng = std::max(1, atoi(e));
    if (s.n0 % s.P != 0) ng = 1;
    while (ng > 1 && (s.b0 % ng != 0 || s.b0 / ng < 8)) --ng;
    s.v2_ng = ng;
    s.v2_gp = s.ln0 / ng;
    if (s.P > B2_MAXSCAT || (int)s.peerR.size() != s.P) return false;
    const ptrdiff_t w = s.v2_w;
    b2fft_geom g;
    if (s.ln0 > 0) {  // n1 pass of one (group, chunk): reads W[gp][n1][n2], stores chunk-major into R_k / S_k
        memset(&g, 0, sizeof g);
        g.kind = B2FFT_C2C;
        g.sign = s.sign;
        g.rank = 1;
        g.dims[0] = b2fft_dim{s.n1, s.n2, s.n2};
        g.vrank = 2;
        g.vdims[0] = b2fft_dim{w, 1, 1};
        g.vdims[1] = b2fft_dim{s.v2_gp, s.n1 * s.n2, s.n1 * s.n2};
        g.in_place = 1;
        s.v2_n1 = b2fft_plan_create(&g);
        if (!s.v2_n1) return false;
        std::vector<long long> ps(s.P), rs(s.P, (long long)w * 2);
        for (int o = 0; o < s.P; ++o) ps[o] = (long long)(o == s.pe ? s.ln1 : s.n1) * w * 2;
        if (b2_plan_enable_scatter(s.v2_n1, s.P, (long long)s.b1, (long long)s.n1 * s.n2 * 2, ps.data(), -1, rs.data(), 1) != 0)
            return false;  // no tile kernel with the scattered store for this n1: round-1 schedule
    }
    if (s.ln1 > 0) {  // n0 pass of one chunk, R_k[n0][ln1][w] -> out[ln1][n0][n2]
        memset(&g, 0, sizeof g);
        g.kind = B2FFT_C2C;
        g.sign = s.sign;
        g.rank = 1;
        g.dims[0] = b2fft_dim{s.n0, s.ln1 * w, s.n2};
        g.vrank = 2;
        g.vdims[0] = b2fft_dim{w, 1, 1};
        g.vdims[1] = b2fft_dim{s.ln1, w, s.n0 * s.n2};
        g.in_place = 0;
        s.v2_n0 = b2fft_plan_create(&g);
        if (!s.v2_n0) return false;
    }
    if (s.P > 1 && !slab_ensure_S(s)) return false;
    bool ok = cudaStreamCreateWithFlags(&s.sb, cudaStreamNonBlocking) == cudaSuccess;
    // opt-in: with the n0 passes on their own stream all n1 passes run first and compete with the whole exchange
    // for HBM -- measured slower on 8 B200 (19.3 against 16.9 ms, profiles/r2_slab_8gpu.md)
    if (getenv("B2FFT_SLAB_V2_SPLIT")) ok = ok && cudaStreamCreateWithFlags(&s.s2, cudaStreamNonBlocking) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&s.v2_e3, cudaEventDisableTiming) == cudaSuccess;
    if (nch > 64) s.v2_flags = false;
    s.v2_e1.assign((size_t)nch * ng, nullptr);
    s.v2_ec.assign(nch, nullptr);
    s.v2_e2.assign(nch, nullptr);
    for (std::vector<cudaEvent_t> *v : {&s.v2_e1, &s.v2_ec, &s.v2_e2})
        for (cudaEvent_t &e : *v) ok = ok && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&s.v2_eb, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreate(&s.v2_ta) == cudaSuccess && cudaEventCreate(&s.v2_tb) == cudaSuccess;
    return ok;
}
// the row pass over the whole slab (contiguous dimension), in -> W; built on first use because in-placeness is
// only known at execute time
static b2fft_plan *slab_v2_row_plan(SlabPlan &s, bool inplace) {
    b2fft_plan *&p = s.v2_row[inplace ? 1 : 0];
    if (p || s.ln0 <= 0) return p;
    b2fft_geom g;
    memset(&g, 0, sizeof g);
    g.kind = B2FFT_C2C;
    g.sign = s.sign;
    g.rank = 1;
    g.dims[0] = b2fft_dim{s.n2f, s.hm, s.hm};
    g.vrank = s.hm > 1 ? 2 : 1;
    g.vdims[0] = b2fft_dim{s.v2_gp * s.n1, s.n2, s.n2};  // one group of planes
    g.vdims[1] = b2fft_dim{s.hm, 1, 1};
    g.in_place = inplace ? 1 : 0;
    p = b2fft_plan_create(&g);
    return p;
}
#endif

// ------------------------------------------------------------------ exchange
// forward: send from `sb` (stage-1 output layout) and receive into `rb` (recv layout);
// reverse: the same messages with the roles swapped (used for natural-order output).
static int slab_exchange(SlabPlan &s, int chunk, cpx_t *sb, cpx_t *rb, bool reverse, stream_t st) {
    std::vector<SlabMsg> sends, recvs;
    slab_messages(s, chunk, sends, recvs);
    if (sends.empty() && recvs.empty()) return 0;
    if (s.P == 1) {
        // one rank: both lists have one entry per plane, in the same order
        for (size_t i = 0; i < sends.size() && i < recvs.size(); ++i) {
            const void *src = reverse ? (const void *)(rb + recvs[i].roff) : (const void *)(sb + sends[i].soff);
            void *dst = reverse ? (void *)(sb + sends[i].soff) : (void *)(rb + recvs[i].roff);
#if B2_CUDA
            B2_CK(cudaMemcpyAsync(dst, src, (size_t)sends[i].cnt * sizeof(cpx_t), cudaMemcpyDeviceToDevice, st));
#else
            (void)st;
            memcpy(dst, src, (size_t)sends[i].cnt * sizeof(cpx_t));
#endif
        }
        return 0;
    }
#if B2_CUDA
    if (!s.comm->nc) {
        set_error("slab exchange: this communicator has no NCCL context (detached test communicator)");
        return -5;
    }
    NcclApi *a = nccl();
    const int dt = sizeof(real_t) == 4 ? 7 : 8;  // ncclFloat32 / ncclFloat64
    B2_NCCL(a->GroupStart());
    const bool skip_self = s.hybrid && !reverse && sb == s.S;  // the n1 pass stored the local block itself
    for (const SlabMsg &m : (reverse ? recvs : sends)) {
        if (skip_self && m.peer == s.pe) continue;
        B2_NCCL(a->Send(reverse ? (const void *)(rb + m.roff) : (const void *)(sb + m.soff), (size_t)m.cnt * 2, dt,
                        m.peer, s.comm->nc, st));
    }
    for (const SlabMsg &m : (reverse ? sends : recvs)) {
        if (skip_self && m.peer == s.pe) continue;
        B2_NCCL(a->Recv(reverse ? (void *)(sb + m.soff) : (void *)(rb + m.roff), (size_t)m.cnt * 2, dt, m.peer,
                        s.comm->nc, st));
    }
    B2_NCCL(a->GroupEnd());
    return 0;
#else
    set_error("slab exchange over more than one rank needs NCCL (emulation build: use the staged test API)");
    return -5;
#endif
}

#if B2_CUDA
// ---- direct exchange through peer memory (CUDA IPC over NVLink) ----------------------------
// Each rank maps every peer's send/receive buffer once, at plan time.  The exchange of a chunk
// is then one strided DMA per peer (cudaMemcpy2DAsync straight from the stage-1 output into the
// peer's receive buffer at its final position): no packing, no SM time, and it overlaps the
// FFTs of the next chunk.  NCCL is used only for the handle all-gather and as a barrier.
static int slab_barrier(SlabPlan &s, cudaStream_t st) {
    if (s.P == 1 || !s.comm->nc) return 0;
    B2_NCCL(nccl()->AllReduce(s.flag, s.flag, 1, 7, 0, s.comm->nc, st));
    return 0;
}
static int slab_setup_p2p(SlabPlan &s) {
    s.p2p = false;
    if (s.P == 1 || !s.comm->nc) return 0;
    NcclApi *a = nccl();
    B2_CK(cudaMalloc(&s.flag, 256));
    B2_CK(cudaMemset(s.flag, 0, 256));
    struct H {
        cudaIpcMemHandle_t s, r, f;
        int ok, has_s, has_f;
    };
    H mine;
    memset(&mine, 0, sizeof mine);
    if (cudaMalloc(&s.F, B2_SLAB_FLAG_BYTES) == cudaSuccess && cudaMemset(s.F, 0, B2_SLAB_FLAG_BYTES) == cudaSuccess &&
        cudaMalloc(&s.epoch_dev, 256) == cudaSuccess && cudaIpcGetMemHandle(&mine.f, s.F) == cudaSuccess)
        mine.has_f = 1;
    cudaGetLastError();
    mine.has_s = s.S != nullptr;  // the v2 schedule has no send buffer
    mine.ok = !getenv("B2FFT_SLAB_SENDRECV") && (!s.S || cudaIpcGetMemHandle(&mine.s, s.S) == cudaSuccess) &&
              cudaIpcGetMemHandle(&mine.r, s.R) == cudaSuccess;
    cudaGetLastError();
    H *dev = nullptr;
    B2_CK(cudaMalloc(&dev, sizeof(H) * s.P));
    B2_CK(cudaMemcpy(dev + s.pe, &mine, sizeof(H), cudaMemcpyHostToDevice));
    B2_NCCL(a->AllGather(dev + s.pe, dev, sizeof(H), 0, s.comm->nc, s.sc));
    B2_CK(cudaStreamSynchronize(s.sc));
    std::vector<H> all(s.P);
    B2_CK(cudaMemcpy(all.data(), dev, sizeof(H) * s.P, cudaMemcpyDeviceToHost));
    cudaFree(dev);
    bool ok = true;
    for (int p = 0; p < s.P; ++p) ok = ok && all[p].ok;
    s.peerS.assign(s.P, nullptr);
    s.peerR.assign(s.P, nullptr);
    s.peerF.assign(s.P, nullptr);
    bool flags_ok = true;
    for (int p = 0; p < s.P; ++p) flags_ok = flags_ok && all[p].has_f;
    float good = 1.0f;
    if (ok)
        for (int p = 0; p < s.P; ++p) {
            if (p == s.pe) {
                s.peerS[p] = s.S;
                s.peerR[p] = s.R;
                s.peerF[p] = s.F;
                continue;
            }
            if (flags_ok && cudaIpcOpenMemHandle((void **)&s.peerF[p], all[p].f, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError();
                good = 0.0f;
                break;
            }
            if ((all[p].has_s && cudaIpcOpenMemHandle((void **)&s.peerS[p], all[p].s, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) ||
                cudaIpcOpenMemHandle((void **)&s.peerR[p], all[p].r, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError();
                good = 0.0f;
                break;
            }
        }
    else
        good = 0.0f;
    // every rank must take the same path: agree on the minimum
    float bad = 1.0f - good;
    B2_CK(cudaMemcpy(s.flag, &bad, sizeof(float), cudaMemcpyHostToDevice));
    B2_NCCL(a->AllReduce(s.flag, s.flag, 1, 7, 0, s.comm->nc, s.sc));
    B2_CK(cudaStreamSynchronize(s.sc));
    B2_CK(cudaMemcpy(&bad, s.flag, sizeof(float), cudaMemcpyDeviceToHost));
    B2_CK(cudaMemset(s.flag, 0, 256));
    s.p2p = (bad == 0.0f);
    // opt-in: measured slower on 8 B200 (profiles/r2_slab_8gpu.md: the 4-byte peer copies cost ~4 ms per execute)
    s.v2_flags = s.p2p && flags_ok && stream_wait_value() != nullptr && getenv("B2FFT_SLAB_V2_FLAGS") != nullptr;
    // One stream per peer (copies to different peers on different copy engines) is opt-in: on 8 B200
    // (profiles/r1_slab_8gpu.md, 2048^3) the shifted schedule p = pe + d on ONE stream keeps every
    // receiver paired with one sender at a time and exposes 6.0 ms of exchange, seven concurrent
    // streams expose 9.0 ms (incast at the receivers); on 2 GPUs the difference is under 2 %.
    const char *ps = getenv("B2FFT_SLAB_PEER_STREAMS");
    if (s.p2p && ps && atoi(ps) > 1) {
        const int ns = std::min(atoi(ps), s.P);  // peer offset d copies on stream d % ns
        s.sx.assign(ns, nullptr);
        s.evx.assign(ns, nullptr);
        for (int d = 0; d < ns; ++d) {
            B2_CK(cudaStreamCreateWithFlags(&s.sx[d], cudaStreamNonBlocking));
            B2_CK(cudaEventCreateWithFlags(&s.evx[d], cudaEventDisableTiming));
        }
        B2_CK(cudaEventCreateWithFlags(&s.evb, cudaEventDisableTiming));
    }
    return 0;
}
// `after`: event the copies must wait for (the chunk's FFTs); with per-peer streams every copy runs on
// its peer's stream and `st` later joins them (slab_join_peers)
static int slab_exchange_p2p(SlabPlan &s, int chunk, bool reverse, cudaStream_t st0, cudaEvent_t after = nullptr,
                             const cpx_t *fwd_src = nullptr) {
    const size_t esz = sizeof(cpx_t);
    if (!fwd_src) fwd_src = s.S;
    for (int d = 0; d < s.P; ++d) {
        const int p = (s.pe + d) % s.P;
        cudaStream_t st = (after && !s.sx.empty()) ? s.sx[d % (int)s.sx.size()] : st0;
        if (after && d < std::max<int>(1, (int)s.sx.size())) B2_CK(cudaStreamWaitEvent(st, after, 0));
        if (!reverse) {
            // my planes [a,b) of this chunk -> peer p's receive buffer R_p[s0 + i0][*][*]
            if (s.hybrid && p == s.pe && fwd_src == s.S) continue;  // the n1 pass stored the local block itself
            const ptrdiff_t ln1p = blk_len(s.n1, s.b1, p);
            const ptrdiff_t a = std::min<ptrdiff_t>(s.ln0, chunk * s.chunk), b = std::min<ptrdiff_t>(s.ln0, a + s.chunk);
            if (ln1p <= 0 || b <= a) continue;
            B2_CK(cudaMemcpy2DAsync(s.peerR[p] + (size_t)((s.s0 + a) * ln1p) * s.n2, (size_t)ln1p * s.n2 * esz,
                                    fwd_src + (size_t)(a * s.n1 + p * s.b1) * s.n2, (size_t)s.n1 * s.n2 * esz,
                                    (size_t)ln1p * s.n2 * esz, (size_t)(b - a), cudaMemcpyDeviceToDevice, st));
        } else {
            // my rows of peer p's planes [a,b) -> peer p's send buffer S_p[i0][s1 + j1][*]
            const ptrdiff_t ln0p = blk_len(s.n0, s.b0, p);
            const ptrdiff_t a = std::min<ptrdiff_t>(ln0p, chunk * s.chunk), b = std::min<ptrdiff_t>(ln0p, a + s.chunk);
            if (s.ln1 <= 0 || b <= a) continue;
            B2_CK(cudaMemcpy2DAsync(s.peerS[p] + (size_t)(a * s.n1 + s.s1) * s.n2, (size_t)s.n1 * s.n2 * esz,
                                    s.R + (size_t)((p * s.b0 + a) * s.ln1) * s.n2, (size_t)s.ln1 * s.n2 * esz,
                                    (size_t)s.ln1 * s.n2 * esz, (size_t)(b - a), cudaMemcpyDeviceToDevice, st));
        }
    }
    return 0;
}
static int slab_join_peers(SlabPlan &s, cudaStream_t st) {
    for (size_t d = 0; d < s.sx.size(); ++d) {
        B2_CK(cudaEventRecord(s.evx[d], s.sx[d]));
        B2_CK(cudaStreamWaitEvent(st, s.evx[d], 0));
    }
    return 0;
}
#endif

// stage 0: local (n1 x n2) FFTs of chunk k, in -> S
static int slab_stage1_chunk(SlabPlan &s, int k, cpx_t *in, stream_t st) {
    const ptrdiff_t a = std::min<ptrdiff_t>(s.ln0, k * s.chunk), b = std::min<ptrdiff_t>(s.ln0, a + s.chunk);
    if (b <= a) return 0;
    const ptrdiff_t cnt = b - a;
    b2fft_plan *p = (s.p1 && cnt == std::min(s.chunk, s.ln0)) ? s.p1 : s.p1_tail;
    if (!p) return -6;
    const size_t off = (size_t)a * s.n1 * s.n2;
    if (s.fused || s.hybrid) {
        std::vector<real_t *> bases(s.P);
        for (int o = 0; o < s.P; ++o) bases[o] = (real_t *)s.peerR[o];
        return b2_plan_execute_scatter(p, in + off, s.S + off, bases.data(), (long long)(s.s0 + a), st);
    }
    return b2fft_execute_async(p, in + off, s.S + off, st);
}

// test hook (also usable by a caller with its own peer mapping): the receive buffers of all ranks as
// seen from this process; switches stage 0 to the fused scatter.  Returns 1 when fused.
extern "C" int b2fft_slab_set_peer_recv(b2fft_plan *p, void *const *recv, int n) {
    if (!p || !p->slab || n != p->slab->P) return -1;
    SlabPlan &s = *p->slab;
    s.peerR.assign(n, nullptr);
    for (int o = 0; o < n; ++o) s.peerR[o] = (cpx_t *)recv[o];
    slab_enable_fused(s);
    slab_describe(s, p);
    return s.fused ? 1 : 0;
}

// ---- staged API: lets a test (or a caller with its own transport) drive the phases ----
//   stage 0: all local (n1 x n2) FFTs, in -> send buffer
//   stage 1: exchange send -> recv buffer (NCCL, or a copy on one rank)
//   stage 2: length-n0 FFTs, recv -> out (transposed output) or in place on recv (natural)
//   stage 3: natural output only: exchange recv -> out
extern "C" int b2fft_slab_stage(b2fft_plan *p, int stage, void *in, void *out) {
    if (!p || !p->slab) return -1;
    DeviceScope caller;
    if (ensure_device_of(p->device) != 0) return -1;
    SlabPlan &s = *p->slab;
    if (!slab_ensure_S(s)) return -3;
    stream_t st = nullptr;
#if B2_CUDA
    st = s.sc;
#endif
    int rc = 0;
    if (s.kind == B2FFT_C2R) {
        // c2r: 0 forward exchange in -> recv (natural input only), 1 backward n0 FFTs, 2 reverse
        // exchange recv -> send, 3 local 2-D c2r send -> out
        if (stage == 0) {
            for (int k = 0; k < s.nchunks && !rc && !s.c2r_tin; ++k) rc = slab_exchange(s, k, (cpx_t *)in, s.R, false, st);
        } else if (stage == 1) {
            if (s.pA) rc = b2fft_execute_async(s.pA, s.c2r_tin ? in : (void *)s.R, s.R, st);
        } else if (stage == 2) {
            for (int k = 0; k < s.nchunks && !rc; ++k) rc = slab_exchange(s, k, s.S, s.R, true, st);
        } else if (stage == 3) {
            if (s.pB) rc = b2fft_execute_async(s.pB, s.S, out, st);
        } else
            rc = -1;
    } else if (stage == 0) {
        for (int k = 0; k < s.nchunks && !rc; ++k) rc = slab_stage1_chunk(s, k, (cpx_t *)in, st);
    } else if (stage == 1) {
        for (int k = 0; k < s.nchunks && !rc && !s.fused; ++k) rc = slab_exchange(s, k, s.S, s.R, false, st);
    } else if (stage == 2) {
        if (s.p2) rc = b2fft_execute_async(s.p2, s.R, s.transposed_out ? out : (void *)s.R, st);
    } else if (stage == 3) {
        if (!s.transposed_out)
            for (int k = 0; k < s.nchunks && !rc; ++k) rc = slab_exchange(s, k, (cpx_t *)out, s.R, true, st);
    } else
        rc = -1;
#if B2_CUDA
    if (!rc) B2_CK(cudaStreamSynchronize(st));
#endif
    return rc;
}
extern "C" void *b2fft_slab_buffer(b2fft_plan *p, int which, size_t *bytes) {
    if (!p || !p->slab) return nullptr;
    SlabPlan &s = *p->slab;
    if (which == 0 && !slab_ensure_S(s)) return nullptr;
    if (bytes) *bytes = (which == 0 ? s.s_elems : s.r_elems) * sizeof(cpx_t);
    return which == 0 ? (void *)s.S : (void *)s.R;
}

#if B2_CUDA
// v2 schedule (see SlabPlan): din / dout are device arrays, W = dout.
//   compute stream sc :  row(g0) | n1(g0,0) .. n1(g0,nch-1) | row(g1) | n1(g1,0) .. n1(g1,nch-1)   (never waits for a peer)
//   copy stream sn    :  per (group, chunk), one 2-D copy per peer straight into the peer's receive buffer; behind the
//                        last group's copies of chunk k, one 4-byte copy per peer: "my part of chunk k has landed"
//   barrier stream sb :  per chunk, after the last group's copies: "every rank's copies of this chunk have landed"
//                        (1-float NCCL all-reduce); the n0 pass of chunk k is queued on sc v2_lag chunks behind n1(k)
//   Opt-in variants, both measured slower on 8 GPUs: B2FFT_SLAB_V2_FLAGS (epoch flags in peer memory + cuStreamWaitValue32
//   instead of the barrier kernels), B2FFT_SLAB_V2_SPLIT (n0 passes on their own stream s2).
static int slab_execute_v2(SlabPlan &s, cpx_t *din, cpx_t *dout) {
    const size_t esz = sizeof(cpx_t);
    const int nch = s.v2_nch, ng = s.v2_ng, P = s.P;
    const size_t w = (size_t)s.v2_w, gp = (size_t)s.v2_gp;
    const bool flags = s.v2_flags && P > 1;
    cudaStream_t sq = s.s2 ? s.s2 : s.sc;  // where the n0 passes run
    StreamWaitValueFn wait_value = flags ? stream_wait_value() : nullptr;
    auto wait_flag = [&](cudaStream_t st, size_t idx, unsigned value) -> int {
        const CUresult r = wait_value((CUstream)st, (CUdeviceptr)(uintptr_t)(s.F + idx), value, CU_STREAM_WAIT_VALUE_GEQ);
        if (r != CUDA_SUCCESS) {
            set_error("cuStreamWaitValue32 failed (%d)", (int)r);
            return 2000 + (int)r;
        }
        return 0;
    };
    cpx_t *W = dout;
    ++s.epoch;
    B2_CK(cudaEventRecord(s.t0, s.sc));
    // nobody may still be reading the receive buffer that this execute is about to overwrite
    if (P > 1 && flags) {
        B2_CK(cudaMemcpyAsync(s.epoch_dev, &s.epoch, sizeof(unsigned), cudaMemcpyHostToDevice, s.sn));
        for (int pe = 0; pe < P; ++pe)
            if (pe != s.pe) {
                int rc = wait_flag(s.sn, (size_t)nch * P + pe, s.epoch - 1);
                if (rc) return rc;
            }
    } else if (P > 1) {
        int rc = slab_barrier(s, s.sb);
        if (rc) return rc;
        B2_CK(cudaEventRecord(s.v2_eb, s.sb));
        B2_CK(cudaStreamWaitEvent(s.sn, s.v2_eb, 0));
    }
    b2fft_plan *row = s.ln0 > 0 ? slab_v2_row_plan(s, din == dout) : nullptr;
    if (s.ln0 > 0 && !row) return -6;
    auto n0_pass = [&](int k) -> int {
        if (flags) {
            B2_CK(cudaStreamWaitEvent(sq, s.v2_e1[(size_t)(ng - 1) * nch + k], 0));  // this rank's own rows of the chunk
            for (int pe = 0; pe < P; ++pe)
                if (pe != s.pe) {
                    int rc = wait_flag(sq, (size_t)k * P + pe, s.epoch);
                    if (rc) return rc;
                }
        } else
            B2_CK(cudaStreamWaitEvent(sq, s.v2_e2[k], 0));
        if (!s.v2_n0) return 0;
        return b2fft_execute_async(s.v2_n0, s.R + (size_t)k * s.n0 * s.ln1 * w, dout + (size_t)k * w, sq);
    };
    const int lag = s.s2 ? 0 : s.v2_lag;
    for (int g = 0; g < ng; ++g) {
        const size_t p0 = (size_t)g * gp;  // first plane of the group
        if (row) {
            int rc = b2fft_execute_async(row, din + p0 * s.n1 * s.n2, W + p0 * s.n1 * s.n2, s.sc);
            if (rc) return rc;
        }
        if (g == 0) B2_CK(cudaEventRecord(s.v2_ta, s.sc));
        for (int k = 0; k < nch; ++k) {
            cpx_t *Sk = s.S ? s.S + (size_t)k * s.ln0 * s.n1 * w + p0 * s.n1 * w : nullptr;
            if (s.v2_n1) {
                cpx_t *wk = W + p0 * s.n1 * s.n2 + (size_t)k * w;
                real_t *bases[B2_MAXSCAT];
                for (int o = 0; o < P; ++o)
                    bases[o] = o == s.pe ? (real_t *)(s.R + (size_t)k * s.n0 * s.ln1 * w + ((size_t)s.s0 + p0) * s.ln1 * w)
                                         : (real_t *)(Sk + (size_t)o * s.b1 * w);
                int rc = b2_plan_execute_scatter(s.v2_n1, wk, wk, bases, 0, s.sc);
                if (rc) return rc;
            }
            cudaEvent_t e1 = s.v2_e1[(size_t)g * nch + k];
            B2_CK(cudaEventRecord(e1, s.sc));
            B2_CK(cudaStreamWaitEvent(s.sn, e1, 0));
            for (int d = 1; d < P; ++d) {  // shifted order: one sender per receiver at a time
                const int pe = (s.pe + d) % P;
                const size_t ln1p = (size_t)blk_len(s.n1, s.b1, pe);
                if (ln1p > 0 && s.ln0 > 0)
                    B2_CK(cudaMemcpy2DAsync(s.peerR[pe] + (size_t)k * s.n0 * ln1p * w + ((size_t)s.s0 + p0) * ln1p * w, ln1p * w * esz,
                                            Sk + (size_t)pe * s.b1 * w, (size_t)s.n1 * w * esz, ln1p * w * esz, gp,
                                            cudaMemcpyDeviceToDevice, s.sn));
                if (flags && g == ng - 1)  // behind the data, same stream: "rank s.pe's part of chunk k is in your receive buffer"
                    B2_CK(cudaMemcpyAsync(s.peerF[pe] + (size_t)k * P + s.pe, s.epoch_dev, sizeof(unsigned), cudaMemcpyDeviceToDevice, s.sn));
            }
            if (g == ng - 1) {
                if (!flags) {
                    B2_CK(cudaEventRecord(s.v2_ec[k], s.sn));
                    B2_CK(cudaStreamWaitEvent(s.sb, s.v2_ec[k], 0));
                    if (P > 1) {
                        int rc = slab_barrier(s, s.sb);
                        if (rc) return rc;
                    }
                    B2_CK(cudaEventRecord(s.v2_e2[k], s.sb));
                }
                if (k >= lag) {
                    int rc = n0_pass(k - lag);
                    if (rc) return rc;
                }
            }
        }
    }
    B2_CK(cudaEventRecord(s.v2_tb, s.sc));
    for (int k = std::max(0, nch - lag); k < nch; ++k) {
        int rc = n0_pass(k);
        if (rc) return rc;
    }
    if (flags)  // "this rank has finished reading its receive buffer for this epoch"
        for (int pe = 0; pe < P; ++pe)
            if (pe != s.pe)
                B2_CK(cudaMemcpyAsync(s.peerF[pe] + (size_t)nch * P + s.pe, s.epoch_dev, sizeof(unsigned), cudaMemcpyDeviceToDevice, sq));
    if (sq != s.sc) {
        B2_CK(cudaEventRecord(s.v2_e3, sq));
        B2_CK(cudaStreamWaitEvent(s.sc, s.v2_e3, 0));
    }
    B2_CK(cudaEventRecord(s.t3, s.sc));
    return 0;
}
#endif

int b2_slab_execute(b2fft_plan_s *p, void *in, void *out) {
    SlabPlan &s = *p->slab;
    if (ensure_device_of(p->device) != 0) return -1;
#if B2_CUDA
    cpx_t *din = (cpx_t *)in, *dout = (cpx_t *)out;
    const bool hin = !is_device_ptr(in), hout = !is_device_ptr(out);
    // The plan works on its own non-blocking streams.  A caller that filled its device arrays with work queued on
    // the (legacy) default stream -- what ordinary plans run on, and what e.g. torch uses -- expects fftw_execute to
    // see that data: order the plan's first stream after everything queued there so far.
    if (!s.ev_in) B2_CK(cudaEventCreateWithFlags(&s.ev_in, cudaEventDisableTiming));
    B2_CK(cudaEventRecord(s.ev_in, 0));
    B2_CK(cudaStreamWaitEvent(s.sc, s.ev_in, 0));
    // (the padded real side of r2c / c2r has 2*n2 reals per row: the same bytes as n2 complex)
    const size_t in_b = (size_t)(s.c2r_tin ? s.ln1 * s.n0 : s.ln0 * s.n1) * s.n2 * sizeof(cpx_t);
    const size_t out_b = (size_t)(s.transposed_out && s.kind != B2FFT_C2R ? s.ln1 * s.n0 : s.ln0 * s.n1) * s.n2 * sizeof(cpx_t);
    if (hin) {
        if (!s.dev_in) s.dev_in = (cpx_t *)dev_malloc(std::max<size_t>(in_b, 16));
        if (!s.dev_in) return -3;
        din = s.dev_in;
        B2_CK(cudaMemcpyAsync(din, in, in_b, cudaMemcpyHostToDevice, s.sc));
    }
    if (hout) {
        // (the v2 schedule also works in this array: it must hold the stage-1 layout too)
        if (!s.dev_out) s.dev_out = (cpx_t *)dev_malloc(std::max<size_t>(std::max(out_b, s.v2 ? in_b : (size_t)0), 16));
        if (!s.dev_out) return -3;
        dout = s.dev_out;
    }
    if (s.v2) {
        int rc = slab_execute_v2(s, din, dout);
        if (rc) return rc;
        if (hout) B2_CK(cudaMemcpyAsync(out, dout, out_b, cudaMemcpyDeviceToHost, s.sc));
        B2_CK(cudaStreamSynchronize(s.sc));
        B2_CK(cudaStreamSynchronize(s.sn));
        B2_CK(cudaStreamSynchronize(s.sb));
        if (s.s2) B2_CK(cudaStreamSynchronize(s.s2));
        // row pass | pipelined n1 passes, exchange and n0 passes | the n0 passes left after the last n1 pass
        cudaEventElapsedTime(&s.ms[0], s.t0, s.v2_ta);
        cudaEventElapsedTime(&s.ms[1], s.v2_ta, s.v2_tb);
        cudaEventElapsedTime(&s.ms[2], s.v2_tb, s.t3);
        return 0;
    }
    if (!slab_ensure_S(s)) return -3;
    B2_CK(cudaEventRecord(s.t0, s.sc));
    if (s.kind == B2FFT_C2R) {
        int rc = 0;
        if (s.p2p) rc = slab_barrier(s, s.sc);  // every rank has finished reading the buffers of its last execute
        if (!s.c2r_tin) {
            for (int k = 0; k < s.nchunks && !rc; ++k)
                rc = s.p2p ? slab_exchange_p2p(s, k, false, s.sc, nullptr, din) : slab_exchange(s, k, din, s.R, false, s.sc);
            if (!rc && s.p2p) rc = slab_barrier(s, s.sc);
        }
        if (!rc && s.pA) rc = b2fft_execute_async(s.pA, s.c2r_tin ? (void *)din : (void *)s.R, s.R, s.sc);
        if (rc) return rc;
        B2_CK(cudaEventRecord(s.t1, s.sc));
        for (int k = 0; k < s.nchunks && !rc; ++k)
            rc = s.p2p ? slab_exchange_p2p(s, k, true, s.sc) : slab_exchange(s, k, s.S, s.R, true, s.sc);
        if (!rc && s.p2p) rc = slab_barrier(s, s.sc);
        if (rc) return rc;
        B2_CK(cudaEventRecord(s.t2, s.sc));
        if (s.pB) rc = b2fft_execute_async(s.pB, s.S, dout, s.sc);
        if (rc) return rc;
        B2_CK(cudaEventRecord(s.t3, s.sc));
        if (hout) B2_CK(cudaMemcpyAsync(out, dout, out_b, cudaMemcpyDeviceToHost, s.sc));
        B2_CK(cudaStreamSynchronize(s.sc));
        cudaEventElapsedTime(&s.ms[0], s.t0, s.t1);
        cudaEventElapsedTime(&s.ms[1], s.t1, s.t2);
        cudaEventElapsedTime(&s.ms[2], s.t2, s.t3);
        return 0;
    }
    if (s.fused) {
        // one stream: [barrier: nobody still reads the receive buffers] row pass, n1 pass + scatter, per
        // chunk; [barrier: every rank's rows have landed]
        int rc = slab_barrier(s, s.sc);
        if (rc) return rc;
        for (int k = 0; k < s.nchunks; ++k) {
            rc = slab_stage1_chunk(s, k, din, s.sc);
            if (rc) return rc;
        }
        rc = slab_barrier(s, s.sc);
        if (rc) return rc;
        B2_CK(cudaEventRecord(s.t1, s.sc));
        B2_CK(cudaEventRecord(s.t2, s.sc));
    } else {
    if (s.p2p) {
        // nobody may still be reading the receive buffer this exchange is about to overwrite
        int rc = slab_barrier(s, s.sn);
        if (rc) return rc;
        if (!s.sx.empty()) {
            B2_CK(cudaEventRecord(s.evb, s.sn));
            for (cudaStream_t x : s.sx) B2_CK(cudaStreamWaitEvent(x, s.evb, 0));
        }
    }
    for (int k = 0; k < s.nchunks; ++k) {
        int rc = slab_stage1_chunk(s, k, din, s.sc);
        if (rc) return rc;
        B2_CK(cudaEventRecord(s.ev[k], s.sc));
        if (s.p2p && !s.sx.empty())
            rc = slab_exchange_p2p(s, k, false, s.sn, s.ev[k]);
        else {
            B2_CK(cudaStreamWaitEvent(s.sn, s.ev[k], 0));
            rc = s.p2p ? slab_exchange_p2p(s, k, false, s.sn) : slab_exchange(s, k, s.S, s.R, false, s.sn);
        }
        if (rc) return rc;
    }
    if (s.p2p) {
        // all pushes into my receive buffer have landed once every rank has passed this point
        int rc = slab_join_peers(s, s.sn);
        if (!rc) rc = slab_barrier(s, s.sn);
        if (rc) return rc;
    }
    B2_CK(cudaEventRecord(s.t1, s.sc));
    B2_CK(cudaEventRecord(s.ev[s.nchunks], s.sn));
    B2_CK(cudaStreamWaitEvent(s.sc, s.ev[s.nchunks], 0));
    B2_CK(cudaEventRecord(s.t2, s.sc));
    }
    if (s.p2) {
        int rc = b2fft_execute_async(s.p2, s.R, s.transposed_out ? (void *)dout : (void *)s.R, s.sc);
        if (rc) return rc;
    }
    if (!s.transposed_out) {
        if (s.p2p) {
            for (int k = 0; k < s.nchunks; ++k) {
                int rc = slab_exchange_p2p(s, k, true, s.sc);
                if (rc) return rc;
            }
            int rc = slab_barrier(s, s.sc);
            if (rc) return rc;
            B2_CK(cudaMemcpyAsync(dout, s.S, (size_t)s.ln0 * s.n1 * s.n2 * sizeof(cpx_t), cudaMemcpyDeviceToDevice, s.sc));
        } else
            for (int k = 0; k < s.nchunks; ++k) {
                int rc = slab_exchange(s, k, dout, s.R, true, s.sc);
                if (rc) return rc;
            }
    }
    B2_CK(cudaEventRecord(s.t3, s.sc));
    if (hout) B2_CK(cudaMemcpyAsync(out, dout, out_b, cudaMemcpyDeviceToHost, s.sc));
    B2_CK(cudaStreamSynchronize(s.sc));
    cudaEventElapsedTime(&s.ms[0], s.t0, s.t1);
    cudaEventElapsedTime(&s.ms[1], s.t1, s.t2);
    cudaEventElapsedTime(&s.ms[2], s.t2, s.t3);
    return 0;
#else
    for (int st = 0; st < 4; ++st) {
        int rc = b2fft_slab_stage(p, st, in, out);
        if (rc) return rc;
    }
    return 0;
#endif
}

extern "C" int b2fft_slab_last_times(const b2fft_plan *p, float *a, float *b, float *c) {
    if (!p || !p->slab) return -1;
    if (a) *a = p->slab->ms[0];
    if (b) *b = p->slab->ms[1];
    if (c) *c = p->slab->ms[2];
    return 0;
}
