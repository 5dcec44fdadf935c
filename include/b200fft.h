/* include/b200fft.h -- the thin C-ABI between the plain-C host side (csrc/fftw_api.c, the
 * drop-in replacement of the reference's fftw/fftw_api.c public layer) and the CUDA engine
 * (csrc/engine.cu: kernel-configuration selector + hand-written sm_100a kernels).
 *
 * Plain pointers and sizes only; no C++ or torch types.  One shared object is built per
 * precision (like the reference: FFTW_SINGLE is a compile-time switch, fftw/fftw3.h:65-82),
 * so "real" below means float in lib/fp32 and double in lib/fp64.
 *
 * What each entry point replaces in the reference:
 *   b2fft_plan_create   <- fftw_mkproblem_dft_d / fftw_mkproblem_rdft2_d_3pointers +
 *                          fftw_mkapiplan (fftw/fftw_api.c:4110, 9358, 317-407): geometry in,
 *                          executable plan out.  The recursive planner (fftw_api.c:14745-15710)
 *                          is replaced by a deterministic selector.
 *   b2fft_execute       <- fftw_execute / fftw_dft_solve / fftw_rdft2_solve
 *                          (fftw/fftw_api.c:428-454, 4589, 10577)
 *   b2fft_plan_destroy  <- fftw_destroy_plan (fftw/fftw_api.c:409-422)
 *   b2fft_malloc/free   <- fftw_malloc / fftw_free (fftw/fftw_api.c:234-256, 14448)
 *   b2fft_slab_*        <- fftw/mpi: local_size_guru (mpi/api.c:248-346), default_block
 *                          (mpi/block.c:39-50), dft-rank-geq2-transposed.c:47-198,
 *                          transpose-alltoall.c:49-104 (MPI_Alltoall -> NCCL over NVLink)
 */
#ifndef B200FFT_H
#define B200FFT_H
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2FFT_MAXRANK 8

enum { B2FFT_C2C = 0, B2FFT_R2C = 1, B2FFT_C2R = 2 };

/* One loop dimension.  Strides are in ELEMENTS of the array they index: complex elements for
 * a complex array, reals for a real array (the public "many" convention, fftw_api.c:655-661;
 * the reference's internal doubling of complex strides, fftw_api.c:842-861, is not exposed). */
typedef struct { ptrdiff_t n, is, os; } b2fft_dim;

typedef struct {
    int kind;                       /* B2FFT_C2C / R2C / C2R */
    int sign;                       /* -1 forward, +1 backward (C2C only; R2C is -1, C2R +1) */
    int rank;                       /* transform dims, outermost first; n = LOGICAL size */
    b2fft_dim dims[B2FFT_MAXRANK];
    int vrank;                      /* batch ("howmany") dims */
    b2fft_dim vdims[B2FFT_MAXRANK];
    int in_place;                   /* in == out at plan time */
    unsigned flags;                 /* FFTW_* planner flags, passed through */
} b2fft_geom;

typedef struct b2fft_plan_s b2fft_plan;

/* NULL when the geometry is unsupported or the device cannot be initialised; the reason is
 * then available from b2fft_last_error(). */
b2fft_plan *b2fft_plan_create(const b2fft_geom *g);
void        b2fft_plan_destroy(b2fft_plan *p);

/* FFTW_MEASURE / PATIENT / EXHAUSTIVE (reference: patience loop fftw/fftw_api.c:351-376, cost by time :15106-15153).
 * cfg = B2FFT_CFG_MEASURE: build the default plan and every alternative configuration the selector knows for this
 * geometry (TMA or round-1 column kernels, pipelined or not, two or three passes of a long axis, the real-transform
 * variants), time each on plan-owned zero-filled device arrays and keep the fastest; cfg >= 0: that configuration,
 * unmeasured (what wisdom replays); cfg = B2FFT_CFG_DEFAULT: b2fft_plan_create.  b2fft_plan_config = the
 * configuration of a plan (a bit mask; 0 = the selector's default), b2fft_plan_measured = 1 when it was timed. */
#define B2FFT_CFG_DEFAULT (-1)
#define B2FFT_CFG_MEASURE (-2)
b2fft_plan *b2fft_plan_create_tuned(const b2fft_geom *g, int cfg);
int         b2fft_plan_config(const b2fft_plan *p);
int         b2fft_plan_measured(const b2fft_plan *p);

/* Synchronous execute.  in/out may each be a host pointer (copied over PCIe inside the call)
 * or a device/managed pointer (transformed in HBM).  Returns 0, or a CUDA error code. */
int b2fft_execute(b2fft_plan *p, void *in, void *out);
/* Asynchronous execute on a caller-provided cudaStream_t (device pointers only). */
int b2fft_execute_async(b2fft_plan *p, void *in, void *out, void *cuda_stream);

/* Introspection: the selected kernel configurations ("passes"). */
int    b2fft_plan_num_passes(const b2fft_plan *p);     /* kernels launched per execute */
int    b2fft_plan_num_hbm_sweeps(const b2fft_plan *p); /* full read+write sweeps over the data */
size_t b2fft_plan_scratch_bytes(const b2fft_plan *p);
int    b2fft_plan_describe(const b2fft_plan *p, char *buf, size_t buflen); /* snprintf-like */
double b2fft_plan_flops(const b2fft_plan *p);          /* 5 N log2 N (c2c), 2.5 N log2 N (real) */
unsigned long long b2fft_launch_count(void);           /* kernels launched so far (process) */

/* Pinned host memory (what fftw_malloc hands out) and device memory helpers. */
void *b2fft_malloc(size_t nbytes);
void  b2fft_free(void *p);
void *b2fft_device_malloc(size_t nbytes);
void  b2fft_device_free(void *p);

int         b2fft_device_count(void);       /* 0 => no usable GPU: plans are NULL */
int         b2fft_set_device(int ordinal);  /* default: $FFTW_CUDA_DEVICE, $LOCAL_RANK or 0 */
const char *b2fft_last_error(void);
void        b2fft_cleanup(void);            /* drops cached twiddle tables / workspaces */
const char *b2fft_precision(void);          /* "fp32" or "fp64" */

/* ---- split-array transforms (separate real / imaginary arrays) ------------------------------
 * Replaces fftw_plan_guru_split_dft / _split_dft_r2c / _split_dft_c2r and fftw_execute_split_dft*
 * (reference fftw/fftw_api.c:1420-1517, 444-454).  The geometry uses the guru convention: strides
 * of every split or real array are in REALS; dims[].n are logical sizes; kind selects c2c (always
 * forward: swap the arrays for backward), r2c (real in, split out) or c2r (split in, real out).
 * Arrays may be host or device pointers.  execute: c2c (ri, ii, ro, io); r2c (in, NULL, ro, io);
 * c2r (ri, ii, out, NULL). */
typedef struct b2fft_split_s b2fft_split;
b2fft_split *b2fft_split_plan_create(const b2fft_geom *g);
void         b2fft_split_plan_destroy(b2fft_split *p);
int          b2fft_split_execute(b2fft_split *p, void *ri, void *ii, void *ro, void *io);
int          b2fft_split_describe(const b2fft_split *p, char *buf, size_t buflen);
double       b2fft_split_flops(const b2fft_split *p);

/* ---- real-to-real transforms (fftw_plan_r2r*, reference fftw/fftw_api.c:1085-1170, 11700-14160) -----
 * kinds[i] is the fftw_r2r_kind of transform dim i (FFTW_R2HC ... FFTW_RODFT11, fftw3.h:117-121);
 * strides in reals; host or device arrays. */
typedef struct b2fft_r2r_s b2fft_r2r;
b2fft_r2r *b2fft_r2r_plan_create(int rank, const b2fft_dim *dims, int vrank, const b2fft_dim *vdims, const int *kinds);
void       b2fft_r2r_plan_destroy(b2fft_r2r *p);
int        b2fft_r2r_execute(b2fft_r2r *p, void *in, void *out);
int        b2fft_r2r_describe(const b2fft_r2r *p, char *buf, size_t buflen);
double     b2fft_r2r_flops(const b2fft_r2r *p);

/* ---- slab-distributed 3-D c2c (fftw_mpi_plan_dft_3d form), one process per GPU --------- */
typedef struct b2fft_comm_s b2fft_comm;

/* ceil(n/P) block distribution, exactly fftw/mpi/block.c:39-50. */
ptrdiff_t b2fft_slab_local_size_3d_transposed(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2,
                                              int n_pes, int my_pe,
                                              ptrdiff_t *local_n0, ptrdiff_t *local_0_start,
                                              ptrdiff_t *local_n1, ptrdiff_t *local_1_start);

/* the same with explicit block sizes (0 = ceil(n/P)): fftw_mpi_local_size_many_transposed / fftw_mpi_plan_many_dft.
 * block0 distributes the plan's input dimension, block1 its post-exchange dimension. */
ptrdiff_t b2fft_slab_local_size_3d_blocks(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, ptrdiff_t block0, ptrdiff_t block1,
                                          int n_pes, int my_pe, ptrdiff_t *local_n0, ptrdiff_t *local_0_start,
                                          ptrdiff_t *local_n1, ptrdiff_t *local_1_start);

int         b2fft_nccl_unique_id(void *id128);                 /* fills 128 bytes */
b2fft_comm *b2fft_comm_create(const void *id128, int n_pes, int my_pe);
b2fft_comm *b2fft_comm_self(void);                             /* single-rank communicator */
void        b2fft_comm_destroy(b2fft_comm *c);
int         b2fft_comm_rank(const b2fft_comm *c);
int         b2fft_comm_size(const b2fft_comm *c);

b2fft_plan *b2fft_slab_plan_dft_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, ptrdiff_t howmany,
                                   b2fft_comm *c, int sign, int transposed_in,
                                   int transposed_out, unsigned flags);
b2fft_plan *b2fft_slab_plan_dft_3d_blocks(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, ptrdiff_t howmany,
                                          ptrdiff_t block0, ptrdiff_t block1, b2fft_comm *c, int sign,
                                          int transposed_in, int transposed_out, unsigned flags);
/* the receive buffers of all ranks as seen from this process (test hook / callers with their own
 * peer mapping): switches the plan to the fused exchange; returns 1 when it did */
int         b2fft_slab_set_peer_recv(b2fft_plan *p, void *const *recv, int n);
/* Real-data forms (fftw_mpi_plan_dft_r2c_3d / _c2r_3d, reference fftw/mpi/api.c:751-858): kind =
 * B2FFT_R2C (real [local_n0][n1][2(n2/2+1)] in, complex out, natural or FFTW_MPI_TRANSPOSED_OUT) or
 * B2FFT_C2R (complex in, natural or FFTW_MPI_TRANSPOSED_IN, real out).  n2 = logical real length. */
b2fft_plan *b2fft_slab_plan_rdft_3d(int kind, ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, b2fft_comm *c,
                                    int transposed_in, int transposed_out, unsigned flags);
/* the same with `howmany` interleaved transforms (tuples of howmany reals / complex per element) */
b2fft_plan *b2fft_slab_plan_many_rdft_3d(int kind, ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, ptrdiff_t howmany,
                                         b2fft_comm *c, int transposed_in, int transposed_out, unsigned flags);
/* ... and with explicit block sizes: block0 over n0 (the distribution of the REAL array and of a natural complex array), block1
 * over n1 (the distribution of the TRANSPOSED complex array); 0 = default.  Replaces the iblock / oblock arguments of
 * fftw_mpi_plan_many_dft_r2c / _c2r (reference fftw/mpi/api.c:790-858). */
b2fft_plan *b2fft_slab_plan_many_rdft_3d_blocks(int kind, ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, ptrdiff_t howmany,
                                                ptrdiff_t block0, ptrdiff_t block1, b2fft_comm *c, int transposed_in,
                                                int transposed_out, unsigned flags);
/* 2-D real data (fftw_mpi_plan_dft_r2c_2d / _c2r_2d): real [local_n0][2(n1/2+1)], complex [local_n0][n1/2+1] or,
 * transposed, [local_n1c][n0] with the n1/2+1 complex columns block-distributed. */
b2fft_plan *b2fft_slab_plan_rdft_2d(int kind, ptrdiff_t n0, ptrdiff_t n1, b2fft_comm *c, int transposed_in,
                                    int transposed_out, unsigned flags);
/* ... with explicit block sizes: block0 over n0, block1 over the n1/2+1 complex columns; 0 = default (the iblock / oblock of
 * fftw_mpi_plan_many_dft_r2c / _c2r with rnk = 2, reference fftw/mpi/api.c:790-836) */
b2fft_plan *b2fft_slab_plan_rdft_2d_blocks(int kind, ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t block0, ptrdiff_t block1, b2fft_comm *c,
                                           int transposed_in, int transposed_out, unsigned flags);
/* Staged execution, for callers that bring their own transport and for the CPU tests of the
 * multi-rank path: stage 0 local (n1 x n2) FFTs in -> send buffer; 1 exchange (NCCL); 2 length-n0
 * FFTs recv buffer -> out; 3 (natural-order output only) exchange back (c2r plans: 0 forward exchange
 * of a natural input, 1 backward n0 FFTs, 2 reverse exchange, 3 local 2-D c2r).  b2fft_slab_buffer returns
 * the send (0) / receive (1) buffer.  b2fft_comm_detached carries (rank, size) without NCCL. */
int         b2fft_slab_stage(b2fft_plan *p, int stage, void *in, void *out);
void       *b2fft_slab_buffer(b2fft_plan *p, int which, size_t *bytes);
b2fft_comm *b2fft_comm_detached(int n_pes, int my_pe);
/* device time (ms) of the last slab execute split into local FFT / exchange / local FFT */
int b2fft_slab_last_times(const b2fft_plan *p, float *ms_fft1, float *ms_exchange, float *ms_fft2);

/* ---- batched 1-D c2c that allocates nothing ----------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  A variant of the engine's
 * two-pass (n = n1 * n2) long 1-D transform whose intermediate array and twiddle table live in a file-scope
 * static __device__ array of the library (csrc/noalloc.cu) instead of plan-owned cudaMalloc'ed buffers.
 * No plan, no cudaMalloc / cudaFree, no allocating library call; asynchronous on `cuda_stream`.
 *   n        power of two, 64 <= n <= B2FFT_NOALLOC_MAX_N
 *   howmany  contiguous transforms (dist n), n * howmany <= B2FFT_NOALLOC_MAX_POINTS, howmany <= 65535
 *   in, out  DEVICE pointers to interleaved complex of the library's precision; in == out is allowed
 *   sign     -1 forward, +1 backward (unnormalised, like fftw_execute)
 * Returns 0, or non-zero with the reason in b2fft_last_error() (sizes beyond the maxima are rejected).
 * Every call shares the one array (its previous contents are never read); calls that overlap on the GPU
 * -- different streams without ordering -- race on it and must be serialised by the caller. */
#define B2FFT_NOALLOC_MAX_POINTS (1 << 20)
#define B2FFT_NOALLOC_MAX_N (1 << 16)
int b2fft_c2c_1d_static_scratch(ptrdiff_t n, ptrdiff_t howmany, const void *in, void *out, int sign, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* B200FFT_H */
