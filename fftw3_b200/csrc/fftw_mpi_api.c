/* csrc/fftw_mpi_api.c -- plain-C host side of the slab-distributed interface: the subset of
 * the reference's fftw/mpi/api.c named by the hot path (fftw_mpi_plan_dft_3d :638-645 ->
 * plan_many_dft :592 -> plan_guru_dft :560-590; local sizes :494-516; execute decl
 * fftw3-mpi.h:192), with MPI_Comm standing for an NCCL communicator (include/fftw3-mpi.h).
 */
#include <stdlib.h>
#include <string.h>

#include "fftw3-mpi.h"
#include "fftw_plan_struct.h"

void fftw_mpi_init(void) {}
void fftw_mpi_cleanup(void) { fftw_cleanup(); }

ptrdiff_t fftw_mpi_local_size_3d_transposed(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, MPI_Comm comm,
                                            ptrdiff_t *local_n0, ptrdiff_t *local_0_start,
                                            ptrdiff_t *local_n1, ptrdiff_t *local_1_start) {
    return b2fft_slab_local_size_3d_transposed(n0, n1, n2, b2fft_comm_size(comm), b2fft_comm_rank(comm), local_n0,
                                               local_0_start, local_n1, local_1_start);
}
ptrdiff_t fftw_mpi_local_size_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, MPI_Comm comm, ptrdiff_t *local_n0,
                                 ptrdiff_t *local_0_start) {
    ptrdiff_t a, b;
    return fftw_mpi_local_size_3d_transposed(n0, n1, n2, comm, local_n0, local_0_start, &a, &b);
}

/* reference fftw/mpi/api.c:348-372 (local_size_many_transposed): rank 3, default blocks, sizes in
 * units of complex elements including the howmany interleaved transforms */
ptrdiff_t fftw_mpi_local_size_many_transposed(int rnk, const ptrdiff_t *n, ptrdiff_t howmany, ptrdiff_t block0,
                                              ptrdiff_t block1, MPI_Comm comm, ptrdiff_t *local_n0,
                                              ptrdiff_t *local_0_start, ptrdiff_t *local_n1,
                                              ptrdiff_t *local_1_start) {
    int P;
    if ((rnk != 3 && rnk != 2) || !n || !comm || howmany < 1) return 0;
    P = b2fft_comm_size(comm);
    if (block0 < 0 || block1 < 0) return 0;
    /* rank 2 = rank 3 with a last dimension of one element (the slab algorithm never looks inside a row) */
    return howmany * b2fft_slab_local_size_3d_blocks(n[0], n[1], rnk == 3 ? n[2] : 1, block0, block1, P,
                                                     b2fft_comm_rank(comm), local_n0, local_0_start, local_n1,
                                                     local_1_start);
}
ptrdiff_t fftw_mpi_local_size_many(int rnk, const ptrdiff_t *n, ptrdiff_t howmany, ptrdiff_t block0, MPI_Comm comm,
                                   ptrdiff_t *local_n0, ptrdiff_t *local_0_start) {
    ptrdiff_t a, b;
    return fftw_mpi_local_size_many_transposed(rnk, n, howmany, block0, FFTW_MPI_DEFAULT_BLOCK, comm, local_n0,
                                               local_0_start, &a, &b);
}

fftw_plan fftw_mpi_plan_many_dft(int rnk, const ptrdiff_t *n, ptrdiff_t howmany, ptrdiff_t block, ptrdiff_t tblock,
                                 FFTW_COMPLEX *in, FFTW_COMPLEX *out, MPI_Comm comm, int sign, unsigned flags) {
    fftw_plan p;
    b2fft_plan *bp;
    int P;
    ptrdiff_t n3[3];
    if ((rnk != 3 && rnk != 2) || !n || !comm) return 0; /* the slab path: 3-D, and 2-D as n2 = 1 */
    n3[0] = n[0];
    n3[1] = n[1];
    n3[2] = rnk == 3 ? n[2] : 1;
    n = n3;
    P = b2fft_comm_size(comm);
    (void)P;
    /* block = block size of the distributed INPUT dimension, tblock = of the distributed OUTPUT dimension
     * (reference fftw/mpi/api.c:604-611); a plan whose output is distributed like its input (no flag, or both)
     * needs the same block on both sides.  More blocks than ranks -> NULL (fftw/mpi/api.c:578-582). */
    if (block < 0 || tblock < 0) return 0;
    {
        const int tin = (flags & FFTW_MPI_TRANSPOSED_IN) != 0, tout = (flags & FFTW_MPI_TRANSPOSED_OUT) != 0;
        const ptrdiff_t nin = n[tin ? 1 : 0];
        ptrdiff_t b1 = tblock;
        if (tin == tout) {
            const ptrdiff_t bi = block ? block : (nin + P - 1) / P, bo = tblock ? tblock : (nin + P - 1) / P;
            if (bi != bo) return 0;
            b1 = 0;
        }
        bp = b2fft_slab_plan_dft_3d_blocks(n[0], n[1], n[2], howmany, block, b1, comm, sign, tin, tout, flags);
    }
    if (!bp) return 0;
    p = (fftw_plan)malloc(sizeof *p);
    if (!p) {
        b2fft_plan_destroy(bp);
        return 0;
    }
    memset(p, 0, sizeof *p);
    p->bp = bp;
    p->in = in;
    p->out = out;
    p->kind = B2FFT_C2C;
    return p;
}

fftw_plan fftw_mpi_plan_dft_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, FFTW_COMPLEX *in, FFTW_COMPLEX *out,
                               MPI_Comm comm, int sign, unsigned flags) {
    ptrdiff_t n[3];
    n[0] = n0;
    n[1] = n1;
    n[2] = n2;
    return fftw_mpi_plan_many_dft(3, n, 1, FFTW_MPI_DEFAULT_BLOCK, FFTW_MPI_DEFAULT_BLOCK, in, out, comm, sign, flags);
}

/* real data (reference fftw/mpi/api.c:751-858): rank 3 with any howmany, rank 2 with howmany 1; explicit block sizes for both */
static fftw_plan mpi_rdft(int kind, int rnk, const ptrdiff_t *n, ptrdiff_t howmany, ptrdiff_t iblock,
                          ptrdiff_t oblock, void *in, void *out, MPI_Comm comm, unsigned flags) {
    fftw_plan p;
    b2fft_plan *bp;
    int P;
    const int tin = (flags & FFTW_MPI_TRANSPOSED_IN) != 0, tout = (flags & FFTW_MPI_TRANSPOSED_OUT) != 0;
    if ((rnk != 3 && rnk != 2) || !n || !comm || howmany < 1 || (rnk == 2 && howmany != 1)) return 0;
    if (iblock < 0 || oblock < 0) return 0;
    P = b2fft_comm_size(comm);
    {
        /* iblock = block size of the distributed INPUT dimension, oblock = of the distributed OUTPUT dimension.  The real
         * array is always distributed over n0; the complex one over dimension 1 (n1 in 3-D, the n1/2+1 complex columns in
         * 2-D) when it is the transposed side, else over n0 too -- and then both sides must use the same block (as in
         * fftw_mpi_plan_many_dft above). */
        const ptrdiff_t rblock = kind == B2FFT_R2C ? iblock : oblock; /* of the real side: over n0 */
        const ptrdiff_t cblock = kind == B2FFT_R2C ? oblock : iblock; /* of the complex side */
        const int ctrans = kind == B2FFT_R2C ? tout : tin;
        ptrdiff_t b1 = cblock;
        if (!ctrans) {
            const ptrdiff_t br = rblock ? rblock : (n[0] + P - 1) / P, bc = cblock ? cblock : (n[0] + P - 1) / P;
            if (br != bc) return 0;
            b1 = 0;
        }
        bp = rnk == 3 ? b2fft_slab_plan_many_rdft_3d_blocks(kind, n[0], n[1], n[2], howmany, rblock, b1, comm, tin, tout, flags)
                      : b2fft_slab_plan_rdft_2d_blocks(kind, n[0], n[1], rblock, b1, comm, tin, tout, flags);
    }
    if (!bp) return 0;
    p = (fftw_plan)malloc(sizeof *p);
    if (!p) {
        b2fft_plan_destroy(bp);
        return 0;
    }
    memset(p, 0, sizeof *p);
    p->bp = bp;
    p->in = in;
    p->out = out;
    p->kind = kind;
    return p;
}
fftw_plan fftw_mpi_plan_many_dft_r2c(int rnk, const ptrdiff_t *n, ptrdiff_t howmany, ptrdiff_t iblock,
                                     ptrdiff_t oblock, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out, MPI_Comm comm,
                                     unsigned flags) {
    return mpi_rdft(B2FFT_R2C, rnk, n, howmany, iblock, oblock, in, out, comm, flags);
}
fftw_plan fftw_mpi_plan_many_dft_c2r(int rnk, const ptrdiff_t *n, ptrdiff_t howmany, ptrdiff_t iblock,
                                     ptrdiff_t oblock, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out, MPI_Comm comm,
                                     unsigned flags) {
    return mpi_rdft(B2FFT_C2R, rnk, n, howmany, iblock, oblock, in, out, comm, flags);
}
fftw_plan fftw_mpi_plan_dft_r2c(int rnk, const ptrdiff_t *n, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out, MPI_Comm comm,
                                unsigned flags) {
    return fftw_mpi_plan_many_dft_r2c(rnk, n, 1, FFTW_MPI_DEFAULT_BLOCK, FFTW_MPI_DEFAULT_BLOCK, in, out, comm, flags);
}
fftw_plan fftw_mpi_plan_dft_c2r(int rnk, const ptrdiff_t *n, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out, MPI_Comm comm,
                                unsigned flags) {
    return fftw_mpi_plan_many_dft_c2r(rnk, n, 1, FFTW_MPI_DEFAULT_BLOCK, FFTW_MPI_DEFAULT_BLOCK, in, out, comm, flags);
}
fftw_plan fftw_mpi_plan_dft_r2c_2d(ptrdiff_t n0, ptrdiff_t n1, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out, MPI_Comm comm,
                                   unsigned flags) {
    ptrdiff_t n[2];
    n[0] = n0;
    n[1] = n1;
    return fftw_mpi_plan_dft_r2c(2, n, in, out, comm, flags);
}
fftw_plan fftw_mpi_plan_dft_c2r_2d(ptrdiff_t n0, ptrdiff_t n1, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out, MPI_Comm comm,
                                   unsigned flags) {
    ptrdiff_t n[2];
    n[0] = n0;
    n[1] = n1;
    return fftw_mpi_plan_dft_c2r(2, n, in, out, comm, flags);
}
fftw_plan fftw_mpi_plan_dft_r2c_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out,
                                   MPI_Comm comm, unsigned flags) {
    ptrdiff_t n[3];
    n[0] = n0;
    n[1] = n1;
    n[2] = n2;
    return fftw_mpi_plan_dft_r2c(3, n, in, out, comm, flags);
}
fftw_plan fftw_mpi_plan_dft_c2r_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out,
                                   MPI_Comm comm, unsigned flags) {
    ptrdiff_t n[3];
    n[0] = n0;
    n[1] = n1;
    n[2] = n2;
    return fftw_mpi_plan_dft_c2r(3, n, in, out, comm, flags);
}

/* 2-D forms (reference fftw/mpi/api.c:488-492, 631-636): [local_n0][n1] in, [local_n1][n0] transposed out */
ptrdiff_t fftw_mpi_local_size_2d_transposed(ptrdiff_t n0, ptrdiff_t n1, MPI_Comm comm, ptrdiff_t *local_n0,
                                            ptrdiff_t *local_0_start, ptrdiff_t *local_n1, ptrdiff_t *local_1_start) {
    return fftw_mpi_local_size_3d_transposed(n0, n1, 1, comm, local_n0, local_0_start, local_n1, local_1_start);
}
ptrdiff_t fftw_mpi_local_size_2d(ptrdiff_t n0, ptrdiff_t n1, MPI_Comm comm, ptrdiff_t *local_n0,
                                 ptrdiff_t *local_0_start) {
    ptrdiff_t a, b;
    return fftw_mpi_local_size_2d_transposed(n0, n1, comm, local_n0, local_0_start, &a, &b);
}
fftw_plan fftw_mpi_plan_dft_2d(ptrdiff_t n0, ptrdiff_t n1, FFTW_COMPLEX *in, FFTW_COMPLEX *out, MPI_Comm comm,
                               int sign, unsigned flags) {
    ptrdiff_t n[2];
    n[0] = n0;
    n[1] = n1;
    return fftw_mpi_plan_many_dft(2, n, 1, FFTW_MPI_DEFAULT_BLOCK, FFTW_MPI_DEFAULT_BLOCK, in, out, comm, sign, flags);
}

void fftw_mpi_execute_dft(fftw_plan p, FFTW_COMPLEX *in, FFTW_COMPLEX *out) { fftw_execute_dft(p, in, out); }
void fftw_mpi_execute_dft_r2c(fftw_plan p, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out) { fftw_execute_dft_r2c(p, in, out); }
void fftw_mpi_execute_dft_c2r(fftw_plan p, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out) { fftw_execute_dft_c2r(p, in, out); }

int fftw_mpi_nccl_unique_id(void *id128) { return b2fft_nccl_unique_id(id128); }
MPI_Comm fftw_mpi_comm_from_nccl(const void *id128, int n_pes, int my_pe) {
    return b2fft_comm_create(id128, n_pes, my_pe);
}
MPI_Comm fftw_mpi_comm_self(void) { return b2fft_comm_self(); }
void fftw_mpi_comm_free(MPI_Comm comm) { b2fft_comm_destroy(comm); }
