/* include/fftw3-mpi.h -- the slab-distributed subset of the reference's MPI interface
 * (fftw/mpi/fftw3-mpi.h:78-79, 113-119, 129-144, 192, 209-215) over NCCL, one process per GPU.
 *
 * Same function names, argument order and layouts as the reference (fftw/doc/mpi.texi:440-516):
 * dimension 0 is block-distributed with block = ceil(n0/P) (fftw/mpi/block.c:39-50); with
 * FFTW_MPI_TRANSPOSED_OUT the result is left as [local_n1][n0][n2], distributed over n1.
 *
 * MPI itself does not exist on the target boxes.  When <mpi.h> is not in use, MPI_Comm is an
 * opaque communicator handle created from an NCCL unique id (fftw_mpi_comm_from_nccl); the id
 * travels between the processes by any side channel (the tests and bench.py use
 * torch.distributed's store).  Define B2FFT_HAVE_MPI before including this header to use a real
 * MPI_Comm type instead (the library then still needs a communicator created here).
 */
#ifndef FFTW3_MPI_H
#define FFTW3_MPI_H

#include "fftw3.h"

#ifdef __cplusplus
extern "C" {
#endif

#ifndef B2FFT_HAVE_MPI
struct b2fft_comm_s;
typedef struct b2fft_comm_s *MPI_Comm;
#endif

#define FFTW_MPI_DEFAULT_BLOCK (0)

/* MPI-specific flags (reference fftw/mpi/fftw3-mpi.h:212-215) */
#define FFTW_MPI_SCRAMBLED_IN   (1U << 27)
#define FFTW_MPI_SCRAMBLED_OUT  (1U << 28)
#define FFTW_MPI_TRANSPOSED_IN  (1U << 29)
#define FFTW_MPI_TRANSPOSED_OUT (1U << 30)

FFTW_EXTERN void fftw_mpi_init(void);
FFTW_EXTERN void fftw_mpi_cleanup(void);

FFTW_EXTERN ptrdiff_t fftw_mpi_local_size_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, MPI_Comm comm,
                                             ptrdiff_t *local_n0, ptrdiff_t *local_0_start);
FFTW_EXTERN ptrdiff_t fftw_mpi_local_size_3d_transposed(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, MPI_Comm comm,
                                                        ptrdiff_t *local_n0, ptrdiff_t *local_0_start,
                                                        ptrdiff_t *local_n1, ptrdiff_t *local_1_start);

/* reference fftw/mpi/fftw3-mpi.h:129-144 (rank 3, default blocks) */
FFTW_EXTERN ptrdiff_t fftw_mpi_local_size_many_transposed(int rnk, const ptrdiff_t *n, ptrdiff_t howmany,
                                                          ptrdiff_t block0, ptrdiff_t block1, MPI_Comm comm,
                                                          ptrdiff_t *local_n0, ptrdiff_t *local_0_start,
                                                          ptrdiff_t *local_n1, ptrdiff_t *local_1_start);
FFTW_EXTERN ptrdiff_t fftw_mpi_local_size_many(int rnk, const ptrdiff_t *n, ptrdiff_t howmany, ptrdiff_t block0,
                                               MPI_Comm comm, ptrdiff_t *local_n0, ptrdiff_t *local_0_start);
/* 2-D: [local_n0][n1] in, [local_n1][n0] with FFTW_MPI_TRANSPOSED_OUT (reference fftw/mpi/fftw3-mpi.h:113-128) */
FFTW_EXTERN ptrdiff_t fftw_mpi_local_size_2d(ptrdiff_t n0, ptrdiff_t n1, MPI_Comm comm, ptrdiff_t *local_n0,
                                             ptrdiff_t *local_0_start);
FFTW_EXTERN ptrdiff_t fftw_mpi_local_size_2d_transposed(ptrdiff_t n0, ptrdiff_t n1, MPI_Comm comm,
                                                        ptrdiff_t *local_n0, ptrdiff_t *local_0_start,
                                                        ptrdiff_t *local_n1, ptrdiff_t *local_1_start);
FFTW_EXTERN fftw_plan fftw_mpi_plan_dft_2d(ptrdiff_t n0, ptrdiff_t n1, FFTW_COMPLEX *in, FFTW_COMPLEX *out,
                                           MPI_Comm comm, int sign, unsigned flags);
FFTW_EXTERN fftw_plan fftw_mpi_plan_dft_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, FFTW_COMPLEX *in,
                                           FFTW_COMPLEX *out, MPI_Comm comm, int sign, unsigned flags);
FFTW_EXTERN fftw_plan fftw_mpi_plan_many_dft(int rnk, const ptrdiff_t *n, ptrdiff_t howmany, ptrdiff_t block,
                                             ptrdiff_t tblock, FFTW_COMPLEX *in, FFTW_COMPLEX *out,
                                             MPI_Comm comm, int sign, unsigned flags);
FFTW_EXTERN void fftw_mpi_execute_dft(fftw_plan p, FFTW_COMPLEX *in, FFTW_COMPLEX *out);

/* real data, rank 3 (reference fftw/mpi/fftw3-mpi.h:161-194): the real array is [local_n0][n1][2(n2/2+1)],
 * the complex one [local_n0][n1][n2/2+1] or, transposed, [local_n1][n0][n2/2+1]; local sizes come from
 * fftw_mpi_local_size_3d(_transposed)(n0, n1, n2/2+1, ...) in complex elements, as in the reference.
 * r2c accepts FFTW_MPI_TRANSPOSED_OUT, c2r FFTW_MPI_TRANSPOSED_IN. */
FFTW_EXTERN fftw_plan fftw_mpi_plan_many_dft_r2c(int rnk, const ptrdiff_t *n, ptrdiff_t howmany, ptrdiff_t iblock,
                                                 ptrdiff_t oblock, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out,
                                                 MPI_Comm comm, unsigned flags);
FFTW_EXTERN fftw_plan fftw_mpi_plan_dft_r2c(int rnk, const ptrdiff_t *n, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out,
                                            MPI_Comm comm, unsigned flags);
/* 2-D real data (reference fftw/mpi/fftw3-mpi.h:168-170, 182-184): real [local_n0][2(n1/2+1)]; complex
 * [local_n0][n1/2+1] or [local_n1][n0] with local_n1 blocks of the n1/2+1 columns
 * (sizes from fftw_mpi_local_size_2d[_transposed](n0, n1/2+1, ...)) */
FFTW_EXTERN fftw_plan fftw_mpi_plan_dft_r2c_2d(ptrdiff_t n0, ptrdiff_t n1, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out,
                                               MPI_Comm comm, unsigned flags);
FFTW_EXTERN fftw_plan fftw_mpi_plan_dft_c2r_2d(ptrdiff_t n0, ptrdiff_t n1, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out,
                                               MPI_Comm comm, unsigned flags);
FFTW_EXTERN fftw_plan fftw_mpi_plan_dft_r2c_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, FFTW_REAL_TYPE *in,
                                               FFTW_COMPLEX *out, MPI_Comm comm, unsigned flags);
FFTW_EXTERN fftw_plan fftw_mpi_plan_many_dft_c2r(int rnk, const ptrdiff_t *n, ptrdiff_t howmany, ptrdiff_t iblock,
                                                 ptrdiff_t oblock, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out,
                                                 MPI_Comm comm, unsigned flags);
FFTW_EXTERN fftw_plan fftw_mpi_plan_dft_c2r(int rnk, const ptrdiff_t *n, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out,
                                            MPI_Comm comm, unsigned flags);
FFTW_EXTERN fftw_plan fftw_mpi_plan_dft_c2r_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2, FFTW_COMPLEX *in,
                                               FFTW_REAL_TYPE *out, MPI_Comm comm, unsigned flags);
FFTW_EXTERN void fftw_mpi_execute_dft_r2c(fftw_plan p, FFTW_REAL_TYPE *in, FFTW_COMPLEX *out);
FFTW_EXTERN void fftw_mpi_execute_dft_c2r(fftw_plan p, FFTW_COMPLEX *in, FFTW_REAL_TYPE *out);

/* ---- extensions: communicator bootstrap without MPI ---- */
FFTW_EXTERN int      fftw_mpi_nccl_unique_id(void *id128);            /* rank 0: fills 128 bytes */
FFTW_EXTERN MPI_Comm fftw_mpi_comm_from_nccl(const void *id128, int n_pes, int my_pe);
FFTW_EXTERN MPI_Comm fftw_mpi_comm_self(void);
FFTW_EXTERN void     fftw_mpi_comm_free(MPI_Comm comm);

#ifdef __cplusplus
}
#endif
#endif /* FFTW3_MPI_H */
