/* oracle/fft_oracle.h -- CPU restatement of the reference DFT hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may build, load or call this.
 * The product (fftw3_b200/) never links or calls anything in oracle/.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function here against
 * golden vectors produced by the UNMODIFIED reference (oracle/_ref, built from
 * /root/reference by oracle/refbuild/Makefile) -- see tests/golden/make_golden.py.
 *
 * All arithmetic is double precision on interleaved (re,im) arrays, the layout of
 * the reference's FFTW_COMPLEX (fftw/fftw3.h:69).  Strides/dists are in units of
 * complex elements for complex arrays and reals for real arrays, exactly as in the
 * public "many" interface (fftw/fftw_api.c:642-736).
 */
#ifndef FFT_ORACLE_H
#define FFT_ORACLE_H
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* rank-r, howmany-batched complex DFT; restates fftw_plan_many_dft + fftw_execute
 * (fftw/fftw_api.c:642-666, 428-431).  sign = -1 forward, +1 backward, unnormalised.
 * inembed/onembed may be NULL (= n).  in may equal out (same strides). returns 0 / -1. */
int oracle_dft_many(int rank, const int *n, int howmany,
                    const double *in, const int *inembed, int istride, int idist,
                    double *out, const int *onembed, int ostride, int odist, int sign);

/* real -> half-complex (n_last/2+1 outputs per row); restates fftw_plan_many_dft_r2c
 * (fftw/fftw_api.c:704-736).  inembed NULL = n (out of place);  the in-place padded
 * layout of fftw_rdft2_pad (fftw/fftw_api.c:774-788) is requested by passing it
 * explicitly.  in (reals) and out (complex) must not overlap. */
int oracle_r2c_many(int rank, const int *n, int howmany,
                    const double *in, const int *inembed, int istride, int idist,
                    double *out, const int *onembed, int ostride, int odist);

/* half-complex -> real, unnormalised; restates fftw_plan_many_dft_c2r
 * (fftw/fftw_api.c:668-702).  The input is NOT destroyed by the oracle. */
int oracle_c2r_many(int rank, const int *n, int howmany,
                    const double *in, const int *inembed, int istride, int idist,
                    double *out, const int *onembed, int ostride, int odist);

/* ---- slab-distributed form (fftw/mpi; specification only in the reference) ---- */

/* restates fftw_mpi_local_size_3d (fftw/mpi/api.c:509-516 -> local_size_guru :248-352)
 * for a communicator of n_pes ranks, as seen by rank my_pe.  Returns the number of
 * complex elements rank my_pe must allocate. */
ptrdiff_t oracle_mpi_local_size_3d(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2,
                                   int n_pes, int my_pe,
                                   ptrdiff_t *local_n0, ptrdiff_t *local_0_start);

/* restates fftw_mpi_local_size_3d_transposed (fftw/mpi/api.c:494-507). */
ptrdiff_t oracle_mpi_local_size_3d_transposed(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2,
                                              int n_pes, int my_pe,
                                              ptrdiff_t *local_n0, ptrdiff_t *local_0_start,
                                              ptrdiff_t *local_n1, ptrdiff_t *local_1_start);

/* The result rank my_pe holds after fftw_mpi_plan_dft_3d(...)+execute on the GLOBAL
 * row-major input `global_in` (n0*n1*n2 complex).  transposed_out != 0 gives the
 * FFTW_MPI_TRANSPOSED_OUT layout [local_n1][n0][n2] (fftw/mpi/dft-rank-geq2-transposed.c:47-70,
 * fftw/doc/mpi.texi:470-509); otherwise [local_n0][n1][n2] (fftw/mpi/dft-rank-geq2.c:40-58).
 * local_out must hold local_n1*n0*n2 (resp. local_n0*n1*n2) complex.  returns 0 / -1. */
int oracle_mpi_dft_3d_local(ptrdiff_t n0, ptrdiff_t n1, ptrdiff_t n2,
                            int n_pes, int my_pe, int transposed_out, int sign,
                            const double *global_in, double *local_out);

#ifdef __cplusplus
}
#endif
#endif
