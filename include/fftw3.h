/* include/fftw3.h -- public C ABI of the B200 FFT engine.
 *
 * This header declares, symbol for symbol and argument for argument, the API that the
 * reference exposes in fftw/fftw3.h:143-463 (Jasonlai0821/fftw3, an FFTW 3.3.x fork), so
 * that a program written against the reference recompiles and relinks against
 * fftw3_b200/lib/{fp32,fp64}/libfftw3.so unchanged.  It is a fresh restatement of the
 * interface, not a copy of the reference file.
 *
 * Conventions inherited from the reference:
 *   - no name mangling: every symbol is fftw_* in both precisions; the precision is the
 *     compile-time switch FFTW_SINGLE (reference fftw/fftw3.h:65-82).  Build the caller
 *     with -DFFTW_SINGLE and link lib/fp32, or without and link lib/fp64.
 *   - complex = FFTW_COMPLEX = FFTW_REAL_TYPE[2], interleaved (re, im) (fftw3.h:69).
 *   - FFTW_FORWARD = exp(-2 pi i jk/n), transforms are unnormalised (fftw3.h:469-470).
 *
 * Pointers handed to plan/execute may be host pointers (staged through pinned memory and
 * copied over PCIe inside fftw_execute) or CUDA device/managed pointers (transformed in
 * HBM with no copies).  fftw_malloc returns pinned host memory.  See INTEGRATION.md.
 */
#ifndef FFTW3_H
#define FFTW3_H

#include <stddef.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef FFTW_TYPE
#define FFTW_TYPE
#  if defined(FFTW_SINGLE)
typedef float FFTW_REAL_TYPE;
#  else
typedef double FFTW_REAL_TYPE;
#  endif
typedef FFTW_REAL_TYPE FFTW_COMPLEX[2];
#endif

#define FFTW_EXTERN extern
#define FFTW_CDECL

/* ---- types (reference fftw3.h:109-141) ------------------------------------------- */
enum fftw_r2r_kind_do_not_use_me {
    FFTW_R2HC = 0, FFTW_HC2R = 1, FFTW_DHT = 2,
    FFTW_REDFT00 = 3, FFTW_REDFT01 = 4, FFTW_REDFT10 = 5, FFTW_REDFT11 = 6,
    FFTW_RODFT00 = 7, FFTW_RODFT01 = 8, FFTW_RODFT10 = 9, FFTW_RODFT11 = 10
};
typedef enum fftw_r2r_kind_do_not_use_me fftw_r2r_kind;

struct fftw_iodim_do_not_use_me   { int n, is, os; };
struct fftw_iodim64_do_not_use_me { ptrdiff_t n, is, os; };
typedef struct fftw_iodim_do_not_use_me   fftw_iodim;
typedef struct fftw_iodim64_do_not_use_me fftw_iodim64;

typedef void (*fftw_write_char_func)(char c, void *);
typedef int  (*fftw_read_char_func)(void *);

typedef struct fftw_plan_s *fftw_plan;

/* shorthands used only inside this header */
#define B2_R FFTW_REAL_TYPE
#define B2_C FFTW_COMPLEX
#define B2_GURU(D)   int rank, const D *dims, int howmany_rank, const D *howmany_dims
#define B2_MANY(I,O) int rank, const int *n, int howmany,                 \
                     I *in, const int *inembed, int istride, int idist,   \
                     O *out, const int *onembed, int ostride, int odist

/* ---- complex DFT: basic, advanced ("many"), guru (reference fftw3.h:143-190) ------ */
FFTW_EXTERN fftw_plan fftw_plan_dft(int rank, const int *n, B2_C *in, B2_C *out, int sign, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_dft_1d(int n, B2_C *in, B2_C *out, int sign, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_dft_2d(int n0, int n1, B2_C *in, B2_C *out, int sign, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_dft_3d(int n0, int n1, int n2, B2_C *in, B2_C *out, int sign, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_many_dft(B2_MANY(B2_C, B2_C), int sign, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru_dft(B2_GURU(fftw_iodim), B2_C *in, B2_C *out, int sign, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru_split_dft(B2_GURU(fftw_iodim), B2_R *ri, B2_R *ii, B2_R *ro, B2_R *io, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru64_dft(B2_GURU(fftw_iodim64), B2_C *in, B2_C *out, int sign, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru64_split_dft(B2_GURU(fftw_iodim64), B2_R *ri, B2_R *ii, B2_R *ro, B2_R *io, unsigned flags);

/* ---- real-input / real-output DFT (reference fftw3.h:196-300) --------------------- */
FFTW_EXTERN fftw_plan fftw_plan_many_dft_r2c(B2_MANY(B2_R, B2_C), unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_dft_r2c(int rank, const int *n, B2_R *in, B2_C *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_dft_r2c_1d(int n, B2_R *in, B2_C *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_dft_r2c_2d(int n0, int n1, B2_R *in, B2_C *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_dft_r2c_3d(int n0, int n1, int n2, B2_R *in, B2_C *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_many_dft_c2r(B2_MANY(B2_C, B2_R), unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_dft_c2r(int rank, const int *n, B2_C *in, B2_R *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_dft_c2r_1d(int n, B2_C *in, B2_R *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_dft_c2r_2d(int n0, int n1, B2_C *in, B2_R *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_dft_c2r_3d(int n0, int n1, int n2, B2_C *in, B2_R *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru_dft_r2c(B2_GURU(fftw_iodim), B2_R *in, B2_C *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru_dft_c2r(B2_GURU(fftw_iodim), B2_C *in, B2_R *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru_split_dft_r2c(B2_GURU(fftw_iodim), B2_R *in, B2_R *ro, B2_R *io, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru_split_dft_c2r(B2_GURU(fftw_iodim), B2_R *ri, B2_R *ii, B2_R *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru64_dft_r2c(B2_GURU(fftw_iodim64), B2_R *in, B2_C *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru64_dft_c2r(B2_GURU(fftw_iodim64), B2_C *in, B2_R *out, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru64_split_dft_r2c(B2_GURU(fftw_iodim64), B2_R *in, B2_R *ro, B2_R *io, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru64_split_dft_c2r(B2_GURU(fftw_iodim64), B2_R *ri, B2_R *ii, B2_R *out, unsigned flags);

/* ---- r2r (ABI-complete, not on the hot path: plans are NULL) (fftw3.h:306-350) ----- */
FFTW_EXTERN fftw_plan fftw_plan_many_r2r(B2_MANY(B2_R, B2_R), const fftw_r2r_kind *kind, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_r2r(int rank, const int *n, B2_R *in, B2_R *out, const fftw_r2r_kind *kind, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_r2r_1d(int n, B2_R *in, B2_R *out, fftw_r2r_kind kind, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_r2r_2d(int n0, int n1, B2_R *in, B2_R *out, fftw_r2r_kind kind0, fftw_r2r_kind kind1, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_r2r_3d(int n0, int n1, int n2, B2_R *in, B2_R *out, fftw_r2r_kind kind0, fftw_r2r_kind kind1, fftw_r2r_kind kind2, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru_r2r(B2_GURU(fftw_iodim), B2_R *in, B2_R *out, const fftw_r2r_kind *kind, unsigned flags);
FFTW_EXTERN fftw_plan fftw_plan_guru64_r2r(B2_GURU(fftw_iodim64), B2_R *in, B2_R *out, const fftw_r2r_kind *kind, unsigned flags);

/* ---- execution (reference fftw3.h:143, 186-190, 296-300, 350) ---------------------- */
FFTW_EXTERN void fftw_execute(const fftw_plan p);
FFTW_EXTERN void fftw_execute_dft(const fftw_plan p, B2_C *in, B2_C *out);
FFTW_EXTERN void fftw_execute_split_dft(const fftw_plan p, B2_R *ri, B2_R *ii, B2_R *ro, B2_R *io);
FFTW_EXTERN void fftw_execute_dft_r2c(const fftw_plan p, B2_R *in, B2_C *out);
FFTW_EXTERN void fftw_execute_dft_c2r(const fftw_plan p, B2_C *in, B2_R *out);
FFTW_EXTERN void fftw_execute_split_dft_r2c(const fftw_plan p, B2_R *in, B2_R *ro, B2_R *io);
FFTW_EXTERN void fftw_execute_split_dft_c2r(const fftw_plan p, B2_R *ri, B2_R *ii, B2_R *out);
FFTW_EXTERN void fftw_execute_r2r(const fftw_plan p, B2_R *in, B2_R *out);

/* ---- lifecycle, threads, wisdom, printing, memory (reference fftw3.h:352-463) ------ */
FFTW_EXTERN void fftw_destroy_plan(fftw_plan p);
FFTW_EXTERN void fftw_forget_wisdom(void);
FFTW_EXTERN void fftw_cleanup(void);
FFTW_EXTERN void fftw_set_timelimit(double t);

FFTW_EXTERN void fftw_plan_with_nthreads(int nthreads);
FFTW_EXTERN int  fftw_init_threads(void);
FFTW_EXTERN void fftw_cleanup_threads(void);
FFTW_EXTERN void fftw_make_planner_thread_safe(void);

FFTW_EXTERN int   fftw_export_wisdom_to_filename(const char *filename);
FFTW_EXTERN void  fftw_export_wisdom_to_file(FILE *output_file);
FFTW_EXTERN char *fftw_export_wisdom_to_string(void);
FFTW_EXTERN void  fftw_export_wisdom(fftw_write_char_func write_char, void *data);
FFTW_EXTERN int   fftw_import_system_wisdom(void);
FFTW_EXTERN int   fftw_import_wisdom_from_filename(const char *filename);
FFTW_EXTERN int   fftw_import_wisdom_from_file(FILE *input_file);
FFTW_EXTERN int   fftw_import_wisdom_from_string(const char *input_string);
FFTW_EXTERN int   fftw_import_wisdom(fftw_read_char_func read_char, void *data);

FFTW_EXTERN void  fftw_fprint_plan(const fftw_plan p, FILE *output_file);
FFTW_EXTERN void  fftw_print_plan(const fftw_plan p);
FFTW_EXTERN char *fftw_sprint_plan(const fftw_plan p);

FFTW_EXTERN void *fftw_malloc(size_t n);
FFTW_EXTERN B2_R *fftw_alloc_real(size_t n);
FFTW_EXTERN B2_C *fftw_alloc_complex(size_t n);
FFTW_EXTERN void  fftw_free(void *p);

FFTW_EXTERN void   fftw_flops(const fftw_plan p, double *add, double *mul, double *fmas);
FFTW_EXTERN double fftw_estimate_cost(const fftw_plan p);
FFTW_EXTERN double fftw_cost(const fftw_plan p);
FFTW_EXTERN int    fftw_alignment_of(B2_R *p);

FFTW_EXTERN const char fftw_version[];
FFTW_EXTERN const char fftw_cc[];
FFTW_EXTERN const char fftw_codelet_optim[];

#undef B2_R
#undef B2_C
#undef B2_GURU
#undef B2_MANY

/* ---- constants (reference fftw3.h:469-499) ----------------------------------------- */
#define FFTW_FORWARD  (-1)
#define FFTW_BACKWARD (+1)
#define FFTW_NO_TIMELIMIT (-1.0)

#define FFTW_MEASURE         (0U)
#define FFTW_DESTROY_INPUT   (1U << 0)
#define FFTW_UNALIGNED       (1U << 1)
#define FFTW_CONSERVE_MEMORY (1U << 2)
#define FFTW_EXHAUSTIVE      (1U << 3)
#define FFTW_PRESERVE_INPUT  (1U << 4)
#define FFTW_PATIENT         (1U << 5)
#define FFTW_ESTIMATE        (1U << 6)
#define FFTW_WISDOM_ONLY     (1U << 21)

/* accepted and ignored (they steer the reference's recursive planner only) */
#define FFTW_ESTIMATE_PATIENT       (1U << 7)
#define FFTW_BELIEVE_PCOST          (1U << 8)
#define FFTW_NO_DFT_R2HC            (1U << 9)
#define FFTW_NO_NONTHREADED         (1U << 10)
#define FFTW_NO_BUFFERING           (1U << 11)
#define FFTW_NO_INDIRECT_OP         (1U << 12)
#define FFTW_ALLOW_LARGE_GENERIC    (1U << 13)
#define FFTW_NO_RANK_SPLITS         (1U << 14)
#define FFTW_NO_VRANK_SPLITS        (1U << 15)
#define FFTW_NO_VRECURSE            (1U << 16)
#define FFTW_NO_SIMD                (1U << 17)
#define FFTW_NO_SLOW                (1U << 18)
#define FFTW_NO_FIXED_RADIX_LARGE_N (1U << 19)
#define FFTW_ALLOW_PRUNING          (1U << 20)

#ifdef __cplusplus
}
#endif
#endif /* FFTW3_H */
