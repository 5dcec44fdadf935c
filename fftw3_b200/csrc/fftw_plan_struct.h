/* csrc/fftw_plan_struct.h -- the opaque fftw_plan of include/fftw3.h (reference: apiplan
 * {pln, prb, sign}, fftw/fftw_api.h:2031-2035) */
#ifndef FFTW_PLAN_STRUCT_H
#define FFTW_PLAN_STRUCT_H
#include "b200fft.h"
struct fftw_plan_s {
    b2fft_plan *bp;
    void *in, *out; /* the arrays given at plan time (reference: problem_dft.ri.., api.h:1874) */
    int kind;
    b2fft_geom g;
    b2fft_split *sp;            /* split-array plan (bp is NULL then) */
    void *ri, *ii, *ro, *io;    /* its arrays at plan time */
    b2fft_r2r *rr;              /* real-to-real plan (bp is NULL then) */
};
#endif
