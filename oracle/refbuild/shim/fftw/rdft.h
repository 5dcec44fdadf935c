/* shim: the fork folded rdft.h into fftw_api.h but fftw/threads still includes it */
#include "fftw/fftw_api.h"
