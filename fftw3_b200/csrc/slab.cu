// csrc/slab.cu -- slab-distributed 3-D c2c (the fftw_mpi_plan_dft_3d form) -- placeholder
// filled in by the multi-GPU milestone.
#include "b200fft.h"
#include "device_rt.h"

struct b2fft_plan_s;
void b2_slab_destroy(b2fft_plan_s *) {}
int b2_slab_execute(b2fft_plan_s *, void *, void *) { return -1; }
