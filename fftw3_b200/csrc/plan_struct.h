// csrc/plan_struct.h -- the opaque b2fft_plan of include/b200fft.h
#pragma once
#include "plan_build.h"

struct SlabPlan;
struct b2fft_plan_s {
    b2::PlanCore pc;
    real_t *scr = nullptr;
    // device staging for host-pointer executes (allocated on first use)
    real_t *dev_in = nullptr, *dev_out = nullptr;
    b2::stream_t stream = nullptr;
    bool own_stream = false;
    SlabPlan *slab = nullptr;
};
void b2_slab_destroy(b2fft_plan_s *p);
int b2_slab_execute(b2fft_plan_s *p, void *in, void *out);
