// csrc/plan_struct.h -- the opaque b2fft_plan of include/b200fft.h
#pragma once
#include <mutex>

#include "plan_build.h"

namespace b2 {
struct Chain;
}

struct SlabPlan;
struct b2fft_plan_s;
// Host-array executes of a batched plan run as a pipeline over chunks of the batch: H2D copy,
// kernels and D2H copy of different chunks overlap on separate streams (PCIe is full duplex).
struct HostPipe {
    int state = 0;  // 0 not tried, 1 ready, -1 not applicable
    int vdim = -1;  // batch dim that is chunked
    long long chunk = 0, nchunks = 0, tail = 0;
    long long in_step = 0, out_step = 0;  // reals between consecutive batch entries
    struct Slot {
        b2fft_plan_s *plan = nullptr, *plan_tail = nullptr;
        real_t *din = nullptr, *dout = nullptr;
        b2::stream_t stream = nullptr;
    } slot[3];
};
struct b2fft_plan_s {
    b2::PlanCore pc;
    real_t *scr = nullptr;
    real_t *wrk = nullptr;  // second work array (Bluestein rows), pc.wrk_reals
    // device staging for host-pointer executes (allocated on first use)
    real_t *dev_in = nullptr, *dev_out = nullptr;
    b2::stream_t stream = nullptr;
    bool own_stream = false;
    int cfg = 0;            // configuration bit mask the plan was built with (B2_CFG_*), 0 = the selector's default
    bool measured = false;  // chosen by timing (FFTW_MEASURE)
    int device = -1;  // the device the plan (its tables, work arrays, streams) lives on; executes make it current
    // host / device classification of the arrays of the last synchronous execute (fftw_execute passes the same
    // two pointers every time: one cudaPointerGetAttributes pair per plan, not per call)
    const void *cls_in = nullptr, *cls_out = nullptr;
    bool cls_din = false, cls_dout = false, cls_valid = false;
    SlabPlan *slab = nullptr;
    HostPipe *pipe = nullptr;
    std::vector<b2::Chain> *chains = nullptr;
    bool needs_aligned = false;     // has steps without a generic-kernel form
    b2fft_plan_s *alt = nullptr;    // the same problem without them, for unaligned arrays
    b2::StepPlan *scat = nullptr;   // the last step re-selected with a scattered store (slab.cu)
    // fftw_execute is re-entrant in the reference, also for one plan from several threads
    // (fftw_api.c:428-454: plans are read-only after awake).  Here a plan owns work arrays and
    // staging buffers, so synchronous executes of the SAME plan are serialised.
    std::mutex exec_mutex;
};
// Fused slab exchange: turn the plan's last step (a column pass writing rows of the transformed
// axis) into one that stores row k into the receive buffer of its owner k / blk.  plane_stride =
// dst stride (reals) of the line dim that numbers the planes; pstride[o] = owner o's plane stride.
// Returns 0 when the step qualifies (tile kernel, column mode, no other epilogue).
// self >= 0: hybrid form, only owner `self` is scattered, the other rows stay in place.
// rstride (optional): per-owner distance between consecutive rows of the block, in reals (default: the step's own
// axis stride); local: all bases are memory of this GPU.
int b2_plan_enable_scatter(b2fft_plan_s *p, int nowners, long long blk, long long plane_stride, const long long *pstride,
                           int self = -1, const long long *rstride = nullptr, int local = 0);
int b2_plan_execute_scatter(b2fft_plan_s *p, void *in, void *out, real_t *const *bases, long long coord0,
                            b2::stream_t stream);
void b2_slab_destroy(b2fft_plan_s *p);
int b2_slab_execute(b2fft_plan_s *p, void *in, void *out);
