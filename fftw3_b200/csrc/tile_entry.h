// csrc/tile_entry.h -- one row of the table of instantiated tile kernels (tile_dispatch.cuh builds the table;
// mixtile.cu contributes the mixed-radix rows from its own translation unit).
#pragma once
#include "device_rt.h"
#include "pass_params.h"

struct TileEntry {
    int n, e, mode, epi, c, threads;
    size_t smem;
    size_t (*smem_for)(int C);
    int (*launch)(const PassParams &, const real_t *, real_t *, long long, long long, int, size_t, b2::stream_t);
    void (*init)();
    // 0: plain tile kernel; 1: persistent, software-pipelined column kernel (colpipe_kernel.cuh); 2: TMA column kernel
    // (coltma_kernel.cuh); 3: mixed-radix tile kernel for a length that is not a power of two (mixtile.cu)
    int pipe;
    const char *sched;  // mixed-radix rows: the radices, for the plan description
};

namespace b2 {
const TileEntry *mix_entries(int *count);       // the mixed-radix rows (mixtile.cu)
const TileEntry *mix_entries_more(int *count);  // ... and mixtile2.cu
const TileEntry *mix_entries_cols(int *count);  // ... and mixtile3.cu
}
