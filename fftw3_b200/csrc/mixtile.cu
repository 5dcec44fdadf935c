// csrc/mixtile.cu -- the first list of mixed-radix tile kernels (the 29 lengths measured on B200, profiles/r2_mix_probe.md;
// kernel and macros: mixtile.cuh; second list: mixtile2.cu)
#include "mixtile.cuh"

namespace b2 {
const TileEntry *mix_entries(int *count) {
    static const TileEntry tab[] = {
#if defined(FFTW_SINGLE)
        // Shapes that measured faster than the ones in the main list below, form by form (profiles/r2_mix_probe.md): listed
        // first, so they are chosen.  Column forms want 128-byte row segments (16 lines) and 20+ points per thread; the
        // row -> column form wants many small CTAs, so it mostly stays with the main list's shape.
        B2_MIX_ONE(250, 10, MODE_COL, 0, 16, 10, 5, 5)   // 0.727 -> 0.466 ms per GiB pass
        B2_MIX_COLS(1000, 20, 16, 10, 10, 10)            // 0.788 -> 0.598; as the passes of n = 10^6: 1.556 -> 1.330
        B2_MIX_ONE(243, 27, MODE_COL, 0, 16, 9, 9, 3)    // 0.727 -> 0.438
        B2_MIX_ONE(243, 27, MODE_COL, 1, 16, 9, 9, 3)    // n = 243 * 128: 0.966 -> 0.920
        B2_MIX_ONE(729, 27, MODE_COL, 0, 16, 9, 9, 9)    // 0.849 -> 0.657
        B2_MIX_ONE(729, 27, MODE_COL, 1, 16, 9, 9, 9)    // n = 729 * 512: 1.383 -> 1.122
        B2_MIX_ROW(5000, 20, 1, 10, 10, 10, 5)           // 0.642 -> 0.545
        B2_MIX_ROW(625, 5, 2, 5, 5, 5, 5)                // 0.626 -> 0.444
#endif
        // 3 * 2^k
        B2_MIX_ALL(96, 24, 64, B2_CC(64, 32), 12, 8)
        B2_MIX_ALL(192, 24, 32, B2_CC(32, 16), 24, 8)
        B2_MIX_ALL(384, 24, 16, B2_CC(16, 8), 8, 8, 6)
        B2_MIX_ALL(768, 24, 8, B2_CC(16, 8), 12, 8, 8)
        B2_MIX_ALL32(1536, 24, 4, 8, 24, 8, 8)
        B2_MIX_ROW(3072, 24, 2, 8, 8, 8, 6)
        // powers of ten and their halves / doubles
        B2_MIX_ALL(100, 10, 24, B2_CC(32, 16), 10, 10)
        B2_MIX_ALL(200, 20, 24, B2_CC(16, 16), 20, 10)
        B2_MIX_ALL(250, 10, 10, B2_CC(8, 8), 10, 5, 5)
        B2_MIX_ALL(400, 20, 12, B2_CC(16, 8), 20, 20)
        B2_MIX_ALL(500, 10, 5, B2_CC(16, 8), 10, 10, 5)
        B2_MIX_ALL(1000, 10, 2, B2_CC(8, 4), 10, 10, 10)
        B2_MIX_ROW(2000, 20, 2, 20, 10, 10)
        B2_MIX_ROW(3000, 30, 2, 30, 10, 10)
        B2_MIX_ROW(4000, 20, 1, 20, 20, 10)
        // powers of 3, 5, 7
        B2_MIX_ALL(81, 9, 28, B2_CC(16, 16), 9, 9)
        B2_MIX_ALL(243, 9, 9, B2_CC(8, 8), 9, 9, 3)
        B2_MIX_ALL(729, 9, 3, B2_CC(8, 4), 9, 9, 9)
        B2_MIX_ROW(2187, 9, 1, 9, 9, 9, 3)
        B2_MIX_ALL(125, 5, 10, B2_CC(8, 8), 5, 5, 5)
        B2_MIX_ROW(625, 25, 10, 5, 5, 5, 5)
        B2_MIX_ROW(3125, 25, 2, 5, 5, 5, 5, 5)
        B2_MIX_ALL(49, 7, 32, B2_CC(16, 16), 7, 7)
        B2_MIX_ALL(343, 7, 5, B2_CC(16, 8), 7, 7, 7)
        B2_MIX_ROW(2401, 7, 1, 7, 7, 7, 7)
        // 1080 = 2^3 3^3 5 (video lines)
        B2_MIX_ROW(1080, 30, 7, 30, 6, 6)
#if defined(FFTW_SINGLE)  // beyond the fp64 row cap (selector.h caps())
        B2_MIX_ROW(5000, 10, 1, 10, 10, 10, 5)
        B2_MIX_ROW(6144, 24, 1, 12, 8, 8, 8)
        B2_MIX_ROW(6561, 27, 1, 9, 9, 9, 9)
#endif
        // real transforms of 2N points on the N-point row shapes with an even E (same shape as the complex row form)
        B2_MIX_REAL(96, 24, 64, 12, 8)
        B2_MIX_REAL(192, 24, 32, 24, 8)
        B2_MIX_REAL(384, 24, 16, 8, 8, 6)
        B2_MIX_REAL(768, 24, 8, 12, 8, 8)
        B2_MIX_REAL(1536, 24, 4, 24, 8, 8)
        B2_MIX_REAL(3072, 24, 2, 8, 8, 8, 6)
        B2_MIX_REAL(100, 10, 24, 10, 10)
        B2_MIX_REAL(200, 20, 24, 20, 10)
        B2_MIX_REAL(250, 10, 10, 10, 5, 5)
        B2_MIX_REAL(400, 20, 12, 20, 20)
        B2_MIX_REAL(500, 10, 5, 10, 10, 5)
        B2_MIX_REAL(1000, 10, 2, 10, 10, 10)
        B2_MIX_REAL(2000, 20, 2, 20, 10, 10)
        B2_MIX_REAL(3000, 30, 2, 30, 10, 10)
        B2_MIX_REAL(4000, 20, 1, 20, 20, 10)
        B2_MIX_REAL(1080, 30, 7, 30, 6, 6)
#if defined(FFTW_SINGLE)
        B2_MIX_REAL(5000, 20, 1, 10, 10, 10, 5)
        B2_MIX_REAL(6144, 24, 1, 12, 8, 8, 8)
        // shapes that measured equal or slower than the ones above (kept selectable with FFTW_B200_MIX_ALT=1, which
        // takes the LAST row listed for a length and form: scripts/gpu_mix_alt_probe.py, gpu_mix_form_probe.py)
        B2_MIX_COLS(500, 20, 16, 10, 10, 5)
        B2_MIX_ONE(250, 10, MODE_COL, 1, 16, 10, 5, 5)
        B2_MIX_ONE(250, 10, MODE_ROW2COL, 0, 16, 10, 5, 5)
        B2_MIX_ONE(243, 27, MODE_ROW2COL, 0, 16, 9, 9, 3)
        B2_MIX_ONE(729, 27, MODE_ROW2COL, 0, 16, 9, 9, 9)
        B2_MIX_ROW(6561, 9, 1, 9, 9, 9, 9)
        B2_MIX_ROW(3125, 5, 1, 5, 5, 5, 5, 5)
#endif
    };
    *count = (int)(sizeof tab / sizeof tab[0]);
    return tab;
}
}  // namespace b2
