// csrc/mixtile3.cu -- the third list of mixed-radix tile kernels: column, column + twiddle and row -> column forms for 22 lengths of
// the second list (mixtile2.cu), so that the other axes of an N-D transform (144^3, 600^3, 1280 x 720, 1920 x 1080 frames ...) and the
// two passes of a long 1-D transform whose factors are such lengths (20736 = 144 x 144, 518400 = 720 x 720) run on tile kernels too.
// Tile widths follow what the first list measured (profiles/r2_mix_probe.md): 16 lines (128-byte row segments in single precision)
// where the CTA then still fits the register file, else 8; more for very short lines.  Not individually measured on the GPU.
#include "mixtile.cuh"

namespace b2 {
const TileEntry *mix_entries_cols(int *count) {
    static const TileEntry tab[] = {
        B2_MIX_COLS(144, 24, B2_CC(32, 32), 12, 12)
        B2_MIX_COLS(288, 24, B2_CC(16, 16), 24, 12)
        B2_MIX_COLS(576, 24, B2_CC(16, 8), 24, 24)
        B2_MIX_COLS(320, 20, B2_CC(16, 16), 20, 4, 4)
        B2_MIX_COLS(640, 20, B2_CC(16, 8), 20, 4, 4, 2)
        B2_MIX_COLS(800, 20, B2_CC(16, 8), 20, 10, 4)
        B2_MIX_COLS(300, 30, B2_CC(16, 16), 30, 10)
        B2_MIX_COLS(600, 30, B2_CC(16, 8), 10, 10, 6)
        B2_MIX_COLS(900, 30, B2_CC(16, 8), 30, 30)
        B2_MIX_COLS(360, 30, B2_CC(16, 16), 6, 6, 10)
        B2_MIX_COLS(720, 30, B2_CC(16, 8), 6, 6, 10, 2)
        B2_MIX_COLS(216, 24, B2_CC(16, 16), 6, 6, 6)
        B2_MIX_COLS(432, 24, B2_CC(16, 8), 24, 6, 3)
        B2_MIX_COLS(864, 24, B2_CC(16, 8), 24, 12, 3)
#if defined(FFTW_SINGLE)  // beyond the fp64 column cap (selector.h caps())
        B2_MIX_COLS(1152, 24, 8, 24, 8, 6)
        B2_MIX_COLS(1280, 20, 8, 20, 4, 4, 4)
        B2_MIX_COLS(1600, 20, 8, 20, 20, 4)
        B2_MIX_COLS(1500, 30, 8, 30, 10, 5)
        B2_MIX_COLS(1800, 30, 8, 30, 10, 6)
        B2_MIX_COLS(1080, 30, 16, 30, 6, 6)
        B2_MIX_COLS(1296, 24, 8, 6, 6, 6, 6)
        B2_MIX_COLS(1728, 24, 8, 12, 12, 12)
#endif
    };
    *count = (int)(sizeof tab / sizeof tab[0]);
    return tab;
}
}  // namespace b2
