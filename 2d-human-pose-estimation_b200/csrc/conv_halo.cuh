// Launch interface of the y-halo conv variant (conv_halo.cu), used by the plan code in conv_gemm.cu.
#pragma once
#include "hpe_common.cuh"
#include <cuda.h>

namespace hpe {

struct HaloParams {
  const float* scale;
  const float* shift;
  int num_tiles, tiles_x, tiles_y;     // 8-wide x 16-tall bricks of one image
  int x_lo, y_lo;                      // box origin = (x0 + box_index - x_lo, y0 - y_lo)
  int base_off_mode;                   // unused (kept for ABI stability of the struct between the two translation units)
};

int halo_launch(int taps_y, int boxes /* 0 = one 10-pixel-wide box holding all three column taps */, int act, const CUtensorMap* tmA, const CUtensorMap* tmW, const CUtensorMap* tmOut,
                const HaloParams& p, int grid, cudaStream_t stream);

}  // namespace hpe
