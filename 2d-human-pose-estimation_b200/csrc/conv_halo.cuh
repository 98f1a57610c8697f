// Launch interface of the y-halo conv variant (conv_halo.cu), used by the plan code in conv_gemm.cu.
#pragma once
#include "hpe_common.cuh"
#include <cuda.h>

namespace hpe {

constexpr int kHaloRows = 32;          // image rows of one brick (8 wide x 32 tall); the A box has kHaloRows + taps_y - 1 rows

struct HaloParams {
  const float* scale;
  const float* shift;
  void* out;                           // y, dense [B, Ho, Wo, 64] bf16
  int num_tiles, tiles_x, tiles_y;     // bricks of one image
  int x_lo, y_lo;                      // box origin = (x0 - x_lo, y0 - y_lo)
  int Wo, Ho;                          // extent of y (the pooled extent with `pool`)
  int pool;                            // != 0 (stem only): MaxPool 3x3/s2/p1 of the ReLU output is taken in the epilogue; bricks are then
                                       // OVERLAPPING, at (6*tx - 1, 30*ty - 1), and produce 3 x 15 pooled pixels each
  int Wc, Hc;                          // pool: extent of the (never written) convolution output
  int reverse, nbatch;                 // reverse != 0: the bricks are walked from the last to the first (hpe_conv_set_schedule); nbatch = B
};

int halo_launch(int taps_y, int act, const CUtensorMap* tmA, const CUtensorMap* tmW, const HaloParams& p, int grid, cudaStream_t stream);

}  // namespace hpe
