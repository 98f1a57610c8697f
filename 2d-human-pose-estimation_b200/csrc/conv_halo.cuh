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
  int pool;                            // != 0 (stem only): MaxPool 3x3/s2/p1 of the ReLU output is taken in the epilogue; tiles are then
                                       // OVERLAPPING 8x16 bricks at (6*tx - 1, 14*ty - 1) producing 3x7 pooled pixels each
  int Wc, Hc;                          // pool: extent of the (never written) convolution output
  int dbg;                             // TEMP experiment bits
  int direct;                          // epilogue writes y with plain 16-byte stores instead of a TMA bulk store
  int Wo, Ho;                          // extent of y (the pooled extent with `pool`)
  void* out;                           // y, dense [B, Ho, Wo, 64] bf16
  long long* prof;                     // TEMP
};

int halo_launch(int taps_y, int boxes /* 0 = one 10-pixel-wide box holding all three column taps */, int act, int pool, const CUtensorMap* tmA, const CUtensorMap* tmW, const CUtensorMap* tmOut,
                const HaloParams& p, int grid, cudaStream_t stream);

}  // namespace hpe
