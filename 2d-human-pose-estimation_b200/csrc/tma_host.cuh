// Host-side TMA descriptor encoding shared by the tcgen05 GEMM kernels (conv_gemm.cu, wgrad_gemm.cu).
#pragma once
#include "hpe_common.cuh"
#include <cuda.h>
#include <mutex>

namespace hpe {

// ---- host side --------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

inline int encode_map(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, const uint32_t* estrides, const char* what) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return fail(HPE_ECUDA, "cuTensorMapEncodeTiled entry point not available");
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), dims, strides_bytes, box, estrides,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(HPE_ECUDA, "cuTensorMapEncodeTiled(%s) failed with CUresult %d (dims %llu,%llu,%llu,%llu box %u,%u,%u,%u)",
                what, (int)r, (unsigned long long)dims[0], (unsigned long long)dims[1],
                (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0),
                box[0], box[1], rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
  return HPE_OK;
}


// (tw, th, nb) brick of `pixels` output pixels that minimises the number of bricks over a Wo x Ho x B grid
// (ties: wider tw)
inline void pick_brick(int pixels, int Wo, int Ho, int B, int max_tw, int* tw, int* th, int* nb) {
  long long best = -1;
  for (int a = 1; a <= pixels; a <<= 1) {
    if (a > max_tw) break;
    for (int b = 1; a * b <= pixels; b <<= 1) {
      const int c = pixels / (a * b);
      const long long tiles = (long long)((Wo + a - 1) / a) * ((Ho + b - 1) / b) * ((B + c - 1) / c);
      if (best < 0 || tiles < best || (tiles == best && a > *tw)) { best = tiles; *tw = a; *th = b; *nb = c; }
    }
  }
}

inline bool pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("HPE_B200_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
  return v != 0;
}

}  // namespace hpe
