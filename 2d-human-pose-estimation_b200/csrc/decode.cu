// Heatmap decode for sm_100a: per-joint first-max argmax + quarter-pixel refinement + inverse affine
// in ONE coalesced, 128-bit-vectorised pass over the heatmaps.
//
// HBM-bound: algorithmic bytes per (b,k) map = H*W*sizeof(T) read + 16 B written.  One warp owns one
// map at a time (grid-stride over maps); every lane keeps UNROLL independent 16-byte streaming loads
// in flight, tracks a private (value,index) best with a strict-greater update over its increasing
// indices, and the warp combines with the order-independent rule "greater value, else smaller index"
// so that ties resolve to the smallest flat index exactly like numpy/torch argmax.
#include "hpe_common.cuh"
#include <limits.h>
#include <math_constants.h>

namespace hpe {

constexpr int kDecodeThreads = 256;
constexpr int kDecodeUnroll  = 8;

__device__ __forceinline__ void upd(float v, int i, float& best, int& bidx) {
  if (v > best) { best = v; bidx = i; }      // NaN never wins (documented deviation, App. A.2)
}

template <bool BF16>
__device__ __forceinline__ float load_elem(const void* base, int i) {
  if (BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[i]);
  return reinterpret_cast<const float*>(base)[i];
}

template <bool BF16>
__global__ void __launch_bounds__(kDecodeThreads, 4)
decode_kernel(const void* __restrict__ hm, int n_maps, int K, int H, int W, int vec_ok,
              const float* __restrict__ affine, int refine,
              float* __restrict__ coords, float* __restrict__ maxvals, int32_t* __restrict__ idx_out,
              float* __restrict__ xyv) {
  constexpr int EPV   = BF16 ? 8 : 4;          // elements per 16-byte vector
  constexpr int ESIZE = BF16 ? 2 : 4;
  const int lane   = threadIdx.x & 31;
  const int warp   = (blockIdx.x * kDecodeThreads + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * kDecodeThreads) >> 5;
  const int HW     = H * W;
  const int nvec   = vec_ok ? HW / EPV : 0;

  for (int map = warp; map < n_maps; map += nwarps) {
    const char* base = reinterpret_cast<const char*>(hm) + (size_t)map * HW * ESIZE;
    float best = -CUDART_INF_F;
    int   bidx = INT_MAX;

    for (int v0 = 0; v0 < nvec; v0 += 32 * kDecodeUnroll) {
      uint4 r[kDecodeUnroll];
#pragma unroll
      for (int u = 0; u < kDecodeUnroll; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < nvec) r[u] = ld_stream_v4(base + (size_t)v * 16);
      }
#pragma unroll
      for (int u = 0; u < kDecodeUnroll; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < nvec) {
          const int e = v * EPV;
          if (BF16) {
            upd(bf16_lo(r[u].x), e + 0, best, bidx); upd(bf16_hi(r[u].x), e + 1, best, bidx);
            upd(bf16_lo(r[u].y), e + 2, best, bidx); upd(bf16_hi(r[u].y), e + 3, best, bidx);
            upd(bf16_lo(r[u].z), e + 4, best, bidx); upd(bf16_hi(r[u].z), e + 5, best, bidx);
            upd(bf16_lo(r[u].w), e + 6, best, bidx); upd(bf16_hi(r[u].w), e + 7, best, bidx);
          } else {
            upd(__uint_as_float(r[u].x), e + 0, best, bidx);
            upd(__uint_as_float(r[u].y), e + 1, best, bidx);
            upd(__uint_as_float(r[u].z), e + 2, best, bidx);
            upd(__uint_as_float(r[u].w), e + 3, best, bidx);
          }
        }
      }
    }
    // scalar tail (also the whole map when the 16-byte path is not usable)
    for (int i = nvec * EPV + lane; i < HW; i += 32) upd(load_elem<BF16>(base, i), i, best, bidx);

    // warp combine: greater value, else smaller index  -> first-max, order independent
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, off);
      const int   oi = __shfl_xor_sync(0xffffffffu, bidx, off);
      if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
    }

    if (lane == 0) {
      if (bidx == INT_MAX) { bidx = 0; best = load_elem<BF16>(base, 0); }   // all -inf / NaN map
      const int py = bidx / W, px = bidx - py * W;
      float x = 0.f, y = 0.f;
      if (best > 0.f) {                          // A.2 mask: coords zeroed when maxval <= 0
        x = (float)px; y = (float)py;
        if (refine && px > 1 && px < W - 1 && py > 1 && py < H - 1) {      // A.3, strictly interior
          const float dx = load_elem<BF16>(base, bidx + 1) - load_elem<BF16>(base, bidx - 1);
          const float dy = load_elem<BF16>(base, bidx + W) - load_elem<BF16>(base, bidx - W);
          x += dx > 0.f ? 0.25f : (dx < 0.f ? -0.25f : 0.f);
          y += dy > 0.f ? 0.25f : (dy < 0.f ? -0.25f : 0.f);
        }
      }
      if (affine) {                              // A.4
        const float* M = affine + (size_t)(map / K) * 6;
        const float xi = fmaf(M[0], x, fmaf(M[1], y, M[2]));
        const float yi = fmaf(M[3], x, fmaf(M[4], y, M[5]));
        x = xi; y = yi;
      }
      if (xyv) {                                 // packed (x, y, maxval) rows: the layout the result gather ships
        xyv[(size_t)map * 3 + 0] = x; xyv[(size_t)map * 3 + 1] = y; xyv[(size_t)map * 3 + 2] = best;
      } else {
        reinterpret_cast<float2*>(coords)[map] = make_float2(x, y);
        maxvals[map] = best;
      }
      if (idx_out) idx_out[map] = bidx;
    }
  }
}

static int decode_launch(const void* hm, int dtype, int B, int K, int H, int W, const float* affine,
                         int refine, float* coords, float* maxvals, int32_t* idx, cudaStream_t st, float* xyv = nullptr) {
  const long long n_maps_ll = (long long)B * K;
  const int n_maps = (int)n_maps_ll;
  const int esize = dtype == HPE_BF16 ? 2 : 4;
  const int epv = 16 / esize;
  const int vec_ok = (((uintptr_t)hm & 15) == 0) && ((H * W) % epv == 0);
  const int warps_per_cta = kDecodeThreads / 32;
  long long ctas = (n_maps_ll + warps_per_cta - 1) / warps_per_cta;
  const long long cap = (long long)num_sms() * 4 * 4;     // 4 resident CTAs/SM, a few waves
  if (ctas > cap) ctas = cap;
  if (dtype == HPE_BF16)
    decode_kernel<true><<<(int)ctas, kDecodeThreads, 0, st>>>(hm, n_maps, K, H, W, vec_ok, affine, refine,
                                                             coords, maxvals, idx, xyv);
  else
    decode_kernel<false><<<(int)ctas, kDecodeThreads, 0, st>>>(hm, n_maps, K, H, W, vec_ok, affine, refine,
                                                              coords, maxvals, idx, xyv);
  return launch_status("decode_kernel");
}

}  // namespace hpe

using namespace hpe;

static int decode_check(const void* hm, int dtype, int B, int K, int H, int W,
                        const float* coords, const float* maxvals) {
  HPE_CHECK_ARG(hm && coords && maxvals, "hpe_decode_heatmaps: null pointer");
  HPE_CHECK_ARG(dtype == HPE_F32 || dtype == HPE_BF16, "hpe_decode_heatmaps: bad dtype %d", dtype);
  HPE_CHECK_ARG(B > 0 && K > 0 && H > 0 && W > 0, "hpe_decode_heatmaps: non-positive size B=%d K=%d H=%d W=%d", B, K, H, W);
  HPE_CHECK_ARG((long long)B * K <= INT_MAX && (long long)H * W <= INT_MAX / 2,
                "hpe_decode_heatmaps: B*K or H*W overflows int32");
  return HPE_OK;
}

extern "C" int hpe_decode_heatmaps(const void* hm, int dtype, int B, int K, int H, int W,
                                   const float* affine, int refine,
                                   float* coords, float* maxvals, int32_t* idx, void* stream) {
  int rc = decode_check(hm, dtype, B, K, H, W, coords, maxvals);
  if (rc) return rc;
  return decode_launch(hm, dtype, B, K, H, W, affine, refine, coords, maxvals, idx, (cudaStream_t)stream);
}

extern "C" int hpe_decode_heatmaps_xyv(const void* hm, int dtype, int B, int K, int H, int W,
                                       const float* affine, int refine, float* xyv, int32_t* idx, void* stream) {
  int rc = decode_check(hm, dtype, B, K, H, W, xyv, xyv);
  if (rc) return rc;
  return decode_launch(hm, dtype, B, K, H, W, affine, refine, nullptr, nullptr, idx, (cudaStream_t)stream, xyv);
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

extern "C" size_t hpe_decode_host_scratch_bytes(int dtype, int B, int K, int H, int W) {
  const size_t esize = dtype == HPE_BF16 ? 2 : 4;
  const size_t maps = (size_t)B * K;
  return align256(maps * H * W * esize) + align256((size_t)B * 24) + align256(maps * 8) +
         align256(maps * 4) + align256(maps * 4);
}

extern "C" int hpe_decode_heatmaps_host(const void* h_hm, int dtype, int B, int K, int H, int W,
                                        const float* h_affine, int refine,
                                        float* h_coords, float* h_maxvals, int32_t* h_idx,
                                        void* d_scratch, size_t scratch_bytes, void* stream) {
  int rc = decode_check(h_hm, dtype, B, K, H, W, h_coords, h_maxvals);
  if (rc) return rc;
  HPE_CHECK_ARG(d_scratch && scratch_bytes >= hpe_decode_host_scratch_bytes(dtype, B, K, H, W),
                "hpe_decode_heatmaps_host: scratch too small (%zu bytes)", scratch_bytes);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t esize = dtype == HPE_BF16 ? 2 : 4;
  const size_t maps = (size_t)B * K;
  char* p = (char*)d_scratch;
  void* d_hm = p;                     p += align256(maps * H * W * esize);
  float* d_aff = (float*)p;           p += align256((size_t)B * 24);
  float* d_coords = (float*)p;        p += align256(maps * 8);
  float* d_max = (float*)p;           p += align256(maps * 4);
  int32_t* d_idx = (int32_t*)p;
  HPE_CUDA_OK(cudaMemcpyAsync(d_hm, h_hm, maps * H * W * esize, cudaMemcpyHostToDevice, st));
  if (h_affine) HPE_CUDA_OK(cudaMemcpyAsync(d_aff, h_affine, (size_t)B * 24, cudaMemcpyHostToDevice, st));
  rc = decode_launch(d_hm, dtype, B, K, H, W, h_affine ? d_aff : nullptr, refine, d_coords, d_max,
                     h_idx ? d_idx : nullptr, st);
  if (rc) return rc;
  HPE_CUDA_OK(cudaMemcpyAsync(h_coords, d_coords, maps * 8, cudaMemcpyDeviceToHost, st));
  HPE_CUDA_OK(cudaMemcpyAsync(h_maxvals, d_max, maps * 4, cudaMemcpyDeviceToHost, st));
  if (h_idx) HPE_CUDA_OK(cudaMemcpyAsync(h_idx, d_idx, maps * 4, cudaMemcpyDeviceToHost, st));
  HPE_CUDA_OK(cudaStreamSynchronize(st));
  return HPE_OK;
}
