// Gaussian target generation + target-weighted joints MSE for sm_100a.
//
// HBM-bound.  Fused mode never materialises the target: a warp owns one (b,k) map, streams the
// prediction with 16-byte loads, evaluates the Gaussian on the fly only where the vector intersects
// the +-3*sigma support, accumulates w^2 (p-g)^2 in fp32 in a fixed order (lane-sequential, then an
// xor-shuffle tree) and optionally writes dL/dpred in the same pass.  Maps whose weight is 0 are not
// read at all (their gradient is written as zeros).  A second single-CTA kernel sums the per-map
// partials in a fixed order, so the loss is run-to-run bit-stable.
#include <mutex>
#include "hpe_common.cuh"
#include <limits.h>

namespace hpe {

constexpr int kLossThreads = 256;
constexpr int kLossUnroll  = 4;

struct GaussSpec {
  int   mu_x, mu_y, t;
  float inv2s2;
  float w;          // effective target weight (A.5)
};

// A.5: weight and rounded centre of one joint
__device__ __forceinline__ GaussSpec make_spec(float x, float y, float v, int H, int W, float sigma, float sx, float sy) {
  GaussSpec g;
  float fx = truncf(x / sx + 0.5f), fy = truncf(y / sy + 0.5f);           // C int() truncation
  fx = fminf(fmaxf(fx, -1.0e6f), 1.0e6f);  fy = fminf(fmaxf(fy, -1.0e6f), 1.0e6f);
  if (!(fx == fx)) fx = -1.0e6f;
  if (!(fy == fy)) fy = -1.0e6f;
  g.mu_x = (int)fx; g.mu_y = (int)fy;
  g.t = (int)(3.0f * sigma);
  g.inv2s2 = 1.0f / (2.0f * sigma * sigma);
  g.w = v > 0.f ? 1.f : 0.f;
  const int ulx = g.mu_x - g.t, uly = g.mu_y - g.t, brx = g.mu_x + g.t + 1, bry = g.mu_y + g.t + 1;
  if (ulx >= W || uly >= H || brx < 0 || bry < 0) g.w = 0.f;
  return g;
}

__device__ __forceinline__ float gauss_at(const GaussSpec& g, int x, int y) {
  const int dx = x - g.mu_x, dy = y - g.mu_y;
  if (abs(dx) > g.t || abs(dy) > g.t) return 0.f;
  const float d2 = (float)(dy * dy) + (float)(dx * dx);
  return expf(-(d2 * g.inv2s2));
}

__device__ __forceinline__ float warp_sum_fixed(float v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

// ---- fused: pred + joints ------------------------------------------------------------------------
template <bool BF16, bool GRAD>
__global__ void __launch_bounds__(kLossThreads, 4)
mse_fused_kernel(const void* __restrict__ pred, const float* __restrict__ joints, int n_maps, int H, int W,
                 int vec_ok, float sigma, float sx, float sy, float* __restrict__ partials,
                 void* __restrict__ grad, float gcoef, float* __restrict__ weight_out) {
  constexpr int EPV   = BF16 ? 8 : 4;
  constexpr int ESIZE = BF16 ? 2 : 4;
  const int lane   = threadIdx.x & 31;
  const int warp   = (blockIdx.x * kLossThreads + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * kLossThreads) >> 5;
  const int HW     = H * W;
  const int nvec   = vec_ok ? HW / EPV : 0;       // vec_ok also guarantees W % EPV == 0
  // This pass is ALU-bound, not latency-bound (r02 measurement: forward-only moved exactly half the bytes per second of the
  // forward+backward variant, bf16 half of fp32, and deeper unrolling made it slower): what counts is instructions per element.
  // Hence no integer division per vector, and vectors outside the +-3 sigma support (94 % of a 64x48 map) take the short
  // path d = p (w == 1 here: maps with w == 0 never reach this loop), one FMA per element.
  constexpr int U  = kLossUnroll;
  const float inv_w = 1.0f / (float)W;            // row of a vector without an integer division (exact: e < 2^22, see below)

  float jx = 0.f, jy = 0.f, jv = 0.f;
  if (warp < n_maps) { jx = __ldg(joints + (size_t)warp * 3); jy = __ldg(joints + (size_t)warp * 3 + 1); jv = __ldg(joints + (size_t)warp * 3 + 2); }
  for (int map = warp; map < n_maps; map += nwarps) {
    const GaussSpec g = make_spec(jx, jy, jv, H, W, sigma, sx, sy);
    if (map + nwarps < n_maps) {                  // next map's joint: its latency hides behind this map's stream
      const size_t nj = (size_t)(map + nwarps) * 3;
      jx = __ldg(joints + nj); jy = __ldg(joints + nj + 1); jv = __ldg(joints + nj + 2);
    }
    const char* base = reinterpret_cast<const char*>(pred) + (size_t)map * HW * ESIZE;
    char* gbase = GRAD ? reinterpret_cast<char*>(grad) + (size_t)map * HW * ESIZE : nullptr;
    if (lane == 0 && weight_out) weight_out[map] = g.w;
    float acc = 0.f;
    if (g.w == 0.f) {
      if (GRAD) {
        for (int v = lane; v < nvec; v += 32) st_stream_v4(gbase + (size_t)v * 16, make_uint4(0, 0, 0, 0));
        for (int i = nvec * EPV + lane; i < HW; i += 32) {
          if (BF16) reinterpret_cast<__nv_bfloat16*>(gbase)[i] = __float2bfloat16(0.f);
          else      reinterpret_cast<float*>(gbase)[i] = 0.f;
        }
      }
      if (lane == 0) partials[map] = 0.f;
      continue;
    }
    // first / one-past-last vector of the rows the Gaussian support touches (one row per vector: W % EPV == 0)
    const int vpr = W / EPV;
    const int v_lo = max(g.mu_y - g.t, 0) * vpr, v_hi = min(g.mu_y + g.t + 1, H) * vpr;
    const float gc = gcoef * g.w;                 // dL/dp = gscale * w^2 (p-g)/n = (w p - w g) * gscale * w / n
    for (int v0 = 0; v0 < nvec; v0 += 32 * U) {
      uint4 r[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < nvec) r[u] = ld_stream_v4(base + (size_t)v * 16);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < nvec) {
          const int e = v * EPV;
          float p[EPV];
          if constexpr (BF16) {
            p[0] = bf16_lo(r[u].x); p[1] = bf16_hi(r[u].x); p[2] = bf16_lo(r[u].y); p[3] = bf16_hi(r[u].y);
            p[4] = bf16_lo(r[u].z); p[5] = bf16_hi(r[u].z); p[6] = bf16_lo(r[u].w); p[7] = bf16_hi(r[u].w);
          } else {
            p[0] = __uint_as_float(r[u].x); p[1] = __uint_as_float(r[u].y);
            p[2] = __uint_as_float(r[u].z); p[3] = __uint_as_float(r[u].w);
          }
          // vectors of rows mu_y - t .. mu_y + t only (two compares); the row / column of a vector is computed inside that
          // window alone: y = floor(e / W) from the fp32 product, corrected by one step -- exact for e < 2^22, no integer division
          bool hit = v >= v_lo && v < v_hi;
          int y = 0, x0 = 0;
          if (hit) {
            y = (int)(((float)e + 0.5f) * inv_w);
            x0 = e - y * W;
            if (x0 < 0) { --y; x0 += W; } else if (x0 >= W) { ++y; x0 -= W; }
            hit = abs(y - g.mu_y) <= g.t && x0 <= g.mu_x + g.t && x0 + EPV - 1 >= g.mu_x - g.t;
          }
          float d[EPV];
          if (!hit) {                                 // g == 0 and w == 1:  w*p - w*0 == p  (same bits as the general form)
#pragma unroll
            for (int j = 0; j < EPV; ++j) { d[j] = p[j]; acc = fmaf(p[j], p[j], acc); }
          } else {
#pragma unroll
            for (int j = 0; j < EPV; ++j) {
              const float gt = gauss_at(g, x0 + j, y);
              d[j] = g.w * p[j] - g.w * gt;
              acc = fmaf(d[j], d[j], acc);
            }
          }
          if (GRAD) {
            uint4 o;
            if constexpr (BF16) {
              o.x = pack_bf16(d[0] * gc, d[1] * gc); o.y = pack_bf16(d[2] * gc, d[3] * gc);
              o.z = pack_bf16(d[4] * gc, d[5] * gc); o.w = pack_bf16(d[6] * gc, d[7] * gc);
            } else {
              o.x = __float_as_uint(d[0] * gc); o.y = __float_as_uint(d[1] * gc);
              o.z = __float_as_uint(d[2] * gc); o.w = __float_as_uint(d[3] * gc);
            }
            st_stream_v4(gbase + (size_t)v * 16, o);
          }
        }
      }
    }
    for (int i = nvec * EPV + lane; i < HW; i += 32) {       // scalar path for odd shapes
      const int y = i / W, x = i - y * W;
      float p;
      if (BF16) p = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[i]);
      else      p = reinterpret_cast<const float*>(base)[i];
      const float d = g.w * p - g.w * gauss_at(g, x, y);
      acc = fmaf(d, d, acc);
      if (GRAD) {
        if (BF16) reinterpret_cast<__nv_bfloat16*>(gbase)[i] = __float2bfloat16(d * gc);
        else      reinterpret_cast<float*>(gbase)[i] = d * gc;
      }
    }
    acc = warp_sum_fixed(acc);
    if (lane == 0) partials[map] = acc;
  }
}

// ---- dense: pred + materialised fp32 target + weight ---------------------------------------------
template <bool BF16, bool GRAD>
__global__ void __launch_bounds__(kLossThreads, 4)
mse_dense_kernel(const void* __restrict__ pred, const float* __restrict__ target, const float* __restrict__ weight,
                 int n_maps, int HW, int vec_ok, float* __restrict__ partials, void* __restrict__ grad, float gcoef) {
  constexpr int ESIZE = BF16 ? 2 : 4;
  const int lane   = threadIdx.x & 31;
  const int warp   = (blockIdx.x * kLossThreads + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * kLossThreads) >> 5;
  const int nvec   = vec_ok ? HW / 4 : 0;           // 4 elements per step: 16 B of target

  for (int map = warp; map < n_maps; map += nwarps) {
    const float w = weight[map];
    const char* pb = reinterpret_cast<const char*>(pred) + (size_t)map * HW * ESIZE;
    const float* tb = target + (size_t)map * HW;
    char* gb = GRAD ? reinterpret_cast<char*>(grad) + (size_t)map * HW * ESIZE : nullptr;
    float acc = 0.f;
    for (int v0 = 0; v0 < nvec; v0 += 32 * kLossUnroll) {
      uint4 t[kLossUnroll]; uint4 p[kLossUnroll];
#pragma unroll
      for (int u = 0; u < kLossUnroll; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < nvec) {
          t[u] = ld_stream_v4(tb + (size_t)v * 4);
          if (BF16) { const uint2 q = __ldg(reinterpret_cast<const uint2*>(pb) + v); p[u].x = q.x; p[u].y = q.y; }
          else      p[u] = ld_stream_v4(pb + (size_t)v * 16);
        }
      }
#pragma unroll
      for (int u = 0; u < kLossUnroll; ++u) {
        const int v = v0 + u * 32 + lane;
        if (v < nvec) {
          float pv[4], d[4];
          if (BF16) { pv[0] = bf16_lo(p[u].x); pv[1] = bf16_hi(p[u].x); pv[2] = bf16_lo(p[u].y); pv[3] = bf16_hi(p[u].y); }
          else      { pv[0] = __uint_as_float(p[u].x); pv[1] = __uint_as_float(p[u].y);
                      pv[2] = __uint_as_float(p[u].z); pv[3] = __uint_as_float(p[u].w); }
          const float tv[4] = {__uint_as_float(t[u].x), __uint_as_float(t[u].y),
                               __uint_as_float(t[u].z), __uint_as_float(t[u].w)};
#pragma unroll
          for (int j = 0; j < 4; ++j) { d[j] = w * pv[j] - w * tv[j]; acc = fmaf(d[j], d[j], acc); }
          if (GRAD) {
            // dL/dp = gscale * w^2 (p - g) / n = (w p - w g) * (gscale * w / n)
            const float c = gcoef * w;
            if (BF16) {
              uint2 o; o.x = pack_bf16(d[0] * c, d[1] * c); o.y = pack_bf16(d[2] * c, d[3] * c);
              reinterpret_cast<uint2*>(gb)[v] = o;
            } else {
              st_stream_v4(gb + (size_t)v * 16, make_uint4(__float_as_uint(d[0] * c), __float_as_uint(d[1] * c),
                                                           __float_as_uint(d[2] * c), __float_as_uint(d[3] * c)));
            }
          }
        }
      }
    }
    for (int i = nvec * 4 + lane; i < HW; i += 32) {
      float p;
      if (BF16) p = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(pb)[i]);
      else      p = reinterpret_cast<const float*>(pb)[i];
      const float d = w * p - w * tb[i];
      acc = fmaf(d, d, acc);
      if (GRAD) {
        const float c = gcoef * w;
        if (BF16) reinterpret_cast<__nv_bfloat16*>(gb)[i] = __float2bfloat16(d * c);
        else      reinterpret_cast<float*>(gb)[i] = d * c;
      }
    }
    acc = warp_sum_fixed(acc);
    if (lane == 0) partials[map] = acc;
  }
}

// ---- A.5 alone: rasterise targets ------------------------------------------------------------------
__global__ void __launch_bounds__(kLossThreads, 4)
generate_target_kernel(const float* __restrict__ joints, int n_maps, int H, int W, int vec_ok,
                       float sigma, float sx, float sy, float* __restrict__ target, float* __restrict__ weight) {
  const int lane   = threadIdx.x & 31;
  const int warp   = (blockIdx.x * kLossThreads + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * kLossThreads) >> 5;
  const int HW     = H * W;
  const int nvec   = vec_ok ? HW / 4 : 0;
  for (int map = warp; map < n_maps; map += nwarps) {
    const GaussSpec g = make_spec(joints[(size_t)map * 3 + 0], joints[(size_t)map * 3 + 1], joints[(size_t)map * 3 + 2], H, W, sigma, sx, sy);
    if (lane == 0) weight[map] = g.w;
    float* tb = target + (size_t)map * HW;
    for (int v = lane; v < nvec; v += 32) {
      const int e = v * 4, y = e / W, x0 = e - y * W;
      float o[4] = {0.f, 0.f, 0.f, 0.f};
      if (g.w != 0.f && abs(y - g.mu_y) <= g.t && x0 <= g.mu_x + g.t && x0 + 3 >= g.mu_x - g.t) {
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] = gauss_at(g, x0 + j, y);
      }
      st_stream_v4(tb + (size_t)v * 4, make_uint4(__float_as_uint(o[0]), __float_as_uint(o[1]),
                                                  __float_as_uint(o[2]), __float_as_uint(o[3])));
    }
    for (int i = nvec * 4 + lane; i < HW; i += 32) {
      const int y = i / W, x = i - y * W;
      tb[i] = g.w != 0.f ? gauss_at(g, x, y) : 0.f;
    }
  }
}

// ---- stage 2: fixed-order sum of the per-map partials ----------------------------------------------
__global__ void __launch_bounds__(1024, 1)
reduce_partials_kernel(const float* __restrict__ partials, int n, double coef, float* __restrict__ loss) {
  __shared__ double sh[1024];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += 1024) acc += (double)partials[i];
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) *loss = (float)(sh[0] * coef);
}

static int loss_grid(long long n_maps) {
  long long ctas = (n_maps + (kLossThreads / 32) - 1) / (kLossThreads / 32);
  const long long cap = (long long)num_sms() * 4 * 4;
  return (int)(ctas > cap ? cap : ctas);
}

}  // namespace hpe

using namespace hpe;

static int loss_check(const char* fn, const void* pred, int dtype, int B, int K, int H, int W,
                      const float* partials, const float* loss) {
  HPE_CHECK_ARG(pred && partials && loss, "%s: null pointer", fn);
  HPE_CHECK_ARG(dtype == HPE_F32 || dtype == HPE_BF16, "%s: bad dtype %d", fn, dtype);
  HPE_CHECK_ARG(B > 0 && K > 0 && H > 0 && W > 0, "%s: non-positive size B=%d K=%d H=%d W=%d", fn, B, K, H, W);
  HPE_CHECK_ARG((long long)B * K <= INT_MAX && (long long)H * W <= INT_MAX / 2, "%s: size overflows int32", fn);
  return HPE_OK;
}

extern "C" int hpe_joints_mse_fused(const void* pred, int dtype, const float* joints,
                                    int B, int K, int H, int W, float sigma, float stride_x, float stride_y,
                                    float* partials, float* loss, void* grad, float grad_scale,
                                    float* weight_out, void* stream) {
  int rc = loss_check("hpe_joints_mse_fused", pred, dtype, B, K, H, W, partials, loss);
  if (rc) return rc;
  HPE_CHECK_ARG(joints, "hpe_joints_mse_fused: null joints");
  HPE_CHECK_ARG(sigma > 0.f && stride_x > 0.f && stride_y > 0.f, "hpe_joints_mse_fused: sigma/stride must be > 0");
  cudaStream_t st = (cudaStream_t)stream;
  const int n_maps = B * K, HW = H * W;
  const int epv = dtype == HPE_BF16 ? 8 : 4;
  const int vec_ok = (((uintptr_t)pred & 15) == 0) && (!grad || ((uintptr_t)grad & 15) == 0) && (W % epv == 0) &&
                     HW <= (1 << 22);      // the kernel's fp32 row computation is exact up to 2^22 elements per map
  const double n = (double)K * B * HW;
  const float gcoef = (float)((double)grad_scale / n);
  const int grid = loss_grid(n_maps);
#define HPE_LAUNCH_FUSED(BF, GR)                                                                         \
  mse_fused_kernel<BF, GR><<<grid, kLossThreads, 0, st>>>(pred, joints, n_maps, H, W, vec_ok, sigma,     \
                                                           stride_x, stride_y, partials, grad, gcoef, weight_out)
  if (dtype == HPE_BF16) { if (grad) HPE_LAUNCH_FUSED(true, true); else HPE_LAUNCH_FUSED(true, false); }
  else                   { if (grad) HPE_LAUNCH_FUSED(false, true); else HPE_LAUNCH_FUSED(false, false); }
#undef HPE_LAUNCH_FUSED
  rc = launch_status("mse_fused_kernel");
  if (rc) return rc;
  reduce_partials_kernel<<<1, 1024, 0, st>>>(partials, n_maps, 0.5 / n, loss);
  return launch_status("reduce_partials_kernel");
}

// ---- allocation-free variant (no counterpart in the reference) -------------------
// The per-map partial sums live in a file-scope device array instead of caller-provided scratch, so a caller that may not call
// cudaMalloc (graph capture, serving loops) needs device memory for inputs and outputs only.  Every GPU that loads the library
// holds the array (kStaticLossMaps = 131072 floats = 512 KB) until exit.  The kernels overwrite partials[0 .. B*K) completely
// (maps with weight 0 store 0) and stage 2 reads exactly that range, so nothing has to be cleared between calls; all calls share
// the array: calls that overlap on one GPU (two streams, two threads) race on it.
namespace hpe {
constexpr int kStaticLossMaps = 1 << 17;
static __device__ __align__(256) float g_static_partials[kStaticLossMaps];
}  // namespace hpe

// Device address of g_static_partials on the CURRENT device.  In host code the array's name is only a handle for
// cudaGetSymbolAddress, never a pointer a kernel may dereference; the address differs per device, so it is looked up once per
// device, with that device current.
static int static_partials_ptr(float** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static float* addr[kMaxDev] = {};
  int dev = 0;
  HPE_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDev) return fail(HPE_EUNSUPPORTED, "hpe_joints_mse_fused_noalloc: device ordinal %d >= %d", dev, kMaxDev);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    float* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, hpe::g_static_partials);
    if (e != cudaSuccess || !ptr)
      return fail(HPE_ECUDA, "hpe_joints_mse_fused_noalloc: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
    addr[dev] = ptr;
  }
  *out = addr[dev];
  return HPE_OK;
}

extern "C" int hpe_static_loss_max_maps(void) { return hpe::kStaticLossMaps; }

extern "C" int hpe_joints_mse_fused_noalloc(const void* pred, int dtype, const float* joints,
                                            int B, int K, int H, int W, float sigma, float stride_x, float stride_y,
                                            float* loss, void* grad, float grad_scale, float* weight_out, void* stream) {
  HPE_CHECK_ARG(B > 0 && K > 0, "hpe_joints_mse_fused_noalloc: non-positive size B=%d K=%d", B, K);
  if ((long long)B * K > hpe::kStaticLossMaps)
    return fail(HPE_EUNSUPPORTED, "hpe_joints_mse_fused_noalloc: B*K = %lld maps exceed the built-in scratch of %d maps; use "
                "hpe_joints_mse_fused with caller-provided partials", (long long)B * K, hpe::kStaticLossMaps);
  float* partials = nullptr;
  const int rc = static_partials_ptr(&partials);
  if (rc) return rc;
  return hpe_joints_mse_fused(pred, dtype, joints, B, K, H, W, sigma, stride_x, stride_y, partials, loss, grad, grad_scale,
                              weight_out, stream);
}

extern "C" int hpe_joints_mse_dense(const void* pred, int dtype, const float* target, const float* weight,
                                    int B, int K, int H, int W,
                                    float* partials, float* loss, void* grad, float grad_scale, void* stream) {
  int rc = loss_check("hpe_joints_mse_dense", pred, dtype, B, K, H, W, partials, loss);
  if (rc) return rc;
  HPE_CHECK_ARG(target && weight, "hpe_joints_mse_dense: null target/weight");
  cudaStream_t st = (cudaStream_t)stream;
  const int n_maps = B * K, HW = H * W;
  const int vec_ok = (((uintptr_t)pred & 15) == 0) && (((uintptr_t)target & 15) == 0) &&
                     (!grad || ((uintptr_t)grad & 15) == 0) && (HW % 4 == 0);
  const double n = (double)K * B * HW;
  const float gcoef = (float)((double)grad_scale / n);
  const int grid = loss_grid(n_maps);
#define HPE_LAUNCH_DENSE(BF, GR)                                                                       \
  mse_dense_kernel<BF, GR><<<grid, kLossThreads, 0, st>>>(pred, target, weight, n_maps, HW, vec_ok,    \
                                                           partials, grad, gcoef)
  if (dtype == HPE_BF16) { if (grad) HPE_LAUNCH_DENSE(true, true); else HPE_LAUNCH_DENSE(true, false); }
  else                   { if (grad) HPE_LAUNCH_DENSE(false, true); else HPE_LAUNCH_DENSE(false, false); }
#undef HPE_LAUNCH_DENSE
  rc = launch_status("mse_dense_kernel");
  if (rc) return rc;
  reduce_partials_kernel<<<1, 1024, 0, st>>>(partials, n_maps, 0.5 / n, loss);
  return launch_status("reduce_partials_kernel");
}

extern "C" int hpe_generate_target(const float* joints, int B, int K, int H, int W,
                                   float sigma, float stride_x, float stride_y,
                                   float* target, float* weight, void* stream) {
  HPE_CHECK_ARG(joints && target && weight, "hpe_generate_target: null pointer");
  HPE_CHECK_ARG(B > 0 && K > 0 && H > 0 && W > 0, "hpe_generate_target: non-positive size");
  HPE_CHECK_ARG((long long)B * K <= INT_MAX && (long long)H * W <= INT_MAX / 2, "hpe_generate_target: size overflows int32");
  HPE_CHECK_ARG(sigma > 0.f && stride_x > 0.f && stride_y > 0.f, "hpe_generate_target: sigma/stride must be > 0");
  const int n_maps = B * K;
  const int vec_ok = (((uintptr_t)target & 15) == 0) && (W % 4 == 0);
  generate_target_kernel<<<loss_grid(n_maps), kLossThreads, 0, (cudaStream_t)stream>>>(
      joints, n_maps, H, W, vec_ok, sigma, stride_x, stride_y, target, weight);
  return launch_status("generate_target_kernel");
}
