// Keypoint metrics on device: consumers of decoded keypoints [N,K,2|3] in image px.
//   oks_kernel            = OKS.compute            (/root/reference/HPE/core/metric/oksap.py:32-74)
//   pck_kernel            = PCK.compute            (/root/reference/HPE/core/metric/pck.py:14-104)
//   keypoint_loss_kernel  = KeypointLoss.forward   (/root/reference/HPE/core/loss/keypoint.py:13-21)
// One warp per instance, lanes over joints, shuffle reductions in a fixed order (run-to-run bit-stable);
// the batch-level sums (PCK counts, KeypointLoss mean) finish in a single-CTA fixed-order second stage.
#include "hpe_common.cuh"

namespace hpe {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__global__ void __launch_bounds__(256)
oks_kernel(const float* __restrict__ gt, const float* __restrict__ pred, const float* __restrict__ areas,
           const float* __restrict__ sigmas, int N, int K, float vis_thres, float eps, float* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= N) return;
  const float area = areas[warp];
  float num = 0.f, cnt = 0.f;
  for (int k = lane; k < K; k += 32) {
    const float* g = gt + (static_cast<size_t>(warp) * K + k) * 3;
    const float* p = pred + (static_cast<size_t>(warp) * K + k) * 2;
    const float dx = p[0] - g[0], dy = p[1] - g[1];
    const float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
    const float s2 = sigmas[k] * 2.f;
    const float denom = __fadd_rn(__fmul_rn(__fmul_rn(2.f, area), __fmul_rn(s2, s2)), eps);
    const float vis = g[2] > vis_thres ? 1.f : 0.f;
    num += expf(-d2 / denom) * vis;
    cnt += vis;
  }
  num = warp_sum(num); cnt = warp_sum(cnt);
  if (lane == 0) out[warp] = num / (cnt + eps);
}

// per-instance (correct, visible) counts; `visible` may be NULL (all visible), `normalizer` NULL (gt bbox) / [1] / [N]
__global__ void __launch_bounds__(256)
pck_kernel(const float* __restrict__ gt, const float* __restrict__ pred, const unsigned char* __restrict__ visible,
           const float* __restrict__ normalizer, int norm_n, int N, int K, float alpha, float eps,
           int* __restrict__ per_inst) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= N) return;
  float L;
  if (normalizer) {
    L = normalizer[norm_n == 1 ? 0 : warp];
  } else {
    float x0 = INFINITY, y0 = INFINITY, x1 = -INFINITY, y1 = -INFINITY;
    for (int k = lane; k < K; k += 32) {
      if (visible && !visible[static_cast<size_t>(warp) * K + k]) continue;
      const float* g = gt + (static_cast<size_t>(warp) * K + k) * 2;
      x0 = fminf(x0, g[0]); x1 = fmaxf(x1, g[0]); y0 = fminf(y0, g[1]); y1 = fmaxf(y1, g[1]);
    }
    x0 = warp_min(x0); y0 = warp_min(y0); x1 = warp_max(x1); y1 = warp_max(y1);
    L = fmaxf(x1 - x0, y1 - y0);                 // nothing visible: -inf - inf = -inf, clamped below
  }
  L = L < eps ? eps : L;                          // torch.clamp(min=eps); NaN stays NaN
  const float thr = __fmul_rn(alpha, L);
  int correct = 0, vis = 0;
  for (int k = lane; k < K; k += 32) {
    const bool v = !visible || visible[static_cast<size_t>(warp) * K + k];
    const float* g = gt + (static_cast<size_t>(warp) * K + k) * 2;
    const float* p = pred + (static_cast<size_t>(warp) * K + k) * 2;
    const float dx = g[0] - p[0], dy = g[1] - p[1];
    const float dist = sqrtf(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
    correct += (v && dist <= thr) ? 1 : 0;
    vis += v ? 1 : 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { correct += __shfl_xor_sync(0xffffffffu, correct, o); vis += __shfl_xor_sync(0xffffffffu, vis, o); }
  if (lane == 0) { per_inst[2 * warp] = correct; per_inst[2 * warp + 1] = vis; }
}

__global__ void __launch_bounds__(256)
pck_reduce_kernel(const int* __restrict__ per_inst, int N, long long* __restrict__ counts) {
  __shared__ long long sc[256], sv[256];
  long long c = 0, v = 0;
  for (int i = threadIdx.x; i < N; i += 256) { c += per_inst[2 * i]; v += per_inst[2 * i + 1]; }
  sc[threadIdx.x] = c; sv[threadIdx.x] = v;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) { sc[threadIdx.x] += sc[threadIdx.x + s]; sv[threadIdx.x] += sv[threadIdx.x + s]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { counts[0] = sc[0]; counts[1] = sv[0]; }
}

__global__ void __launch_bounds__(256)
keypoint_loss_kernel(const float* __restrict__ pred, int pred_stride, const float* __restrict__ gt, int gt_stride,
                     const float* __restrict__ mask, const float* __restrict__ area, const float* __restrict__ sigmas,
                     int N, int K, float* __restrict__ per_inst) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= N) return;
  const float a = area[warp] + 1e-9f;
  float nz = 0.f, acc = 0.f;
  for (int k = lane; k < K; k += 32) {
    const float* p = pred + (static_cast<size_t>(warp) * K + k) * pred_stride;
    const float* g = gt + (static_cast<size_t>(warp) * K + k) * gt_stride;
    const float m = mask[static_cast<size_t>(warp) * K + k];
    const float dx = p[0] - g[0], dy = p[1] - g[1];
    const float d = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
    const float s2 = 2.f * sigmas[k];
    const float e = d / (s2 * s2 * a * 2.f);
    acc += (1.f - expf(-e)) * m;
    nz += m != 0.f ? 1.f : 0.f;
  }
  acc = warp_sum(acc); nz = warp_sum(nz);
  if (lane == 0) per_inst[warp] = static_cast<float>(K) / (nz + 1e-9f) * acc;
}

__global__ void __launch_bounds__(256)
mean_reduce_kernel(const float* __restrict__ v, int N, double denom, float* __restrict__ out) {
  __shared__ double s[256];
  double a = 0.0;
  for (int i = threadIdx.x; i < N; i += 256) a += static_cast<double>(v[i]);
  s[threadIdx.x] = a;
  __syncthreads();
  for (int k = 128; k > 0; k >>= 1) { if (threadIdx.x < k) s[threadIdx.x] += s[threadIdx.x + k]; __syncthreads(); }
  if (threadIdx.x == 0) out[0] = static_cast<float>(s[0] / denom);
}

}  // namespace hpe
using namespace hpe;

static inline int warps_grid(int N) { return (N * 32 + 255) / 256; }

extern "C" int hpe_oks(const float* gt, const float* pred, const float* areas, const float* sigmas, int N, int K,
                       float in_vis_thres, float eps, float* out, void* stream) {
  HPE_CHECK_ARG(gt && pred && areas && sigmas && out, "hpe_oks: null pointer");
  HPE_CHECK_ARG(N > 0 && K > 0, "hpe_oks: non-positive size");
  oks_kernel<<<warps_grid(N), 256, 0, (cudaStream_t)stream>>>(gt, pred, areas, sigmas, N, K, in_vis_thres, eps, out);
  return launch_status("oks_kernel");
}

extern "C" int hpe_pck(const float* gt, const float* pred, const unsigned char* visible, const float* normalizer, int norm_n,
                       int N, int K, float alpha, float eps, int32_t* scratch, long long* counts, void* stream) {
  HPE_CHECK_ARG(gt && pred && scratch && counts, "hpe_pck: null pointer");
  HPE_CHECK_ARG(N > 0 && K > 0, "hpe_pck: non-positive size");
  HPE_CHECK_ARG(!normalizer || norm_n == 1 || norm_n == N, "hpe_pck: normalizer must have 1 or N elements");
  pck_kernel<<<warps_grid(N), 256, 0, (cudaStream_t)stream>>>(gt, pred, visible, normalizer, norm_n, N, K, alpha, eps, scratch);
  int rc = launch_status("pck_kernel");
  if (rc) return rc;
  pck_reduce_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(scratch, N, counts);
  return launch_status("pck_reduce_kernel");
}

extern "C" int hpe_keypoint_loss(const float* pred, int pred_stride, const float* gt, int gt_stride, const float* mask,
                                 const float* area, const float* sigmas, int N, int K, float* scratch, float* loss,
                                 void* stream) {
  HPE_CHECK_ARG(pred && gt && mask && area && sigmas && scratch && loss, "hpe_keypoint_loss: null pointer");
  HPE_CHECK_ARG(N > 0 && K > 0 && pred_stride >= 2 && gt_stride >= 2, "hpe_keypoint_loss: bad size/stride");
  keypoint_loss_kernel<<<warps_grid(N), 256, 0, (cudaStream_t)stream>>>(pred, pred_stride, gt, gt_stride, mask, area, sigmas, N, K, scratch);
  int rc = launch_status("keypoint_loss_kernel");
  if (rc) return rc;
  mean_reduce_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(scratch, N, (double)N * K, loss);
  return launch_status("mean_reduce_kernel");
}
