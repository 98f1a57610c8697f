// Fused device path of the reference's detection-style pose loss:
// ComputeLoss.__call__ (/root/reference/HPE/loss.py:349-448) with VarifocalLoss (loss.py:9-27),
// HybridKeypointLoss = wing-OKS + 0.5 L1 + 0.1 smoothness (loss.py:41-84, 171-222), BCE visibility,
// point2box_xywh (loss.py:277-315) and CIoU bbox_iou (loss.py:317-347).
//
// The reference loops over positives in Python, allocates dense zero targets per level and syncs the host
// three times (`.item()`); here one call = a handful of stream-ordered launches, no host sync:
//   closs_targets_kernel  warp per target row: positive cell, keypoint terms, robust box + CIoU -> soft label
//   closs_level_sums      fixed-order reduction of the per-target partials (deterministic)
//   closs_cls_kernel      varifocal class term over ALL cells (the bandwidth-bound part) + its gradient
//   closs_kptgrad_kernel  analytic gradient of the keypoint / visibility terms at the positive cells
//   closs_final_kernel    normalise by num_pos and assemble {total, cls, kpt, vis}
// Soft labels of duplicate rows in one cell: the LAST row wins (CPU index_put order), made deterministic with a
// 64-bit atomicMax on (row+1, iou bits).
#include "hpe_common.cuh"

namespace hpe {

constexpr int kCLMaxLevels = 4;
constexpr int kCLPart = 5;                // per-target partials: wing, l1, mask, smooth, bce_vis
constexpr int kCLWarps = 4;               // warps per CTA in the per-target kernels
constexpr int kClsThreads = 256;

struct CLLevel {
  const float* pred; float* grad;
  unsigned long long* tmap;               // [B*H*W*nc] packed (row+1, iou bits); 0 = background
  int* cellmask;                          // [B*H*W]
  float* partials;                        // [n*kCLPart]
  float* cls_partials;                    // [cls_grid]
  int H, W, cls_grid;
};
struct CLParams {
  CLLevel lv[kCLMaxLevels];
  int L, B, nc, K, n;
  const float* targets; const float* sigmas;
  int* num_pos;                           // [1] unique positive cells over all levels
  float* level_sums;                      // [L*kCLPart]
  float* out;                             // [4]
  float grad_scale;
};

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float softplus_f(float x) { return fmaxf(x, 0.f) + log1pf(expf(-fabsf(x))); }
__device__ __forceinline__ float sigm(float x) { return 1.f / (1.f + expf(-x)); }

// CIoU of (cx,cy,w,h) boxes, loss.py:317-347
__device__ float ciou_xywh(float ax, float ay, float aw, float ah, float bx, float by, float bw, float bh) {
  const float eps = 1e-7f;
  const float a_x1 = ax - aw / 2, a_x2 = ax + aw / 2, a_y1 = ay - ah / 2, a_y2 = ay + ah / 2;
  const float b_x1 = bx - bw / 2, b_x2 = bx + bw / 2, b_y1 = by - bh / 2, b_y2 = by + bh / 2;
  const float inter = fmaxf(fminf(a_x2, b_x2) - fmaxf(a_x1, b_x1), 0.f) * fmaxf(fminf(a_y2, b_y2) - fmaxf(a_y1, b_y1), 0.f);
  const float uni = aw * ah + bw * bh - inter + eps;
  const float iou = inter / uni;
  const float cw = fmaxf(a_x2, b_x2) - fminf(a_x1, b_x1), ch = fmaxf(a_y2, b_y2) - fminf(a_y1, b_y1);
  const float c2 = cw * cw + ch * ch + eps;
  const float rx = a_x1 + a_x2 - b_x1 - b_x2, ry = a_y1 + a_y2 - b_y1 - b_y2;
  const float rho2 = (rx * rx + ry * ry) / 4.f;
  const float dv = atanf(bw / (bh + eps)) - atanf(aw / (ah + eps));
  const float v = static_cast<float>(4.0 / (3.141592 * 3.141592)) * dv * dv;
  const float alpha = v / (v - iou + (1.f + eps));
  const float r = iou - (rho2 / c2 + v * alpha);
  return isnan(r) ? 0.f : r;
}

struct TargetGeom { int b, c, gx, gy; float bx, by, bw, bh; };
__device__ __forceinline__ TargetGeom target_geom(const float* row, int H, int W) {
  TargetGeom g;
  g.b = static_cast<int>(row[0]); g.c = static_cast<int>(row[1]);
  g.bx = row[2] * W; g.by = row[3] * H; g.bw = row[4] * W; g.bh = row[5] * H;
  g.gx = min(max(static_cast<int>(g.bx), 0), W - 1);
  g.gy = min(max(static_cast<int>(g.by), 0), H - 1);
  return g;
}

__global__ void __launch_bounds__(kCLWarps * 32)
closs_targets_kernel(const CLParams P, int l) {
  extern __shared__ float sm[];                        // per warp: px[K], py[K], sel[K], 4 result slots
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x * kCLWarps + wib;
  if (t >= P.n) return;
  const CLLevel& lv = P.lv[l];
  const int K = P.K, H = lv.H, W = lv.W, C = P.nc + 3 * K;
  float* px = sm + wib * (3 * K + 4); float* py = px + K; float* sel = py + K; float* res = sel + K;
  const float* row = P.targets + static_cast<size_t>(t) * (6 + 3 * K);
  const TargetGeom g = target_geom(row, H, W);
  const size_t cell = (static_cast<size_t>(g.b) * H + g.gy) * W + g.gx;
  const size_t plane = static_cast<size_t>(H) * W;
  const float* pb = lv.pred + static_cast<size_t>(g.b) * C * plane + static_cast<size_t>(g.gy) * W + g.gx;
  if (lane == 0 && atomicExch(&lv.cellmask[cell], 1) == 0) atomicAdd(P.num_pos, 1);
  const float area = fmaxf(g.bw * g.bh, 1.f);
  float s_wing = 0.f, s_l1 = 0.f, s_m = 0.f, s_sm = 0.f, s_bce = 0.f;
  for (int k = lane; k < K; k += 32) {
    const float x = pb[(P.nc + 3 * k) * plane], y = pb[(P.nc + 3 * k + 1) * plane], cf = pb[(P.nc + 3 * k + 2) * plane];
    const float tx = row[6 + 3 * k] * W - static_cast<float>(g.gx), ty = row[7 + 3 * k] * H - static_cast<float>(g.gy);
    const float m = row[8 + 3 * k] > 0.f ? 1.f : 0.f;
    px[k] = x; py[k] = y; sel[k] = m;
    const float dx = x - tx, dy = y - ty;
    const float d = dx * dx + dy * dy;
    const float s2 = 2.f * P.sigmas[k];
    const float e = d / (s2 * s2 * area * 2.f + 1e-6f);
    const float wing = e < 2.f ? 10.f * logf(1.f + e / 10.f) : e - 2.f + 10.f * logf(1.f + 2.f / 10.f);
    s_wing += wing * m; s_l1 += (fabsf(dx) + fabsf(dy)) * m; s_m += m;
    s_bce += fmaxf(cf, 0.f) - cf * m + log1pf(expf(-fabsf(cf)));
  }
  __syncwarp();
  for (int k = lane; k < K; k += 32)
    if (k >= 1) { const float ax = px[k] - px[k - 1], ay = py[k] - py[k - 1]; s_sm += ax * ax + ay * ay; }
  s_wing = wsum(s_wing); s_l1 = wsum(s_l1); s_m = wsum(s_m); s_sm = wsum(s_sm); s_bce = wsum(s_bce);
  // ---- robust box from the predicted keypoints (loss.py:277-315)
  const int nvis = static_cast<int>(s_m);
  if (nvis < 2) { for (int k = lane; k < K; k += 32) sel[k] = 1.f; }
  __syncwarp();
  const int cnt = nvis < 2 ? K : nvis;
  int lo_rank = 0, hi_rank = cnt - 1;
  if (cnt > 4) { int trim = static_cast<int>(cnt * 0.1); if (trim < 1) trim = 1; lo_rank = trim; hi_rank = cnt - trim - 1; }
  for (int k = lane; k < K; k += 32) {
    if (sel[k] == 0.f) continue;
    int rx = 0, ry = 0;
    for (int j = 0; j < K; ++j) {
      if (sel[j] == 0.f) continue;
      rx += (px[j] < px[k] || (px[j] == px[k] && j < k)) ? 1 : 0;
      ry += (py[j] < py[k] || (py[j] == py[k] && j < k)) ? 1 : 0;
    }
    if (rx == lo_rank) res[0] = px[k];
    if (rx == hi_rank) res[1] = px[k];
    if (ry == lo_rank) res[2] = py[k];
    if (ry == hi_rank) res[3] = py[k];
  }
  __syncwarp();
  if (lane == 0) {
    const float x0 = res[0], x1 = res[1], y0 = res[2], y1 = res[3];
    const float w = fmaxf(x1 - x0, 1.f), h = fmaxf(y1 - y0, 1.f);
    float iou = ciou_xywh((x0 + x1) * 0.5f, (y0 + y1) * 0.5f, w, h, g.bx - static_cast<float>(g.gx), g.by - static_cast<float>(g.gy), g.bw, g.bh);
    iou = fminf(fmaxf(iou, 0.f), 1.f);
    atomicMax(&lv.tmap[cell * P.nc + g.c], (static_cast<unsigned long long>(t + 1) << 32) | __float_as_uint(iou));
    float* o = lv.partials + static_cast<size_t>(t) * kCLPart;
    o[0] = s_wing; o[1] = s_l1; o[2] = s_m; o[3] = s_sm; o[4] = s_bce;
  }
}

// level_sums[l] = fixed-order sums of the per-target partials (one CTA, one warp per partial column)
__global__ void __launch_bounds__(kCLMaxLevels * kCLPart * 32)
closs_level_sums_kernel(const CLParams P) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (w >= P.L * kCLPart) return;
  const int l = w / kCLPart, j = w % kCLPart;
  double a = 0.0;
  for (int t = lane; t < P.n; t += 32) a += static_cast<double>(P.lv[l].partials[static_cast<size_t>(t) * kCLPart + j]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
  if (lane == 0) P.level_sums[w] = static_cast<float>(a);
}

__global__ void __launch_bounds__(kClsThreads)
closs_cls_kernel(const CLParams P, int l) {
  const CLLevel& lv = P.lv[l];
  const int H = lv.H, W = lv.W, C = P.nc + 3 * P.K;
  const long long plane = static_cast<long long>(H) * W, total = static_cast<long long>(P.B) * P.nc * plane;
  const float gs = lv.grad ? P.grad_scale / static_cast<float>(max(*P.num_pos, 1)) : 0.f;
  float acc = 0.f;
  for (long long i = static_cast<long long>(blockIdx.x) * kClsThreads + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * kClsThreads) {
    const long long hw = i % plane; const int c = static_cast<int>((i / plane) % P.nc); const long long b = i / (plane * P.nc);
    const size_t off = (static_cast<size_t>(b) * C + c) * plane + hw;
    const float p = __ldg(lv.pred + off);
    const unsigned long long tv = lv.tmap[(static_cast<size_t>(b) * plane + hw) * P.nc + c];
    const float t = __uint_as_float(static_cast<unsigned>(tv & 0xFFFFFFFFu));       // 0 bits = 0.0f background
    const float s = sigm(p);
    if (t > 0.f) {
      acc += (fmaxf(p, 0.f) - p * t + log1pf(expf(-fabsf(p)))) * t;
      if (lv.grad) lv.grad[off] = gs * t * (s - t);
    } else {
      const float sp = softplus_f(p);
      acc += sp * 0.75f * s * s;
      if (lv.grad) lv.grad[off] = gs * 0.75f * s * s * (s + 2.f * sp * (1.f - s));
    }
  }
  __shared__ float red[kClsThreads / 32];
  acc = wsum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f;
    for (int i = 0; i < kClsThreads / 32; ++i) a += red[i];
    lv.cls_partials[blockIdx.x] = a;
  }
}

__global__ void __launch_bounds__(kCLWarps * 32)
closs_kptgrad_kernel(const CLParams P, int l) {
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x * kCLWarps + wib;
  if (t >= P.n) return;
  const CLLevel& lv = P.lv[l];
  const int K = P.K, H = lv.H, W = lv.W, C = P.nc + 3 * K;
  const float* row = P.targets + static_cast<size_t>(t) * (6 + 3 * K);
  const TargetGeom g = target_geom(row, H, W);
  const size_t plane = static_cast<size_t>(H) * W;
  const size_t base = static_cast<size_t>(g.b) * C * plane + static_cast<size_t>(g.gy) * W + g.gx;
  const float* pb = lv.pred + base; float* gb = lv.grad + base;
  const float inv_pos = P.grad_scale / static_cast<float>(max(*P.num_pos, 1));
  const float V = fmaxf(P.level_sums[l * kCLPart + 2], 1.f);
  const float w_kpt = 10.f * inv_pos, w_vis = 5.f * inv_pos;
  const float area = fmaxf(g.bw * g.bh, 1.f);
  for (int k = lane; k < K; k += 32) {
    const float x = pb[(P.nc + 3 * k) * plane], y = pb[(P.nc + 3 * k + 1) * plane], cf = pb[(P.nc + 3 * k + 2) * plane];
    const float tx = row[6 + 3 * k] * W - static_cast<float>(g.gx), ty = row[7 + 3 * k] * H - static_cast<float>(g.gy);
    const float m = row[8 + 3 * k] > 0.f ? 1.f : 0.f;
    const float dx = x - tx, dy = y - ty;
    const float s2 = 2.f * P.sigmas[k];
    const float den = s2 * s2 * area * 2.f + 1e-6f;
    const float e = (dx * dx + dy * dy) / den;
    const float dwing = e < 2.f ? 1.f / (1.f + e / 10.f) : 1.f;
    float gx = m * (dwing * 2.f * dx / den + 0.5f * (dx > 0.f ? 1.f : (dx < 0.f ? -1.f : 0.f))) / V;
    float gy = m * (dwing * 2.f * dy / den + 0.5f * (dy > 0.f ? 1.f : (dy < 0.f ? -1.f : 0.f))) / V;
    float sx = 0.f, sy = 0.f;                            // smoothness: d/dp_k of sum (p_k - p_{k-1})^2 / n
    if (k >= 1)    { sx += x - pb[(P.nc + 3 * (k - 1)) * plane]; sy += y - pb[(P.nc + 3 * (k - 1) + 1) * plane]; }
    if (k < K - 1) { sx -= pb[(P.nc + 3 * (k + 1)) * plane] - x; sy -= pb[(P.nc + 3 * (k + 1) + 1) * plane] - y; }
    gx += 0.1f * 2.f * sx / static_cast<float>(P.n);
    gy += 0.1f * 2.f * sy / static_cast<float>(P.n);
    atomicAdd(&gb[(P.nc + 3 * k) * plane], w_kpt * gx);
    atomicAdd(&gb[(P.nc + 3 * k + 1) * plane], w_kpt * gy);
    atomicAdd(&gb[(P.nc + 3 * k + 2) * plane], w_vis * (sigm(cf) - m));
  }
}

__global__ void closs_final_kernel(const CLParams P) {
  if (threadIdx.x != 0) return;
  const float npos = static_cast<float>(max(*P.num_pos, 1));
  float cls = 0.f, kpt = 0.f, vis = 0.f;
  for (int l = 0; l < P.L; ++l) {
    double a = 0.0;
    for (int i = 0; i < P.lv[l].cls_grid; ++i) a += static_cast<double>(P.lv[l].cls_partials[i]);
    cls += static_cast<float>(a);
    if (P.n > 0) {
      const float* s = P.level_sums + l * kCLPart;
      const float V = fmaxf(s[2], 1.f);
      const float sm = P.K > 1 ? s[3] / static_cast<float>(P.n) : 0.f;
      kpt += s[0] / V + 0.5f * (s[1] / V) + 0.1f * sm;
      vis += s[4];
    }
  }
  cls = cls / npos * 1.f; kpt = kpt / npos * 10.f; vis = vis / npos * 5.f;
  P.out[0] = cls + kpt + vis; P.out[1] = cls; P.out[2] = kpt; P.out[3] = vis;
}

static inline size_t align256(size_t x) { return (x + 255) & ~size_t(255); }
static inline int cls_grid_for(long long total) {
  const long long want = (total + kClsThreads - 1) / kClsThreads, cap = (long long)num_sms() * 8;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace hpe
using namespace hpe;

extern "C" size_t hpe_compute_loss_workspace_bytes(int n_levels, int B, const int* Hs, const int* Ws, int nc, int n_targets) {
  if (n_levels < 1 || n_levels > kCLMaxLevels || B <= 0 || nc <= 0 || n_targets < 0 || !Hs || !Ws) return 0;
  size_t bytes = 256 + align256(sizeof(float) * kCLMaxLevels * kCLPart);
  for (int l = 0; l < n_levels; ++l) {
    const size_t cells = (size_t)B * Hs[l] * Ws[l];
    bytes += align256(cells * nc * 8) + align256(cells * 4) + align256((size_t)(n_targets > 0 ? n_targets : 1) * kCLPart * 4) +
             align256((size_t)num_sms() * 8 * 4);
  }
  return bytes;
}

extern "C" int hpe_compute_loss(const float* const* preds, const int* Hs, const int* Ws, int n_levels, int B, int nc, int nkpts,
                                const float* targets, int n_targets, const float* sigmas, void* workspace, size_t workspace_bytes,
                                float* loss_out, float* const* grads, float grad_scale, void* stream) {
  HPE_CHECK_ARG(preds && Hs && Ws && sigmas && workspace && loss_out, "hpe_compute_loss: null pointer");
  HPE_CHECK_ARG(n_levels >= 1 && n_levels <= kCLMaxLevels, "hpe_compute_loss: 1..%d levels", kCLMaxLevels);
  HPE_CHECK_ARG(B > 0 && nc > 0 && nkpts > 0 && n_targets >= 0, "hpe_compute_loss: bad sizes");
  HPE_CHECK_ARG(n_targets == 0 || targets, "hpe_compute_loss: null targets");
  const size_t need = hpe_compute_loss_workspace_bytes(n_levels, B, Hs, Ws, nc, n_targets);
  HPE_CHECK_ARG(workspace_bytes >= need, "hpe_compute_loss: workspace %zu < %zu bytes", workspace_bytes, need);
  cudaStream_t st = (cudaStream_t)stream;
  CLParams P = {};
  P.L = n_levels; P.B = B; P.nc = nc; P.K = nkpts; P.n = n_targets; P.targets = targets; P.sigmas = sigmas;
  P.out = loss_out; P.grad_scale = grad_scale;
  char* w = (char*)workspace;
  P.num_pos = (int*)w; w += 256;
  P.level_sums = (float*)w; w += align256(sizeof(float) * kCLMaxLevels * kCLPart);
  const int C = nc + 3 * nkpts;
  for (int l = 0; l < n_levels; ++l) {
    HPE_CHECK_ARG(preds[l] && Hs[l] > 0 && Ws[l] > 0, "hpe_compute_loss: bad level %d", l);
    CLLevel& lv = P.lv[l];
    const size_t cells = (size_t)B * Hs[l] * Ws[l];
    lv.pred = preds[l]; lv.grad = grads ? grads[l] : nullptr; lv.H = Hs[l]; lv.W = Ws[l];
    lv.tmap = (unsigned long long*)w; w += align256(cells * nc * 8);
    lv.cellmask = (int*)w; w += align256(cells * 4);
    lv.partials = (float*)w; w += align256((size_t)(n_targets > 0 ? n_targets : 1) * kCLPart * 4);
    lv.cls_partials = (float*)w; w += align256((size_t)num_sms() * 8 * 4);
    lv.cls_grid = cls_grid_for((long long)cells * nc);
    if (grads) HPE_CHECK_ARG(grads[l], "hpe_compute_loss: null grad for level %d", l);
  }
  // tmap + cellmask + num_pos + level_sums are zero-initialised every call (the reference allocates zero targets, loss.py:364-365)
  HPE_CUDA_OK(cudaMemsetAsync(workspace, 0, (size_t)(w - (char*)workspace), st));
  if (grads)
    for (int l = 0; l < n_levels; ++l)
      HPE_CUDA_OK(cudaMemsetAsync(grads[l], 0, (size_t)B * C * Hs[l] * Ws[l] * 4, st));
  if (n_targets > 0) {
    const int grid = (n_targets + kCLWarps - 1) / kCLWarps;
    const size_t smem = (size_t)kCLWarps * (3 * nkpts + 4) * 4;
    if (smem > 48 * 1024) return fail(HPE_EUNSUPPORTED, "hpe_compute_loss: %d keypoints exceed the per-warp staging budget", nkpts);
    for (int l = 0; l < n_levels; ++l) {
      closs_targets_kernel<<<grid, kCLWarps * 32, smem, st>>>(P, l);
      int rc = launch_status("closs_targets_kernel"); if (rc) return rc;
    }
    closs_level_sums_kernel<<<1, kCLMaxLevels * kCLPart * 32, 0, st>>>(P);
    int rc = launch_status("closs_level_sums_kernel"); if (rc) return rc;
  }
  for (int l = 0; l < n_levels; ++l) {
    closs_cls_kernel<<<P.lv[l].cls_grid, kClsThreads, 0, st>>>(P, l);
    int rc = launch_status("closs_cls_kernel"); if (rc) return rc;
  }
  if (grads && n_targets > 0) {
    const int grid = (n_targets + kCLWarps - 1) / kCLWarps;
    for (int l = 0; l < n_levels; ++l) {
      closs_kptgrad_kernel<<<grid, kCLWarps * 32, 0, st>>>(P, l);
      int rc = launch_status("closs_kptgrad_kernel"); if (rc) return rc;
    }
  }
  closs_final_kernel<<<1, 32, 0, st>>>(P);
  return launch_status("closs_final_kernel");
}
