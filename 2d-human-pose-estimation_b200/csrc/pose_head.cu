// Device path of the reference's OWN post-processing for its pose head:
//   * pose_decode_kernel  = PoseHead._inference_pose + kpts_decode + make_anchors
//                           (/root/reference/HPE/models/nn/modules/head.py:185-211, HPE/models/utils.py:15-27):
//                           level concat + anchor add + stride scale + sigmoid in ONE coalesced pass;
//   * pose_nms_kernel     = non_max_suppression_pose (/root/reference/HPE/test.py:69-185) incl. the greedy
//                           torchvision.ops.nms it calls: confidence filter, first-max label, keypoint min/max box,
//                           class offset 7680, stable descending sort, greedy suppression, max_det -- one CTA per
//                           image, no host sync, no Python loop.
// Arithmetic is written with explicit round-to-nearest intrinsics (no FMA contraction) so box coordinates,
// areas and IoU decisions are bit-identical to the fp32 PyTorch/torchvision CPU path.
#include "hpe_common.cuh"

namespace hpe {

constexpr int kMaxLevels = 4;

struct DecodeLevels {
  const void* ptr[kMaxLevels];
  int H[kMaxLevels], W[kMaxLevels], a0[kMaxLevels];     // a0 = first anchor index of the level
  float stride[kMaxLevels];
  int n;
};

__device__ __forceinline__ float sigmoid_f(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
// accurate variant used where the value feeds a threshold decision downstream (class scores)
__device__ __forceinline__ float sigmoid_acc(float x) { return 1.f / (1.f + expf(-x)); }

template <bool BF16>
__global__ void __launch_bounds__(256)
pose_decode_kernel(const DecodeLevels lv, int B, int C, int A, int nc, int ndim, float* __restrict__ out) {
  const long long total = static_cast<long long>(B) * C * A;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int a = static_cast<int>(i % A);
    const int c = static_cast<int>((i / A) % C);
    const int b = static_cast<int>(i / (static_cast<long long>(A) * C));
    int l = 0;
#pragma unroll
    for (int j = 1; j < kMaxLevels; ++j) if (j < lv.n && a >= lv.a0[j]) l = j;
    const int la = a - lv.a0[l], hw = lv.H[l] * lv.W[l];
    const size_t off = (static_cast<size_t>(b) * C + c) * hw + la;
    float v;
    if (BF16) v = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(lv.ptr[l])[off]);
    else      v = __ldg(reinterpret_cast<const float*>(lv.ptr[l]) + off);
    float r;
    if (c < nc) {
      r = sigmoid_acc(v);
    } else {
      const int d = (c - nc) % ndim;
      if (d == 2) r = sigmoid_acc(v);
      else {
        const float anchor = static_cast<float>(d == 0 ? la % lv.W[l] : la / lv.W[l]) + 0.5f;
        r = __fmul_rn(__fadd_rn(__fsub_rn(v, 0.5f), anchor), lv.stride[l]);      // (raw - 0.5 + anchor) * stride
      }
    }
    out[i] = r;
  }
}

// ---- NMS -------------------------------------------------------------------------------------------
// scratch per image: A records of {x1, y1, x2, y2, score, label}
constexpr int kNmsRec = 6;
constexpr int kNmsThreads = 256;

__global__ void __launch_bounds__(kNmsThreads)
pose_nms_kernel(const float* __restrict__ pred, int C, int A, int nc, int nkpts, int kdim, float conf_thres,
                double iou_thres, int max_det, int n_pow2, float* __restrict__ scratch,
                float* __restrict__ out_boxes, float* __restrict__ out_scores, int* __restrict__ out_labels,
                float* __restrict__ out_kpts, int* __restrict__ out_count) {
  extern __shared__ unsigned long long keys[];                 // n_pow2 sort keys, then the suppression bitmask
  unsigned int* supp = reinterpret_cast<unsigned int*>(keys + n_pow2);
  __shared__ int s_n, s_next;
  const int b = blockIdx.x, tid = threadIdx.x;
  const float* p = pred + static_cast<size_t>(b) * C * A;
  float* rec = scratch + static_cast<size_t>(b) * A * kNmsRec;
  if (tid == 0) s_n = 0;
  for (int i = tid; i < n_pow2; i += kNmsThreads) keys[i] = ~0ull;
  __syncthreads();
  // 1. confidence filter + first-max label (test.py:100-137); key = (descending score, ascending anchor)
  for (int a = tid; a < A; a += kNmsThreads) {
    float best = p[a]; int lab = 0;
    for (int c = 1; c < nc; ++c) { const float v = p[static_cast<size_t>(c) * A + a]; if (v > best) { best = v; lab = c; } }
    if (best > conf_thres) {
      const int slot = atomicAdd(&s_n, 1);
      keys[slot] = (static_cast<unsigned long long>(0xFFFFFFFFu - __float_as_uint(best)) << 32) | static_cast<unsigned>(a);
      rec[static_cast<size_t>(a) * kNmsRec + 4] = best;
      rec[static_cast<size_t>(a) * kNmsRec + 5] = __int_as_float(lab);
    }
  }
  __syncthreads();
  const int n = s_n;
  if (n == 0) { if (tid == 0) out_count[b] = 0; return; }
  // 2. bitonic sort of the candidate keys (ascending key = descending score, ties by anchor index = stable)
  int m = 1; while (m < n) m <<= 1;
  for (int k = 2; k <= m; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < m; i += kNmsThreads) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long x = keys[i], y = keys[ixj];
          const bool up = (i & k) == 0;
          if ((x > y) == up) { keys[i] = y; keys[ixj] = x; }
        }
      }
      __syncthreads();
    }
  // 3. keypoint min/max box of every candidate (test.py:155-163)
  for (int i = tid; i < n; i += kNmsThreads) {
    const int a = static_cast<int>(keys[i] & 0xFFFFFFFFu);
    float x1 = INFINITY, y1 = INFINITY, x2 = -INFINITY, y2 = -INFINITY;
    for (int k = 0; k < nkpts; ++k) {
      const float x = p[static_cast<size_t>(nc + k * kdim) * A + a], y = p[static_cast<size_t>(nc + k * kdim + 1) * A + a];
      x1 = fminf(x1, x); x2 = fmaxf(x2, x); y1 = fminf(y1, y); y2 = fmaxf(y2, y);
    }
    float* r = rec + static_cast<size_t>(a) * kNmsRec;
    r[0] = x1; r[1] = y1; r[2] = x2; r[3] = y2;
  }
  for (int i = tid; i < (n + 31) / 32; i += kNmsThreads) supp[i] = 0u;
  __syncthreads();
  // 4. greedy suppression in sorted order (torchvision nms_kernel_impl); one iteration per KEPT detection
  int pos = 0, kept = 0;                        // every thread runs the same loop: `kept` is replicated
  while (true) {
    if (tid == 0) s_next = n;
    __syncthreads();
    for (int w = (pos >> 5) + tid; w < (n + 31) / 32; w += kNmsThreads) {
      unsigned int free_bits = ~supp[w];
      if (w == (pos >> 5)) free_bits &= ~0u << (pos & 31);
      if (w == (n >> 5)) free_bits &= (1u << (n & 31)) - 1u;
      if (free_bits) { atomicMin(&s_next, w * 32 + __ffs(free_bits) - 1); break; }   // later words of this thread are larger
    }
    __syncthreads();
    const int cur = s_next;
    if (cur >= n) break;
    const int ai = static_cast<int>(keys[cur] & 0xFFFFFFFFu);
    const float* ri = rec + static_cast<size_t>(ai) * kNmsRec;
    const float ci = __fmul_rn(static_cast<float>(__float_as_int(ri[5])), 7680.f);
    const float ix1 = __fadd_rn(ri[0], ci), iy1 = __fadd_rn(ri[1], ci), ix2 = __fadd_rn(ri[2], ci), iy2 = __fadd_rn(ri[3], ci);
    const float iarea = __fmul_rn(__fsub_rn(ix2, ix1), __fsub_rn(iy2, iy1));
    if (tid == 0) {
      float* ob = out_boxes + (static_cast<size_t>(b) * max_det + kept) * 4;
      ob[0] = ri[0]; ob[1] = ri[1]; ob[2] = ri[2]; ob[3] = ri[3];
      out_scores[static_cast<size_t>(b) * max_det + kept] = ri[4];
      out_labels[static_cast<size_t>(b) * max_det + kept] = __float_as_int(ri[5]);
    }
    for (int k = tid; k < nkpts * 3; k += kNmsThreads) {
      const int kp = k / 3, d = k - kp * 3;
      out_kpts[(static_cast<size_t>(b) * max_det + kept) * nkpts * 3 + k] =
          d < kdim ? p[static_cast<size_t>(nc + kp * kdim + d) * A + ai] : 1.f;           // kdim == 2: conf = 1 (test.py:146-150)
    }
    if (++kept >= max_det) break;
    for (int j = cur + 1 + tid; j < n; j += kNmsThreads) {
      if (supp[j >> 5] & (1u << (j & 31))) continue;
      const float* rj = rec + static_cast<size_t>(keys[j] & 0xFFFFFFFFu) * kNmsRec;
      const float cj = __fmul_rn(static_cast<float>(__float_as_int(rj[5])), 7680.f);
      const float jx1 = __fadd_rn(rj[0], cj), jy1 = __fadd_rn(rj[1], cj), jx2 = __fadd_rn(rj[2], cj), jy2 = __fadd_rn(rj[3], cj);
      const float jarea = __fmul_rn(__fsub_rn(jx2, jx1), __fsub_rn(jy2, jy1));
      const float w = fmaxf(0.f, __fsub_rn(fminf(ix2, jx2), fmaxf(ix1, jx1)));
      const float h = fmaxf(0.f, __fsub_rn(fminf(iy2, jy2), fmaxf(iy1, jy1)));
      const float inter = __fmul_rn(w, h);
      const float ovr = __fdiv_rn(inter, __fsub_rn(__fadd_rn(iarea, jarea), inter));
      if (static_cast<double>(ovr) > iou_thres) atomicOr(&supp[j >> 5], 1u << (j & 31));
    }
    pos = cur + 1;
    __syncthreads();
  }
  if (tid == 0) out_count[b] = kept;
}

}  // namespace hpe
using namespace hpe;

extern "C" int hpe_pose_decode(const void* const* levels, const int* Hs, const int* Ws, const float* strides, int n_levels,
                               int dtype, int B, int nc, int nkpts, int ndim, float* out, void* stream) {
  HPE_CHECK_ARG(levels && Hs && Ws && strides && out, "hpe_pose_decode: null pointer");
  HPE_CHECK_ARG(n_levels >= 1 && n_levels <= kMaxLevels, "hpe_pose_decode: 1..%d levels (got %d)", kMaxLevels, n_levels);
  HPE_CHECK_ARG(B > 0 && nc > 0 && nkpts > 0 && (ndim == 2 || ndim == 3), "hpe_pose_decode: bad B/nc/nkpts/ndim");
  HPE_CHECK_ARG(dtype == HPE_F32 || dtype == HPE_BF16, "hpe_pose_decode: bad dtype %d", dtype);
  DecodeLevels lv = {};
  int A = 0;
  for (int l = 0; l < n_levels; ++l) {
    HPE_CHECK_ARG(levels[l] && Hs[l] > 0 && Ws[l] > 0, "hpe_pose_decode: bad level %d", l);
    lv.ptr[l] = levels[l]; lv.H[l] = Hs[l]; lv.W[l] = Ws[l]; lv.stride[l] = strides[l]; lv.a0[l] = A;
    A += Hs[l] * Ws[l];
  }
  lv.n = n_levels;
  const int C = nc + nkpts * ndim;
  const long long total = (long long)B * C * A;
  const long long want = (total + 255) / 256, cap = (long long)num_sms() * 16;
  const int grid = (int)(want < cap ? want : cap);
  if (dtype == HPE_BF16) pose_decode_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(lv, B, C, A, nc, ndim, out);
  else                   pose_decode_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(lv, B, C, A, nc, ndim, out);
  return launch_status("pose_decode_kernel");
}

extern "C" size_t hpe_pose_nms_scratch_elems(int B, int A) {
  if (B <= 0 || A <= 0) return 0;
  return (size_t)B * A * kNmsRec;
}

extern "C" int hpe_pose_nms(const float* pred, int B, int C, int A, int nc, int nkpts, float conf_thres, double iou_thres,
                            int max_det, float* scratch, float* boxes, float* scores, int32_t* labels, float* keypoints,
                            int32_t* count, void* stream) {
  HPE_CHECK_ARG(pred && scratch && boxes && scores && labels && keypoints && count, "hpe_pose_nms: null pointer");
  HPE_CHECK_ARG(B > 0 && A > 0 && nc > 0 && nkpts > 0 && max_det > 0, "hpe_pose_nms: non-positive size");
  const int nk = C - nc;
  HPE_CHECK_ARG(nk > 0 && nk % nkpts == 0 && (nk / nkpts == 2 || nk / nkpts == 3), "hpe_pose_nms: C=%d is not nc + nkpts*{2,3}", C);
  int n_pow2 = 1; while (n_pow2 < A) n_pow2 <<= 1;
  const size_t smem = (size_t)n_pow2 * 8 + (size_t)(A + 31) / 32 * 4 + 16;
  if (smem > 200 * 1024) return fail(HPE_EUNSUPPORTED, "hpe_pose_nms: %d anchors need %zu B of shared memory (max 16384 anchors)", A, smem);
  // the opt-in is a per-DEVICE function attribute: set it on every call that needs it (a cached flag would skip a second GPU)
  if (smem > 48 * 1024)
    HPE_CUDA_OK(cudaFuncSetAttribute((const void*)pose_nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  pose_nms_kernel<<<B, kNmsThreads, smem, (cudaStream_t)stream>>>(pred, C, A, nc, nkpts, nk / nkpts, conf_thres, iou_thres, max_det,
                                                                 n_pow2, scratch, boxes, scores, labels, keypoints, count);
  return launch_status("pose_nms_kernel");
}
