// Training-mode kernels of the reference's real pose model beyond the ResNet path: the neck and head of DINOv3Pose
// (FeatureAdaptor + SPPF + PAN + PoseHead, /root/reference/HPE/models/pose.py:47-65 with the backbone frozen) are built from
// ConvBlock = Conv2d + BatchNorm2d(batch statistics) + SiLU (HPE/models/nn/modules/block.py:20-45), a depthwise 3x3 ConvBlock,
// MaxPool2d(5, 1, 2) (block.py:102-114) and bilinear 2x up-sampling (block.py:132-151).  The convolutions run on the tcgen05
// kernels (forward, dgrad, wgrad); this file holds the bandwidth-bound rest, for ANY channel count that is a multiple of 8
// (the neck's 192 / 384 / 576 / 768 / 1152 / 1536 are not the powers of two train_ops.cu is specialised for):
//   bnact_*            BatchNorm2d forward statistics / apply / backward with act in {none, ReLU, SiLU}
//   dwconv3_wgrad_*    weight gradient of a depthwise 3x3 convolution
//   maxpool5_idx / bwd  MaxPool2d(5, 1, 2) with the recorded arg-max (first-max tie rule) and its backward
//   upsample2x_bwd     backward of F.interpolate(mode='bilinear', align_corners=False)
// Every reduction is two-stage with a fixed order (no atomics): gradients are bit-reproducible run to run.
#include "hpe_common.cuh"
#include <math.h>

namespace hpe {

__device__ __forceinline__ void nt_unpack8(const uint4& v, float* f) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 nt_pack8(const float* f) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}
// act: 0 none, 1 ReLU, 2 SiLU
__device__ __forceinline__ float nt_act(float u, int act) {
  if (act == 1) return fmaxf(u, 0.f);
  if (act == 2) return u / (1.f + __expf(-u));
  return u;
}
__device__ __forceinline__ float nt_dact(float u, int act) {
  if (act == 1) return u > 0.f ? 1.f : 0.f;
  // SiLU'(u) = s (1 + u (1 - s)), s = sigmoid(u); the reciprocal is the 1-ulp MUFU approximation: the IEEE division cost ~8 issue
  // slots per element and made the backward reduction compute-bound at the 1-2 blocks per SM these small tensors give
  if (act == 2) { const float s = __fdividef(1.f, 1.f + __expf(-u)); return s * fmaf(u, 1.f - s, 1.f); }
  return 1.f;
}

constexpr int kNtSlices = 256;           // max row slices of the two-stage reductions (partials [slice][2][C])

// Stage 1 of a per-channel reduction over the M rows of an [M][C] bf16 tensor.  Block = 8 rows x 32 channel vectors (a warp
// reads 512 contiguous bytes of one row); grid = (ceil(Cv / 32), kNtSlices).
// MODE 0: s0 = sum z, s1 = sum z^2.   MODE 1: g = dy * act'(z*scale + shift); s0 = sum g, s1 = sum g * (z - mean) * rstd.
template <int MODE>
__global__ void __launch_bounds__(256)
bnact_reduce_kernel(const __nv_bfloat16* __restrict__ z, const __nv_bfloat16* __restrict__ dy, long long M, int C, int act,
                    const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ mean,
                    const float* __restrict__ rstd, float* __restrict__ part) {
  __shared__ float red[8][32][17];
  const int Cv = C >> 3, tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int vec = blockIdx.x * 32 + tx;
  float a0[8], a1[8], sc[8], sh[8], mu[8], rs[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) { a0[e] = 0.f; a1[e] = 0.f; sc[e] = 1.f; sh[e] = 0.f; mu[e] = 0.f; rs[e] = 1.f; }
  if (vec < Cv) {
    if (MODE == 1) {
#pragma unroll
      for (int e = 0; e < 8; ++e) { sc[e] = scale[vec * 8 + e]; sh[e] = shift[vec * 8 + e]; mu[e] = mean[vec * 8 + e]; rs[e] = rstd[vec * 8 + e]; }
    }
    const long long per = (M + gridDim.y - 1) / gridDim.y;
    const long long r0 = static_cast<long long>(blockIdx.y) * per, r1 = r0 + per < M ? r0 + per : M;
    // four rows per pass, every load issued before the first use (with one row in flight per thread and 1-2 blocks per SM the
    // kernel ran at the latency of a single load: 58 launches of the backward reduction took 1.33 ms at B = 64); the rows are
    // still added in the order r0 + ty, + 8, + 16, ...: same bits as before
    constexpr int U = 4;
    for (long long r = r0 + ty; r < r1; r += 8 * U) {
      uint4 vz[U], vd[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long rr = r + 8 * u;
        if (rr < r1) {
          vz[u] = __ldg(reinterpret_cast<const uint4*>(z) + rr * Cv + vec);
          if (MODE == 1) vd[u] = __ldg(reinterpret_cast<const uint4*>(dy) + rr * Cv + vec);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (r + 8 * u >= r1) break;
        float fz[8];
        nt_unpack8(vz[u], fz);
        if (MODE == 0) {
#pragma unroll
          for (int e = 0; e < 8; ++e) { a0[e] += fz[e]; a1[e] = fmaf(fz[e], fz[e], a1[e]); }
        } else {
          float fd[8];
          nt_unpack8(vd[u], fd);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float g = fd[e] * nt_dact(fmaf(fz[e], sc[e], sh[e]), act);
            a0[e] += g; a1[e] = fmaf(g, (fz[e] - mu[e]) * rs[e], a1[e]);
          }
        }
      }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) { red[ty][tx][e] = a0[e]; red[ty][tx][8 + e] = a1[e]; }
  __syncthreads();
  // thread (tx, comp = ty*2 + {0,1}) adds the 8 row groups in order
  for (int comp = ty * 2; comp < ty * 2 + 2; ++comp) {
    if (vec < Cv) {
      float s = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r) s += red[r][tx][comp];
      part[(static_cast<size_t>(blockIdx.y) * 2 + (comp >> 3)) * C + vec * 8 + (comp & 7)] = s;
    }
  }
}

// Stage 2.  MODE 0: batch statistics -> scale = gamma*rstd, shift = beta - mean*scale, saved mean / rstd, running statistics
// (momentum, unbiased variance) as torch.nn.BatchNorm2d does in training mode.
// MODE 1: dbeta, dgamma and the coefficients of dz = A*g + Bz*z + C0.
template <int MODE>
__global__ void __launch_bounds__(256)
bnact_finish_kernel(const float* __restrict__ part, int nslices, long long M, int C, const float* __restrict__ gamma,
                    const float* __restrict__ beta, float eps, float momentum, float* running_mean, float* running_var,
                    float* __restrict__ o0, float* __restrict__ o1, float* __restrict__ o2, float* __restrict__ o3,
                    float* __restrict__ coef, int frozen) {
  // a warp per channel: lane l adds slices l, l + 32, ... in order, then a fixed shuffle tree (deterministic; one thread walking
  // up to 256 dependent loads per channel made these 116 tiny launches ~9 us each)
  const int lane = threadIdx.x & 31, c = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (c >= C) return;
  float s0 = 0.f, s1 = 0.f;
  for (int s = lane; s < nslices; s += 32) { s0 += part[(static_cast<size_t>(s) * 2) * C + c]; s1 += part[(static_cast<size_t>(s) * 2 + 1) * C + c]; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { s0 += __shfl_xor_sync(0xffffffffu, s0, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o); }
  if (lane != 0) return;
  const float inv = 1.f / static_cast<float>(M);
  if (MODE == 1) {
    o0[c] = s0; o1[c] = s1;                                    // dbeta, dgamma
    const float rs = o3[c], mu = o2[c], A = gamma[c] * rs;     // saved rstd, mean
    // frozen statistics (the module is in eval mode: running mean / variance are constants): dz = gamma * rstd * g
    coef[c] = A; coef[C + c] = frozen ? 0.f : -A * rs * s1 * inv; coef[2 * C + c] = frozen ? 0.f : A * (mu * rs * s1 - s0) * inv;
    return;
  }
  const float mu = s0 * inv, var = fmaxf(s1 * inv - mu * mu, 0.f), rs = rsqrtf(var + eps), sc = gamma[c] * rs;
  o0[c] = sc; o1[c] = beta[c] - mu * sc; o2[c] = mu; o3[c] = rs;
  if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mu;
  if (running_var) {
    const float unbiased = M > 1 ? var * static_cast<float>(M) / static_cast<float>(M - 1) : var;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
}

// y = act(z * scale + shift)
__global__ void __launch_bounds__(256)
bnact_apply_kernel(const __nv_bfloat16* __restrict__ z, const float* __restrict__ scale, const float* __restrict__ shift,
                   int act, long long nvec, int Cv, __nv_bfloat16* __restrict__ y) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(i % Cv) * 8;
    float f[8];
    nt_unpack8(__ldg(reinterpret_cast<const uint4*>(z) + i), f);
    const float4 s0 = __ldg(reinterpret_cast<const float4*>(scale + c0)), s1 = __ldg(reinterpret_cast<const float4*>(scale + c0 + 4));
    const float4 h0 = __ldg(reinterpret_cast<const float4*>(shift + c0)), h1 = __ldg(reinterpret_cast<const float4*>(shift + c0 + 4));
    const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w}, sh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = nt_act(fmaf(f[e], sc[e], sh[e]), act);
    reinterpret_cast<uint4*>(y)[i] = nt_pack8(f);
  }
}

// dz = A*g + Bz*z + C0 with g = dy * act'(z*scale + shift)
__global__ void __launch_bounds__(256)
bnact_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ z, const float* __restrict__ coef,
                       const float* __restrict__ scale, const float* __restrict__ shift, int act, int C, long long nvec, int Cv,
                       __nv_bfloat16* __restrict__ dz) {
  // two vectors per pass (four 16-byte loads in flight per thread), per-channel coefficients as 16-byte loads (C % 8 == 0 and the
  // arrays are 16-byte aligned: torch allocations / scratch + a multiple of 4 floats); same arithmetic per element as before
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i0 = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i0 < nvec; i0 += 2 * stride) {
    uint4 vd[2], vz[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long i = i0 + u * stride;
      if (i < nvec) { vd[u] = __ldg(reinterpret_cast<const uint4*>(dy) + i); vz[u] = __ldg(reinterpret_cast<const uint4*>(z) + i); }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long i = i0 + u * stride;
      if (i >= nvec) break;
      const int c0 = static_cast<int>(i % Cv) * 8;
      float fd[8], fz[8], o[8], k[5][8];
      nt_unpack8(vd[u], fd);
      nt_unpack8(vz[u], fz);
      const float* src[5] = {scale + c0, shift + c0, coef + c0, coef + C + c0, coef + 2 * C + c0};
#pragma unroll
      for (int a = 0; a < 5; ++a) {
        const float4 lo = __ldg(reinterpret_cast<const float4*>(src[a])), hi = __ldg(reinterpret_cast<const float4*>(src[a]) + 1);
        k[a][0] = lo.x; k[a][1] = lo.y; k[a][2] = lo.z; k[a][3] = lo.w; k[a][4] = hi.x; k[a][5] = hi.y; k[a][6] = hi.z; k[a][7] = hi.w;
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float g = fd[e] * nt_dact(fmaf(fz[e], k[0][e], k[1][e]), act);
        o[e] = fmaf(k[2][e], g, fmaf(k[3][e], fz[e], k[4][e]));
      }
      reinterpret_cast<uint4*>(dz)[i] = nt_pack8(o);
    }
  }
}

// ---- depthwise 3x3 (pad 1, stride 1) weight gradient: dw[tap][c] = sum_{b,y,x} dz[b,y,x,c] * x[b,y+ky-1,x+kx-1,c] --------------
// stage 1: grid = (ceil(Cv / 32), kNtSlices) over the B*H rows of the image stack; block = 8 row groups x 32 channel vectors
__global__ void __launch_bounds__(256)
dwconv3_wgrad_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dz, int B, int H, int W, int C,
                     float* __restrict__ part /* [slice][9][C] */) {
  __shared__ float red[8][32][9];
  const int Cv = C >> 3, tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int vec = blockIdx.x * 32 + tx;
  float acc[9][8];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[t][e] = 0.f;
  const long long rows = static_cast<long long>(B) * H;
  const long long per = (rows + gridDim.y - 1) / gridDim.y;
  const long long r0 = static_cast<long long>(blockIdx.y) * per, r1 = r0 + per < rows ? r0 + per : rows;
  if (vec < Cv) {
    for (long long r = r0 + ty; r < r1; r += 8) {
      const int y = static_cast<int>(r % H);
      const long long b = r / H;
      for (int xo = 0; xo < W; ++xo) {
        float g[8];
        nt_unpack8(__ldg(reinterpret_cast<const uint4*>(dz) + ((b * H + y) * W + xo) * Cv + vec), g);
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
          const int iy = y + ky - 1;
          if (iy < 0 || iy >= H) continue;
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const int ix = xo + kx - 1;
            if (ix < 0 || ix >= W) continue;
            float v[8];
            nt_unpack8(__ldg(reinterpret_cast<const uint4*>(x) + ((b * H + iy) * W + ix) * Cv + vec), v);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[ky * 3 + kx][e] = fmaf(g[e], v[e], acc[ky * 3 + kx][e]);
          }
        }
      }
    }
  }
  // fixed-order sum of the 8 row groups, one channel at a time (8 passes keep the shared buffer small)
  for (int e = 0; e < 8; ++e) {
#pragma unroll
    for (int t = 0; t < 9; ++t) red[ty][tx][t] = acc[t][e];
    __syncthreads();
    for (int t = ty; t < 9; t += 8) {
      if (vec < Cv) {
        float s = 0.f;
#pragma unroll
        for (int r = 0; r < 8; ++r) s += red[r][tx][t];
        part[(static_cast<size_t>(blockIdx.y) * 9 + t) * C + vec * 8 + e] = s;
      }
    }
    __syncthreads();
  }
}
// stage 2: dw_torch[c][0][ky][kx] = sum over slices
__global__ void __launch_bounds__(256)
dwconv3_wgrad_finish_kernel(const float* __restrict__ part, int nslices, int C, float* __restrict__ dw) {
  const int i = blockIdx.x * 256 + threadIdx.x;          // (c, tap)
  if (i >= C * 9) return;
  const int c = i / 9, t = i % 9;
  float s = 0.f;
  for (int sl = 0; sl < nslices; ++sl) s += part[(static_cast<size_t>(sl) * 9 + t) * C + c];
  dw[i] = s;
}

__device__ __forceinline__ bool nt_bf16_gt(uint16_t a, uint16_t b) {
  return __uint_as_float(static_cast<uint32_t>(a) << 16) > __uint_as_float(static_cast<uint32_t>(b) << 16);
}

// ---- MaxPool2d(5, 1, 2) in training mode: the forward also records, per output element, which window position (wy * 5 + wx)
// holds the FIRST maximum (row-major scan of the window, strict >, as ATen's max_pool2d); the backward gathers
// dx[p] = sum of dy[q] over the <= 25 windows q around p whose recorded arg-max is p (fixed order, no atomics).
__global__ void __launch_bounds__(256)
maxpool5_idx_kernel(const __nv_bfloat16* __restrict__ x, int B, int H, int W, int C, __nv_bfloat16* __restrict__ y,
                    uint8_t* __restrict__ idx) {
  const int Cv = C / 8;
  const long long total = static_cast<long long>(B) * H * W * Cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(i % Cv);
    const int qx = static_cast<int>((i / Cv) % W);
    const int qy = static_cast<int>((i / (static_cast<long long>(Cv) * W)) % H);
    const long long b = i / (static_cast<long long>(Cv) * W * H);
    const uint4* xb = reinterpret_cast<const uint4*>(x) + b * H * W * Cv + cv;
    uint16_t best[8]; uint8_t bi[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { best[e] = 0xFF80; bi[e] = 255; }
#pragma unroll
    for (int wy = 0; wy < 5; ++wy) {
      const int iy = qy + wy - 2;
      if (iy < 0 || iy >= H) continue;
#pragma unroll
      for (int wx = 0; wx < 5; ++wx) {
        const int ix = qx + wx - 2;
        if (ix < 0 || ix >= W) continue;
        const uint4 v = __ldg(xb + (static_cast<long long>(iy) * W + ix) * Cv);
        const uint16_t* h = reinterpret_cast<const uint16_t*>(&v);
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (bi[e] == 255 || nt_bf16_gt(h[e], best[e])) { best[e] = h[e]; bi[e] = static_cast<uint8_t>(wy * 5 + wx); }
      }
    }
    uint4 o; uint16_t* oh = reinterpret_cast<uint16_t*>(&o);
    uint2 oi; uint8_t* ob = reinterpret_cast<uint8_t*>(&oi);
#pragma unroll
    for (int e = 0; e < 8; ++e) { oh[e] = best[e]; ob[e] = bi[e]; }
    reinterpret_cast<uint4*>(y)[i] = o;
    reinterpret_cast<uint2*>(idx)[i] = oi;
  }
}
__global__ void __launch_bounds__(256)
maxpool5_bwd_kernel(const uint8_t* __restrict__ idx, const __nv_bfloat16* __restrict__ dy, int B, int H, int W, int C,
                    __nv_bfloat16* __restrict__ dx) {
  const int Cv = C / 8;
  const long long total = static_cast<long long>(B) * H * W * Cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(i % Cv);
    const int px = static_cast<int>((i / Cv) % W);
    const int py = static_cast<int>((i / (static_cast<long long>(Cv) * W)) % H);
    const long long b = i / (static_cast<long long>(Cv) * W * H);
    const uint2* ib = reinterpret_cast<const uint2*>(idx) + b * H * W * Cv + cv;
    const uint4* gb = reinterpret_cast<const uint4*>(dy) + b * H * W * Cv + cv;
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
    for (int wy = 0; wy < 5; ++wy) {
      const int qy = py - (wy - 2);                      // window q whose position (wy, wx) is p
      if (qy < 0 || qy >= H) continue;
#pragma unroll
      for (int wx = 0; wx < 5; ++wx) {
        const int qx = px - (wx - 2);
        if (qx < 0 || qx >= W) continue;
        const long long o = (static_cast<long long>(qy) * W + qx) * Cv;
        const uint2 k = __ldg(ib + o);
        const uint8_t* kb = reinterpret_cast<const uint8_t*>(&k);
        float g[8];
        nt_unpack8(__ldg(gb + o), g);
#pragma unroll
        for (int e = 0; e < 8; ++e) if (kb[e] == wy * 5 + wx) acc[e] += g[e];
      }
    }
    reinterpret_cast<uint4*>(dx)[i] = nt_pack8(acc);
  }
}

// ---- bilinear up-sampling backward (align_corners=False): dx[iy,ix] = sum over the outputs that read it of weight * dy -------
__device__ __forceinline__ void nt_src(int o, int in_size, float scale, int& i0, int& i1, float& l1) {
  float s = (static_cast<float>(o) + 0.5f) * scale - 0.5f;       // ATen area_pixel_compute_source_index (cubic = false)
  if (s < 0.f) s = 0.f;
  i0 = static_cast<int>(s);
  i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
  l1 = s - static_cast<float>(i0);
}
__global__ void __launch_bounds__(256)
upsample_bilinear_bwd_kernel(const __nv_bfloat16* __restrict__ dy, int B, int Hi, int Wi, int C, int Ho, int Wo,
                             __nv_bfloat16* __restrict__ dx) {
  const int Cv = C / 8;
  const float sy = static_cast<float>(Hi) / Ho, sx = static_cast<float>(Wi) / Wo;
  const long long total = static_cast<long long>(B) * Hi * Wi * Cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(i % Cv);
    const int ix = static_cast<int>((i / Cv) % Wi);
    const int iy = static_cast<int>((i / (static_cast<long long>(Cv) * Wi)) % Hi);
    const long long b = i / (static_cast<long long>(Cv) * Wi * Hi);
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    // outputs whose source rows / columns can include (iy, ix): a window of the output grid around (iy / sy, ix / sx)
    const int oy_lo = max(0, static_cast<int>(floorf((iy - 1) / sy)) - 1), oy_hi = min(Ho - 1, static_cast<int>(ceilf((iy + 1) / sy)) + 1);
    const int ox_lo = max(0, static_cast<int>(floorf((ix - 1) / sx)) - 1), ox_hi = min(Wo - 1, static_cast<int>(ceilf((ix + 1) / sx)) + 1);
    for (int oy = oy_lo; oy <= oy_hi; ++oy) {
      int y0, y1; float ly;
      nt_src(oy, Hi, sy, y0, y1, ly);
      const float wy = (y0 == iy ? 1.f - ly : 0.f) + (y1 == iy ? ly : 0.f);
      if (wy == 0.f) continue;
      for (int ox = ox_lo; ox <= ox_hi; ++ox) {
        int x0, x1; float lx;
        nt_src(ox, Wi, sx, x0, x1, lx);
        const float wx = (x0 == ix ? 1.f - lx : 0.f) + (x1 == ix ? lx : 0.f);
        if (wx == 0.f) continue;
        float g[8];
        nt_unpack8(__ldg(reinterpret_cast<const uint4*>(dy) + ((b * Ho + oy) * Wo + ox) * Cv + cv), g);
        const float w = wy * wx;
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(w, g[e], acc[e]);
      }
    }
    reinterpret_cast<uint4*>(dx)[i] = nt_pack8(acc);
  }
}

static inline int nt_grid(long long n) {
  long long g = (n + 255) / 256;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  return static_cast<int>(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace hpe

using namespace hpe;

// row slices of a reduction over M rows: ~128 rows per block, at least 16 (small maps) and at most kNtSlices
static int nt_slices(long long M) {
  long long n = (M + 127) / 128;
  return (int)(n < 16 ? 16 : (n > kNtSlices ? kNtSlices : n));
}

static int nt_shape_ok(long long M, int C, const char* fn) {
  HPE_CHECK_ARG(M > 0 && C >= 8 && C % 8 == 0 && C <= 4096, "%s: C=%d must be a multiple of 8 in [8, 4096] and M > 0", fn, C);
  return HPE_OK;
}

extern "C" size_t hpe_bnact_scratch_elems(int C) { return C > 0 ? (size_t)kNtSlices * 2 * C + 3 * (size_t)C : 0; }

extern "C" int hpe_bnact_stats(const void* z, long long M, int C, const float* gamma, const float* beta, float eps, float momentum,
                               float* running_mean, float* running_var, float* scratch, float* scale, float* shift, float* mean,
                               float* rstd, void* stream) {
  int rc = nt_shape_ok(M, C, "hpe_bnact_stats");
  if (rc) return rc;
  HPE_CHECK_ARG(z && gamma && beta && scratch && scale && shift && mean && rstd, "hpe_bnact_stats: null pointer");
  HPE_CHECK_ARG(((uintptr_t)z & 15) == 0, "hpe_bnact_stats: misaligned pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const int ns = nt_slices(M);
  const dim3 grid((C / 8 + 31) / 32, ns);
  bnact_reduce_kernel<0><<<grid, 256, 0, st>>>((const __nv_bfloat16*)z, nullptr, M, C, 0, nullptr, nullptr, nullptr, nullptr, scratch);
  rc = launch_status("bnact_reduce_kernel<0>");
  if (rc) return rc;
  bnact_finish_kernel<0><<<(C + 7) / 8, 256, 0, st>>>(scratch, ns, M, C, gamma, beta, eps, momentum, running_mean, running_var,
                                                        scale, shift, mean, rstd, nullptr, 0);
  return launch_status("bnact_finish_kernel<0>");
}

extern "C" int hpe_bnact_apply(const void* z, long long M, int C, const float* scale, const float* shift, int act, void* y, void* stream) {
  int rc = nt_shape_ok(M, C, "hpe_bnact_apply");
  if (rc) return rc;
  HPE_CHECK_ARG(z && scale && shift && y && act >= 0 && act <= 2, "hpe_bnact_apply: bad argument");
  HPE_CHECK_ARG((((uintptr_t)z | (uintptr_t)y | (uintptr_t)scale | (uintptr_t)shift) & 15) == 0, "hpe_bnact_apply: misaligned pointer");
  const long long nvec = M * (C / 8);
  bnact_apply_kernel<<<nt_grid(nvec), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)z, scale, shift, act, nvec, C / 8, (__nv_bfloat16*)y);
  return launch_status("bnact_apply_kernel");
}

extern "C" int hpe_bnact_backward(const void* dy, const void* z, long long M, int C, const float* mean, const float* rstd,
                                  const float* gamma, const float* scale, const float* shift, int act, int frozen_stats, float* scratch,
                                  float* dbeta, float* dgamma, void* dz, void* stream) {
  int rc = nt_shape_ok(M, C, "hpe_bnact_backward");
  if (rc) return rc;
  HPE_CHECK_ARG(dy && z && mean && rstd && gamma && scale && shift && scratch && dbeta && dgamma && dz && act >= 0 && act <= 2,
                "hpe_bnact_backward: bad argument");
  HPE_CHECK_ARG((((uintptr_t)dy | (uintptr_t)z | (uintptr_t)dz | (uintptr_t)scale | (uintptr_t)shift | (uintptr_t)scratch) & 15) == 0,
                "hpe_bnact_backward: misaligned pointer (tensors, scale / shift and scratch must be 16-byte aligned)");
  cudaStream_t st = (cudaStream_t)stream;
  float* coef = scratch + (size_t)kNtSlices * 2 * C;
  const int ns = nt_slices(M);
  const dim3 grid((C / 8 + 31) / 32, ns);
  bnact_reduce_kernel<1><<<grid, 256, 0, st>>>((const __nv_bfloat16*)z, (const __nv_bfloat16*)dy, M, C, act, scale, shift, mean, rstd, scratch);
  rc = launch_status("bnact_reduce_kernel<1>");
  if (rc) return rc;
  bnact_finish_kernel<1><<<(C + 7) / 8, 256, 0, st>>>(scratch, ns, M, C, gamma, nullptr, 0.f, 0.f, nullptr, nullptr, dbeta, dgamma,
                                                        const_cast<float*>(mean), const_cast<float*>(rstd), coef, frozen_stats ? 1 : 0);
  rc = launch_status("bnact_finish_kernel<1>");
  if (rc) return rc;
  const long long nvec = M * (C / 8);
  bnact_bwd_apply_kernel<<<nt_grid(nvec), 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)z, coef, scale, shift, act, C, nvec,
                                                       C / 8, (__nv_bfloat16*)dz);
  return launch_status("bnact_bwd_apply_kernel");
}

extern "C" size_t hpe_dwconv3_wgrad_scratch_elems(int C) { return C > 0 ? (size_t)kNtSlices * 9 * C : 0; }

extern "C" int hpe_dwconv3_wgrad(const void* x, const void* dz, int B, int H, int W, int C, float* scratch, float* dw_torch, void* stream) {
  HPE_CHECK_ARG(x && dz && scratch && dw_torch && B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0, "hpe_dwconv3_wgrad: bad argument");
  HPE_CHECK_ARG((((uintptr_t)x | (uintptr_t)dz) & 15) == 0, "hpe_dwconv3_wgrad: misaligned pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const int ns = nt_slices((long long)B * H * 32);          // ~4 image rows per block (the channel groups alone are 1-3 blocks)
  const dim3 grid((C / 8 + 31) / 32, ns);
  dwconv3_wgrad_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dz, B, H, W, C, scratch);
  int rc = launch_status("dwconv3_wgrad_kernel");
  if (rc) return rc;
  dwconv3_wgrad_finish_kernel<<<(C * 9 + 255) / 256, 256, 0, st>>>(scratch, ns, C, dw_torch);
  return launch_status("dwconv3_wgrad_finish_kernel");
}

extern "C" int hpe_maxpool5_train(const void* x, int B, int H, int W, int C, void* y, uint8_t* idx, void* stream) {
  HPE_CHECK_ARG(x && y && idx && B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0, "hpe_maxpool5_train: bad argument");
  HPE_CHECK_ARG((((uintptr_t)x | (uintptr_t)y) & 15) == 0 && ((uintptr_t)idx & 7) == 0, "hpe_maxpool5_train: misaligned pointer");
  maxpool5_idx_kernel<<<nt_grid((long long)B * H * W * (C / 8)), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, B, H, W, C,
                                                                                               (__nv_bfloat16*)y, idx);
  return launch_status("maxpool5_idx_kernel");
}

extern "C" int hpe_maxpool5_backward(const uint8_t* idx, const void* dy, int B, int H, int W, int C, void* dx, void* stream) {
  HPE_CHECK_ARG(idx && dy && dx && B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0, "hpe_maxpool5_backward: bad argument");
  HPE_CHECK_ARG((((uintptr_t)dy | (uintptr_t)dx) & 15) == 0 && ((uintptr_t)idx & 7) == 0, "hpe_maxpool5_backward: misaligned pointer");
  maxpool5_bwd_kernel<<<nt_grid((long long)B * H * W * (C / 8)), 256, 0, (cudaStream_t)stream>>>(idx, (const __nv_bfloat16*)dy, B, H, W, C,
                                                                                               (__nv_bfloat16*)dx);
  return launch_status("maxpool5_bwd_kernel");
}

extern "C" int hpe_upsample_bilinear_backward(const void* dy, int B, int Hi, int Wi, int C, int Ho, int Wo, void* dx, void* stream) {
  HPE_CHECK_ARG(dy && dx && B > 0 && Hi > 0 && Wi > 0 && Ho > 0 && Wo > 0 && C > 0 && C % 8 == 0, "hpe_upsample_bilinear_backward: bad argument");
  HPE_CHECK_ARG((((uintptr_t)dy | (uintptr_t)dx) & 15) == 0, "hpe_upsample_bilinear_backward: misaligned pointer");
  upsample_bilinear_bwd_kernel<<<nt_grid((long long)B * Hi * Wi * (C / 8)), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dy, B, Hi, Wi, C, Ho, Wo,
                                                                                                         (__nv_bfloat16*)dx);
  return launch_status("upsample_bilinear_bwd_kernel");
}
