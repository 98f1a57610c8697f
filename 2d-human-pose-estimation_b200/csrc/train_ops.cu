// Bandwidth-bound kernels of the training step (NHWC bf16 activations, fp32 statistics / parameters):
// BatchNorm batch statistics + apply (+residual, ReLU), BatchNorm backward (two passes), MaxPool with
// saved argmax + its backward, heatmap-gradient layout change, SGD momentum step.
// All are plain vectorised (16-byte) grid-stride kernels: one pass over the tensor each, fp32 math.
#include "hpe_common.cuh"
#include <math.h>
#include <cooperative_groups.h>

namespace hpe {

__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

// ---- per-channel reductions over the M = B*H*W rows of an [M][C] tensor --------------------------------
// MODE 0: out0 = sum z, out1 = sum z^2                       (BatchNorm batch statistics)
// MODE 1: g = relu ? (y > 0 ? dy : 0) : dy;  out0 = sum g,  out1 = sum g * (z - mean) * rstd   (BN backward)
//         With y == nullptr and msc/msh given, the ReLU mask is recomputed as fmaf(z, msc, msh) > 0 -- the forward's own
//         expression for units without a residual -- so the backward never reads y (2 of its 6 bytes per element).
// 256 threads = (256 / (C/8)) rows x (C/8) 16-byte channel vectors; 4 rows in flight per thread.
// Deterministic: every block writes its partial sums to part[blockIdx.x][2C]; bn_collect_kernel adds the
// partials in a fixed order (no atomics), so the forward pass is bit-reproducible run to run.
template <int MODE>
__global__ void __launch_bounds__(256)
bn_reduce_kernel(const __nv_bfloat16* __restrict__ z, const __nv_bfloat16* __restrict__ dy,
                 const __nv_bfloat16* __restrict__ y, const float* __restrict__ mean, const float* __restrict__ rstd,
                 int relu, long long M, int C, float* __restrict__ part,
                 const float* __restrict__ msc = nullptr, const float* __restrict__ msh = nullptr,
                 const uint8_t* __restrict__ bits = nullptr) {
  __shared__ float red[256][17];
  const int cv = C >> 3, rpp = 256 / cv;
  const int vec = threadIdx.x % cv, r0 = threadIdx.x / cv;
  float a0[8], a1[8], mu[8], rs[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) { a0[e] = 0.f; a1[e] = 0.f; mu[e] = 0.f; rs[e] = 1.f; }
  float ks[8], kh[8];
  // ReLU mask source: `bits` (one byte per 16-byte vector, bit e = y[e] > 0, written by the forward's bn_apply: 1/16 of the
  // bytes of y), else recomputed from z (units without a residual), else read from y
  const bool bmask = MODE == 1 && relu && bits != nullptr;
  const bool zmask = MODE == 1 && relu && !bmask && y == nullptr;
  if (MODE == 1) {
#pragma unroll
    for (int e = 0; e < 8; ++e) { mu[e] = mean[vec * 8 + e]; rs[e] = rstd[vec * 8 + e]; }
#pragma unroll
    for (int e = 0; e < 8; ++e) { ks[e] = zmask ? msc[vec * 8 + e] : 0.f; kh[e] = zmask ? msh[vec * 8 + e] : 0.f; }
  }
  const long long step = static_cast<long long>(gridDim.x) * rpp;
  if (r0 < rpp) {
    constexpr int U = MODE == 0 ? 8 : 4;
    for (long long row = static_cast<long long>(blockIdx.x) * rpp + r0; row < M; row += U * step) {
      uint4 vz[U], vd[U], vy[U];
      uint32_t vb[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long r = row + u * step;
        if (r < M) {
          const size_t off = static_cast<size_t>(r) * cv + vec;
          vz[u] = __ldg(reinterpret_cast<const uint4*>(z) + off);
          if (MODE == 1) {
            vd[u] = __ldg(reinterpret_cast<const uint4*>(dy) + off);
            if (bmask) vb[u] = __ldg(bits + off);
            else if (relu && !zmask) vy[u] = __ldg(reinterpret_cast<const uint4*>(y) + off);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (row + u * step < M) {
          float fz[8]; unpack8(vz[u], fz);
          if (MODE == 0) {
#pragma unroll
            for (int e = 0; e < 8; ++e) { a0[e] += fz[e]; a1[e] = fmaf(fz[e], fz[e], a1[e]); }
          } else {
            float fd[8], fy[8]; unpack8(vd[u], fd);
            if (bmask) {
#pragma unroll
              for (int e = 0; e < 8; ++e) fy[e] = ((vb[u] >> e) & 1u) ? 1.f : 0.f;
            } else if (relu && !zmask) unpack8(vy[u], fy);
            if (zmask) {
#pragma unroll
              for (int e = 0; e < 8; ++e) fy[e] = fmaf(fz[e], ks[e], kh[e]);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const float g = (relu && !(fy[e] > 0.f)) ? 0.f : fd[e];
              a0[e] += g; a1[e] = fmaf(g, (fz[e] - mu[e]) * rs[e], a1[e]);
            }
          }
        }
      }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) { red[threadIdx.x][e] = a0[e]; red[threadIdx.x][8 + e] = a1[e]; }
  __syncthreads();
  // thread t < cv*16 handles (vec = t / 16, component = t % 16): sum over the rpp rows
  for (int t = threadIdx.x; t < cv * 16; t += 256) {
    const int v = t >> 4, comp = t & 15;
    float s = 0.f;
    for (int r = 0; r < rpp; ++r) s += red[r * cv + v][comp];
    part[static_cast<size_t>(blockIdx.x) * 2 * C + (comp < 8 ? 0 : C) + v * 8 + (comp & 7)] = s;
  }
}

// Fixed-order sum of the per-block partials, then
// MODE 0: BatchNorm finalize (scale/shift/mean/rstd + running statistics);
// MODE 1: dbeta = sum g, dgamma = sum g*xhat, and the per-channel coefficients of
//         dz = A*g + Bz*z + C0  with  A = gamma*rstd, Bz = -gamma*rstd^2*dgamma/M, C0 = A*(mean*rstd*dgamma - dbeta)/M.
// block = 32 channels x 32 slices of the partial list.
template <int MODE>
__global__ void __launch_bounds__(1024)
bn_collect_kernel(const float* __restrict__ part, int nparts, long long M, int C,
                  const float* __restrict__ gamma, const float* __restrict__ beta, float eps, float momentum,
                  float* running_mean, float* running_var, float* __restrict__ o0, float* __restrict__ o1,
                  float* __restrict__ o2, float* __restrict__ o3, float* __restrict__ coef) {
  __shared__ float sm[2][32][33];
  const int cl = threadIdx.x & 31, r = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cl;
  float s0 = 0.f, s1 = 0.f;
  if (c < C)
#pragma unroll 8
    for (int b = r; b < nparts; b += 32) {      // independent loads: unrolled so ~8 L2 round trips overlap (was latency-bound)
      s0 += part[static_cast<size_t>(b) * 2 * C + c];
      s1 += part[static_cast<size_t>(b) * 2 * C + C + c];
    }
  sm[0][r][cl] = s0; sm[1][r][cl] = s1;
  __syncthreads();
  if (r != 0 || c >= C) return;
  s0 = 0.f; s1 = 0.f;
#pragma unroll
  for (int k = 0; k < 32; ++k) { s0 += sm[0][k][cl]; s1 += sm[1][k][cl]; }
  const float inv = 1.f / static_cast<float>(M);
  if (MODE == 1) {
    o0[c] = s0; o1[c] = s1;                      // dbeta, dgamma
    const float rs = o3[c], mu = o2[c];          // saved rstd, mean
    const float A = gamma[c] * rs;
    coef[c] = A; coef[C + c] = -A * rs * s1 * inv; coef[2 * C + c] = A * (mu * rs * s1 - s0) * inv;
    return;
  }
  const float mu = s0 * inv;
  const float var = fmaxf(s1 * inv - mu * mu, 0.f);
  const float rs = rsqrtf(var + eps);
  const float sc = gamma[c] * rs;
  o0[c] = sc; o1[c] = beta[c] - mu * sc; o2[c] = mu; o3[c] = rs;
  if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mu;
  if (running_var) {
    const float unbiased = M > 1 ? var * static_cast<float>(M) / static_cast<float>(M - 1) : var;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
}

// y = act(z * scale + shift (+ res)).  The grid stride (gridDim.x * 256 vectors) is a multiple of cv (a power of two <= 256),
// so a thread stays on ONE channel vector.  Its coefficients sit in shared memory (read with LDS.128 at the point of use):
// keeping them in registers cost occupancy, and these passes are bound by bytes in flight per SM (measured: 96 registers ->
// 2 blocks/SM -> 49 KB in flight was slower although it read fewer bytes).  4 independent 16-byte loads per input per thread.
__global__ void __launch_bounds__(256)
bn_apply_kernel(const __nv_bfloat16* __restrict__ z, const float* __restrict__ scale, const float* __restrict__ shift,
                const __nv_bfloat16* __restrict__ res, int relu, long long nvec, int cv, __nv_bfloat16* __restrict__ y,
                uint8_t* __restrict__ bits) {
  constexpr int U = 4;
  __shared__ __align__(16) float sk[2][2048];
  for (int c = threadIdx.x; c < cv * 8; c += 256) { sk[0][c] = __ldg(scale + c); sk[1][c] = __ldg(shift + c); }
  __syncthreads();
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const int c0 = static_cast<int>(threadIdx.x % cv) * 8;        // == (tid % cv) * 8 since 256 % cv == 0
  for (long long i = tid; i < nvec; i += U * stride) {
    uint4 vz[U], vr[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      if (j < nvec) {
        vz[u] = __ldg(reinterpret_cast<const uint4*>(z) + j);
        if (res) vr[u] = __ldg(reinterpret_cast<const uint4*>(res) + j);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      if (j >= nvec) break;
      float f[8], sc[8], sh[8]; unpack8(vz[u], f);
      *reinterpret_cast<float4*>(sc) = *reinterpret_cast<const float4*>(&sk[0][c0]);
      *reinterpret_cast<float4*>(sc + 4) = *reinterpret_cast<const float4*>(&sk[0][c0 + 4]);
      *reinterpret_cast<float4*>(sh) = *reinterpret_cast<const float4*>(&sk[1][c0]);
      *reinterpret_cast<float4*>(sh + 4) = *reinterpret_cast<const float4*>(&sk[1][c0 + 4]);
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] = fmaf(f[e], sc[e], sh[e]);
      if (res) {
        float r[8]; unpack8(vr[u], r);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] += r[e];
      }
      if (bits) {                                  // ReLU mask of this vector for the backward pass (bit e = pre-activation > 0)
        uint32_t m = 0;
#pragma unroll
        for (int e = 0; e < 8; ++e) m |= (f[e] > 0.f ? 1u : 0u) << e;
        bits[j] = static_cast<uint8_t>(m);
      }
      if (relu) {
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
      }
      reinterpret_cast<uint4*>(y)[j] = pack8(f);
    }
  }
}

// dz = A*g + Bz*z + C0 (coefficients from bn_collect_kernel<1>);  optionally g_out = g (masked gradient, for the shortcut)
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                    const __nv_bfloat16* __restrict__ z, const float* __restrict__ coef, int C,
                    int relu, long long nvec, int cv, __nv_bfloat16* __restrict__ dz, __nv_bfloat16* __restrict__ g_out,
                    const float* __restrict__ msc, const float* __restrict__ msh, const uint8_t* __restrict__ bits) {
  constexpr int U = 2;
  const bool bmask = relu && bits != nullptr;   // ReLU mask from the forward's bit mask (see bn_reduce_kernel)
  const bool zmask = relu && !bmask && y == nullptr;      // ... or recomputed from z
  __shared__ __align__(16) float sk[5][2048];   // A, Bz, C0, mask scale, mask shift
  for (int c = threadIdx.x; c < C; c += 256) {
    sk[0][c] = __ldg(coef + c); sk[1][c] = __ldg(coef + C + c); sk[2][c] = __ldg(coef + 2 * C + c);
    sk[3][c] = zmask ? __ldg(msc + c) : 0.f; sk[4][c] = zmask ? __ldg(msh + c) : 0.f;
  }
  __syncthreads();
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const int c0 = static_cast<int>(threadIdx.x % cv) * 8;       // constant per thread: the stride is a multiple of cv
  for (long long i = tid; i < nvec; i += U * stride) {
    uint4 vd[U], vz[U], vy[U];
    uint32_t vb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      vy[u] = make_uint4(0, 0, 0, 0); vb[u] = 0;
      if (j < nvec) {
        vd[u] = __ldg(reinterpret_cast<const uint4*>(dy) + j);
        vz[u] = __ldg(reinterpret_cast<const uint4*>(z) + j);
        if (bmask) vb[u] = __ldg(bits + j);
        else if (relu && !zmask) vy[u] = __ldg(reinterpret_cast<const uint4*>(y) + j);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long j = i + u * stride;
      if (j >= nvec) break;
      float fd[8], fz[8], fy[8], o[8], k[8];
      unpack8(vd[u], fd); unpack8(vz[u], fz);
      if (bmask) {
#pragma unroll
        for (int e = 0; e < 8; ++e) fy[e] = ((vb[u] >> e) & 1u) ? 1.f : 0.f;
      } else if (relu && !zmask) unpack8(vy[u], fy);
      if (zmask) {
        float ks[8];
        *reinterpret_cast<float4*>(ks) = *reinterpret_cast<const float4*>(&sk[3][c0]);
        *reinterpret_cast<float4*>(ks + 4) = *reinterpret_cast<const float4*>(&sk[3][c0 + 4]);
        *reinterpret_cast<float4*>(k) = *reinterpret_cast<const float4*>(&sk[4][c0]);
        *reinterpret_cast<float4*>(k + 4) = *reinterpret_cast<const float4*>(&sk[4][c0 + 4]);
#pragma unroll
        for (int e = 0; e < 8; ++e) fy[e] = fmaf(fz[e], ks[e], k[e]);
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) fd[e] = (relu && !(fy[e] > 0.f)) ? 0.f : fd[e];     // g
      *reinterpret_cast<float4*>(k) = *reinterpret_cast<const float4*>(&sk[1][c0]);
      *reinterpret_cast<float4*>(k + 4) = *reinterpret_cast<const float4*>(&sk[1][c0 + 4]);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = k[e] * fz[e];
      *reinterpret_cast<float4*>(k) = *reinterpret_cast<const float4*>(&sk[2][c0]);
      *reinterpret_cast<float4*>(k + 4) = *reinterpret_cast<const float4*>(&sk[2][c0 + 4]);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] += k[e];                    // == fmaf(Bz, z, C0)
      *reinterpret_cast<float4*>(k) = *reinterpret_cast<const float4*>(&sk[0][c0]);
      *reinterpret_cast<float4*>(k + 4) = *reinterpret_cast<const float4*>(&sk[0][c0 + 4]);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = fmaf(k[e], fd[e], o[e]);
      reinterpret_cast<uint4*>(dz)[j] = pack8(o);
      if (g_out) reinterpret_cast<uint4*>(g_out)[j] = pack8(fd);
    }
  }
}

__device__ __forceinline__ bool bf16_gt(uint16_t a, uint16_t b) {
  return __uint_as_float(static_cast<uint32_t>(a) << 16) > __uint_as_float(static_cast<uint32_t>(b) << 16);
}

// MaxPool 3x3 s2 p1 forward that also records which window position (ky*3+kx) won (first max in scan order,
// torch semantics), 8 channels per thread
__global__ void __launch_bounds__(256)
maxpool_idx_kernel(const __nv_bfloat16* __restrict__ x, int B, int H, int W, int C, __nv_bfloat16* __restrict__ y,
                   uint8_t* __restrict__ idx) {
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1, Cv = C / 8;
  const long long total = static_cast<long long>(B) * Ho * Wo * Cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(i % Cv);
    const int ox = static_cast<int>((i / Cv) % Wo);
    const int oy = static_cast<int>((i / (static_cast<long long>(Cv) * Wo)) % Ho);
    const int b  = static_cast<int>(i / (static_cast<long long>(Cv) * Wo * Ho));
    uint16_t best[8]; uint8_t bi[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { best[e] = 0xFF80; bi[e] = 0; }      // -inf
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int iy = 2 * oy - 1 + ky;
      if (iy < 0 || iy >= H) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int ix = 2 * ox - 1 + kx;
        if (ix < 0 || ix >= W) continue;
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(x + ((static_cast<size_t>(b) * H + iy) * W + ix) * C) + cv);
        const uint16_t* h = reinterpret_cast<const uint16_t*>(&v);
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (bf16_gt(h[e], best[e])) { best[e] = h[e]; bi[e] = static_cast<uint8_t>(ky * 3 + kx); }
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

// dx[b,iy,ix,c] = sum of dy over the (<= 4) windows that contain the pixel and whose argmax is this pixel
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ idx, int B, int H, int W, int C,
                   __nv_bfloat16* __restrict__ dx) {
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1, Cv = C / 8;
  const long long total = static_cast<long long>(B) * H * W * Cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(i % Cv);
    const int ix = static_cast<int>((i / Cv) % W);
    const int iy = static_cast<int>((i / (static_cast<long long>(Cv) * W)) % H);
    const int b  = static_cast<int>(i / (static_cast<long long>(Cv) * W * H));
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    // windows along one axis: even pixel 2a -> (o=a, k=1); odd pixel 2a+1 -> (o=a, k=2), (o=a+1, k=0)
    const int ny = (iy & 1) ? 2 : 1, nx = (ix & 1) ? 2 : 1;
    for (int jy = 0; jy < ny; ++jy) {
      const int oy = (iy >> 1) + jy, ky = (iy & 1) ? (jy ? 0 : 2) : 1;
      if (oy >= Ho) continue;
      for (int jx = 0; jx < nx; ++jx) {
        const int ox = (ix >> 1) + jx, kx = (ix & 1) ? (jx ? 0 : 2) : 1;
        if (ox >= Wo) continue;
        const size_t o = ((static_cast<size_t>(b) * Ho + oy) * Wo + ox) * Cv + cv;
        const uint2 vi = __ldg(reinterpret_cast<const uint2*>(idx) + o);
        const uint4 vd = __ldg(reinterpret_cast<const uint4*>(dy) + o);
        const uint8_t* pi = reinterpret_cast<const uint8_t*>(&vi);
        float fd[8]; unpack8(vd, fd);
        const uint8_t want = static_cast<uint8_t>(ky * 3 + kx);
#pragma unroll
        for (int e = 0; e < 8; ++e) if (pi[e] == want) acc[e] += fd[e];
      }
    }
    reinterpret_cast<uint4*>(dx)[i] = pack8(acc);
  }
}

// dL/dheatmap [B,K,H,W] (f32 | bf16, NCHW) -> [B,H,W,64] bf16 NHWC, joints zero-padded (the final conv's dgrad /
// wgrad operand); also dbias[k] = sum over b,y,x
template <bool BF16>
__global__ void __launch_bounds__(256)
hm_grad_nhwc_kernel(const void* __restrict__ dhm, int B, int K, int HW, __nv_bfloat16* __restrict__ out) {
  const long long total = static_cast<long long>(B) * HW;
  for (long long base = static_cast<long long>(blockIdx.x) * blockDim.x; base < total;
       base += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long i = base + threadIdx.x;
    if (i >= total) continue;
    const int b = static_cast<int>(i / HW), pix = static_cast<int>(i % HW);
    for (int k0 = 0; k0 < 64; k0 += 8) {
      float f[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int k = k0 + e;
        float v = 0.f;
        if (k < K) {
          const size_t off = (static_cast<size_t>(b) * K + k) * HW + pix;
          v = BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(dhm)[off]) : reinterpret_cast<const float*>(dhm)[off];
        }
        f[e] = v;
      }
      reinterpret_cast<uint4*>(out + static_cast<size_t>(i) * 64 + k0)[0] = pack8(f);
    }
  }
}

// dbias[k] = sum over b, y, x of dL/dheatmap[b][k]: one block per joint, a warp per map (coalesced, fixed lane order + shuffle
// tree), maps b = warp, warp + 8, ... added in order, the eight warps' sums added in order -> deterministic (round 1 used
// shared + global fp32 atomics here, the last run-to-run varying reduction of the training step besides the wgrad split-K)
template <bool BF16>
__global__ void __cluster_dims__(8, 1, 1) __launch_bounds__(512)
hm_dbias_kernel(const void* __restrict__ dhm, int B, int K, int HW, float* __restrict__ dbias) {
  // one 8-CTA cluster per joint (8K CTAs ~ one per SM for K = 17): CTA r sums the maps b = r*16 + warp (mod 128), a warp per map
  // with 16-byte loads; the cluster's eight totals are added by CTA 0 through distributed shared memory in rank order
  __shared__ float red[16];
  __shared__ float total;
  cooperative_groups::cluster_group cl = cooperative_groups::this_cluster();
  const unsigned r = cl.block_rank();
  const int k = blockIdx.x >> 3, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int EPV = BF16 ? 8 : 4;
  const bool vec = HW % EPV == 0;
  float acc = 0.f;
  for (int b = static_cast<int>(r) * 16 + warp; b < B; b += 128) {
    const size_t off = (static_cast<size_t>(b) * K + k) * HW;
    float a = 0.f;
    if (vec) {
      const uint4* src = reinterpret_cast<const uint4*>(BF16 ? static_cast<const void*>(reinterpret_cast<const __nv_bfloat16*>(dhm) + off)
                                                             : static_cast<const void*>(reinterpret_cast<const float*>(dhm) + off));
      const int nv = HW / EPV;
      for (int i = lane; i < nv; i += 128) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = (i + u * 32 < nv) ? __ldg(src + i + u * 32) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (BF16) { float f[8]; unpack8(v[u], f); a += ((f[0] + f[1]) + (f[2] + f[3])) + ((f[4] + f[5]) + (f[6] + f[7])); }
          else a += (__uint_as_float(v[u].x) + __uint_as_float(v[u].y)) + (__uint_as_float(v[u].z) + __uint_as_float(v[u].w));
        }
      }
    } else {
      for (int i = lane; i < HW; i += 32)
        a += BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(dhm)[off + i]) : reinterpret_cast<const float*>(dhm)[off + i];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    acc += a;
  }
  if (lane == 0) red[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) { float s = 0.f; for (int w = 0; w < 16; ++w) s += red[w]; total = s; }
  cl.sync();
  if (r == 0 && threadIdx.x == 0) {
    float s = 0.f;
    for (unsigned j = 0; j < 8; ++j) s += *cl.map_shared_rank(&total, j);
    dbias[k] = s;
  }
  cl.sync();                               // peers stay resident until CTA 0 has read their totals
}

// SGD with momentum (torch.optim.SGD, dampening 0, no nesterov):  g' = g*grad_scale + wd*p; m = mu*m + g'; p -= lr*m
__global__ void __launch_bounds__(256)
sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, long long n, float lr, float mu,
           float wd, float grad_scale) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float pi = p[i];
    const float gi = fmaf(g[i], grad_scale, wd * pi);
    const float mi = fmaf(mu, m[i], gi);
    m[i] = mi;
    p[i] = pi - lr * mi;
  }
}

// ---- AdamW + gradient-norm clipping (the reference's working trainer: optim.AdamW + clip_grad_norm_(10), HPE/train.py:52-55, 136) --
// Sum of squares of a flat fp32 buffer, deterministic: fixed per-thread order, shuffle tree, one partial per block, then a
// single-block fixed-order finish (no atomics) -> the clip coefficient, hence the whole update, is run-to-run bit-stable.
__global__ void __launch_bounds__(256)
sumsq_partial_kernel(const float* __restrict__ g, long long n, float* __restrict__ part) {
  __shared__ float red[8];
  float a = 0.f;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) { const float v = g[i]; a = fmaf(v, v, a); }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) a += __shfl_xor_sync(0xffffffffu, a, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) { float s = 0.f; for (int w = 0; w < 8; ++w) s += red[w]; part[blockIdx.x] = s; }
}
__global__ void __launch_bounds__(256)
sumsq_finish_kernel(const float* __restrict__ part, int nparts, float* __restrict__ out, int accumulate) {
  __shared__ double red[256];
  double a = 0.0;
  for (int i = threadIdx.x; i < nparts; i += 256) a += static_cast<double>(part[i]);
  red[threadIdx.x] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int i = 0; i < 256; ++i) s += red[i];
    out[0] = (accumulate ? out[0] : 0.f) + static_cast<float>(s);
  }
}
// torch.optim.AdamW (decoupled weight decay, no amsgrad), with torch.nn.utils.clip_grad_norm_ folded in:
//   clip = min(1, max_norm / (sqrt(sumsq) + 1e-6));  g = grad * grad_scale * clip;  p *= 1 - lr*wd
//   m = b1*m + (1-b1)*g;  v = b2*v + (1-b2)*g*g;  p -= (lr / bc1) * m / (sqrt(v) / sqrt(bc2) + eps)
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, long long n,
             float lr, float b1, float b2, float eps, float wd, float bc1, float rsqrt_bc2, float grad_scale,
             const float* __restrict__ sumsq, float max_norm) {
  float clip = 1.f;
  if (sumsq != nullptr && max_norm > 0.f) {
    // the norm is of the gradient AFTER grad_scale (the averaged gradient), like clip_grad_norm_ on DDP-averaged .grad
    const float norm = sqrtf(__ldg(sumsq)) * fabsf(grad_scale);
    clip = fminf(1.f, max_norm / (norm + 1e-6f));
  }
  const float gs = grad_scale * clip, step = lr / bc1, decay = 1.f - lr * wd;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float gi = g[i] * gs;
    const float mi = fmaf(b1, m[i], (1.f - b1) * gi);
    const float vi = fmaf(b2, v[i], (1.f - b2) * gi * gi);
    m[i] = mi; v[i] = vi;
    const float denom = fmaf(sqrtf(vi), rsqrt_bc2, eps);
    p[i] = fmaf(-step, __fdiv_rn(mi, denom), p[i] * decay);
  }
}

static inline int ew_grid(long long n, int per_block = 256) {
  long long g = (n + per_block - 1) / per_block;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  return static_cast<int>(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace hpe

using namespace hpe;

static int bn_shape_ok(long long M, int C, const char* fn) {
  HPE_CHECK_ARG(M > 0 && C >= 64 && C <= 2048 && (C & (C - 1)) == 0, "%s: C=%d must be a power of two in [64, 2048] and M > 0", fn, C);
  return HPE_OK;
}
constexpr int kBnMaxParts = 1184;           // 8 blocks per SM on B200; bounds the scratch
static int reduce_grid(long long M, int C) {
  const int rpp = 256 / (C / 8);
  long long g = (M + (long long)rpp * 8 - 1) / ((long long)rpp * 8);
  return (int)(g < 1 ? 1 : (g > kBnMaxParts ? kBnMaxParts : g));
}

// per-block partials [kBnMaxParts][2C] followed by the three backward coefficient arrays [3C]
extern "C" size_t hpe_bn_scratch_elems(int C) { return C > 0 ? (size_t)kBnMaxParts * 2 * C + 3 * (size_t)C : 0; }

extern "C" int hpe_bn_stats(const void* z, long long M, int C, const float* gamma, const float* beta, float eps,
                            float momentum, float* running_mean, float* running_var, float* scratch, float* scale,
                            float* shift, float* mean, float* rstd, void* stream) {
  int rc = bn_shape_ok(M, C, "hpe_bn_stats");
  if (rc) return rc;
  HPE_CHECK_ARG(z && gamma && beta && scratch && scale && shift && mean && rstd, "hpe_bn_stats: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = reduce_grid(M, C);
  bn_reduce_kernel<0><<<grid, 256, 0, st>>>((const __nv_bfloat16*)z, nullptr, nullptr, nullptr, nullptr, 0, M, C, scratch);
  bn_collect_kernel<0><<<(C + 31) / 32, 1024, 0, st>>>(scratch, grid, M, C, gamma, beta, eps, momentum, running_mean,
                                                       running_var, scale, shift, mean, rstd, nullptr);
  return launch_status("bn_stats");
}

extern "C" int hpe_bn_apply_bits(const void* z, long long M, int C, const float* scale, const float* shift, const void* residual,
                                 int relu, void* y, uint8_t* bits, void* stream) {
  int rc = bn_shape_ok(M, C, "hpe_bn_apply");
  if (rc) return rc;
  HPE_CHECK_ARG(z && scale && shift && y, "hpe_bn_apply: null pointer");
  const long long nvec = M * (C / 8);
  bn_apply_kernel<<<ew_grid(nvec), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)z, scale, shift,
                                                                  (const __nv_bfloat16*)residual, relu, nvec, C / 8,
                                                                  (__nv_bfloat16*)y, bits);
  return launch_status("bn_apply_kernel");
}

extern "C" int hpe_bn_apply(const void* z, long long M, int C, const float* scale, const float* shift, const void* residual,
                            int relu, void* y, void* stream) {
  return hpe_bn_apply_bits(z, M, C, scale, shift, residual, relu, y, nullptr, stream);
}

extern "C" int hpe_bn_backward(const void* dy, const void* y, const void* z, long long M, int C, const float* mean,
                               const float* rstd, const float* gamma, int relu, float* scratch, float* dbeta,
                               float* dgamma, void* dz, void* g_out, void* stream) {
  HPE_CHECK_ARG(y || !relu, "hpe_bn_backward: the ReLU mask needs y (or use hpe_bn_backward_zmask)");
  return hpe_bn_backward_zmask(dy, y, z, M, C, mean, rstd, gamma, relu, nullptr, nullptr, scratch, dbeta, dgamma, dz, g_out, stream);
}

static int bn_backward_impl(const void* dy, const void* y, const uint8_t* bits, const void* z, long long M, int C, const float* mean,
                            const float* rstd, const float* gamma, int relu, const float* scale, const float* shift,
                            float* scratch, float* dbeta, float* dgamma, void* dz, void* g_out, void* stream) {
  int rc = bn_shape_ok(M, C, "hpe_bn_backward");
  if (rc) return rc;
  HPE_CHECK_ARG(dy && z && mean && rstd && gamma && scratch && dbeta && dgamma && dz && (y || bits || !relu || (scale && shift)),
                "hpe_bn_backward: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = reduce_grid(M, C);
  bn_reduce_kernel<1><<<grid, 256, 0, st>>>((const __nv_bfloat16*)z, (const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, mean,
                                            rstd, relu, M, C, scratch, scale, shift, bits);
  float* coef = scratch + (size_t)kBnMaxParts * 2 * C;
  bn_collect_kernel<1><<<(C + 31) / 32, 1024, 0, st>>>(scratch, grid, M, C, gamma, nullptr, 0.f, 0.f, nullptr, nullptr,
                                                       dbeta, dgamma, const_cast<float*>(mean), const_cast<float*>(rstd), coef);
  const long long nvec = M * (C / 8);
  bn_bwd_apply_kernel<<<ew_grid(nvec), 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, (const __nv_bfloat16*)z,
                                                    coef, C, relu, nvec, C / 8, (__nv_bfloat16*)dz, (__nv_bfloat16*)g_out, scale, shift, bits);
  return launch_status("bn_backward");
}

extern "C" int hpe_bn_backward_zmask(const void* dy, const void* y, const void* z, long long M, int C, const float* mean,
                                     const float* rstd, const float* gamma, int relu, const float* scale, const float* shift,
                                     float* scratch, float* dbeta, float* dgamma, void* dz, void* g_out, void* stream) {
  return bn_backward_impl(dy, y, nullptr, z, M, C, mean, rstd, gamma, relu, scale, shift, scratch, dbeta, dgamma, dz, g_out, stream);
}

extern "C" int hpe_bn_backward_bits(const void* dy, const uint8_t* bits, const void* z, long long M, int C, const float* mean,
                                    const float* rstd, const float* gamma, float* scratch, float* dbeta, float* dgamma, void* dz,
                                    void* g_out, void* stream) {
  HPE_CHECK_ARG(bits, "hpe_bn_backward_bits: null mask");
  return bn_backward_impl(dy, nullptr, bits, z, M, C, mean, rstd, gamma, 1, nullptr, nullptr, scratch, dbeta, dgamma, dz, g_out, stream);
}

extern "C" int hpe_maxpool3x3s2_train(const void* x, int B, int H, int W, int C, void* y, uint8_t* idx, void* stream) {
  HPE_CHECK_ARG(x && y && idx, "hpe_maxpool3x3s2_train: null pointer");
  HPE_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0, "hpe_maxpool3x3s2_train: C must be a positive multiple of 8");
  const long long total = (long long)B * ((H - 1) / 2 + 1) * ((W - 1) / 2 + 1) * (C / 8);
  maxpool_idx_kernel<<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, B, H, W, C, (__nv_bfloat16*)y, idx);
  return launch_status("maxpool_idx_kernel");
}

extern "C" int hpe_maxpool3x3s2_backward(const void* dy, const uint8_t* idx, int B, int H, int W, int C, void* dx, void* stream) {
  HPE_CHECK_ARG(dy && idx && dx, "hpe_maxpool3x3s2_backward: null pointer");
  HPE_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0, "hpe_maxpool3x3s2_backward: C must be a positive multiple of 8");
  const long long total = (long long)B * H * W * (C / 8);
  maxpool_bwd_kernel<<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dy, idx, B, H, W, C, (__nv_bfloat16*)dx);
  return launch_status("maxpool_bwd_kernel");
}

extern "C" int hpe_heatmap_grad_nhwc(const void* dhm, int dtype, int B, int K, int H, int W, void* out, float* dbias, void* stream) {
  HPE_CHECK_ARG(dhm && out, "hpe_heatmap_grad_nhwc: null pointer");
  HPE_CHECK_ARG(B > 0 && K > 0 && K <= 64 && H > 0 && W > 0, "hpe_heatmap_grad_nhwc: need 0 < K <= 64");
  HPE_CHECK_ARG(dtype == HPE_F32 || dtype == HPE_BF16, "hpe_heatmap_grad_nhwc: bad dtype %d", dtype);
  cudaStream_t st = (cudaStream_t)stream;
  const long long total = (long long)B * H * W;
  const int grid = ew_grid(total);
  if (dtype == HPE_BF16) hm_grad_nhwc_kernel<true><<<grid, 256, 0, st>>>(dhm, B, K, H * W, (__nv_bfloat16*)out);
  else                   hm_grad_nhwc_kernel<false><<<grid, 256, 0, st>>>(dhm, B, K, H * W, (__nv_bfloat16*)out);
  int rc = launch_status("hm_grad_nhwc_kernel");
  if (rc || !dbias) return rc;
  if (dtype == HPE_BF16) hm_dbias_kernel<true><<<K * 8, 512, 0, st>>>(dhm, B, K, H * W, dbias);
  else                   hm_dbias_kernel<false><<<K * 8, 512, 0, st>>>(dhm, B, K, H * W, dbias);
  return launch_status("hm_dbias_kernel");
}

extern "C" int hpe_sgd_step(float* param, const float* grad, float* momentum_buf, long long n, float lr, float momentum,
                            float weight_decay, float grad_scale, void* stream) {
  HPE_CHECK_ARG(param && grad && momentum_buf && n > 0, "hpe_sgd_step: null pointer or empty");
  sgd_kernel<<<ew_grid(n), 256, 0, (cudaStream_t)stream>>>(param, grad, momentum_buf, n, lr, momentum, weight_decay, grad_scale);
  return launch_status("sgd_kernel");
}

extern "C" size_t hpe_grad_sumsq_scratch_elems(void) { return (size_t)num_sms() * 16; }

extern "C" int hpe_grad_sumsq(const float* grad, long long n, float* scratch, float* out, int accumulate, void* stream) {
  HPE_CHECK_ARG(grad && scratch && out && n > 0, "hpe_grad_sumsq: null pointer or empty");
  const int grid = ew_grid(n);
  sumsq_partial_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(grad, n, scratch);
  int rc = launch_status("sumsq_partial_kernel");
  if (rc) return rc;
  sumsq_finish_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(scratch, grid, out, accumulate);
  return launch_status("sumsq_finish_kernel");
}

extern "C" int hpe_adamw_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n, float lr, float beta1,
                              float beta2, float eps, float weight_decay, int step, float grad_scale, const float* grad_sumsq,
                              float max_norm, void* stream) {
  HPE_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && n > 0, "hpe_adamw_step: null pointer or empty");
  HPE_CHECK_ARG(step >= 1 && beta1 >= 0.f && beta1 < 1.f && beta2 >= 0.f && beta2 < 1.f, "hpe_adamw_step: step must be >= 1 and betas in [0,1)");
  const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  adamw_kernel<<<ew_grid(n), 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay,
                                                            (float)bc1, (float)(1.0 / sqrt(bc2)), grad_scale, grad_sumsq, max_norm);
  return launch_status("adamw_kernel");
}
