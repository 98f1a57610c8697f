// Bandwidth-bound NHWC bf16 kernels around the tcgen05 GEMMs of the reference's real pose model
// (DINOv3Pose = ConvNeXt backbone + FeatureAdaptor + SPPF + PAN + PoseHead):
//   dwconv_kernel<KS>        depthwise KSxKS conv (pad KS/2, stride 1), two fused epilogues:
//                              LN   : + bias -> LayerNorm over C (eps, gamma, beta)   = ConvNeXt Block.dwconv + .norm
//                                     (/root/reference/HPE/thirdparty/dinov3/dinov3/models/convnext.py:57-58, 73-76)
//                              BNACT: * scale + shift -> SiLU                          = ConvBlock(k=3, g=C) of the head's
//                                     class branch (/root/reference/HPE/models/nn/modules/head.py:143-144, block.py:20-45)
//   layernorm_nhwc_kernel    per-pixel LayerNorm over C (channels_first LayerNorm of the stem / downsample layers,
//                            convnext.py:112-117, 157-167)
//   patchify4_kernel         fp32 NCHW image -> bf16 [B,H/4,W/4,64] rows (c,ky,kx) of the 4x4/s4 stem conv (48 real + 16 zero)
//   maxpool5_kernel          MaxPool2d(5, 1, 2) of SPPF (block.py:102-114), reading/writing channel slices of the concat buffer
//   upsample_bilinear_kernel F.interpolate(mode='bilinear', align_corners=False) of PAN (block.py:132-151) into a concat slice
// All tensors are NHWC bf16 with an explicit channel pitch (elements between consecutive pixels), so torch.cat along
// channels never materialises: producers write into their slice.
#include "hpe_common.cuh"

namespace hpe {

__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

constexpr int kDwRX = 4;          // output pixels along x per thread (register sliding window)
constexpr int kDwThreads = 128;

// grid: (ceil(W / (kDwRX * px_per_cta)), H, B); block: Cv = C/8 channel groups x px_per_cta pixel groups.
// LN epilogue: conv results of the CTA's pixels go through shared memory so one warp can normalise a pixel.
template <int KS, bool LN>
__global__ void __launch_bounds__(kDwThreads)
dwconv_kernel(const __nv_bfloat16* __restrict__ x, int in_pitch, const float* __restrict__ wt /* [KS*KS][C] */,
              const float* __restrict__ a /* LN: bias | BNACT: scale */, const float* __restrict__ b2 /* LN: gamma | BNACT: shift */,
              const float* __restrict__ b3 /* LN: beta */, float eps, int H, int W, int C,
              __nv_bfloat16* __restrict__ y, int out_pitch, int groups_per_cta) {
  extern __shared__ float tile[];                  // LN: [groups_per_cta * kDwRX][C] fp32
  constexpr int R = KS / 2;
  const int Cv = C / 8;
  const int cv = threadIdx.x % Cv, pg = threadIdx.x / Cv;
  const int b = blockIdx.z, oy = blockIdx.y;
  const int ox0 = (blockIdx.x * groups_per_cta + pg) * kDwRX;
  const bool active = pg < groups_per_cta && ox0 < W;
  float acc[kDwRX][8];
#pragma unroll
  for (int r = 0; r < kDwRX; ++r)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[r][e] = 0.f;
  if (active) {
    for (int ky = 0; ky < KS; ++ky) {
      const int iy = oy + ky - R;
      if (iy < 0 || iy >= H) continue;
      const __nv_bfloat16* rowp = x + (static_cast<size_t>(b) * H + iy) * W * in_pitch + cv * 8;
#pragma unroll
      for (int ix_rel = 0; ix_rel < kDwRX + KS - 1; ++ix_rel) {
        const int ix = ox0 + ix_rel - R;
        if (ix < 0 || ix >= W) continue;
        float v[8];
        unpack8(__ldg(reinterpret_cast<const uint4*>(rowp + static_cast<size_t>(ix) * in_pitch)), v);
#pragma unroll
        for (int r = 0; r < kDwRX; ++r) {
          const int kx = ix_rel - r;
          if (kx < 0 || kx >= KS) continue;
          const float4 w0 = __ldg(reinterpret_cast<const float4*>(wt + static_cast<size_t>(ky * KS + kx) * C + cv * 8));
          const float4 w1 = __ldg(reinterpret_cast<const float4*>(wt + static_cast<size_t>(ky * KS + kx) * C + cv * 8 + 4));
          acc[r][0] = fmaf(v[0], w0.x, acc[r][0]); acc[r][1] = fmaf(v[1], w0.y, acc[r][1]);
          acc[r][2] = fmaf(v[2], w0.z, acc[r][2]); acc[r][3] = fmaf(v[3], w0.w, acc[r][3]);
          acc[r][4] = fmaf(v[4], w1.x, acc[r][4]); acc[r][5] = fmaf(v[5], w1.y, acc[r][5]);
          acc[r][6] = fmaf(v[6], w1.z, acc[r][6]); acc[r][7] = fmaf(v[7], w1.w, acc[r][7]);
        }
      }
    }
  }
  if (!LN) {
    if (!active) return;
    float sc[8], sh[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { sc[e] = __ldg(a + cv * 8 + e); sh[e] = __ldg(b2 + cv * 8 + e); }
#pragma unroll
    for (int r = 0; r < kDwRX; ++r) {
      const int ox = ox0 + r;
      if (ox >= W) break;
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) { const float t = fmaf(acc[r][e], sc[e], sh[e]); o[e] = t / (1.f + __expf(-t)); }
      *reinterpret_cast<uint4*>(y + ((static_cast<size_t>(b) * H + oy) * W + ox) * out_pitch + cv * 8) = pack8(o);
    }
    return;
  }
  // ---- LayerNorm epilogue
  if (active) {
#pragma unroll
    for (int r = 0; r < kDwRX; ++r)
#pragma unroll
      for (int e = 0; e < 8; ++e) tile[static_cast<size_t>(pg * kDwRX + r) * C + cv * 8 + e] = acc[r][e] + __ldg(a + cv * 8 + e);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = kDwThreads / 32;
  for (int pl = warp; pl < groups_per_cta * kDwRX; pl += nwarps) {
    const int ox = blockIdx.x * groups_per_cta * kDwRX + pl;
    if (ox >= W) continue;
    const float* t = tile + static_cast<size_t>(pl) * C;
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += t[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s / C;
    float q = 0.f;
    for (int c = lane; c < C; c += 32) { const float d = t[c] - mean; q += d * d; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q / C + eps);
    __nv_bfloat16* yo = y + ((static_cast<size_t>(b) * H + oy) * W + ox) * out_pitch;
    for (int c8 = lane; c8 < Cv; c8 += 32) {
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) { const int c = c8 * 8 + e; o[e] = (t[c] - mean) * rstd * __ldg(b2 + c) + __ldg(b3 + c); }
      *reinterpret_cast<uint4*>(yo + c8 * 8) = pack8(o);
    }
  }
}

// warp per pixel; C multiple of 8
__global__ void __launch_bounds__(256)
layernorm_nhwc_kernel(const __nv_bfloat16* __restrict__ x, int in_pitch, const float* __restrict__ gamma,
                      const float* __restrict__ beta, float eps, long long M, int C, __nv_bfloat16* __restrict__ y, int out_pitch) {
  const long long pix = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (pix >= M) return;
  const int Cv = C / 8;
  const __nv_bfloat16* xp = x + pix * in_pitch;
  float s = 0.f, q = 0.f;
  for (int c8 = lane; c8 < Cv; c8 += 32) {
    float v[8]; unpack8(__ldg(reinterpret_cast<const uint4*>(xp + c8 * 8)), v);
#pragma unroll
    for (int e = 0; e < 8; ++e) s += v[e];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s / C;
  for (int c8 = lane; c8 < Cv; c8 += 32) {
    float v[8]; unpack8(__ldg(reinterpret_cast<const uint4*>(xp + c8 * 8)), v);
#pragma unroll
    for (int e = 0; e < 8; ++e) { const float d = v[e] - mean; q += d * d; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q / C + eps);
  __nv_bfloat16* yp = y + pix * out_pitch;
  for (int c8 = lane; c8 < Cv; c8 += 32) {
    float v[8], o[8]; unpack8(__ldg(reinterpret_cast<const uint4*>(xp + c8 * 8)), v);
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = (v[e] - mean) * rstd * __ldg(gamma + c8 * 8 + e) + __ldg(beta + c8 * 8 + e);
    *reinterpret_cast<uint4*>(yp + c8 * 8) = pack8(o);
  }
}

// one thread per (output pixel, 8-element group of the 64-wide row); k = c*16 + ky*4 + kx  (Conv2d weight.reshape(Cout, 48))
__global__ void __launch_bounds__(256)
patchify4_kernel(const float* __restrict__ img, int B, int H, int W, __nv_bfloat16* __restrict__ out) {
  const int Ho = H / 4, Wo = W / 4;
  const long long total = static_cast<long long>(B) * Ho * Wo * 8;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % 8);
    const long long pix = i / 8;
    const int ox = static_cast<int>(pix % Wo), oy = static_cast<int>((pix / Wo) % Ho), b = static_cast<int>(pix / (static_cast<long long>(Wo) * Ho));
    float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (g < 6) {                                     // k0 = g*8: c = g/2, ky = (g%2)*2 + {0,1}, kx = 0..3
      const int c = g >> 1, ky0 = (g & 1) * 2;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(img + ((static_cast<size_t>(b) * 3 + c) * H + (4 * oy + ky0 + j)) * W + 4 * ox));
        o[4 * j] = v.x; o[4 * j + 1] = v.y; o[4 * j + 2] = v.z; o[4 * j + 3] = v.w;
      }
    }
    *reinterpret_cast<uint4*>(out + pix * 64 + g * 8) = pack8(o);
  }
}

__device__ __forceinline__ uint32_t bmax2(uint32_t p, uint32_t q) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&p), *reinterpret_cast<__nv_bfloat162*>(&q));
  return *reinterpret_cast<uint32_t*>(&r);
}

__global__ void __launch_bounds__(256)
maxpool5_kernel(const __nv_bfloat16* __restrict__ x, int in_pitch, int B, int H, int W, int C,
                __nv_bfloat16* __restrict__ y, int out_pitch) {
  const int Cv = C / 8;
  const long long total = static_cast<long long>(B) * H * W * Cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(i % Cv);
    const int ox = static_cast<int>((i / Cv) % W), oy = static_cast<int>((i / (static_cast<long long>(Cv) * W)) % H);
    const int b = static_cast<int>(i / (static_cast<long long>(Cv) * W * H));
    const uint32_t ninf = 0xFF80FF80u;
    uint4 m = make_uint4(ninf, ninf, ninf, ninf);
    for (int ky = -2; ky <= 2; ++ky) {
      const int iy = oy + ky;
      if (iy < 0 || iy >= H) continue;
      for (int kx = -2; kx <= 2; ++kx) {
        const int ix = ox + kx;
        if (ix < 0 || ix >= W) continue;
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(x + ((static_cast<size_t>(b) * H + iy) * W + ix) * in_pitch + cv * 8));
        m.x = bmax2(m.x, v.x); m.y = bmax2(m.y, v.y); m.z = bmax2(m.z, v.z); m.w = bmax2(m.w, v.w);
      }
    }
    *reinterpret_cast<uint4*>(y + ((static_cast<size_t>(b) * H + oy) * W + ox) * out_pitch + cv * 8) = m;
  }
}

// torch upsample_bilinear2d, align_corners=False: src = max((dst + 0.5) * (in/out) - 0.5, 0)
__global__ void __launch_bounds__(256)
upsample_bilinear_kernel(const __nv_bfloat16* __restrict__ x, int in_pitch, int B, int Hi, int Wi, int C, int Ho, int Wo,
                         __nv_bfloat16* __restrict__ y, int out_pitch) {
  const int Cv = C / 8;
  const float sy = static_cast<float>(Hi) / Ho, sx = static_cast<float>(Wi) / Wo;
  const long long total = static_cast<long long>(B) * Ho * Wo * Cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(i % Cv);
    const int ox = static_cast<int>((i / Cv) % Wo), oy = static_cast<int>((i / (static_cast<long long>(Cv) * Wo)) % Ho);
    const int b = static_cast<int>(i / (static_cast<long long>(Cv) * Wo * Ho));
    const float fy = fmaxf((oy + 0.5f) * sy - 0.5f, 0.f), fx = fmaxf((ox + 0.5f) * sx - 0.5f, 0.f);
    const int y0 = static_cast<int>(fy), x0 = static_cast<int>(fx);
    const int y1 = y0 + (y0 < Hi - 1 ? 1 : 0), x1 = x0 + (x0 < Wi - 1 ? 1 : 0);
    const float ly = fy - y0, lx = fx - x0, hy = 1.f - ly, hx = 1.f - lx;
    const __nv_bfloat16* base = x + static_cast<size_t>(b) * Hi * Wi * in_pitch + cv * 8;
    float p00[8], p01[8], p10[8], p11[8], o[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(base + (static_cast<size_t>(y0) * Wi + x0) * in_pitch)), p00);
    unpack8(__ldg(reinterpret_cast<const uint4*>(base + (static_cast<size_t>(y0) * Wi + x1) * in_pitch)), p01);
    unpack8(__ldg(reinterpret_cast<const uint4*>(base + (static_cast<size_t>(y1) * Wi + x0) * in_pitch)), p10);
    unpack8(__ldg(reinterpret_cast<const uint4*>(base + (static_cast<size_t>(y1) * Wi + x1) * in_pitch)), p11);
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = hy * (hx * p00[e] + lx * p01[e]) + ly * (hx * p10[e] + lx * p11[e]);
    *reinterpret_cast<uint4*>(y + ((static_cast<size_t>(b) * Ho + oy) * Wo + ox) * out_pitch + cv * 8) = pack8(o);
  }
}

}  // namespace hpe
using namespace hpe;

static inline int ew_grid(long long total) {
  const long long want = (total + 255) / 256, cap = (long long)num_sms() * 16;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}
#define HPE_PITCH_OK(p, C) ((p) >= (C) && (p) % 8 == 0)

extern "C" int hpe_dwconv(const void* x, int in_pitch, int B, int H, int W, int C, int ksize, int mode, const float* w_taps,
                          const float* p0, const float* p1, const float* p2, float eps, void* y, int out_pitch, void* stream) {
  HPE_CHECK_ARG(x && y && w_taps && p0 && p1, "hpe_dwconv: null pointer");
  HPE_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0, "hpe_dwconv: C must be a positive multiple of 8");
  HPE_CHECK_ARG(HPE_PITCH_OK(in_pitch, C) && HPE_PITCH_OK(out_pitch, C), "hpe_dwconv: bad channel pitch");
  HPE_CHECK_ARG((((uintptr_t)x | (uintptr_t)y | (uintptr_t)w_taps) & 15) == 0, "hpe_dwconv: misaligned pointer");
  HPE_CHECK_ARG(mode == 0 || (mode == 1 && p2), "hpe_dwconv: mode 0 = scale/shift+SiLU, 1 = bias+LayerNorm (needs beta)");
  const int Cv = C / 8;
  if (Cv > kDwThreads) return fail(HPE_EUNSUPPORTED, "hpe_dwconv: C=%d exceeds %d", C, kDwThreads * 8);
  const int groups = kDwThreads / Cv;                       // pixel groups per CTA
  const size_t smem = mode == 1 ? (size_t)groups * kDwRX * C * 4 : 0;
  dim3 grid((W + groups * kDwRX - 1) / (groups * kDwRX), H, B);
  cudaStream_t st = (cudaStream_t)stream;
  const __nv_bfloat16* xi = (const __nv_bfloat16*)x; __nv_bfloat16* yo = (__nv_bfloat16*)y;
  if (ksize == 7 && mode == 1)      dwconv_kernel<7, true><<<grid, kDwThreads, smem, st>>>(xi, in_pitch, w_taps, p0, p1, p2, eps, H, W, C, yo, out_pitch, groups);
  else if (ksize == 3 && mode == 0) dwconv_kernel<3, false><<<grid, kDwThreads, smem, st>>>(xi, in_pitch, w_taps, p0, p1, p2, eps, H, W, C, yo, out_pitch, groups);
  else if (ksize == 7 && mode == 0) dwconv_kernel<7, false><<<grid, kDwThreads, smem, st>>>(xi, in_pitch, w_taps, p0, p1, p2, eps, H, W, C, yo, out_pitch, groups);
  else if (ksize == 3 && mode == 1) dwconv_kernel<3, true><<<grid, kDwThreads, smem, st>>>(xi, in_pitch, w_taps, p0, p1, p2, eps, H, W, C, yo, out_pitch, groups);
  else return fail(HPE_EUNSUPPORTED, "hpe_dwconv: kernel size %d not in {3,7}", ksize);
  return launch_status("dwconv_kernel");
}

extern "C" int hpe_layernorm_nhwc(const void* x, int in_pitch, long long M, int C, const float* gamma, const float* beta,
                                  float eps, void* y, int out_pitch, void* stream) {
  HPE_CHECK_ARG(x && y && gamma && beta, "hpe_layernorm_nhwc: null pointer");
  HPE_CHECK_ARG(M > 0 && C > 0 && C % 8 == 0 && HPE_PITCH_OK(in_pitch, C) && HPE_PITCH_OK(out_pitch, C), "hpe_layernorm_nhwc: bad size/pitch");
  const long long blocks = (M * 32 + 255) / 256;
  layernorm_nhwc_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, in_pitch, gamma, beta, eps, M, C,
                                                                          (__nv_bfloat16*)y, out_pitch);
  return launch_status("layernorm_nhwc_kernel");
}

extern "C" int hpe_patchify4(const float* img, int B, int H, int W, void* out, void* stream) {
  HPE_CHECK_ARG(img && out, "hpe_patchify4: null pointer");
  HPE_CHECK_ARG(B > 0 && H > 0 && W > 0 && H % 4 == 0 && W % 4 == 0, "hpe_patchify4: H and W must be multiples of 4");
  HPE_CHECK_ARG((((uintptr_t)img | (uintptr_t)out) & 15) == 0, "hpe_patchify4: misaligned pointer");
  patchify4_kernel<<<ew_grid((long long)B * (H / 4) * (W / 4) * 8), 256, 0, (cudaStream_t)stream>>>(img, B, H, W, (__nv_bfloat16*)out);
  return launch_status("patchify4_kernel");
}

extern "C" int hpe_maxpool5(const void* x, int in_pitch, int B, int H, int W, int C, void* y, int out_pitch, void* stream) {
  HPE_CHECK_ARG(x && y, "hpe_maxpool5: null pointer");
  HPE_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0 && HPE_PITCH_OK(in_pitch, C) && HPE_PITCH_OK(out_pitch, C), "hpe_maxpool5: bad size/pitch");
  maxpool5_kernel<<<ew_grid((long long)B * H * W * (C / 8)), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, in_pitch, B, H, W, C,
                                                                                           (__nv_bfloat16*)y, out_pitch);
  return launch_status("maxpool5_kernel");
}

extern "C" int hpe_upsample_bilinear(const void* x, int in_pitch, int B, int Hi, int Wi, int C, int Ho, int Wo, void* y, int out_pitch,
                                     void* stream) {
  HPE_CHECK_ARG(x && y, "hpe_upsample_bilinear: null pointer");
  HPE_CHECK_ARG(B > 0 && Hi > 0 && Wi > 0 && Ho > 0 && Wo > 0 && C > 0 && C % 8 == 0 && HPE_PITCH_OK(in_pitch, C) && HPE_PITCH_OK(out_pitch, C),
                "hpe_upsample_bilinear: bad size/pitch");
  upsample_bilinear_kernel<<<ew_grid((long long)B * Ho * Wo * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, in_pitch, B, Hi, Wi, C, Ho, Wo, (__nv_bfloat16*)y, out_pitch);
  return launch_status("upsample_bilinear_kernel");
}
