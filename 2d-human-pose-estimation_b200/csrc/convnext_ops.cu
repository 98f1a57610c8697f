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
#include "ptx_sm100.cuh"
#include "tma_host.cuh"

namespace hpe {

__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

constexpr int kDwThreads = 256;
constexpr int kDwRX = 4;          // output pixels along x per thread (register sliding window)

// Work unit = "pixel group": kDwRX consecutive output pixels of one row x all C channels, handled by Cv = C/8 threads
// (8 channels each).  Groups are numbered over the flattened (b, y, x-group) space and a CTA takes G = 256 / Cv
// consecutive ones, so every lane is busy whatever W is and the groups of one CTA are vertically adjacent (their input
// rows are shared through L1).  Per filter row a thread keeps the KS x 8 weights in registers and slides over
// kDwRX + KS - 1 input pixels (one 16-byte load each).
// LN epilogue: two fixed-order reductions over the Cv threads of a pixel through shared memory (mean, then centred
// variance); the normalised values never leave registers before the bf16 store.
template <int KS, bool LN>
__global__ void __launch_bounds__(kDwThreads, 2)
dwconv_kernel(const __nv_bfloat16* __restrict__ x, int in_pitch, const float* __restrict__ wt /* [KS*KS][C] */,
              const float* __restrict__ a /* LN: bias | BNACT: scale */, const float* __restrict__ b2 /* LN: gamma | BNACT: shift */,
              const float* __restrict__ b3 /* LN: beta */, float eps, int H, int W, int C, long long n_groups,
              __nv_bfloat16* __restrict__ y, int out_pitch, int G, int red_lanes) {
  extern __shared__ float smem_f[];                // LN: part[G*kDwRX][Cv], stat[G*kDwRX]
  constexpr int R = KS / 2;
  constexpr int RX = kDwRX;
  const int Cv = C / 8;
  const int cv = threadIdx.x % Cv, pg = threadIdx.x / Cv;
  const int gx = (W + RX - 1) / RX;
  const long long gid = static_cast<long long>(blockIdx.x) * G + pg;
  const bool active = pg < G && gid < n_groups;
  const int xg = static_cast<int>(gid % gx), oy = static_cast<int>((gid / gx) % H);
  const long long b = gid / (static_cast<long long>(gx) * H);
  const int ox0 = xg * RX;
  float acc[RX][8];
#pragma unroll
  for (int r = 0; r < RX; ++r)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[r][e] = 0.f;
  if (active) {
#pragma unroll 1
    for (int ky = 0; ky < KS; ++ky) {
      const int iy = oy + ky - R;
      if (iy < 0 || iy >= H) continue;
      float wk[KS][8];
#pragma unroll
      for (int kx = 0; kx < KS; ++kx) {
        const float4 w0 = __ldg(reinterpret_cast<const float4*>(wt + static_cast<size_t>(ky * KS + kx) * C + cv * 8));
        const float4 w1 = __ldg(reinterpret_cast<const float4*>(wt + static_cast<size_t>(ky * KS + kx) * C + cv * 8 + 4));
        wk[kx][0] = w0.x; wk[kx][1] = w0.y; wk[kx][2] = w0.z; wk[kx][3] = w0.w;
        wk[kx][4] = w1.x; wk[kx][5] = w1.y; wk[kx][6] = w1.z; wk[kx][7] = w1.w;
      }
      const __nv_bfloat16* rowp = x + (static_cast<size_t>(b) * H + iy) * W * in_pitch + cv * 8;
#pragma unroll
      for (int ix_rel = 0; ix_rel < RX + KS - 1; ++ix_rel) {
        const int ix = ox0 + ix_rel - R;
        if (ix < 0 || ix >= W) continue;
        float v[8];
        unpack8(__ldg(reinterpret_cast<const uint4*>(rowp + static_cast<size_t>(ix) * in_pitch)), v);
#pragma unroll
        for (int r = 0; r < RX; ++r) {
          const int kx = ix_rel - r;
          if (kx < 0 || kx >= KS) continue;
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[r][e] = fmaf(v[e], wk[kx][e], acc[r][e]);
        }
      }
    }
  }
  __nv_bfloat16* yo = y + ((static_cast<size_t>(b) * H + oy) * W + ox0) * out_pitch + cv * 8;
  if (!LN) {
    if (!active) return;
    float sc[8], sh[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { sc[e] = __ldg(a + cv * 8 + e); sh[e] = __ldg(b2 + cv * 8 + e); }
#pragma unroll
    for (int r = 0; r < RX; ++r) {
      if (ox0 + r < W) {
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) { const float t = fmaf(acc[r][e], sc[e], sh[e]); o[e] = eps < 0.f ? t : t / (1.f + __expf(-t)); }   // eps < 0: no activation (training: raw conv output)
        *reinterpret_cast<uint4*>(yo + static_cast<size_t>(r) * out_pitch) = pack8(o);
      }
    }
    return;
  }
  // ---- LayerNorm over C for each of the group's RX pixels
  float* part = smem_f;                              // [G*RX][Cv]
  float* stat = smem_f + G * RX * Cv;                // [G*RX]
  const int pairs = G * RX;
  const int rp = threadIdx.x / red_lanes, rs = threadIdx.x % red_lanes;     // reducer: pixel slot, sub-lane
  if (active) {
#pragma unroll
    for (int e = 0; e < 8; ++e) { const float bias = __ldg(a + cv * 8 + e);
#pragma unroll
      for (int r = 0; r < RX; ++r) acc[r][e] += bias; }
#pragma unroll
    for (int r = 0; r < RX; ++r)
      part[(pg * RX + r) * Cv + cv] = ((acc[r][0] + acc[r][1]) + (acc[r][2] + acc[r][3])) + ((acc[r][4] + acc[r][5]) + (acc[r][6] + acc[r][7]));
  }
  __syncthreads();
  {
    float s = 0.f;
    if (rp < pairs) for (int c = rs; c < Cv; c += red_lanes) s += part[rp * Cv + c];
    for (int o = red_lanes >> 1; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (rp < pairs && rs == 0) stat[rp] = s / C;
  }
  __syncthreads();
  float mean[RX];
#pragma unroll
  for (int r = 0; r < RX; ++r) mean[r] = active ? stat[pg * RX + r] : 0.f;
  __syncthreads();                                   // stat and part are reused for the variance
  if (active) {
#pragma unroll
    for (int r = 0; r < RX; ++r) {
      float q = 0.f;
#pragma unroll
      for (int e = 0; e < 8; ++e) { acc[r][e] -= mean[r]; q = fmaf(acc[r][e], acc[r][e], q); }
      part[(pg * RX + r) * Cv + cv] = q;
    }
  }
  __syncthreads();
  {
    float s = 0.f;
    if (rp < pairs) for (int c = rs; c < Cv; c += red_lanes) s += part[rp * Cv + c];
    for (int o = red_lanes >> 1; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (rp < pairs && rs == 0) stat[rp] = rsqrtf(s / C + eps);
  }
  __syncthreads();
  if (!active) return;
  float gam[8], bet[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) { gam[e] = __ldg(b2 + cv * 8 + e); bet[e] = __ldg(b3 + cv * 8 + e); }
#pragma unroll
  for (int r = 0; r < RX; ++r) {
    if (ox0 + r < W) {
      const float rstd = stat[pg * RX + r];
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = fmaf(acc[r][e] * rstd, gam[e], bet[e]);
      *reinterpret_cast<uint4*>(yo + static_cast<size_t>(r) * out_pitch) = pack8(o);
    }
  }
}

// ---- FMA-bound formulation of the 7x7 depthwise conv (the ConvNeXt block's 0.34 GFLOP/img run on the fp32 CUDA cores).
// CTA = one image tile (th x tw outputs, halo 3) x one 64-channel slice, staged ONCE into shared memory as fp32 with a zero
// halo; a lane owns a channel PAIR and keeps all 49 x 2 filter weights in registers; a warp computes 4 x 4 output blocks:
// 100 conflict-free 8-byte shared loads feed 1568 FMAs with no bounds checks and no weight reloads.
constexpr int kDtTY = 4, kDtRX = 4, kDtSlice = 64, kDtMaxWarps = 6;   // 2 CTAs/SM: one stages its tile while the other computes
__global__ void __launch_bounds__(kDtMaxWarps * 32, 2)
dwconv7_tiled_kernel(const __nv_bfloat16* __restrict__ x, int in_pitch, const float* __restrict__ wt /* [49][C] */,
                     const float* __restrict__ bias, int H, int W, int C, int th, int tw, int tiles_x, int tiles_y, int n_slices,
                     __nv_bfloat16* __restrict__ y, int out_pitch) {
  extern __shared__ float tile[];                  // [(th+6)][(tw+6)][64]
  int t = blockIdx.x;
  const int slice = t % n_slices; t /= n_slices;
  const int tx = t % tiles_x; t /= tiles_x;
  const int ty = t % tiles_y; const int b = t / tiles_y;
  const int x0 = tx * tw, y0 = ty * th, c0 = slice * kDtSlice;
  const int pw = tw + 6, ph = th + 6;
  // filter weights first: their global loads overlap the tile staging below instead of adding a latency round trip after it
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int ch = c0 + 2 * lane;
  const bool chok = ch < C;
  float2 w[49];
#pragma unroll
  for (int k = 0; k < 49; ++k) w[k] = chok ? __ldg(reinterpret_cast<const float2*>(wt + static_cast<size_t>(k) * C + ch)) : make_float2(0.f, 0.f);
  const float2 bs = chok ? __ldg(reinterpret_cast<const float2*>(bias + ch)) : make_float2(0.f, 0.f);
  // stage: 8 x (8 channels = 16 B) per pixel; 8 independent loads in flight per thread (a 22 x 18 tile is ~16 loads per thread:
  // the staging phase is a chain of global-load round trips and was as long as the FMA phase with 4 in flight)
  constexpr int kStageU = 8;
  const int n_items = ph * pw * 8;
  // (row, col) of a thread's item advance by blockDim / 8 pixels per step: tracked incrementally -- the two integer divisions
  // per item were ~40 of the staging loop's ~60 instructions (ncu: only 52 % of the kernel's instructions were FFMAs)
  const int ppstep = blockDim.x >> 3;               // pixels per step (blockDim is a multiple of 32)
  int srow = (threadIdx.x >> 3) / pw, scol = (threadIdx.x >> 3) % pw;
  const int sc = c0 + (threadIdx.x & 7) * 8;
  const __nv_bfloat16* xb = x + static_cast<size_t>(b) * H * W * in_pitch + sc;
  for (int i0 = threadIdx.x; i0 < n_items; i0 += kStageU * blockDim.x) {
    uint4 v[kStageU];
#pragma unroll
    for (int j = 0; j < kStageU; ++j) {
      const int i = i0 + j * blockDim.x;
      const int iy = y0 - 3 + srow, ix = x0 - 3 + scol;
      v[j] = make_uint4(0u, 0u, 0u, 0u);
      if (i < n_items && iy >= 0 && iy < H && ix >= 0 && ix < W && sc < C)
        v[j] = __ldg(reinterpret_cast<const uint4*>(xb + (static_cast<size_t>(iy) * W + ix) * in_pitch));
      scol += ppstep;
      while (scol >= pw) { scol -= pw; ++srow; }
    }
#pragma unroll
    for (int j = 0; j < kStageU; ++j) {
      const int i = i0 + j * blockDim.x;
      if (i < n_items) {
        float f[8]; unpack8(v[j], f);
        float4* d = reinterpret_cast<float4*>(tile + static_cast<size_t>(i >> 3) * kDtSlice + (i & 7) * 8);
        d[0] = make_float4(f[0], f[1], f[2], f[3]); d[1] = make_float4(f[4], f[5], f[6], f[7]);
      }
    }
  }
  __syncthreads();
  const int nbx = tw / kDtRX, nblk = (th / kDtTY) * nbx;
  for (int blk = warp; blk < nblk; blk += nwarps) {
    const int by = (blk / nbx) * kDtTY, bx = (blk % nbx) * kDtRX;
    float2 acc[kDtTY][kDtRX];
#pragma unroll
    for (int ry = 0; ry < kDtTY; ++ry)
#pragma unroll
      for (int rx = 0; rx < kDtRX; ++rx) acc[ry][rx] = bs;
    const float* tb = tile + (static_cast<size_t>(by) * pw + bx) * kDtSlice + 2 * lane;
#pragma unroll
    for (int iy = 0; iy < kDtTY + 6; ++iy) {
#pragma unroll
      for (int ix = 0; ix < kDtRX + 6; ++ix) {
        const float2 v = *reinterpret_cast<const float2*>(tb + (static_cast<size_t>(iy) * pw + ix) * kDtSlice);
#pragma unroll
        for (int ry = 0; ry < kDtTY; ++ry) {
          const int ky = iy - ry;
          if (ky < 0 || ky > 6) continue;
#pragma unroll
          for (int rx = 0; rx < kDtRX; ++rx) {
            const int kx = ix - rx;
            if (kx < 0 || kx > 6) continue;
            // packed fp32x2 FMA: the same two IEEE fmas in ONE issue slot (a 3-register FFMA issues every other cycle per
            // scheduler on sm_100; the scalar form ran this kernel at 26 % of the fp32 peak)
            acc[ry][rx] = __ffma2_rn(v, w[ky * 7 + kx], acc[ry][rx]);
          }
        }
      }
    }
    if (chok) {
#pragma unroll
      for (int ry = 0; ry < kDtTY; ++ry) {
        const int oy = y0 + by + ry;
        if (oy >= H) continue;
#pragma unroll
        for (int rx = 0; rx < kDtRX; ++rx) {
          const int ox = x0 + bx + rx;
          if (ox < W)
            *reinterpret_cast<uint32_t*>(y + ((static_cast<size_t>(b) * H + oy) * W + ox) * out_pitch + ch) = pack_bf16(acc[ry][rx].x, acc[ry][rx].y);
        }
      }
    }
  }
}

// ---- The same, fed by TMA and software-pipelined (round 2).  ncu of the kernel above (C = 384 @16x12, B = 256): fp32 FMA pipe
// active 42 % of the time, the rest staging (global load -> bf16->fp32 convert -> st.shared, `long_scoreboard` / `barrier`
// stalls) that two co-resident CTAs overlap only by chance.  Here a PERSISTENT CTA owns one 64-channel slice (its 49 x 2 filter
// weights stay in registers across tiles) and walks spatial tiles; the halo'd bf16 tile {64 ch, tw+6, th+6} arrives as ONE TMA
// box (zero padding = out-of-bounds fill, no staging instructions at all) into a two-deep ring, so tile i+1 loads while tile i
// computes; the inner loop converts bf16 pairs on the ALU pipe (two shifts per 8-byte... per 4-byte load) beside the packed FMAs.
__global__ void __launch_bounds__(kDtMaxWarps * 32, 2)
dwconv7_tma_kernel(const __grid_constant__ CUtensorMap tmX, const float* __restrict__ wt /* [49][C] */, const float* __restrict__ bias,
                   int H, int W, int C, int th, int tw, int tiles_x, int tiles_y, int n_slices, int n_spatial, int tile_bytes,
                   __nv_bfloat16* __restrict__ y, int out_pitch, float2* __restrict__ stats) {
  using namespace ptx;
  extern __shared__ uint8_t dsm_raw[];
  uint8_t* dsm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(dsm_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full_bar[2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int slice = blockIdx.x % n_slices, sp0 = blockIdx.x / n_slices, sp_step = gridDim.x / n_slices;
  const int c0 = slice * kDtSlice, ch = c0 + 2 * lane;
  const bool chok = ch < C;
  if (threadIdx.x == 0) { prefetch_tmap(&tmX); mbar_init(&full_bar[0], 1); mbar_init(&full_bar[1], 1); fence_mbar_init(); }
  float2 w[49];
#pragma unroll
  for (int k = 0; k < 49; ++k) w[k] = chok ? __ldg(reinterpret_cast<const float2*>(wt + static_cast<size_t>(k) * C + ch)) : make_float2(0.f, 0.f);
  const float2 bs = chok ? __ldg(reinterpret_cast<const float2*>(bias + ch)) : make_float2(0.f, 0.f);
  __syncthreads();
  const int pw = tw + 6;
  auto issue = [&](int sp, int buf) {                 // one elected thread: the halo'd tile of spatial tile `sp`
    const int tx = sp % tiles_x, ty = (sp / tiles_x) % tiles_y, b = sp / (tiles_x * tiles_y);
    mbar_arrive_expect_tx(&full_bar[buf], tile_bytes);
    tma_load_4d(dsm + buf * ((tile_bytes + 1023) & ~1023), &tmX, &full_bar[buf], c0, tx * tw - 3, ty * th - 3, b);
  };
  if (threadIdx.x == 0 && sp0 < n_spatial) issue(sp0, 0);
  const int nbx = tw / kDtRX, nblk = (th / kDtTY) * nbx;
  int it = 0;
  for (int sp = sp0; sp < n_spatial; sp += sp_step, ++it) {
    const int buf = it & 1;
    if (threadIdx.x == 0 && sp + sp_step < n_spatial) issue(sp + sp_step, buf ^ 1);     // that buffer was drained before the last __syncthreads
    mbar_wait(&full_bar[buf], (it >> 1) & 1);
    const int tx = sp % tiles_x, ty = (sp / tiles_x) % tiles_y, b = sp / (tiles_x * tiles_y);
    const int x0 = tx * tw, y0 = ty * th;
    const uint32_t tbase = smem_u32(dsm + buf * ((tile_bytes + 1023) & ~1023));
    for (int blk = warp; blk < nblk; blk += nwarps) {
      const int by = (blk / nbx) * kDtTY, bx = (blk % nbx) * kDtRX;
      float2 acc[kDtTY][kDtRX];
#pragma unroll
      for (int ry = 0; ry < kDtTY; ++ry)
#pragma unroll
        for (int rx = 0; rx < kDtRX; ++rx) acc[ry][rx] = bs;
#pragma unroll
      for (int iy = 0; iy < kDtTY + 6; ++iy) {
#pragma unroll
        for (int ix = 0; ix < kDtRX + 6; ++ix) {
          // pixel row p of the box = 128 B (64 bf16), 16-byte chunks XOR-swizzled with (p & 7) by TMA (SWIZZLE_128B); a warp reads the
          // whole row (lane -> channel pair 2 lane), i.e. every chunk once: conflict-free
          const int pix = (by + iy) * pw + bx + ix;
          uint32_t u;
          asm volatile("ld.shared.u32 %0, [%1];" : "=r"(u) : "r"(tbase + pix * 128 + ((((lane >> 2) ^ (pix & 7)) << 4) | ((lane & 3) << 2))));
          const float2 v = make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
#pragma unroll
          for (int ry = 0; ry < kDtTY; ++ry) {
            const int ky = iy - ry;
            if (ky < 0 || ky > 6) continue;
#pragma unroll
            for (int rx = 0; rx < kDtRX; ++rx) {
              const int kx = ix - rx;
              if (kx < 0 || kx > 6) continue;
              acc[ry][rx] = __ffma2_rn(v, w[ky * 7 + kx], acc[ry][rx]);
            }
          }
        }
      }
      if (chok) {
#pragma unroll
        for (int ry = 0; ry < kDtTY; ++ry) {
          const int oy = y0 + by + ry;
          if (oy >= H) continue;
#pragma unroll
          for (int rx = 0; rx < kDtRX; ++rx) {
            const int ox = x0 + bx + rx;
            if (ox < W)
              *reinterpret_cast<uint32_t*>(y + ((static_cast<size_t>(b) * H + oy) * W + ox) * out_pitch + ch) = pack_bf16(acc[ry][rx].x, acc[ry][rx].y);
          }
        }
      }
      if (stats != nullptr) {
        // LayerNorm statistics of the STORED (bf16-rounded) values: per pixel the sum and the sum of squares over this slice's 64
        // channels, i.e. over the warp's 32 lanes.  Reduce-scatter: at offset 16 / 8 / 4 / 2 a lane keeps one half of its
        // pixels and trades the other, so after four steps it holds ONE pixel (index lane >> 1), then one plain butterfly step:
        // 16 shuffles per quantity for the block's 16 pixels instead of 80.
        float sv[kDtTY * kDtRX], qv[kDtTY * kDtRX];
#pragma unroll
        for (int i = 0; i < kDtTY * kDtRX; ++i) {
          const uint32_t pk = pack_bf16(acc[i / kDtRX][i % kDtRX].x, acc[i / kDtRX][i % kDtRX].y);
          const float a = chok ? bf16_lo(pk) : 0.f, c2 = chok ? bf16_hi(pk) : 0.f;
          sv[i] = a + c2; qv[i] = fmaf(a, a, c2 * c2);
        }
#pragma unroll
        for (int off = 16, n = kDtTY * kDtRX; n > 1; off >>= 1, n >>= 1) {
          const bool hi = (lane & off) != 0;
#pragma unroll
          for (int i = 0; i < n / 2; ++i) {
            const float ks = hi ? sv[i + n / 2] : sv[i], ts = hi ? sv[i] : sv[i + n / 2];
            const float kq = hi ? qv[i + n / 2] : qv[i], tq = hi ? qv[i] : qv[i + n / 2];
            sv[i] = ks + __shfl_xor_sync(0xffffffffu, ts, off);
            qv[i] = kq + __shfl_xor_sync(0xffffffffu, tq, off);
          }
        }
        sv[0] += __shfl_xor_sync(0xffffffffu, sv[0], 1);
        qv[0] += __shfl_xor_sync(0xffffffffu, qv[0], 1);
        const int pi = lane >> 1, oy = y0 + by + pi / kDtRX, ox = x0 + bx + pi % kDtRX;
        if ((lane & 1) == 0 && oy < H && ox < W)
          stats[((static_cast<size_t>(b) * H + oy) * W + ox) * n_slices + slice] = make_float2(sv[0], qv[0]);
      }
    }
    __syncthreads();                                  // every warp is done with this buffer: it may be refilled two tiles later
  }
}

// A pixel is normalised by a group of `gsz` lanes (power of two >= C/8, at most 32; narrow tensors put 2-8 pixels in a
// warp); the pixel's channels stay in registers (C <= 1024: 4 x 16 B per lane), so the kernel is one read + one write and
// in-place safe.
__global__ void __launch_bounds__(256)
layernorm_nhwc_kernel(const __nv_bfloat16* __restrict__ x, int in_pitch, const float* __restrict__ gamma,
                      const float* __restrict__ beta, float eps, long long M, int C, int gsz,
                      __nv_bfloat16* __restrict__ y, int out_pitch) {
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long pix = tid / gsz;
  const int sub = static_cast<int>(tid % gsz);
  const bool ok = pix < M;
  const int Cv = C / 8;
  const __nv_bfloat16* xp = x + (ok ? pix : 0) * in_pitch;
  float v[4][8];
  float s = 0.f, q = 0.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c8 = sub + gsz * j;
    if (ok && c8 < Cv) {
      unpack8(__ldg(reinterpret_cast<const uint4*>(xp + c8 * 8)), v[j]);
#pragma unroll
      for (int e = 0; e < 8; ++e) s += v[j][e];
    }
  }
  for (int o = gsz >> 1; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s / C;
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (ok && sub + gsz * j < Cv) {
#pragma unroll
      for (int e = 0; e < 8; ++e) { v[j][e] -= mean; q = fmaf(v[j][e], v[j][e], q); }
    }
  for (int o = gsz >> 1; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rstd = rsqrtf(q / C + eps);
  __nv_bfloat16* yp = y + (ok ? pix : 0) * out_pitch;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c8 = sub + gsz * j;
    if (ok && c8 < Cv) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + c8 * 8)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + c8 * 8 + 4));
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + c8 * 8)), b1 = __ldg(reinterpret_cast<const float4*>(beta + c8 * 8 + 4));
      float o[8];
      o[0] = fmaf(v[j][0] * rstd, g0.x, b0.x); o[1] = fmaf(v[j][1] * rstd, g0.y, b0.y);
      o[2] = fmaf(v[j][2] * rstd, g0.z, b0.z); o[3] = fmaf(v[j][3] * rstd, g0.w, b0.w);
      o[4] = fmaf(v[j][4] * rstd, g1.x, b1.x); o[5] = fmaf(v[j][5] * rstd, g1.y, b1.y);
      o[6] = fmaf(v[j][6] * rstd, g1.z, b1.z); o[7] = fmaf(v[j][7] * rstd, g1.w, b1.w);
      *reinterpret_cast<uint4*>(yp + c8 * 8) = pack8(o);
    }
  }
}

// one thread per (output pixel, 8-element group of the 64-wide row); k = c*16 + ky*4 + kx  (Conv2d weight.reshape(Cout, 48)).
// uint8 input = the reference's `imgs.float() / 255.0` (HPE/train.py:117-120) fused in.
__device__ __forceinline__ float4 load_px4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 load_px4(const unsigned char* p) {
  const uchar4 u = __ldg(reinterpret_cast<const uchar4*>(p));
  return make_float4(__fdiv_rn(static_cast<float>(u.x), 255.f), __fdiv_rn(static_cast<float>(u.y), 255.f),
                     __fdiv_rn(static_cast<float>(u.z), 255.f), __fdiv_rn(static_cast<float>(u.w), 255.f));
}
template <typename T>
__global__ void __launch_bounds__(256)
patchify4_kernel(const T* __restrict__ img, int B, int H, int W, __nv_bfloat16* __restrict__ out) {
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
        const float4 v = load_px4(img + ((static_cast<size_t>(b) * 3 + c) * H + (4 * oy + ky0 + j)) * W + 4 * ox);
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
  const int G = kDwThreads / Cv;                            // pixel groups per CTA
  const long long n_groups = (long long)B * H * ((W + kDwRX - 1) / kDwRX);
  int red_lanes = 1;                                        // lanes cooperating on one pixel's reduction (power of two)
  while (red_lanes < 32 && G * kDwRX * red_lanes * 2 <= kDwThreads && red_lanes * 2 <= Cv) red_lanes <<= 1;
  const size_t smem = mode == 1 ? ((size_t)G * kDwRX * Cv + (size_t)G * kDwRX) * 4 : 0;
  const long long grid = (n_groups + G - 1) / G;
  if (grid > 0x7fffffffLL) return fail(HPE_EUNSUPPORTED, "hpe_dwconv: too many pixel groups");
  cudaStream_t st = (cudaStream_t)stream;
  const __nv_bfloat16* xi = (const __nv_bfloat16*)x; __nv_bfloat16* yo = (__nv_bfloat16*)y;
#define HPE_DW_LAUNCH(KS, LNM) dwconv_kernel<KS, LNM><<<(unsigned)grid, kDwThreads, smem, st>>>(xi, in_pitch, w_taps, p0, p1, p2, eps, H, W, C, n_groups, yo, out_pitch, G, red_lanes)
  if (ksize == 7 && mode == 1)      HPE_DW_LAUNCH(7, true);
  else if (ksize == 7 && mode == 0) HPE_DW_LAUNCH(7, false);
  else if (ksize == 3 && mode == 0) HPE_DW_LAUNCH(3, false);
  else if (ksize == 3 && mode == 1) HPE_DW_LAUNCH(3, true);
  else return fail(HPE_EUNSUPPORTED, "hpe_dwconv: kernel size %d not in {3,7}", ksize);
#undef HPE_DW_LAUNCH
  return launch_status("dwconv_kernel");
}

static int dwconv7_bias_impl(const void* x, int in_pitch, int B, int H, int W, int C, const float* w_taps, const float* bias,
                            void* y, int out_pitch, float* rowstats, void* stream) {
  HPE_CHECK_ARG(x && y && w_taps && bias, "hpe_dwconv7_bias: null pointer");
  HPE_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0, "hpe_dwconv7_bias: C must be a positive multiple of 8");
  HPE_CHECK_ARG(HPE_PITCH_OK(in_pitch, C) && HPE_PITCH_OK(out_pitch, C), "hpe_dwconv7_bias: bad channel pitch");
  HPE_CHECK_ARG((((uintptr_t)x | (uintptr_t)y | (uintptr_t)w_taps | (uintptr_t)bias) & 15) == 0, "hpe_dwconv7_bias: misaligned pointer");
  auto tile_of = [](int n) { int t = (n + 3) / 4 * 4; return t > 16 ? 16 : t; };     // (16+6)*(12+6)*256 B = 99 KB -> 2 CTAs/SM
  // tiles as even as possible: n = 24 -> two tiles of 12, not 16 + 8
  auto even_tile = [&](int n) { const int full = tile_of(n); const int cnt = (n + full - 1) / full; return ((n + cnt - 1) / cnt + 3) / 4 * 4; };
  int th = even_tile(H), tw = even_tile(W);
  if (th * tw > 16 * 12) { if (tw > 12) tw = 12; else th = 12; }
  const int tiles_y = (H + th - 1) / th, tiles_x = (W + tw - 1) / tw, n_slices = (C + kDtSlice - 1) / kDtSlice;
  const int nblk = (th / kDtTY) * (tw / kDtRX);
  const int warps = nblk <= kDtMaxWarps ? nblk : kDtMaxWarps;
  {
    // TMA-fed persistent variant (default; HPE_B200_DWCONV_TMA=0 keeps the staged kernel): x must be 16-byte aligned with a pitch
    // that is a multiple of 8 channels (checked above)
    static int use_tma = -1;
    if (use_tma < 0) { const char* e = getenv("HPE_B200_DWCONV_TMA"); use_tma = (e && e[0] == '0') ? 0 : 1; }
    const long long n_spatial = (long long)B * tiles_y * tiles_x;
    if (use_tma && n_spatial < 0x7fffffffLL) {
      CUtensorMap tm;
      const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
      const uint64_t str[3] = {(uint64_t)in_pitch * 2, (uint64_t)W * in_pitch * 2, (uint64_t)H * W * in_pitch * 2};
      const uint32_t box[4] = {64, (uint32_t)(tw + 6), (uint32_t)(th + 6), 1};
      const uint32_t ones[4] = {1, 1, 1, 1};
      int rc = encode_map(&tm, x, 4, dims, str, box, ones, "dwconv7 x");
      if (rc) return rc;
      const int tile_bytes = (th + 6) * (tw + 6) * 128;
      const size_t smem = 2 * (size_t)((tile_bytes + 1023) & ~1023) + 1024;
      HPE_CUDA_OK(cudaFuncSetAttribute((const void*)dwconv7_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      // persistent grid: 2 CTAs per SM, a multiple of the slice count, every CTA at least one spatial tile
      long long per_slice = (long long)num_sms() * 2 / n_slices;
      if (per_slice < 1) per_slice = 1;
      if (per_slice > n_spatial) per_slice = n_spatial;
      const int grid = (int)(per_slice * n_slices);
      dwconv7_tma_kernel<<<grid, warps * 32, smem, (cudaStream_t)stream>>>(tm, w_taps, bias, H, W, C, th, tw, tiles_x, tiles_y, n_slices,
                                                                         (int)n_spatial, tile_bytes, (__nv_bfloat16*)y, out_pitch,
                                                                         reinterpret_cast<float2*>(rowstats));
      return launch_status("dwconv7_tma_kernel");
    }
  }
  if (rowstats) return fail(HPE_EUNSUPPORTED, "hpe_dwconv7_bias_stats: LayerNorm statistics come from the TMA-fed kernel only (HPE_B200_DWCONV_TMA=0 is set)");
  const size_t smem = (size_t)(th + 6) * (tw + 6) * kDtSlice * 4;
  // per-DEVICE function attribute: set on every call that needs it
  if (smem > 48 * 1024)
    HPE_CUDA_OK(cudaFuncSetAttribute((const void*)dwconv7_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long grid = (long long)B * tiles_y * tiles_x * n_slices;
  if (grid > 0x7fffffffLL) return fail(HPE_EUNSUPPORTED, "hpe_dwconv7_bias: too many tiles");
  dwconv7_tiled_kernel<<<(unsigned)grid, warps * 32, smem, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, in_pitch, w_taps, bias, H, W, C, th, tw,
                                                                                 tiles_x, tiles_y, n_slices, (__nv_bfloat16*)y, out_pitch);
  return launch_status("dwconv7_tiled_kernel");
}

extern "C" int hpe_dwconv7_bias(const void* x, int in_pitch, int B, int H, int W, int C, const float* w_taps, const float* bias,
                               void* y, int out_pitch, void* stream) {
  return dwconv7_bias_impl(x, in_pitch, B, H, W, C, w_taps, bias, y, out_pitch, nullptr, stream);
}

extern "C" size_t hpe_dwconv7_stats_elems(int B, int H, int W, int C) {
  return (B > 0 && H > 0 && W > 0 && C > 0) ? (size_t)B * H * W * ((C + kDtSlice - 1) / kDtSlice) * 2 : 0;
}

extern "C" int hpe_dwconv7_bias_stats(const void* x, int in_pitch, int B, int H, int W, int C, const float* w_taps, const float* bias,
                                     void* y, int out_pitch, float* rowstats, void* stream) {
  HPE_CHECK_ARG(rowstats && ((uintptr_t)rowstats & 7) == 0, "hpe_dwconv7_bias_stats: rowstats must be a non-null, 8-byte aligned pointer");
  return dwconv7_bias_impl(x, in_pitch, B, H, W, C, w_taps, bias, y, out_pitch, rowstats, stream);
}

extern "C" int hpe_layernorm_nhwc(const void* x, int in_pitch, long long M, int C, const float* gamma, const float* beta,
                                  float eps, void* y, int out_pitch, void* stream) {
  HPE_CHECK_ARG(x && y && gamma && beta, "hpe_layernorm_nhwc: null pointer");
  HPE_CHECK_ARG(M > 0 && C > 0 && C % 8 == 0 && HPE_PITCH_OK(in_pitch, C) && HPE_PITCH_OK(out_pitch, C), "hpe_layernorm_nhwc: bad size/pitch");
  if (C > 1024) return fail(HPE_EUNSUPPORTED, "hpe_layernorm_nhwc: C=%d exceeds 1024", C);
  HPE_CHECK_ARG((((uintptr_t)gamma | (uintptr_t)beta) & 15) == 0, "hpe_layernorm_nhwc: gamma/beta must be 16-byte aligned");
  // lanes per pixel: the smallest power of two with <= 4 sixteen-byte vectors per lane (C = 384 -> 16 lanes x 3 vectors, two
  // pixels per warp).  Fewer, busier lanes put 3-4 independent loads in flight per lane instead of 1-2: the kernel is
  // latency-bound (75 MB in 36 us at C = 384 before), not bandwidth-bound.
  int gsz = 4;
  while (gsz < 32 && gsz * 4 < C / 8) gsz <<= 1;
  const long long blocks = (M * gsz + 255) / 256;
  layernorm_nhwc_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, in_pitch, gamma, beta, eps, M, C, gsz,
                                                                          (__nv_bfloat16*)y, out_pitch);
  return launch_status("layernorm_nhwc_kernel");
}

extern "C" int hpe_patchify4(const float* img, int B, int H, int W, void* out, void* stream) {
  HPE_CHECK_ARG(img && out, "hpe_patchify4: null pointer");
  HPE_CHECK_ARG(B > 0 && H > 0 && W > 0 && H % 4 == 0 && W % 4 == 0, "hpe_patchify4: H and W must be multiples of 4");
  HPE_CHECK_ARG((((uintptr_t)img | (uintptr_t)out) & 15) == 0, "hpe_patchify4: misaligned pointer");
  patchify4_kernel<float><<<ew_grid((long long)B * (H / 4) * (W / 4) * 8), 256, 0, (cudaStream_t)stream>>>(img, B, H, W, (__nv_bfloat16*)out);
  return launch_status("patchify4_kernel");
}

extern "C" int hpe_patchify4_u8(const unsigned char* img, int B, int H, int W, void* out, void* stream) {
  HPE_CHECK_ARG(img && out, "hpe_patchify4_u8: null pointer");
  HPE_CHECK_ARG(B > 0 && H > 0 && W > 0 && H % 4 == 0 && W % 4 == 0, "hpe_patchify4_u8: H and W must be multiples of 4");
  HPE_CHECK_ARG((((uintptr_t)img & 3) | ((uintptr_t)out & 15)) == 0, "hpe_patchify4_u8: misaligned pointer");
  patchify4_kernel<unsigned char><<<ew_grid((long long)B * (H / 4) * (W / 4) * 8), 256, 0, (cudaStream_t)stream>>>(img, B, H, W, (__nv_bfloat16*)out);
  return launch_status("patchify4_kernel<u8>");
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
