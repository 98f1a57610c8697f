// "y-halo" implicit-GEMM convolution for the 64-output-channel layers whose generic kernel is bound by operand delivery from
// L2 (profiles/: one TMA box per filter tap moves 7x-27x the algorithmic bytes through the L2->SM fabric):
//   * the 7x7/s2 stem on the space-to-depth image (4 row taps, K = 64 per tap, Cout = 64), optionally with the
//     MaxPool2d(3, 2, 1) that follows it taken in the epilogue (hpe_conv_desc.pool),
//   * 3x3 / stride 1 / 64 -> 64 (ResNet layer1 conv2; also its backward-data conv).
// An M tile is an 8-wide x 32-tall brick of ONE image.  One halo'd TMA box {64 ch, 8 (+2) px, 32 + taps_y - 1 rows} holds
// every filter-row shift of the brick: a shift by one image row is a shift of 8 (10) shared-memory rows, so each row tap
// reads the SAME box through a different descriptor start; the 3x3's column taps are 128-byte shifts of the start
// (SWIZZLE_128B is a function of the absolute shared-memory address, so such starts work with base offset 0).
//
// TWO IMAGE ROWS PER TMEM LANE.  Measured on B200 (in-kernel ablation, profiles/): a 128 x 64 x 16 tcgen05.mma from shared-memory
// operands takes ~67 cycles and a 128 x 128 x 16 one ~90, so N = 64 MMAs leave most of the tensor pipe idle.  The kernel
// therefore makes N = 128 out of Cout = 64: lane m = (r, x) of the accumulator holds image row 2r in columns 0..63 and row
// 2r+1 in columns 64..127.  With the A descriptor's 8-row-group stride set to TWO image rows, box row 2r + t' is operand row
// r of "tap t'", and
//     out(2r)   = sum_t  X(2r + t)     * W_t          out(2r+1) = sum_t' X(2r + t') * W_{t'-1}
// so tap t' = 1..taps_y-1 is ONE N = 128 MMA against the stacked weights [W_t' ; W_t'-1] (the weight taps are kept in
// descending order in shared memory, so every such pair is contiguous), and only t' = 0 / t' = taps_y are N = 64 MMAs on one
// column half (both rows thus add their taps in the same order: a pixel's value does not depend on its row's parity inside
// the brick).  Stem: 12 N=128 + 8 N=64 MMAs per 256 pixels instead of 32 N=64; 3x3: 24 + 24 instead of 72.
//
// Warp roles: warp 0 TMA producer, warps 1-2 tcgen05.mma issuers (even / odd tiles, one accumulator each), warps 3-10 two
// epilogue groups (tcgen05.ld -> scale/shift -> ReLU|SiLU -> bf16 -> swizzled smem -> [max-pool] -> coalesced 16-byte stores).
#include "hpe_common.cuh"
#include "ptx_sm100.cuh"
#include "tma_host.cuh"
#include "conv_halo.cuh"

namespace hpe {
using namespace ptx;

constexpr int kHThreads = 352, kHEpiWarp0 = 3;
constexpr int kHN = 64;                       // Cout
constexpr int kHRows = 32;                    // image rows per brick
constexpr int kHSBBytes = kHRows * 8 * 128;   // one epilogue group's staging buffer: 256 pixels x 64 bf16

__device__ __forceinline__ uint4 h_lds_v4(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
// (explicit ld.shared: pointers derived from the re-aligned dynamic shared-memory base are generic to the compiler, and
//  generic loads may not be moved across the st.shared of the staging loop -- measured 1100 -> 670 cycles per tile)
__device__ __forceinline__ float4 h_lds_f4(uint32_t addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void h_sts_v4(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t h_relu2(uint32_t v) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&v), __nv_bfloat162(__float2bfloat16(0.f), __float2bfloat16(0.f)));
  return *reinterpret_cast<uint32_t*>(&r);
}
__device__ __forceinline__ uint32_t h_max2(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}

// Walk of the tiles first, first + step, ... as (tx, ty, b) without a division per tile.
struct TileWalk {
  int tx, ty, b, s_x, s_y, s_b;
  __device__ __forceinline__ TileWalk(int first, int step, int tiles_x, int tiles_y) {
    const int per_img = tiles_x * tiles_y;
    b = first / per_img; ty = (first % per_img) / tiles_x; tx = (first % per_img) % tiles_x;
    s_b = step / per_img; s_y = (step % per_img) / tiles_x; s_x = (step % per_img) % tiles_x;
  }
  __device__ __forceinline__ void next(int tiles_x, int tiles_y) {
    tx += s_x; if (tx >= tiles_x) { tx -= tiles_x; ++ty; }
    ty += s_y; if (ty >= tiles_y) { ty -= tiles_y; ++b; }
    b += s_b;
  }
};

// TAPS_Y filter rows, TX filter columns inside one (8 + TX - 1)-pixel-wide box.
// POOL (stem, ReLU): the epilogue takes MaxPool2d(3, 2, 1) of the brick while it is in shared memory and stores 3 x 15 pooled
// pixels; bricks overlap by one output row/column (origin (6*tx - 1, 30*ty - 1)) so every pooling window lies inside one brick.
// Convolution positions outside the image are staged as 0, which is the identity of max over ReLU outputs (each window holds
// at least one real position), so the result equals maxpool(relu(bn(conv))) of the unfused pair bit for bit.
template <int TAPS_Y, int ACT, int kHStages, int TX, bool POOL>
__global__ void __launch_bounds__(kHThreads, 1)
conv_yhalo_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const HaloParams p) {
  constexpr int kTaps = TAPS_Y * TX;
  constexpr int kBW = 8 + TX - 1;                                  // box width in pixels
  constexpr int kBoxRows = (kHRows + TAPS_Y - 1) * kBW;            // smem rows (128 B each) of one halo'd box
  constexpr int kABytes = (kBoxRows * 128 + 1023) / 1024 * 1024;
  constexpr int kWBytes = kTaps * kHN * 128;                       // [column tap][filter row, descending][64 rows][64 k] bf16
  constexpr int kOffW = kHStages * kABytes;
  constexpr int kOffSB = kOffW + kWBytes;
  constexpr int kOffSS = kOffSB + 2 * kHSBBytes;
  constexpr int kOffBar = kOffSS + 2 * kHN * 4;
  constexpr uint32_t kIdesc128 = make_idesc_bf16_m128(2 * kHN), kIdesc64 = make_idesc_bf16_m128(kHN);
  constexpr int kAccCols = 2 * kHN;                                // columns 0..63: image row 2r, 64..127: image row 2r + 1
  constexpr int kTmemCols = 2 * kAccCols;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem; uint8_t* sW = smem + kOffW; uint8_t* sSB = smem + kOffSB;
  float* sScale = reinterpret_cast<float*>(smem + kOffSS); float* sShift = sScale + kHN;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* full_bar = bars;                    // [kHStages]
  uint64_t* empty_bar = bars + kHStages;        // [kHStages]
  uint64_t* tfull_bar = bars + 2 * kHStages;    // [2]
  uint64_t* tempty_bar = tfull_bar + 2;         // [2]
  uint64_t* wfull_bar = tempty_bar + 2;         // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA); prefetch_tmap(&tmW);
    for (int s = 0; s < kHStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], 4); }
    mbar_init(wfull_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) { tmem_alloc(tmem_slot, kTmemCols); tmem_relinquish(); }
  for (int i = threadIdx.x; i < kHN; i += kHThreads) { sScale[i] = __ldg(p.scale + i); sShift[i] = __ldg(p.shift + i); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);      // warp-uniform for the compiler
  griddep_launch_dependents();
  griddep_wait();

  if (warp == 0) {
    if (lane == 0) {
      // weights: resident for the whole kernel (they do not depend on the previous layer, but TMA reads global memory
      // written by the pack kernel earlier in the stream: keep it after griddep_wait).  Packed order: tap = ky * TX + kx.
      mbar_arrive_expect_tx(wfull_bar, kWBytes);
      for (int ky = 0; ky < TAPS_Y; ++ky)
        for (int kx = 0; kx < TX; ++kx)
          tma_load_2d(sW + (kx * TAPS_Y + (TAPS_Y - 1 - ky)) * kHN * 128, &tmW, wfull_bar, (ky * TX + kx) * 64, 0);
      int stage = 0; uint32_t phase = 0;
      TileWalk w(blockIdx.x, gridDim.x, p.tiles_x, p.tiles_y);
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        // reversed walk: mirroring every coordinate of the mixed-radix tile index = walking the index backwards
        const int wtx = p.reverse ? p.tiles_x - 1 - w.tx : w.tx, wty = p.reverse ? p.tiles_y - 1 - w.ty : w.ty;
        const int wb = p.reverse ? p.nbatch - 1 - w.b : w.b;
        const int x0 = POOL ? 6 * wtx - 1 : wtx * 8, y0 = POOL ? 30 * wty - 1 : wty * kHRows;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[stage], kBoxRows * 128);      // bytes the box delivers (a stage is padded to 1 KB)
        tma_load_4d(sA + stage * kABytes, &tmA, &full_bar[stage], 0, x0 - p.x_lo, y0 - p.y_lo, wb);
        if (++stage == kHStages) { stage = 0; phase ^= 1; }
        w.next(p.tiles_x, p.tiles_y);
      }
    }
  } else if (warp == 1 || warp == 2) {
    // TWO issuing warps: warp 1 takes the even tiles (accumulator 0), warp 2 the odd ones (accumulator 1), so the issue paths
    // of consecutive tiles overlap; the tensor pipe serialises the execution.  The whole warp walks the tile loop (uniform
    // control flow keeps the stage / descriptor arithmetic in uniform registers); one elected lane issues.
    mbar_wait(wfull_bar, 0);
    int stage = 0; uint32_t phase = 0; int it = 0;
    const uint32_t sA0 = smem_u32(sA), sW0 = smem_u32(sW);
    // descriptor high words: 8-row-group stride (A: TWO image rows of the box; B: 1024 B) | version | SWIZZLE_128B
    constexpr uint32_t a_hi = ((2 * kBW * 128) >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t b_hi = (1024 >> 4) | (1u << 14) | (2u << 29);
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
      const int acc = it & 1;
      if (acc != warp - 1) {
        // The other issuer's tile.  Its full barrier is still observed: with an odd number of stages a warp that only looked at
        // its own tiles would see every OTHER phase of a stage, and a parity wait cannot tell phase n from phase n + 2 (a wait
        // for tile i + 6 could pass on tile i's completion while tile i + 3's box is still in flight).
        mbar_wait(&full_bar[stage], phase);
        if (++stage == kHStages) { stage = 0; phase ^= 1; }
        continue;
      }
      mbar_wait(&tempty_bar[acc], ((it >> 1) & 1) ^ 1);
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * kAccCols;
      const uint32_t a_stage = sA0 + stage * kABytes;
#pragma unroll
      for (int kx = 0; kx < TX; ++kx) {
        // Both image rows of a lane accumulate their taps in the SAME order (W_0, W_1, ...), so a pixel's value does not depend
        // on the parity of its row inside the brick (the pooled stem's bricks start on odd rows, the plain one's on even rows).
        const uint32_t a0 = (((a_stage + kx * 128) & 0x3FFFFu) >> 4) | (1u << 16);                       // box rows 2r + 0
        const uint32_t w0 = (((sW0 + (kx * TAPS_Y + TAPS_Y - 1) * kHN * 128) & 0x3FFFFu) >> 4) | (1u << 16);   // W_0
        // t' = 0 feeds the even image rows only: N = 64 on columns 0..63 (the tile's first MMA initialises them)
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16_elect_lo(tmem_d, a0 + 2 * k, a_hi, w0 + 2 * k, b_hi, kIdesc64, (kx | k) != 0);
        // t' = 1 .. TAPS_Y-1: N = 128 against [W_t' ; W_t'-1].  Columns 64..127 are still uninitialised at the tile's first such
        // step, which is therefore issued as two N = 64 halves (the upper one overwriting).
#pragma unroll
        for (int t = 1; t < TAPS_Y; ++t) {
          const uint32_t a_lo = (((a_stage + (t * kBW + kx) * 128) & 0x3FFFFu) >> 4) | (1u << 16);
          const uint32_t b_lo = (((sW0 + (kx * TAPS_Y + (TAPS_Y - 1 - t)) * kHN * 128) & 0x3FFFFu) >> 4) | (1u << 16);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (kx == 0 && t == 1 && k == 0) {
              umma_bf16_elect_lo(tmem_d, a_lo, a_hi, b_lo, b_hi, kIdesc64, true);
              umma_bf16_elect_lo(tmem_d + kHN, a_lo, a_hi, b_lo + ((kHN * 128) >> 4), b_hi, kIdesc64, false);
            } else {
              umma_bf16_elect_lo(tmem_d, a_lo + 2 * k, a_hi, b_lo + 2 * k, b_hi, kIdesc128, true);
            }
          }
        }
        // t' = TAPS_Y feeds the odd image rows only (W_{TAPS_Y-1}): N = 64 on columns 64..127
        const uint32_t a1 = (((a_stage + (TAPS_Y * kBW + kx) * 128) & 0x3FFFFu) >> 4) | (1u << 16);
        const uint32_t w1 = (((sW0 + (kx * TAPS_Y) * kHN * 128) & 0x3FFFFu) >> 4) | (1u << 16);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16_elect_lo(tmem_d + kHN, a1 + 2 * k, a_hi, w1 + 2 * k, b_hi, kIdesc64, true);
      }
      umma_commit_elect(&empty_bar[stage]); umma_commit_elect(&tfull_bar[acc]);
      __syncwarp();
      if (++stage == kHStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp >= kHEpiWarp0) {
    // Two independent epilogue groups of 4 warps: group g drains accumulator g (tiles blockIdx.x + (2k + g) * gridDim.x)
    // through its own staging buffer, so the tcgen05.ld -> math -> store chains of consecutive tiles overlap.
    const int ew = warp & 3, grp = (warp - kHEpiWarp0) >> 2;
    const int m = ew * 32 + lane, px = m & 7, rp = m >> 3;          // TMEM lane = (row pair rp, column px) of the brick
    const uint32_t sb = smem_u32(sSB + grp * kHSBBytes);
    const uint32_t ss = smem_u32(sScale), hs = smem_u32(sShift);
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + grp * kAccCols;
    uint4* const out = reinterpret_cast<uint4*>(p.out);
    TileWalk w(blockIdx.x + grp * gridDim.x, 2 * gridDim.x, p.tiles_x, p.tiles_y);
    int it = 0;
    for (int tile = blockIdx.x + grp * gridDim.x; tile < p.num_tiles; tile += 2 * gridDim.x, ++it) {
      const int wtx = p.reverse ? p.tiles_x - 1 - w.tx : w.tx, wty = p.reverse ? p.tiles_y - 1 - w.ty : w.ty;
      const int wb = p.reverse ? p.nbatch - 1 - w.b : w.b;
      const int x0 = POOL ? 6 * wtx - 1 : wtx * 8, y0 = POOL ? 30 * wty - 1 : wty * kHRows;
      mbar_wait(&tfull_bar[grp], it & 1);
      tc_fence_after();
#pragma unroll
      for (int h = 0; h < 2; ++h) {                                  // h = 0: image row 2 rp (columns 0..63), h = 1: row 2 rp + 1
        uint32_t v[64];
        tmem_ld_x32(taddr + h * kHN, v);
        tmem_ld_x32(taddr + h * kHN + 32, v + 32);
        tmem_ld_wait();
        if (h == 1) {                                                // accumulator drained: hand it back before anything else
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[grp]);
        } else {
          // the previous tile's reads of the staging buffer are done (without the pool a warp re-reads only its own rows)
          if (POOL) named_bar_sync(1 + grp, 128); else __syncwarp();
        }
        const int srow = (2 * rp + h) * 8 + px;                      // staging row = pixel index inside the brick, y-major
        // POOL: a convolution position outside the image is staged as 0
        const bool inside = !POOL || (static_cast<unsigned>(x0 + px) < static_cast<unsigned>(p.Wc) &&
                                      static_cast<unsigned>(y0 + 2 * rp + h) < static_cast<unsigned>(p.Hc));
        const uint32_t rowp = sb + srow * 128;
        // scale/shift fetched one 8-channel chunk ahead of the st.shared of the previous one
        float4 nsa = h_lds_f4(ss), nsb = h_lds_f4(ss + 16), nha = h_lds_f4(hs), nhb = h_lds_f4(hs + 16);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 sa = nsa, sb4 = nsb, ha = nha, hb = nhb;
          if (j < 7) {
            nsa = h_lds_f4(ss + (j + 1) * 32); nsb = h_lds_f4(ss + (j + 1) * 32 + 16);
            nha = h_lds_f4(hs + (j + 1) * 32); nhb = h_lds_f4(hs + (j + 1) * 32 + 16);
          }
          const float sc[8] = {sa.x, sa.y, sa.z, sa.w, sb4.x, sb4.y, sb4.z, sb4.w};
          const float sh[8] = {ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w};
          // packed fp32x2 FMAs (same bits as fmaf per element) and ReLU on the packed bf16 pairs (relu(round(x)) == round(relu(x))):
          // 12 instead of 20 arithmetic instructions per 8 channels in an epilogue that shares its schedulers with the MMA issuers
          float o[8];
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            const float2 r = __ffma2_rn(make_float2(__uint_as_float(v[j * 8 + e]), __uint_as_float(v[j * 8 + e + 1])),
                                        make_float2(sc[e], sc[e + 1]), make_float2(sh[e], sh[e + 1]));
            o[e] = r.x; o[e + 1] = r.y;
          }
          if (ACT == HPE_ACT_SILU) {
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] = o[e] / (1.f + __expf(-o[e]));
          }
          uint4 pk = make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
          if (ACT == HPE_ACT_RELU) { pk.x = h_relu2(pk.x); pk.y = h_relu2(pk.y); pk.z = h_relu2(pk.z); pk.w = h_relu2(pk.w); }
          if (POOL && !inside) pk = make_uint4(0u, 0u, 0u, 0u);
          h_sts_v4(rowp + ((j ^ px) << 4), pk);                      // srow & 7 == px
        }
      }
      if (POOL) {
        named_bar_sync(3 + grp, 128);                                // the whole brick is staged
#pragma unroll
        for (int o = m; o < 45 * 8; o += 128) {                      // (pooled pixel, 16-byte channel chunk)
          const int pp = o >> 3, j = o & 7, ppy = pp / 3, ppx = pp - 3 * ppy;
          uint4 mx = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
          for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
              const int r = (2 * ppy + ky) * 8 + 2 * ppx + kx;
              const uint4 q = h_lds_v4(sb + r * 128 + ((j ^ (r & 7)) << 4));
              mx.x = h_max2(mx.x, q.x); mx.y = h_max2(mx.y, q.y); mx.z = h_max2(mx.z, q.z); mx.w = h_max2(mx.w, q.w);
            }
          const int oy = 15 * wty + ppy, ox = 3 * wtx + ppx;
          if (oy < p.Ho && ox < p.Wo) out[((static_cast<size_t>(wb) * p.Ho + oy) * p.Wo + ox) * 8 + j] = mx;
        }
      } else {
        // Staging buffer -> global memory with plain 16-byte stores, 4 pixels (512 contiguous bytes of an image row) per
        // instruction: no proxy fence, no bulk-store wait, no block barrier -- the warp staged image rows 8 ew .. 8 ew + 7 itself.
        __syncwarp();
        const int ox = x0 + (lane >> 3), oy = y0 + 8 * ew, c = lane & 7;
        uint4* const o0 = out + ((static_cast<size_t>(wb) * p.Ho + oy) * p.Wo + ox) * 8 + c;
        const uint32_t s0 = sb + (ew * 64 + (lane >> 3)) * 128;
#pragma unroll
        for (int i = 0; i < 16; ++i) {                               // brick row 8 ew + (i >> 1), columns 4 (i & 1) .. + 3
          const int r = i * 4 + (lane >> 3);                         // (staging row - 64 ew): r & 7 == 4 (i & 1) + (lane >> 3)
          const uint4 q = h_lds_v4(s0 + i * 512 + ((c ^ (r & 7)) << 4));
          if (oy + (i >> 1) < p.Ho && ox + 4 * (i & 1) < p.Wo) o0[((i >> 1) * p.Wo + 4 * (i & 1)) * 8] = q;
        }
      }
      w.next(p.tiles_x, p.tiles_y);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, kTmemCols); }
}

template <int TAPS_Y, int kHStages, int TX>
static size_t halo_smem_bytes() {
  return (size_t)kHStages * (((kHRows + TAPS_Y - 1) * (8 + TX - 1) * 128 + 1023) / 1024 * 1024) + (size_t)TAPS_Y * TX * kHN * 128 +
         2 * kHSBBytes + 2 * kHN * 4 + (2 * kHStages + 5) * 8 + 16 + 1024;
}

typedef void (*HaloFn)(const CUtensorMap, const CUtensorMap, const HaloParams);

int halo_launch(int taps_y, int act, const CUtensorMap* tmA, const CUtensorMap* tmW, const HaloParams& p, int grid, cudaStream_t stream) {
  HaloFn fn = nullptr; size_t smem = 0;
  if (p.pool) {
    if (taps_y != 4 || act != HPE_ACT_RELU) return fail(HPE_EUNSUPPORTED, "halo_launch: the fused max-pool exists for the ReLU stem only");
    smem = halo_smem_bytes<4, 3, 1>();
    fn = conv_yhalo_kernel<4, HPE_ACT_RELU, 3, 1, true>;
  } else if (taps_y == 4) {                      // stem: 35 KB per 256-pixel brick, 3 bricks in flight hide the HBM latency of the image read
    smem = halo_smem_bytes<4, 3, 1>();
    fn = act == HPE_ACT_RELU ? conv_yhalo_kernel<4, HPE_ACT_RELU, 3, 1, false>
         : (act == HPE_ACT_SILU ? conv_yhalo_kernel<4, HPE_ACT_SILU, 3, 1, false> : conv_yhalo_kernel<4, HPE_ACT_NONE, 3, 1, false>);
  } else if (taps_y == 3) {                      // 3x3: one 10-pixel-wide box (43 KB) per brick + 72 KB of resident weights
    smem = halo_smem_bytes<3, 2, 3>();
    fn = act == HPE_ACT_RELU ? conv_yhalo_kernel<3, HPE_ACT_RELU, 2, 3, false>
         : (act == HPE_ACT_SILU ? conv_yhalo_kernel<3, HPE_ACT_SILU, 2, 3, false> : conv_yhalo_kernel<3, HPE_ACT_NONE, 2, 3, false>);
  } else {
    return fail(HPE_EUNSUPPORTED, "halo_launch: no kernel for %d row taps", taps_y);
  }
  cudaError_t e = cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return fail(HPE_ECUDA, "cudaFuncSetAttribute(halo smem=%zu) failed: %s", smem, cudaGetErrorString(e));
  HaloParams pp = p;
  void* args[3] = {(void*)tmA, (void*)tmW, &pp};
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kHThreads); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  e = cudaLaunchKernelExC(&cfg, (const void*)fn, args);
  if (e != cudaSuccess) return fail(HPE_ECUDA, "launch of conv_yhalo_kernel failed: %s", cudaGetErrorString(e));
  return launch_status("conv_yhalo_kernel");
}
}  // namespace hpe
