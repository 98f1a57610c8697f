// Implicit-GEMM convolution for sm_100a: tcgen05.mma (accumulators in TMEM) fed by TMA, with the folded
// BatchNorm scale/shift (+ residual add) (+ ReLU | SiLU) fused into the epilogue.
//
// GEMM view:  D[M = output pixels][N = Cout] = sum_taps  A_tap[M][Cin] * W_tap[N][Cin]^T
//   * activations are NHWC bf16; an M tile is a (nb x th x tw) brick of 128 output pixels, so for every
//     filter tap the A operand is ONE 4-D TMA box {64 ch, tw, th, nb} of the input shifted by the tap
//     offset -- zero padding comes from TMA out-of-bounds fill, stride 2 from TMA element strides, and
//     no im2col buffer ever exists.  The box lands in shared memory as 128 rows x 128 B with the 128-byte
//     swizzle, which is exactly the K-major SWIZZLE_128B operand layout tcgen05.mma consumes.
//   * weights are pre-packed [rows = Cout][K = taps*Cin] bf16 and loaded by 2-D TMA boxes {64, BLOCK_N}.
//   * ConvTranspose2d(k4,s2,p1) = four 2x2 sub-pixel convolutions (one per output parity class), each a
//     GEMM with 4 taps whose output tile is written through a strided TMA store map (no zero insertion).
//   * the 7x7/s2 stem runs on a space-to-depth copy of the image (16 ch per 2x2 pixel block): with an
//     overlapping-stride TMA map one box row = 4 adjacent blocks = 64 contiguous bf16, so the stem is a
//     4-tap GEMM with K = 4 x 64.
//
//   * a 1x1 conv may take a SECOND input (Cin2): its k-blocks follow the first input's in the same accumulation, which is how
//     a Bottleneck's projection shortcut disappears into conv3 (hpe_conv_run2).
//   * HEAD2: the last deconv stage keeps its bf16 output tile in shared memory and feeds it to a second tcgen05 GEMM against
//     the joint weights -- deconv3 + final_layer in one kernel (hpe_conv_run_head2).
//
// Persistent, warp-specialised CTA (1 per SM, 352 threads):
//   warp 0  TMA producer (A + W operand ring, STAGES deep); the whole warp runs the loop, one elected lane issues
//   warp 1  TMEM allocator + tcgen05.mma issuer (elected lane; double-buffered accumulators)
//   warp 2  residual-tile TMA producer (into the epilogue staging ring) / loads the fused head's joint weights
//   warps 3-10  epilogue (two warps per TMEM lane quarter, each owning 32 of a chunk's 64 columns):
//               tcgen05.ld -> scale/shift/residual/activation -> bf16 -> swizzled smem -> TMA store
// PAIR kernels (256- and 192-wide tiles): the two CTAs of a cluster share ONE tcgen05.mma.cta_group::2 stream with M = 256;
// each CTA loads its own A brick and half of the weight tile and drains its own 128 TMEM lanes.
#include "hpe_common.cuh"
#include "ptx_sm100.cuh"
#include "tma_host.cuh"
#include "conv_halo.cuh"
#include <stdlib.h>
#include <string.h>

namespace hpe {
using namespace ptx;

constexpr int kConvThreads   = 352;                   // 3 control warps + 8 epilogue warps
constexpr int kConvThreads2  = 608;                   // EPI2: 3 control warps + two groups of 8 epilogue warps
constexpr int kConvThreadsC  = 640;                   // CHAIN: 4 control warps (+ the second GEMM's issuer) + two groups of 8 epilogue warps
constexpr int kEpiThreads    = 256;                   // epilogue threads of ONE group
constexpr int kEpiWarp0      = 3;
constexpr int kMaxCout       = 2048;
constexpr int kMaxBlockN     = 256;                   // scale/shift of ONE N tile staged in shared memory
constexpr int kTileM         = 128;
constexpr int kKBlock        = 64;                    // bf16 elements per k-block = one 128-byte swizzle row
constexpr int kABytes        = kTileM * kKBlock * 2;  // 16 KB
constexpr int kSBBytes       = kTileM * 64 * 2;       // 128 rows x 64 bf16

struct OutMaps { CUtensorMap m[4]; };

struct ConvParams {
  int kind, taps, kc_per_tap, in_stride;
  int kc2, in_stride2;          // second A source (fused projection shortcut): its k-blocks follow the first source's
  int tw, th, nb;
  int tiles_x, tiles_y, tiles_b, n_tiles_n, n_classes, num_tiles;   // PAIR kernels: num_tiles counts tile PAIRS
  int Ho, Wo, B, Cout;          // Ho,Wo = extent of the tile grid (deconv: input-resolution grid per class)
  int cout_total;               // HPE_CONV_FINAL: channel count of the NCHW tensor y is a slice of
  int act, has_res;
  const float* scale;
  const float* shift;
  float* hm_out;                // HPE_CONV_FINAL / fused head: fp32 NCHW heatmaps
  int head2_k;                  // fused head (HEAD2): number of joints K <= 32 of the second GEMM
  const float* bias2;           // fused head: conv bias [K]
  const float* scale2;          // CHAIN: folded-BN scale / shift [N2] of the chained 1x1 conv
  const float* shift2;
  int res_prefetch;             // != 0: the residual producer prefetches the NEXT tile's residual boxes into L2 while it serves this one
  int chain_defer;              // CHAIN: the chained conv's output chunks of tile i are drained after this many chunks of tile i+1 (0 = right behind tile i)
  int reverse;                  // != 0: the work items are walked from the last to the first (hpe_conv_set_schedule)
  unsigned long long l2_out, l2_res;     // != 0: L2 eviction-priority policy word of the output stores / the residual loads
  unsigned long long l2_in;              // policy word of the A-operand loads (kL2EvictNormal unless hpe_conv_set_schedule says otherwise)
  const float2* rowstats;       // HPE_ACT_GELU_LN: per input pixel and 64-channel slice (sum, sum of squares) of the stored input
  int ln_slices;                //   slices per pixel;  ln_inv_c = 1 / channels, ln_eps: the LayerNorm folded into this 1x1 conv
  float ln_inv_c, ln_eps;
};

struct TileCoord { int n_tile, cls, x0, y0, b0; };

// CHAIN kernels: entry q of a CTA's chunk sequence.  Tile i contributes kCT chunks of the layer's own output and kC2 chunks of the
// chained conv's output (c >= kCT); the latter need the second GEMM over ALL of tile i's chunks, so they are walked `defer` chunks
// into tile i+1 (right behind the last tile's chunks for the last tile) -- by then their accumulator has long been committed and
// no epilogue group waits for it.  defer = 0 is the plain order [tile i | its chained chunks].  The residual producer, the second
// GEMM's issuer and both epilogue groups all walk this one sequence; false = past the end.
template <int kCT, int kC2>
__device__ __forceinline__ bool chain_seq(int q, int my_tiles, int defer, int& it, int& c) {
  if (q < kCT) { it = 0; c = q; return my_tiles > 0; }
  const int q2 = q - kCT, j = q2 / (kCT + kC2), pos = q2 - j * (kCT + kC2);
  if (j + 1 >= my_tiles) { it = my_tiles - 1; c = kCT + pos; return pos < kC2; }        // behind the last tile: its chained chunks
  if (pos < defer) { it = j + 1; c = pos; }
  else if (pos < defer + kC2) { it = j; c = kCT + pos - defer; }
  else { it = j + 1; c = pos - kC2; }
  return true;
}

__device__ __forceinline__ TileCoord tile_coord(const ConvParams& p, int tile) {
  TileCoord t;
  t.n_tile = tile % p.n_tiles_n;  int r = tile / p.n_tiles_n;
  t.cls = r % p.n_classes;        r /= p.n_classes;
  t.x0 = (r % p.tiles_x) * p.tw;  r /= p.tiles_x;
  t.y0 = (r % p.tiles_y) * p.th;  r /= p.tiles_y;
  t.b0 = r * p.nb;
  return t;
}

// input-pixel offset of filter tap `t` (deconv: depends on the output parity class)
__device__ __forceinline__ void tap_offset(const ConvParams& p, int t, int cls, int& dx, int& dy) {
  if (p.kind == HPE_CONV_3x3)          { dy = t / 3 - 1; dx = t % 3 - 1; }
  else if (p.kind == HPE_CONV_STEM7x7) { dy = t - 2;     dx = 0; }
  else if (p.kind == HPE_CONV_DECONV4) {
    const int py = cls >> 1, px = cls & 1, ty = t >> 1, tx = t & 1;
    dy = ty ? (py ? 1 : -1) : 0;
    dx = tx ? (px ? 1 : -1) : 0;
  } else if (p.kind == HPE_CONV_DGRAD3x3S2) {           // class (py,px): 1 tap on an even axis, 2 on an odd one
    const int py = cls >> 1, px = cls & 1, nx = 1 + px, ty = t / nx, tx = t - ty * nx;
    dy = (py && ty) ? 1 : 0;
    dx = (px && tx) ? 1 : 0;
  } else if (p.kind == HPE_CONV_4x4S2) { dy = (t >> 2) - 1; dx = (t & 3) - 1; }
  else if (p.kind == HPE_CONV_2x2S2) { dy = t >> 1; dx = t & 1; }
  else { dx = 0; dy = 0; }
}
// number of 64-wide k-blocks of a tile (depends on the parity class only for the stride-2 3x3 dgrad)
__device__ __forceinline__ int kb_count(const ConvParams& p, int cls) {
  if (p.kind == HPE_CONV_DGRAD3x3S2) return (1 + (cls >> 1)) * (1 + (cls & 1)) * p.kc_per_tap;
  return p.taps * p.kc_per_tap + p.kc2;
}

// NSB = epilogue staging buffers (16 KB each): 4 with a residual (they double as its prefetch ring), else 3 -- the bytes
// saved go to a deeper operand ring, which is what bounds the deep-K layers (ring depth / hand-off round trip).
// EPI2 (two epilogue groups): NSB counts the buffers of BOTH groups and each group has its own scale/shift table.
// KBR (CHAIN kernels): the layer's whole weight tile (KBR k-blocks of BLOCK_N x 64) stays RESIDENT and the ring carries A only.
// NH (CHAIN kernels): 256-wide halves of the layer's output a CTA computes per pixel tile (2 = Cout 512); then the chained conv's
// [N2][256 NH] weights do not fit and their 64-wide K slices stream through a ring of kW2Slots buffers instead.
constexpr int kW2Slots = 3;
// AH (3x3 / stride 1, Cin = 128): the A operand is NOT ringed.  Per pixel tile (8 x 16 brick of one image) two halo'd boxes
// {64 ch, 10 px, 18 rows} (one per 64-channel block) land in a double-buffered region and every filter tap reads them through a
// shifted descriptor start (conv_halo.cu's trick); the ring carries the weight tiles only.
constexpr int kAHBoxBytes = (18 * 10 * 128 + 1023) / 1024 * 1024;      // one halo box, padded to the 1024-byte swizzle atom
template <int BLOCK_N, int STAGES, bool PAIR = false, int NSB = 4, bool HEAD2 = false, bool EPI2 = false, int CHAIN = 0, int KBR = 0, int NH = 1, bool AH = false>
struct SmemLayout {
  // fused head: this CTA's 16 rows x 256 K of the joint weights; CHAIN: the chained 1x1 conv's [N2 rows][256 K] weights
  static constexpr int kW2Bytes    = HEAD2 ? 4 * 16 * 128 : (NH == 1 ? 4 : kW2Slots) * CHAIN * 128;
  static constexpr int kBBytes     = (PAIR ? BLOCK_N / 2 : BLOCK_N) * kKBlock * 2;    // a CTA of a pair holds half of B's rows
  static constexpr int kStageBytes = (AH ? 0 : kABytes) + (KBR ? 0 : kBBytes);
  static constexpr int kOffA   = 0;
  static constexpr int kOffB   = AH ? 4 * kAHBoxBytes : STAGES * kABytes;
  static constexpr int kOffSB  = kOffB + (KBR ? KBR : STAGES) * kBBytes;
  static constexpr int kOffSS  = kOffSB + NSB * kSBBytes;             // float scale[kMaxBlockN], shift[kMaxBlockN] (x 2 groups with EPI2)
  static constexpr int kOffSS2 = kOffSS + ((EPI2 || NH >= 2) ? 4 : 2) * kMaxBlockN * 4;      // CHAIN: float scale2[128], shift2[128]
  static constexpr int kOffW2  = kOffSS2 + (CHAIN ? 1024 : 0);                  // 1024-byte aligned
  static constexpr int kOffBar = kOffW2 + kW2Bytes;
  static constexpr int kNumBar = 2 * STAGES + 4 + 4 * NSB + 7 + 2 * kW2Slots;       // + g2done[NSB], staged[NSB] (CHAIN) + w2full, s2full, d2full[2], a2empty[2], wres + w2 ring full/empty
  static constexpr int kTotal  = kOffBar + kNumBar * 8 + 16;
  static constexpr int kDynamic = kTotal;             // the dynamic segment starts 1024-byte aligned (checked in the kernel)
  static_assert(kTotal <= 232448, "exceeds the 227 KB of shared memory a CTA may use");
};

__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void sts_v4(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ uint32_t relu_bf16x2(uint32_t v) {
  const __nv_bfloat162 z = __floats2bfloat162_rn(0.f, 0.f);
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&v), z);
  return *reinterpret_cast<uint32_t*>(&r);
}

// work item -> this CTA's tile.  PAIR: work item = (n_tile, class, pair of M tiles); CTA `rank` of the 2-CTA cluster takes
// M tile 2*pair + rank.  An odd tile count leaves the last peer an M tile past the batch: its TMA loads zero-fill and its
// stores are clipped.
template <bool PAIR>
__device__ __forceinline__ int work_tile(const ConvParams& p, int work, int rank) {
  if (p.reverse) work = p.num_tiles - 1 - work;
  if (!PAIR) return work;
  const int per_m = p.n_tiles_n * p.n_classes;
  return ((work / per_m) * 2 + rank) * per_m + work % per_m;
}

// Exact-form GELU 0.5 x (1 + erf(x / sqrt 2)) with erf from Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7, far below the bf16
// rounding of the output): branch-free, 5 FMAs + one MUFU.RCP + one MUFU.EX2.  libdevice's erff evaluates two divergent
// branches per warp (~30 instructions per element), which made the 4C-wide ConvNeXt pwconv1 epilogue the layer's bottleneck.
__device__ __forceinline__ float gelu_erf(float x) {
  const float z = fabsf(x) * 0.70710678118654752f;
  float t, ex;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.f)));          // MUFU.RCP (1 ulp): ample here
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"(-1.4426950408889634f * z * z));      // MUFU.EX2
  float poly = fmaf(1.061405429f, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  const float e = poly * t * ex;                                        // 1 - erf(z), z >= 0
  const float half_erfc = 0.5f * e;                                    // Phi(-|x|)
  return x * (x >= 0.f ? 1.f - half_erfc : half_erfc);
}

// SiLU x * sigmoid(x) with MUFU.EX2 + MUFU.RCP (the IEEE division of `x / (1 + expf(-x))` costs ~10 instructions per element);
// both approximations are accurate to ~1 ulp of fp32, far below the bf16 rounding of the output
__device__ __forceinline__ float silu_fast(float x) {
  float ex, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"(-1.4426950408889634f * x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f + ex));
  return x * r;
}

// One epilogue warp's share of a staged chunk: 32 accumulator columns of tile row `row` -> scale/shift (+ residual read back from
// the chunk buffer) -> activation -> bf16 -> the chunk buffer's 128-byte-swizzled row.  (Same arithmetic as the generic epilogue
// below; used by the chained kernel's two-group epilogue.)
template <int ACT>
__device__ __forceinline__ void epi_chunk_math(const uint32_t (&v)[32], uint32_t ss, uint32_t sh, uint32_t rowp, int half, int row, bool add_res) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float4 s0 = lds_f4(ss + j * 32), s1 = lds_f4(ss + j * 32 + 16);
    const float4 h0 = lds_f4(sh + j * 32), h1 = lds_f4(sh + j * 32 + 16);
    float2 o01 = __ffma2_rn(make_float2(__uint_as_float(v[j * 8 + 0]), __uint_as_float(v[j * 8 + 1])), make_float2(s0.x, s0.y), make_float2(h0.x, h0.y));
    float2 o23 = __ffma2_rn(make_float2(__uint_as_float(v[j * 8 + 2]), __uint_as_float(v[j * 8 + 3])), make_float2(s0.z, s0.w), make_float2(h0.z, h0.w));
    float2 o45 = __ffma2_rn(make_float2(__uint_as_float(v[j * 8 + 4]), __uint_as_float(v[j * 8 + 5])), make_float2(s1.x, s1.y), make_float2(h1.x, h1.y));
    float2 o67 = __ffma2_rn(make_float2(__uint_as_float(v[j * 8 + 6]), __uint_as_float(v[j * 8 + 7])), make_float2(s1.z, s1.w), make_float2(h1.z, h1.w));
    const uint32_t slot = rowp + (((half * 4 + j) ^ (row & 7)) << 4);     // 128-byte swizzle
    if (add_res) {
      const uint4 rr = lds_v4(slot);
      o01 = __fadd2_rn(o01, make_float2(bf16_lo(rr.x), bf16_hi(rr.x))); o23 = __fadd2_rn(o23, make_float2(bf16_lo(rr.y), bf16_hi(rr.y)));
      o45 = __fadd2_rn(o45, make_float2(bf16_lo(rr.z), bf16_hi(rr.z))); o67 = __fadd2_rn(o67, make_float2(bf16_lo(rr.w), bf16_hi(rr.w)));
    }
    float o[8] = {o01.x, o01.y, o23.x, o23.y, o45.x, o45.y, o67.x, o67.y};
    if (ACT == HPE_ACT_SILU) {
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = silu_fast(o[e]);
    }
    uint4 pk = make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
    if (ACT == HPE_ACT_RELU) { pk.x = relu_bf16x2(pk.x); pk.y = relu_bf16x2(pk.y); pk.z = relu_bf16x2(pk.z); pk.w = relu_bf16x2(pk.w); }
    sts_v4(slot, pk);
  }
}

// ACT: HPE_ACT_*; RES: residual add; FINAL: fp32 NCHW heatmap epilogue (HPE_CONV_FINAL);
// PAIR: cta_group::2 -- a 2-CTA cluster computes a 256 x BLOCK_N tile with ONE tcgen05.mma stream issued by the leader:
// each CTA loads its own 128-pixel A brick and HALF of the weight tile, so a CTA ingests 32 KB instead of 48 KB per k-block
// (BLOCK_N = 256) through the L2->SM path that bounds the deep-K layers.  Each CTA drains its own 128 TMEM lanes.
// HEAD2 (pair kernels, Cout == BLOCK_N == 256): the layer's bf16 output tile is NOT stored; its four staged 64-channel chunks are
// exactly the SWIZZLE_128B K-major A operand of a second GEMM against the K-joint 1x1 weights (16 rows per CTA, resident in
// smem), accumulated into the first 32 columns of the accumulator just drained, and written as fp32 NCHW heatmaps: deconv3 +
// final_layer of the App. A.1 head in one kernel (saves writing and re-reading the 403 MB deconv3 output at B=256).
// EPI2: TWO independent epilogue groups of 8 warps (608 threads).  Group g drains accumulator g (the tiles with it & 1 == g) through
// its own staging buffers, residual ring, scale/shift table, named barrier and TMA-store stream, so the tcgen05.ld -> residual
// wait -> math -> fence -> barrier -> store chains of consecutive tiles overlap (in-kernel profile of the single lock-step group on
// the residual 1x1 layers: 8.2 k cycles per tile against 2 k of MMA with the schedulers 26 % busy -- every warp waited at the same
// points).  The group leader waits for ALL of its bulk stores to have left shared memory at the top of a chunk (wait_group.read 0),
// which makes 3 buffers per group enough for a residual lead of 2 chunks (2 buffers without a residual).
// CHAIN = N2 (64 | 128; 256-wide single-CTA tiles of a 1x1 conv with Cout = 256): the NEXT layer's 1x1 conv (256 -> N2, folded BN,
// same activation) runs on the output tile while it is staged for its own TMA store -- each 64-channel chunk the epilogue stages
// is at once the source of the bulk store of y AND a K = 64 slice of the SWIZZLE_128B A operand of a second tcgen05 GEMM against the
// resident [N2][256] weights, accumulated over the tile's four chunks into a separate TMEM accumulator and written as a second
// bf16 NHWC tensor.  conv3 (+ shortcut + ReLU) of a Bottleneck and conv1 of the next one in one kernel: the 403 MB block output
// of layer1 is written once and read once (as the next residual) instead of twice.  The main accumulator is single-buffered
// (256 + N2 TMEM columns); with K = 64..128 its 4-8 MMAs are a small bubble in an HBM-bound tile.
// KBR = the layer's k-blocks per tile (1 | 2, CHAIN only): its weight tile is loaded once and stays resident, the ring carries the A
// bricks only, and the shared memory saved goes to a deeper chunk-buffer ring (the chained kernel was bound by the residual ring:
// 4 buffers / (HBM latency + processing + drain) = one chunk per ~1.85 k cycles).
// NH = 2 (CHAIN = 128, Cout = 512: layer2): the CTA computes BOTH 256-wide halves of a pixel tile one after the other through the
// single main accumulator (weights streamed as usual), eight real chunks feed the chained GEMM (K = 512), whose weight slices
// [128][64] stream through a small ring loaded by the issuer warp itself.
template <int BLOCK_N, int STAGES, int ACT, bool RES, bool FINAL, bool PAIR = false, bool HEAD2 = false, bool EPI2 = false, int CHAIN = 0, int KBR = 0, int NH = 1, bool AH = false>
__global__ void __launch_bounds__(CHAIN ? kConvThreadsC : (EPI2 ? kConvThreads2 : kConvThreads), 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2,
                 const __grid_constant__ CUtensorMap tmW, const __grid_constant__ OutMaps tmOut,
                 const __grid_constant__ CUtensorMap tmRes, const ConvParams p) {
  // staging buffer q % kNumSB is rewritten by chunk q; the leader confirms "stores <= j-2 have left smem" at the start of
  // chunk j and everyone meets at chunk j's named barrier, so at the start of chunk q only stores <= q-3 are known done:
  // 3 buffers is the minimum.  With a residual the buffers double as its prefetch ring (4, lead 2).
  // EPI2 (leader waits for stores <= j-1): a non-leader thread at chunk q knows stores <= q-2 are done -> 2 buffers; with a
  // residual the buffer of chunk q-1 is handed to the producer at the top of chunk q -> 3 buffers give a lead of 2.
  static_assert(!HEAD2 || (PAIR && BLOCK_N == 256 && !RES && !FINAL), "fused head: pair kernel, 256 channels, no residual");
  static_assert(!EPI2 || (!FINAL && !HEAD2), "two epilogue groups: bf16 NHWC TMA-store path only");
  static_assert(!CHAIN || ((CHAIN == 64 || CHAIN == 128) && (BLOCK_N == 256 || BLOCK_N == 128) && !PAIR && !HEAD2 && !EPI2 && !FINAL), "chained 1x1: single-CTA tiles");
  // CHAIN with 128-wide tiles (Cout 512 as FOUR quarters per pixel tile): the main accumulator is double-buffered again (2 x 128 +
  // 2 x 128 chained columns), so the MMAs of quarter i+1 overlap the epilogue of quarter i, and a 32 KB ring stage is one k-block
  constexpr bool kAcc2x = CHAIN != 0 && BLOCK_N == 128;
  constexpr int kGroups = EPI2 ? 2 : 1;
  // per group; the fp32 heatmap epilogue stores from registers: all smem to the ring.  CHAIN = 128: its 64 KB of resident weights
  // leave room for 3 buffers only (residual lead 1)
  static_assert(!(CHAIN == 128 && NH == 1 && !RES), "chained 128-wide conv behind 256 channels: residual variant only (shared-memory budget)");
  static_assert((KBR != 0) == (CHAIN != 0 && NH == 1) && KBR <= 2, "resident weights: the 256-channel chained kernels (1-2 k-blocks per tile)");
  static_assert(NH == 1 || ((NH == 2 || NH == 4) && CHAIN == 128 && NH * BLOCK_N == 512), "several N tiles per pixel tile: the 512 -> 128 chain of layer2");
  // CHAIN: one ring for both groups, as deep as fits (the residual rides it with a lead of kNumSB - 2 chunks)
  // RING6: the 256-wide pair kernel without a residual with SIX ring stages -- the deep-K layers are bound by bytes in flight per
  // SM -- paid for by a 2-buffer staging ring under the `wait_group.read 0` protocol (the leader waits for the previous chunk's
  // store to have left shared memory; these layers' epilogues have slack)
  constexpr bool kRing6 = !RES && !HEAD2 && !EPI2 && !FINAL && CHAIN == 0 && STAGES == 6 && ((PAIR && BLOCK_N == 256) || (!PAIR && BLOCK_N == 128));
  constexpr int kNumSB = FINAL ? 0 : (CHAIN ? ((RES && CHAIN == 64) ? 6 : 4) : ((EPI2 || kRing6) ? (RES ? 3 : 2) : ((RES || HEAD2) ? 4 : 3)));
  constexpr int kResLead = 2;                     // the residual of chunk q + 2 is requested while chunk q is processed
  constexpr int kLead = CHAIN ? kNumSB - 2 : kResLead;
  constexpr int kChunks2 = CHAIN / 64;            // 64-column chunks of the chained conv's output
  static_assert(!AH || (BLOCK_N == 128 && !RES && !FINAL && !HEAD2 && !EPI2 && CHAIN == 0), "halo'd A operand: 128-wide tiles of a plain 3x3 (single CTA or cta_group::2)");
  using L = SmemLayout<BLOCK_N, STAGES, PAIR, kGroups * kNumSB, HEAD2, EPI2, CHAIN, KBR, NH, AH>;
  static_assert(!PAIR || (BLOCK_N % 32 == 0 && !FINAL), "pair kernels: N multiple of 32, bf16 NHWC epilogue");
  constexpr int kChunks   = BLOCK_N >= 64 ? BLOCK_N / 64 : 1;
  constexpr int kTmemCols = CHAIN ? 512 : 2 * BLOCK_N <= 32 ? 32 : (2 * BLOCK_N <= 64 ? 64 : (2 * BLOCK_N <= 128 ? 128 : (2 * BLOCK_N <= 256 ? 256 : 512)));  // power of two
  constexpr uint32_t kIdesc = PAIR ? make_idesc_bf16_m256(BLOCK_N) : make_idesc_bf16_m128(BLOCK_N);
  const int rank  = PAIR ? static_cast<int>(cluster_ctarank()) : 0;
  const int work0 = PAIR ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int wstep = PAIR ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if (smem_u32(smem) & 1023u) __trap();         // SWIZZLE_128B operand tiles need 1024-byte alignment
  uint8_t* sA  = smem + L::kOffA;
  uint8_t* sB  = smem + L::kOffB;
  uint8_t* sSB = smem + L::kOffSB;
  float*   sScale = reinterpret_cast<float*>(smem + L::kOffSS);
  float*   sShift = sScale + kMaxBlockN;
  uint64_t* bars       = reinterpret_cast<uint64_t*>(smem + L::kOffBar);
  uint64_t* full_bar   = bars;
  uint64_t* empty_bar  = bars + STAGES;
  uint64_t* tfull_bar  = bars + 2 * STAGES;
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;
  uint64_t* sbfull_bar = bars + 2 * STAGES + 4;
  uint64_t* sbempty_bar = bars + 2 * STAGES + 4 + kGroups * kNumSB;
  uint64_t* g2_bar     = bars + 2 * STAGES + 4 + 2 * kGroups * kNumSB;       // CHAIN: the second GEMM's MMAs on staging buffer b have completed
  uint64_t* staged_bar = g2_bar + kGroups * kNumSB;                // CHAIN: chunk buffer b is staged (its group's leader, after the group barrier)
  uint64_t* w2full_bar = staged_bar + kGroups * kNumSB;            // fused head / CHAIN: second-GEMM weights landed (own CTA)
  uint64_t* s2full_bar = w2full_bar + 1;                           // fused head: both CTAs staged their A' tile (leader's barrier)
  uint64_t* d2full_bar = w2full_bar + 2;                           // fused head: second GEMM done (multicast commit); CHAIN: [2], per chained accumulator
  uint64_t* a2empty_bar = w2full_bar + 4;                          // CHAIN: [2] chained accumulator drained
  uint64_t* wres_bar   = w2full_bar + 6;                           // KBR: the resident weight tile has landed
  uint64_t* w2rfull_bar = w2full_bar + 7;                          // NH = 2: [kW2Slots] a chained-weight K slice has landed
  uint64_t* w2rempty_bar = w2rfull_bar + kW2Slots;                 // NH = 2: [kW2Slots] ... has been consumed by its MMAs
  uint8_t*  sW2 = smem + L::kOffW2;
  float*    sScale2 = reinterpret_cast<float*>(smem + L::kOffSS2);   // CHAIN
  float*    sShift2 = sScale2 + 128;
  uint32_t* tmem_slot  = reinterpret_cast<uint32_t*>(bars + L::kNumBar);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA); prefetch_tmap(&tmW);
    if (p.kc2) prefetch_tmap(&tmA2);
    if (!FINAL) for (int c = 0; c < p.n_classes; ++c) prefetch_tmap(&tmOut.m[c]);
    if (RES) prefetch_tmap(&tmRes);
    // PAIR: the leader's full barrier collects both producers' arrivals, its tempty both CTAs' epilogue warps
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], PAIR ? 2 : 1); mbar_init(&empty_bar[s], 1); }
    // CHAIN: both epilogue groups drain the one main accumulator
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], (PAIR || CHAIN ? 2 : 1) * kEpiThreads / 32); }
    for (int b = 0; b < kGroups * kNumSB; ++b) { mbar_init(&sbfull_bar[b], 1); mbar_init(&sbempty_bar[b], 1); }
    for (int b = 0; b < kGroups * kNumSB; ++b) { mbar_init(&g2_bar[b], 1); mbar_init(&staged_bar[b], 1); }
    mbar_init(w2full_bar, 1); mbar_init(s2full_bar, 2); mbar_init(d2full_bar, 1); mbar_init(d2full_bar + 1, 1);
    mbar_init(&a2empty_bar[0], CHAIN ? 8 * kChunks2 : 1); mbar_init(&a2empty_bar[1], CHAIN ? 8 * kChunks2 : 1);
    mbar_init(wres_bar, 1);
    for (int b = 0; b < kW2Slots; ++b) { mbar_init(&w2rfull_bar[b], (AH && PAIR) ? 2 : 1); mbar_init(&w2rempty_bar[b], 1); }   // AH: the leader's barrier collects both CTAs' boxes
    fence_mbar_init();
  }
  if (warp == 1) {
    if constexpr (PAIR) { tmem_alloc_2sm(tmem_slot, kTmemCols); tmem_relinquish_2sm(); }
    else                { tmem_alloc(tmem_slot, kTmemCols); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (PAIR) cluster_sync_all();       // the peer's barriers are initialised before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // PDL: everything above (barrier init, TMEM alloc, descriptor prefetch, scale/shift staging -- all
  // independent of the previous layer's output) may overlap the previous kernel's tail.
  griddep_launch_dependents();
  griddep_wait();

  if (warp == 0) {
    // ===================== operand producer =====================
    // The WHOLE warp runs this loop converged and one elected lane issues (elect.sync inside the asm): every operand is
    // warp-uniform, so the barrier/TMA instructions take uniform registers directly.  Issued from a divergent
    // `if (lane == 0)` the same loop compiled to ~190 instructions per k-block (R2UR.BROADCAST + BRA.U.ANY re-broadcast loops
    // around each UTMALDG, an integer division and the tap switch) -- measured ~540 cycles per k-block with loads, MMAs and
    // epilogue all disabled, i.e. slower than the 512-cycle MMA work of a 128x256x64 block.
    int stage = 0; uint32_t phase = 0;
    if constexpr (AH) {
      // halo'd A: two boxes per tile into buffer it & 1 (barriers: w2rfull = landed, w2rempty = its MMAs have retired), then the
      // tile's 18 weight tiles through the ring
      // cta_group::2: each CTA loads the boxes of ITS M tile and half of every weight tile; all bytes are accounted on the
      // leader's barriers (the ones its MMA warp waits on), the peer adds a remote arrival
      int it = 0;
      for (int work = work0; work < p.num_tiles; work += wstep, ++it) {
        const TileCoord t = tile_coord(p, work_tile<PAIR>(p, work, rank));
        const int buf = it & 1;
        const int wrow = t.n_tile * BLOCK_N + (PAIR ? rank * (BLOCK_N / 2) : 0);
        mbar_wait(&w2rempty_bar[buf], ((it >> 1) & 1) ^ 1);
        if (lane == 0) {
          if constexpr (PAIR) {
            const uint32_t afull0 = mapa_shared(smem_u32(&w2rfull_bar[buf]), 0);
            if (rank == 0) mbar_arrive_expect_tx(&w2rfull_bar[buf], 2 * 2 * 18 * 10 * 128);
            else           mbar_arrive_cluster(afull0);
            tma_load_4d_2sm(sA + (buf * 2 + 0) * kAHBoxBytes, &tmA, afull0, 0, t.x0 - 1, t.y0 - 1, t.b0);
            tma_load_4d_2sm(sA + (buf * 2 + 1) * kAHBoxBytes, &tmA, afull0, 64, t.x0 - 1, t.y0 - 1, t.b0);
          } else {
            mbar_arrive_expect_tx(&w2rfull_bar[buf], 2 * 18 * 10 * 128);
            tma_load_4d(sA + (buf * 2 + 0) * kAHBoxBytes, &tmA, &w2rfull_bar[buf], 0, t.x0 - 1, t.y0 - 1, t.b0);
            tma_load_4d(sA + (buf * 2 + 1) * kAHBoxBytes, &tmA, &w2rfull_bar[buf], 64, t.x0 - 1, t.y0 - 1, t.b0);
          }
        }
        __syncwarp();
        for (int kb = 0; kb < 18; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          const uint32_t fb = smem_u32(&full_bar[stage]);
          if constexpr (PAIR) {
            const uint32_t full0 = mapa_shared(fb, 0);
            if (rank == 0) mbar_arrive_expect_tx_elect(fb, 2 * L::kBBytes);
            else           mbar_arrive_cluster_elect(full0);
            tma_load_b_2sm_elect(full0, smem_u32(sB + stage * L::kBBytes), &tmW, kb * kKBlock, wrow);
          } else {
            tma_load_b_elect(fb, L::kBBytes, smem_u32(sB + stage * L::kBBytes), &tmW, kb * kKBlock, wrow);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    } else {
    if constexpr (KBR != 0) {                     // the whole [BLOCK_N][64 KBR] weight tile, once
      if (lane == 0) {
        mbar_arrive_expect_tx(wres_bar, KBR * L::kBBytes);
        for (int kb = 0; kb < KBR; ++kb) tma_load_2d(sB + kb * L::kBBytes, &tmW, wres_bar, kb * kKBlock, 0);
      }
      __syncwarp();
    }
    for (int work = work0; work < p.num_tiles; work += wstep) {
      const TileCoord t = tile_coord(p, work_tile<PAIR>(p, work, rank));
     for (int nh = 0; nh < NH; ++nh) {              // NH = 2: both 256-wide halves of the pixel tile (A is loaded again for the second)
      const int wrow = t.cls * p.Cout + (t.n_tile * NH + nh) * BLOCK_N + (PAIR ? rank * (BLOCK_N / 2) : 0);
      const int num_kb = kb_count(p, t.cls), num_kb1 = num_kb - p.kc2;
      const int bx = t.x0 * p.in_stride, by = t.y0 * p.in_stride;
      int tap = 0, kc = 0, dx, dy;
      tap_offset(p, 0, t.cls, dx, dy);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);          // own barrier: the multicast commit frees the stage in both CTAs
        const uint32_t dA = smem_u32(sA + stage * kABytes), dB = smem_u32(sB + stage * L::kBBytes);
        const uint32_t fb = smem_u32(&full_bar[stage]);
        const CUtensorMap* mA = &tmA;
        int c0 = kc * kKBlock, cx = bx + dx, cy = by + dy;
        if (kb >= num_kb1) {       // second source (1x1, own stride): the block input feeding the projection shortcut
          mA = &tmA2; c0 = (kb - num_kb1) * kKBlock; cx = t.x0 * p.in_stride2; cy = t.y0 * p.in_stride2;
        }
        if constexpr (PAIR) {
          // both CTAs' bytes are accounted on the LEADER's barrier (the one its MMA warp waits on)
          const uint32_t full0 = mapa_shared(fb, 0);
          if (rank == 0) mbar_arrive_expect_tx_elect(fb, 2 * L::kStageBytes);
          else           mbar_arrive_cluster_elect(full0);
          tma_load_ab_2sm_elect(full0, dA, mA, c0, cx, cy, t.b0, dB, &tmW, kb * kKBlock, wrow, p.l2_in);
        } else if constexpr (KBR != 0) {
          mbar_arrive_expect_tx_elect(fb, kABytes);
          tma_load_4d_elect(dA, mA, fb, c0, cx, cy, t.b0, p.l2_in);
        } else {
          tma_load_ab_elect(fb, L::kStageBytes, dA, mA, c0, cx, cy, t.b0, dB, &tmW, kb * kKBlock, wrow, p.l2_in);
        }
        if (++kc == p.kc_per_tap) { kc = 0; ++tap; tap_offset(p, tap, t.cls, dx, dy); }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
     }
    }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // whole warp converged, one elected lane issues; descriptors as 32-bit halves (constant high word, low word = stage
    // base + immediate), see umma_bf16_elect_lo.  PAIR: only the leader CTA issues (for both CTAs' TMEM).
    if (rank == 0) {
      constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);     // SBO = 1024 B, version 1, SWIZZLE_128B
      int stage = 0; uint32_t phase = 0; int it = 0;
      if constexpr (KBR != 0) mbar_wait(wres_bar, 0);
      if constexpr (AH) {
        // k-block kb = tap * 2 + channel block: the A rows of tap (ky, kx) are the box rows (y + ky) * 10 + x + kx, i.e. the box
        // read from row offset ky * 10 + kx with a 10-row (1280-byte) stride between the 8-row groups
        constexpr uint32_t kDescHiA = (1280u >> 4) | (1u << 14) | (2u << 29);
        for (int work = work0; work < p.num_tiles; work += wstep, ++it) {
          const int acc = it & 1, buf = it & 1;
          mbar_wait(&tempty_bar[acc], ((it >> 1) & 1) ^ 1);
          mbar_wait(&w2rfull_bar[buf], (it >> 1) & 1);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
          for (int kb = 0; kb < 18; ++kb) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const int tap = kb >> 1, ky = tap / 3, kx = tap - 3 * ky;
            const uint32_t a_addr = smem_u32(sA + (buf * 2 + (kb & 1)) * kAHBoxBytes) + (ky * 10 + kx) * 128;
            const uint32_t a_lo = ((a_addr & 0x3FFFFu) >> 4) | (1u << 16);
            const uint32_t b_lo = ((smem_u32(sB + stage * L::kBBytes) & 0x3FFFFu) >> 4) | (1u << 16);
#pragma unroll
            for (int k = 0; k < kKBlock / 16; ++k) {
              if constexpr (PAIR) umma_bf16_2sm_elect_lo(tmem_d, a_lo + 2 * k, kDescHiA, b_lo + 2 * k, kDescHi, kIdesc, (kb | k) != 0);
              else                umma_bf16_elect_lo(tmem_d, a_lo + 2 * k, kDescHiA, b_lo + 2 * k, kDescHi, kIdesc, (kb | k) != 0);
            }
            if constexpr (PAIR) umma_commit_2sm_elect(&empty_bar[stage], 3); else umma_commit_elect(&empty_bar[stage]);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          // the boxes (of both CTAs) may be overwritten once these MMAs retire; the accumulator is ready
          if constexpr (PAIR) { umma_commit_2sm_elect(&w2rempty_bar[buf], 3); umma_commit_2sm_elect(&tfull_bar[acc], 3); }
          else                { umma_commit_elect(&w2rempty_bar[buf]); umma_commit_elect(&tfull_bar[acc]); }
        }
      } else
      for (int work = work0; work < p.num_tiles; work += wstep) for (int nh = 0; nh < NH; ++nh, ++it) {     // it counts (tile, half)
        const int acc = (CHAIN && !kAcc2x) ? 0 : (it & 1);     // CHAIN (256-wide): one main accumulator (columns 256.. hold the chained conv's)
        mbar_wait(&tempty_bar[acc], ((CHAIN && !kAcc2x) ? (it & 1) : ((it >> 1) & 1)) ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
        const int num_kb = kb_count(p, tile_coord(p, work_tile<PAIR>(p, work, 0)).cls);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_lo = ((smem_u32(sA + stage * kABytes) & 0x3FFFFu) >> 4) | (1u << 16);
          const uint32_t b_lo = ((smem_u32(sB + (KBR ? kb : stage) * L::kBBytes) & 0x3FFFFu) >> 4) | (1u << 16);
#pragma unroll
          for (int k = 0; k < kKBlock / 16; ++k) {    // UMMA_K = 16 bf16 = 32 B -> +2 in the (addr>>4) field
            if constexpr (PAIR) umma_bf16_2sm_elect_lo(tmem_d, a_lo + 2 * k, kDescHi, b_lo + 2 * k, kDescHi, kIdesc, (kb | k) != 0);
            else                umma_bf16_elect_lo(tmem_d, a_lo + 2 * k, kDescHi, b_lo + 2 * k, kDescHi, kIdesc, (kb | k) != 0);
          }
          // frees the smem stage (in both CTAs of a pair) when these MMAs retire
          if constexpr (PAIR) umma_commit_2sm_elect(&empty_bar[stage], 3); else umma_commit_elect(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        // accumulator ready for the epilogue (of both CTAs)
        if constexpr (PAIR) umma_commit_2sm_elect(&tfull_bar[acc], 3); else umma_commit_elect(&tfull_bar[acc]);
      }
    }
  } else if (warp == 2) {
    // ===================== residual producer =====================
    if (HEAD2 && lane == 0) {        // fused head: this CTA's 16 joint-weight rows (4 k-blocks of 16 x 128 B), loaded once
      mbar_arrive_expect_tx(w2full_bar, L::kW2Bytes);
      for (int kb = 0; kb < 4; ++kb) tma_load_2d(sW2 + kb * 2048, &tmRes, w2full_bar, kb * kKBlock, rank * 16);
    }
    if (CHAIN && NH == 1 && lane == 0) {        // chained 1x1: [N2 rows][256 K] as 4 k-blocks of N2 x 128 B, loaded once (map in tmOut.m[2])
      mbar_arrive_expect_tx(w2full_bar, L::kW2Bytes);
      for (int kb = 0; kb < 4; ++kb) tma_load_2d(sW2 + kb * (CHAIN * 128), &tmOut.m[2], w2full_bar, kb * kKBlock, 0);
    }
    auto res_load = [&](void* dst, uint64_t* bar, int c0, int c1, int c2, int c3) {
      if (p.l2_res) tma_load_4d_hint(dst, &tmRes, bar, c0, c1, c2, c3, p.l2_res); else tma_load_4d(dst, &tmRes, bar, c0, c1, c2, c3);
    };
    if (RES && lane == 0) {
      if constexpr (EPI2) {
        // two rings (one per epilogue group, tiles it & 1 == g): one thread serves both, polling so that neither ring waits
        // behind the other's full buffers
        int wk[2] = {work0, work0 + wstep}, ch[2] = {0, 0}, rr[2] = {0, 0};
        TileCoord tg[2];
#pragma unroll
        for (int g = 0; g < 2; ++g) if (wk[g] < p.num_tiles) tg[g] = tile_coord(p, work_tile<PAIR>(p, wk[g], rank));
        while (wk[0] < p.num_tiles || wk[1] < p.num_tiles) {
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            if (wk[g] >= p.num_tiles) continue;
            const int b = g * kNumSB + rr[g] % kNumSB;
            if (!mbar_test_wait(&sbempty_bar[b], (rr[g] / kNumSB) & 1)) continue;
            mbar_arrive_expect_tx(&sbfull_bar[b], kSBBytes);
            res_load(sSB + b * kSBBytes, &sbfull_bar[b], tg[g].n_tile * BLOCK_N + ch[g] * 64, tg[g].x0, tg[g].y0, tg[g].b0);
            ++rr[g];
            if (++ch[g] == kChunks) {
              ch[g] = 0; wk[g] += 2 * wstep;
              if (wk[g] < p.num_tiles) tg[g] = tile_coord(p, work_tile<PAIR>(p, wk[g], rank));
            }
          }
        }
      } else if constexpr (CHAIN != 0) {
        // one ring in the order of chain_seq: the chained conv's output chunks take a buffer too, but carry no residual
        const int my_tiles = work0 < p.num_tiles ? (p.num_tiles - 1 - work0) / wstep + 1 : 0;
        TileCoord t = {};
        int it = 0, c = 0, it_cur = -1;
        for (int r = 0; chain_seq<NH * kChunks, kChunks2>(r, my_tiles, p.chain_defer, it, c); ++r) {
          const int b = r % kNumSB;
          mbar_wait(&sbempty_bar[b], (r / kNumSB) & 1);     // the epilogue leaders hand buffers over with a lead of kLead chunks
          if (c >= NH * kChunks) { mbar_arrive(&sbfull_bar[b]); continue; }
          if (it != it_cur) {
            it_cur = it; t = tile_coord(p, work_tile<false>(p, work0 + it * wstep, 0));
            if (p.res_prefetch && it + 1 < my_tiles) {       // (experiment, off) the next tile's residual boxes into L2 a tile ahead
              const TileCoord tn = tile_coord(p, work_tile<false>(p, work0 + (it + 1) * wstep, 0));
              for (int cn = 0; cn < NH * kChunks; ++cn) tma_prefetch_l2_4d(&tmRes, tn.n_tile * NH * BLOCK_N + cn * 64, tn.x0, tn.y0, tn.b0);
            }
          }
          mbar_arrive_expect_tx(&sbfull_bar[b], kSBBytes);
          res_load(sSB + b * kSBBytes, &sbfull_bar[b], t.n_tile * NH * BLOCK_N + c * 64, t.x0, t.y0, t.b0);
        }
      } else {
        int r = 0;
        for (int work = work0; work < p.num_tiles; work += wstep) {
          const TileCoord t = tile_coord(p, work_tile<PAIR>(p, work, rank));
          if (p.res_prefetch && work + wstep < p.num_tiles) {
            // the ring's lead (2-4 chunks) is shorter than the HBM latency once two epilogue groups drain it: pull the next tile's
            // residual into L2 a whole tile ahead, so the ring's own loads are L2 hits
            const TileCoord tn = tile_coord(p, work_tile<PAIR>(p, work + wstep, rank));
            for (int c = 0; c < kChunks; ++c) tma_prefetch_l2_4d(&tmRes, tn.n_tile * BLOCK_N + c * 64, tn.x0, tn.y0, tn.b0);
          }
          for (int c = 0; c < kChunks; ++c, ++r) {
            const int b = r % kNumSB;
            mbar_wait(&sbempty_bar[b], (r / kNumSB) & 1);   // the epilogue leader hands buffers over with a lead of kResLead chunks
            mbar_arrive_expect_tx(&sbfull_bar[b], kSBBytes);
            res_load(sSB + b * kSBBytes, &sbfull_bar[b], t.n_tile * BLOCK_N + c * 64, t.x0, t.y0, t.b0);
          }
        }
      }
    }
  } else if (CHAIN != 0 && warp == 3) {
    // ===================== CHAIN: issuer of the second GEMM =====================
    // One thread issues every chunk's K = 64 slice in chunk order (so the accumulation order over the 256 input channels is
    // the separate conv1 launch's), as soon as the chunk's group has staged it.
    if constexpr (CHAIN != 0) {
      if (lane == 0) {
        constexpr int kCT = NH * kChunks;                        // real chunks per tile (4 of y per 256-wide half)
        constexpr uint32_t kIdesc2 = make_idesc_bf16_m128(CHAIN);
        constexpr int kSlice = CHAIN * 128;                      // bytes of one [N2][64] K slice of the chained weights
        if constexpr (NH == 1) mbar_wait(w2full_bar, 0);
        static_assert(NH == 1 || BLOCK_N * NH == 512, "");
        int q = 0, it = 0;
        // NH = 2: the K slices stream through kW2Slots buffers, requested kW2Slots - 1 real chunks ahead
        int w_req = 0, w_use = 0;                                // slices requested / consumed so far (over the whole kernel)
        const int my_tiles = work0 < p.num_tiles ? (p.num_tiles - 1 - work0) / wstep + 1 : 0;
        const int w_total = my_tiles * kCT;
        auto request = [&]() {
          const int slot = w_req % kW2Slots;
          mbar_wait(&w2rempty_bar[slot], ((w_req / kW2Slots) & 1) ^ 1);
          mbar_arrive_expect_tx(&w2rfull_bar[slot], kSlice);
          tma_load_2d(sW2 + slot * kSlice, &tmOut.m[2], &w2rfull_bar[slot], (w_req % kCT) * kKBlock, 0);
          ++w_req;
        };
        if constexpr (NH >= 2) { while (w_req < kW2Slots - 1 && w_req < w_total) request(); }
        int c = 0;
        for (; chain_seq<kCT, kChunks2>(q, my_tiles, p.chain_defer, it, c); ++q) {
          const int sb = q % kNumSB, a2 = it & 1;
          if (c == 0) mbar_wait(&a2empty_bar[a2], ((it >> 1) & 1) ^ 1);      // tile it-2's chained output has left accumulator a2
          if (NH >= 2 && c < kCT && w_req < w_total) request();
          mbar_wait(&staged_bar[sb], (q / kNumSB) & 1);
          if (c < kCT) {
            const uint8_t* wsl = sW2 + c * kSlice;
            if constexpr (NH >= 2) {
              const int slot = w_use % kW2Slots;
              mbar_wait(&w2rfull_bar[slot], (w_use / kW2Slots) & 1);
              wsl = sW2 + slot * kSlice;
            }
            tc_fence_after();
            const uint64_t adesc = make_sw128_kmajor_desc(sSB + sb * kSBBytes);
            const uint64_t bdesc = make_sw128_kmajor_desc(wsl);
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(tmem_base + 256 + a2 * CHAIN, adesc + 2 * k, bdesc + 2 * k, kIdesc2, (c | k) != 0);
            if constexpr (NH >= 2) { umma_commit(&w2rempty_bar[w_use % kW2Slots]); ++w_use; }
          }
          umma_commit(&g2_bar[sb]);                              // every chunk commits: a buffer's phase sequence stays uniform
          if (c == kCT - 1) umma_commit(&d2full_bar[a2]);
        }
      }
    }
  } else if (CHAIN != 0) {
    // ===================== CHAIN: two epilogue groups, chunk-interleaved =====================
    // Entry q of this CTA's chunk sequence (chain_seq) goes to group q & 1 (warps 4-11 / 12-19): both groups drain the SAME
    // tile, so two tcgen05.ld -> residual wait -> math -> fence -> barrier -> store chains are in flight (a single group ran the
    // chained kernel at 11.6 k cycles per tile against ~7.9 k of HBM time).  The buffers form ONE ring in chunk order; with an
    // even buffer count, or the residual hand-over two chunks back, a buffer is always recycled by the group that stored it.
    if constexpr (CHAIN != 0) {
      constexpr int kCT = NH * kChunks;
      const int grp = (warp - 4) >> 3, ew = warp & 3, half = ((warp - 4) >> 2) & 1, row = ew * 32 + lane;
      const int atid = threadIdx.x - 128, etid = atid - grp * kEpiThreads;
      const bool leader = etid == 0;
      const uint32_t bar_id = 1 + grp;
      float* const scl = sScale; float* const shf = sScale + NH * BLOCK_N;      // (NH = 2: a 4 KB table region)
      for (int n = atid; n < NH * BLOCK_N; n += 2 * kEpiThreads) { scl[n] = n < p.Cout ? __ldg(p.scale + n) : 0.f; shf[n] = n < p.Cout ? __ldg(p.shift + n) : 0.f; }
      if (atid < CHAIN) { sScale2[atid] = __ldg(p.scale2 + atid); sShift2[atid] = __ldg(p.shift2 + atid); }
      named_bar_sync(3, 2 * kEpiThreads);
      if (RES && atid == 0) for (int b = 0; b < kLead; ++b) mbar_arrive(&sbempty_bar[b]);
      const uint32_t taddr0 = tmem_base + (static_cast<uint32_t>(ew * 32) << 16);
      const int my_tiles = work0 < p.num_tiles ? (p.num_tiles - 1 - work0) / wstep + 1 : 0;
      TileCoord tce = {}, tco = {};                              // coordinates of the even / odd tile in flight (a deferred chained chunk
      int ite = -1, ito = -1, it = 0, c = 0;                     // belongs to the tile BEFORE the one whose chunks surround it)
      for (int q = grp; chain_seq<kCT, kChunks2>(q, my_tiles, p.chain_defer, it, c); q += 2) {
        if (it & 1) { if (ito != it) { tco = tile_coord(p, work_tile<false>(p, work0 + it * wstep, 0)); ito = it; } }
        else        { if (ite != it) { tce = tile_coord(p, work_tile<false>(p, work0 + it * wstep, 0)); ite = it; } }
        const TileCoord t = (it & 1) ? tco : tce;
        const bool e2 = c >= kCT;
        const int sb = q % kNumSB, a2 = it & 1;
        const int nh = e2 ? 0 : c / kChunks, cc = e2 ? c - kCT : c - nh * kChunks;     // 256-wide half, chunk inside it
        if (leader) {
          bulk_wait_read<0>();                                   // this group's stores (chunks q-2, q-4, ...) have left shared memory
          if (q >= 2) mbar_wait(&g2_bar[(q - 2) % kNumSB], ((q - 2) / kNumSB) & 1);     // ... and the second GEMM has read chunk q-2
          if (RES) mbar_arrive(&sbempty_bar[(q + kLead) % kNumSB]);                     // = chunk q-2's buffer: to the residual producer
        }
        const int ih = it * NH + nh, a1 = kAcc2x ? (ih & 1) : 0;              // main-accumulator use number / buffer
        if (e2) mbar_wait(&d2full_bar[a2], (it >> 1) & 1); else mbar_wait(&tfull_bar[a1], kAcc2x ? ((ih >> 1) & 1) : (ih & 1));
        tc_fence_after();
        uint32_t v[32];
        tmem_ld_x32(e2 ? taddr0 + 256 + a2 * CHAIN + cc * 64 + half * 32 : taddr0 + a1 * BLOCK_N + cc * 64 + half * 32, v);
        if (RES) mbar_wait(&sbfull_bar[sb], (q / kNumSB) & 1);
        tmem_ld_wait();
        if (e2 || cc + 2 >= kChunks) {                           // this group's last read of the main (half) / of the chained accumulator
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(e2 ? &a2empty_bar[a2] : &tempty_bar[a1]);
        }
        const int n0 = (e2 ? cc : c) * 64 + half * 32;
        epi_chunk_math<ACT>(v, smem_u32((e2 ? sScale2 : scl) + n0), smem_u32((e2 ? sShift2 : shf) + n0),
                            smem_u32(sSB + sb * kSBBytes) + row * 128, half, row, RES && !e2);
        fence_proxy_async_smem();                                // generic-proxy writes -> visible to TMA and to the tensor core
        named_bar_sync(bar_id, kEpiThreads);
        if (leader) {
          if (e2)            tma_store_4d(&tmOut.m[1], sSB + sb * kSBBytes, cc * 64, t.x0, t.y0, t.b0);
          else if (p.l2_out) tma_store_4d_hint(&tmOut.m[0], sSB + sb * kSBBytes, c * 64, t.x0, t.y0, t.b0, p.l2_out);
          else               tma_store_4d(&tmOut.m[0], sSB + sb * kSBBytes, c * 64, t.x0, t.y0, t.b0);
          bulk_commit();
          mbar_arrive(&staged_bar[sb]);
        }
      }
      if (leader) bulk_wait_all();
    }
  } else {
    // ===================== epilogue (8 warps per group: TMEM lane quarter = warp & 3, column half = ((warp - 3) >> 2) & 1) ====
    const int grp = EPI2 ? (warp - kEpiWarp0) >> 3 : 0;          // EPI2: group g drains accumulator g (tiles it & 1 == g)
    const int ew = warp & 3;                      // TMEM lane quarter this warp may access
    const int half = ((warp - kEpiWarp0) >> 2) & 1;     // which 32 of the chunk's 64 columns
    const int row = ew * 32 + lane;               // tile row = TMEM lane = pixel inside the brick
    const int etid = threadIdx.x - kEpiWarp0 * 32 - grp * kEpiThreads;
    const bool leader = etid == 0;
    // per-group resources
    sSB += grp * kNumSB * kSBBytes; sbfull_bar += grp * kNumSB; sbempty_bar += grp * kNumSB;
    sScale += grp * 2 * kMaxBlockN; sShift += grp * 2 * kMaxBlockN;
    const uint32_t bar_id = 1 + grp;
    if (RES && leader) for (int b = 0; b < kLead; ++b) mbar_arrive(&sbempty_bar[b]);
    int q = 0, staged_n = -1;
    // HPE_ACT_GELU_LN: this row's LayerNorm partial sums of the tile `w` (first kLnPf slices into registers; ln_st = null: no pixel)
    constexpr int kLnPf = 8;
    float2 ln_pf[kLnPf];
    const float2* ln_st = nullptr;
    auto ln_fetch = [&](int w) {
      ln_st = nullptr;
#pragma unroll
      for (int i = 0; i < kLnPf; ++i) ln_pf[i] = make_float2(0.f, 0.f);
      if (w >= p.num_tiles) return;
      const TileCoord tn = tile_coord(p, work_tile<PAIR>(p, w, rank));
      const int xi = row % p.tw, yi = (row / p.tw) % p.th, bi = row / (p.tw * p.th);
      const int x = tn.x0 + xi, y = tn.y0 + yi, b = tn.b0 + bi;
      if (x < p.Wo && y < p.Ho && b < p.B) {
        ln_st = p.rowstats + (static_cast<size_t>(b * p.Ho + y) * p.Wo + x) * p.ln_slices;
#pragma unroll
        for (int i = 0; i < kLnPf; ++i) if (i < p.ln_slices) ln_pf[i] = __ldg(ln_st + i);
      }
    };
    if constexpr (ACT == HPE_ACT_GELU_LN) ln_fetch(work0 + grp * wstep);
    for (int work = work0 + grp * wstep, it = grp; work < p.num_tiles; work += kGroups * wstep, it += kGroups) {
      const TileCoord t = tile_coord(p, work_tile<PAIR>(p, work, rank));
      const int acc = it & 1;
      if (t.n_tile != staged_n) {
        // this N tile's folded-BN scale/shift -> smem (channels past Cout: 0).  Every thread of the group finished reading the
        // previous tile's values before the named barrier of its last chunk, so the table is free to overwrite here.
        if (etid < BLOCK_N) {
          const int n = t.n_tile * BLOCK_N + etid;
          sScale[etid] = n < p.Cout ? __ldg(p.scale + n) : 0.f; sShift[etid] = n < p.Cout ? __ldg(p.shift + n) : 0.f;
        }
        staged_n = t.n_tile;
        named_bar_sync(bar_id, kEpiThreads);
      }
      // HPE_ACT_GELU_LN: LayerNorm over the INPUT channels folded into this 1x1 conv.  With W' = gamma * W the layer computes
      //   gelu( LN(x) W^T + b ) = gelu( rstd_r * (x W'^T - mu_r * colsum(W')) + (b + W beta) ),
      // so the A operand is the raw tensor and the epilogue applies a per-ROW (mu, rstd) beside the per-column tables
      // (scale = colsum(W'), shift = b + W beta).  The row's statistics come from the producer's partial sums.
      // The partial sums of a tile are requested one tile ahead (ln_fetch below) and consumed here, so their latency hides
      // behind a whole tile of epilogue work (fetched at the point of use they cost the single-group 192-wide kernel +33 %).
      float2 ln_nmu = make_float2(0.f, 0.f), ln_rstd = make_float2(1.f, 1.f);
      if constexpr (ACT == HPE_ACT_GELU_LN) {
        if (ln_st != nullptr) {
          float s = 0.f, q = 0.f;
#pragma unroll
          for (int i = 0; i < kLnPf; ++i) { s += ln_pf[i].x; q += ln_pf[i].y; }
          for (int i = kLnPf; i < p.ln_slices; ++i) { const float2 v2 = __ldg(ln_st + i); s += v2.x; q += v2.y; }
          const float mu = s * p.ln_inv_c, rs = rsqrtf(fmaxf(fmaf(-mu, mu, q * p.ln_inv_c), 0.f) + p.ln_eps);
          ln_nmu = make_float2(-mu, -mu); ln_rstd = make_float2(rs, rs);
        }
        ln_fetch(work + kGroups * wstep);
      }
      mbar_wait(&tfull_bar[acc], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + acc * BLOCK_N;

      if constexpr (FINAL) {
        // fp32 NCHW heatmaps straight from registers: for a fixed joint the warp's 32 rows are
        // consecutive pixels of the brick -> 64-byte (tw=16) contiguous runs.  Column groups of 32
        // alternate between the two warps of a lane quarter.
        const int xi = row % p.tw, yi = (row / p.tw) % p.th, bi = row / (p.tw * p.th);
        const int x = t.x0 + xi, y = t.y0 + yi, b = t.b0 + bi;
        const bool ok = x < p.Wo && y < p.Ho && b < p.B;
        const size_t plane = static_cast<size_t>(p.Ho) * p.Wo;
        for (int g = half; g * 32 < p.Cout; g += 2) {
          uint32_t v[32];
          tmem_ld_x32(taddr + g * 32, v);
          tmem_ld_wait();
          if (ok) {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const int k = g * 32 + e;
              if (k < p.Cout) {
                const float o = fmaf(__uint_as_float(v[e]), sScale[k], sShift[k]);
                p.hm_out[(static_cast<size_t>(b) * p.cout_total + k) * plane + static_cast<size_t>(y) * p.Wo + x] = o;
              }
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      } else {
        for (int c = 0; c < kChunks; ++c, ++q) {
          const int sb = HEAD2 ? c : q % kNumSB;                    // fused head: the tile's four chunks stay staged (A' operand)
          const uint32_t stg = smem_u32(sSB + sb * kSBBytes);
          if (!HEAD2 && leader) {
            bulk_wait_read<(EPI2 || kRing6) ? 0 : 1>();             // stores <= q-2 (EPI2 / RING6: <= q-1) have left shared memory
            if (RES) mbar_arrive(&sbempty_bar[(q + kLead) % kNumSB]);   // hand the buffer of chunk q + lead to warp 2
          }
          uint32_t v[32];
          tmem_ld_x32(taddr + c * 64 + half * 32, v);
          if (RES) mbar_wait(&sbfull_bar[sb], (q / kNumSB) & 1);
          tmem_ld_wait();
          if (!HEAD2 && c == kChunks - 1) {                         // accumulator drained -> MMA may reuse it
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if constexpr (PAIR) mbar_arrive_cluster(mapa_shared(smem_u32(&tempty_bar[acc]), 0));   // the leader's barrier
              else                mbar_arrive(&tempty_bar[acc]);
            }
          }
          const int n0 = c * 64 + half * 32;                        // first tile-local channel of this warp's 32 columns
          const uint32_t ss = smem_u32(sScale + n0), sh = smem_u32(sShift + n0);
          const uint32_t rowp = stg + row * 128;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 s0 = lds_f4(ss + j * 32), s1 = lds_f4(ss + j * 32 + 16);
            const float4 h0 = lds_f4(sh + j * 32), h1 = lds_f4(sh + j * 32 + 16);
            // packed fp32x2 arithmetic (FFMA2 / FADD2: two independent IEEE operations per instruction, same bits as the scalar
            // form) -- the epilogue is issue-bound (two epilogue warps per scheduler, in-kernel profile in profiles/)
            float2 o01 = __ffma2_rn(make_float2(__uint_as_float(v[j * 8 + 0]), __uint_as_float(v[j * 8 + 1])), make_float2(s0.x, s0.y), make_float2(h0.x, h0.y));
            float2 o23 = __ffma2_rn(make_float2(__uint_as_float(v[j * 8 + 2]), __uint_as_float(v[j * 8 + 3])), make_float2(s0.z, s0.w), make_float2(h0.z, h0.w));
            float2 o45 = __ffma2_rn(make_float2(__uint_as_float(v[j * 8 + 4]), __uint_as_float(v[j * 8 + 5])), make_float2(s1.x, s1.y), make_float2(h1.x, h1.y));
            float2 o67 = __ffma2_rn(make_float2(__uint_as_float(v[j * 8 + 6]), __uint_as_float(v[j * 8 + 7])), make_float2(s1.z, s1.w), make_float2(h1.z, h1.w));
            if constexpr (ACT == HPE_ACT_GELU_LN) {     // rstd * (acc - mu * colsum) + shift
              o01 = __ffma2_rn(__ffma2_rn(make_float2(s0.x, s0.y), ln_nmu, make_float2(__uint_as_float(v[j * 8 + 0]), __uint_as_float(v[j * 8 + 1]))), ln_rstd, make_float2(h0.x, h0.y));
              o23 = __ffma2_rn(__ffma2_rn(make_float2(s0.z, s0.w), ln_nmu, make_float2(__uint_as_float(v[j * 8 + 2]), __uint_as_float(v[j * 8 + 3]))), ln_rstd, make_float2(h0.z, h0.w));
              o45 = __ffma2_rn(__ffma2_rn(make_float2(s1.x, s1.y), ln_nmu, make_float2(__uint_as_float(v[j * 8 + 4]), __uint_as_float(v[j * 8 + 5]))), ln_rstd, make_float2(h1.x, h1.y));
              o67 = __ffma2_rn(__ffma2_rn(make_float2(s1.z, s1.w), ln_nmu, make_float2(__uint_as_float(v[j * 8 + 6]), __uint_as_float(v[j * 8 + 7]))), ln_rstd, make_float2(h1.z, h1.w));
            }
            const uint32_t slot = rowp + (((half * 4 + j) ^ (row & 7)) << 4);     // 128-byte swizzle
            if (RES) {
              const uint4 rr = lds_v4(slot);
              o01 = __fadd2_rn(o01, make_float2(bf16_lo(rr.x), bf16_hi(rr.x))); o23 = __fadd2_rn(o23, make_float2(bf16_lo(rr.y), bf16_hi(rr.y)));
              o45 = __fadd2_rn(o45, make_float2(bf16_lo(rr.z), bf16_hi(rr.z))); o67 = __fadd2_rn(o67, make_float2(bf16_lo(rr.w), bf16_hi(rr.w)));
            }
            float o[8] = {o01.x, o01.y, o23.x, o23.y, o45.x, o45.y, o67.x, o67.y};
            if (ACT == HPE_ACT_SILU) {
#pragma unroll
              for (int e = 0; e < 8; ++e) o[e] = silu_fast(o[e]);
            }
            if (ACT == HPE_ACT_GELU || ACT == HPE_ACT_GELU_LN) {         // exact (erf) GELU = nn.GELU() default (dinov3 convnext.py:60)
#pragma unroll
              for (int e = 0; e < 8; ++e) o[e] = gelu_erf(o[e]);
            }
            uint4 pk = make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
            if (ACT == HPE_ACT_RELU) {        // relu(round(x)) == round(relu(x)): do it on packed bf16x2
              pk.x = relu_bf16x2(pk.x); pk.y = relu_bf16x2(pk.y); pk.z = relu_bf16x2(pk.z); pk.w = relu_bf16x2(pk.w);
            }
            sts_v4(slot, pk);
          }
          if constexpr (HEAD2) {
            if (c == kChunks - 1) {                                  // all four chunks staged: one fence + barrier per tile
              fence_proxy_async_smem();                              // generic-proxy writes -> visible to the tensor core
              tc_fence_before();                                     // our tcgen05.ld of this accumulator precede its reuse below
              named_bar_sync(bar_id, kEpiThreads);
            }
          } else {
            fence_proxy_async_smem();                                // generic-proxy writes -> visible to TMA
            named_bar_sync(bar_id, kEpiThreads);
            if (leader) {
              if (p.l2_out) tma_store_4d_hint(&tmOut.m[t.cls], sSB + sb * kSBBytes, t.n_tile * BLOCK_N + c * 64, t.x0, t.y0, t.b0, p.l2_out);
              else          tma_store_4d(&tmOut.m[t.cls], sSB + sb * kSBBytes, t.n_tile * BLOCK_N + c * 64, t.x0, t.y0, t.b0);
              bulk_commit();
            }
          }
        }
        if constexpr (HEAD2) {
          // ---- second GEMM: heatmaps[256 px (both CTAs)][32] = A'[256][256] * W2[32][256]^T into columns 0-31 of `acc` ----
          if (leader) {
            if (it == 0) mbar_wait(w2full_bar, 0);                   // own joint weights landed (implied for the peer by its arrive)
            mbar_arrive_cluster_release(mapa_shared(smem_u32(s2full_bar), 0));
            if (rank == 0) {
              mbar_wait_cluster_acquire(s2full_bar, it & 1);         // both CTAs staged their A' tile
              tc_fence_after();
              constexpr uint32_t kIdesc2 = make_idesc_bf16_m256(32);
              const uint32_t tmem_d2 = tmem_base + acc * BLOCK_N;
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                const uint64_t adesc = make_sw128_kmajor_desc(sSB + c * kSBBytes);
                const uint64_t bdesc = make_sw128_kmajor_desc(sW2 + c * 2048);
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_bf16_2sm(tmem_d2, adesc + 2 * k, bdesc + 2 * k, kIdesc2, (c | k) != 0);
              }
              umma_commit_2sm(d2full_bar, 3);
            }
          }
          mbar_wait(d2full_bar, it & 1);
          tc_fence_after();
          if (half == 0) {
            uint32_t v[32];
            tmem_ld_x32(taddr, v);
            tmem_ld_wait();
            const int xi = row % p.tw, yi = (row / p.tw) % p.th, bi = row / (p.tw * p.th);
            const int x = t.x0 + xi, y = t.y0 + yi, b = t.b0 + bi;
            if (x < p.Wo && y < p.Ho && b < p.B) {
              const int up = p.n_classes == 4 ? 2 : 1;               // deconv: tile = one parity class of the 2x-upsampled output
              const int oy = up * y + (up == 2 ? (t.cls >> 1) : 0), ox = up * x + (up == 2 ? (t.cls & 1) : 0);
              const size_t plane = static_cast<size_t>(up * p.Ho) * (up * p.Wo);
              float* dst = p.hm_out + static_cast<size_t>(b) * p.head2_k * plane + static_cast<size_t>(oy) * (up * p.Wo) + ox;
#pragma unroll
              for (int e = 0; e < 32; ++e)
                if (e < p.head2_k) dst[e * plane] = __uint_as_float(v[e]) + __ldg(p.bias2 + e);
            }
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&tempty_bar[acc]), 0));     // accumulator free (leader's barrier)
        }
      }
    }
    if (!HEAD2 && leader) bulk_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if constexpr (PAIR) cluster_sync_all();       // no CTA leaves while its peer may still signal its barriers / read its smem
  if (warp == 1) {
    tc_fence_after();
    if constexpr (PAIR) tmem_dealloc_2sm(tmem_base, kTmemCols); else tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ---- auxiliary bandwidth kernels -------------------------------------------------------------------
// fp32 NCHW image -> bf16 space-to-depth [B, H/2, W/2 + 4, 16]; channel = (py*2+px)*4 + c (c==3 -> 0),
// 2 zero columns of left padding and 2 of right padding (the stem's x padding is physical, its y padding
// is TMA out-of-bounds fill).
constexpr int kS2DPad = 4;
// two horizontally adjacent pixels of one channel row; uint8 input = the reference's `imgs.float() / 255.0`
// (/root/reference/HPE/train.py:117-120, HPE/test.py:16-29) fused into the layout change
__device__ __forceinline__ float2 load_px2(const float* p) { return __ldg(reinterpret_cast<const float2*>(p)); }
__device__ __forceinline__ float2 load_px2(const unsigned char* p) {
  const uchar2 u = __ldg(reinterpret_cast<const uchar2*>(p));
  return make_float2(__fdiv_rn(static_cast<float>(u.x), 255.f), __fdiv_rn(static_cast<float>(u.y), 255.f));
}
template <typename T>
__global__ void __launch_bounds__(256)
preprocess_kernel(const T* __restrict__ img, int B, int H, int W, __nv_bfloat16* __restrict__ out) {
  const int Hs = H / 2, Ws = W / 2, Wp = Ws + kS2DPad;
  const long long total = static_cast<long long>(B) * Hs * Wp;
  const int lane = threadIdx.x & 31;
  // whole warps walk the loop: the output is linear in i (32 bytes per thread), and the warp trades its 16-byte pieces by shuffle
  // so that each of its two store instructions covers 512 contiguous bytes (not 32 half-written sectors)
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i - lane < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int xp = static_cast<int>(i % Wp);
    const int ys = static_cast<int>((i / Wp) % Hs);
    const int b  = static_cast<int>(i / (static_cast<long long>(Wp) * Hs));
    const int xs = xp - 2;
    uint4 lo = make_uint4(0, 0, 0, 0), hi = make_uint4(0, 0, 0, 0);
    if (i < total && xs >= 0 && xs < Ws) {
      float v[2][2][3];
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int py = 0; py < 2; ++py) {
          const float2 f = load_px2(img + ((static_cast<size_t>(b) * 3 + c) * H + (2 * ys + py)) * W + 2 * xs);
          v[py][0][c] = f.x; v[py][1][c] = f.y;
        }
      lo = make_uint4(pack_bf16(v[0][0][0], v[0][0][1]), pack_bf16(v[0][0][2], 0.f),
                      pack_bf16(v[0][1][0], v[0][1][1]), pack_bf16(v[0][1][2], 0.f));
      hi = make_uint4(pack_bf16(v[1][0][0], v[1][0][1]), pack_bf16(v[1][0][2], 0.f),
                      pack_bf16(v[1][1][0], v[1][1][1]), pack_bf16(v[1][1][2], 0.f));
    }
    uint4* o = reinterpret_cast<uint4*>(out) + (i - lane) * 2;            // the warp's first pixel
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int src = j * 16 + (lane >> 1);                                // piece j * 32 + lane = half (lane & 1) of that lane's pixel
      uint4 a, c2;
      a.x = __shfl_sync(0xffffffffu, lo.x, src); a.y = __shfl_sync(0xffffffffu, lo.y, src);
      a.z = __shfl_sync(0xffffffffu, lo.z, src); a.w = __shfl_sync(0xffffffffu, lo.w, src);
      c2.x = __shfl_sync(0xffffffffu, hi.x, src); c2.y = __shfl_sync(0xffffffffu, hi.y, src);
      c2.z = __shfl_sync(0xffffffffu, hi.z, src); c2.w = __shfl_sync(0xffffffffu, hi.w, src);
      if (i - lane + src < total) o[j * 32 + lane] = (lane & 1) ? c2 : a;
    }
  }
}

// uint8 image, W % 8 == 0: one thread = FOUR adjacent space-to-depth pixels (xp = 4g .. 4g+3, i.e. source columns 8g-4 .. 8g+3 of two
// image rows) from 4-byte loads; same bf16 values as preprocess_kernel<unsigned char> (x / 255 in fp32, rounded to bf16), computed
// as x * fp32(1/255): the fp32 products differ from the quotients in the last bit for 126 of the 256 inputs, but no x / 255 lies
// within 7.7e-6 (relative) of a bf16 rounding boundary, so the bf16 results are the same for ALL 256 inputs (checked exhaustively
// in tests/test_oracle_cpu.py and on the device in tests/test_model_gpu.py).  96 IEEE divisions per thread had made the kernel
// compute-bound: 94 us under ncu against 40 us for the fp32 image, which moves 1.8x the bytes.
__global__ void __launch_bounds__(256)
preprocess_u8x4_kernel(const unsigned char* __restrict__ img, int B, int H, int W, __nv_bfloat16* __restrict__ out) {
  constexpr float kInv255 = 1.0f / 255.0f;
  __shared__ uint4 stage_all[8][256];                                      // per warp: 32 lanes x 8 pieces of 16 bytes
  const int lane = threadIdx.x & 31;
  uint4* const stage = stage_all[threadIdx.x >> 5];
  const int Hs = H / 2, Ws = W / 2, Wp = Ws + kS2DPad, G = Wp / 4;         // W % 8 == 0 -> Wp % 4 == 0
  const long long total = static_cast<long long>(B) * Hs * G;
  // whole warps walk the loop (i - lane < total): the lanes past the end load nothing and store nothing, but take part in the exchange
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i - lane < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const bool live = i < total;
    const int g  = static_cast<int>(i % G);
    const int ys = static_cast<int>((i / G) % Hs);
    const int b  = live ? static_cast<int>(i / (static_cast<long long>(G) * Hs)) : 0;
    const int sx0 = 8 * g - 4;                                             // first source column of the group (multiple of 4)
    float v[3][2][8];                                                      // [c][py][8 source columns]
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int py = 0; py < 2; ++py) {
        const unsigned char* row = img + ((static_cast<size_t>(b) * 3 + c) * H + (2 * ys + py)) * W;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int sx = sx0 + 4 * h;
          uchar4 u = make_uchar4(0, 0, 0, 0);
          if (live && sx >= 0 && sx < W) u = __ldg(reinterpret_cast<const uchar4*>(row + sx));
          v[c][py][4 * h + 0] = static_cast<float>(u.x) * kInv255; v[c][py][4 * h + 1] = static_cast<float>(u.y) * kInv255;
          v[c][py][4 * h + 2] = static_cast<float>(u.z) * kInv255; v[c][py][4 * h + 3] = static_cast<float>(u.w) * kInv255;
        }
      }
    // The thread's 8 sixteen-byte pieces (128 contiguous bytes; the output is LINEAR in i: group i starts at byte 128 i) go through
    // shared memory so that each store instruction of the warp covers 512 contiguous bytes instead of 32 half-written sectors
    // 128 bytes apart.  Piece k of lane l sits at 16-byte slot l * 8 + (k ^ (l & 7)): conflict-free both ways.
#pragma unroll
    for (int k = 0; k < 4; ++k) {                                          // s2d pixel xp = 4g + k = source columns 2k, 2k+1 of the group
      // channel = (py*2+px)*4 + c (c == 3 -> 0); columns outside the image were loaded as 0 (the physical x padding)
      stage[lane * 8 + ((2 * k) ^ (lane & 7))] =
          make_uint4(pack_bf16(v[0][0][2 * k], v[1][0][2 * k]), pack_bf16(v[2][0][2 * k], 0.f),
                     pack_bf16(v[0][0][2 * k + 1], v[1][0][2 * k + 1]), pack_bf16(v[2][0][2 * k + 1], 0.f));
      stage[lane * 8 + ((2 * k + 1) ^ (lane & 7))] =
          make_uint4(pack_bf16(v[0][1][2 * k], v[1][1][2 * k]), pack_bf16(v[2][1][2 * k], 0.f),
                     pack_bf16(v[0][1][2 * k + 1], v[1][1][2 * k + 1]), pack_bf16(v[2][1][2 * k + 1], 0.f));
    }
    __syncwarp();
    uint4* o = reinterpret_cast<uint4*>(out) + (i - lane) * 8;            // the warp's first group
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int owner = j * 4 + (lane >> 3), k = lane & 7;                // piece j * 32 + lane of the warp's 256
      const uint4 piece = stage[owner * 8 + (k ^ (owner & 7))];
      if (i - lane + owner < total) o[j * 32 + lane] = piece;
    }
    __syncwarp();
  }
}

__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}

// MaxPool 3x3 s2 p1 on NHWC bf16, 8 channels (16 B) per thread
__global__ void __launch_bounds__(256)
maxpool_kernel(const __nv_bfloat16* __restrict__ x, int B, int H, int W, int C, __nv_bfloat16* __restrict__ y) {
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1, Cv = C / 8;
  const long long total = static_cast<long long>(B) * Ho * Wo * Cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(i % Cv);
    const int ox = static_cast<int>((i / Cv) % Wo);
    const int oy = static_cast<int>((i / (static_cast<long long>(Cv) * Wo)) % Ho);
    const int b  = static_cast<int>(i / (static_cast<long long>(Cv) * Wo * Ho));
    const uint32_t ninf = 0xFF80FF80u;    // (-inf, -inf) in bf16
    uint4 m = make_uint4(ninf, ninf, ninf, ninf);
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int iy = 2 * oy - 1 + ky;
      if (iy < 0 || iy >= H) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int ix = 2 * ox - 1 + kx;
        if (ix < 0 || ix >= W) continue;
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(x + ((static_cast<size_t>(b) * H + iy) * W + ix) * C) + cv);
        m.x = bf16x2_max(m.x, v.x); m.y = bf16x2_max(m.y, v.y); m.z = bf16x2_max(m.z, v.z); m.w = bf16x2_max(m.w, v.w);
      }
    }
    reinterpret_cast<uint4*>(y + ((static_cast<size_t>(b) * Ho + oy) * Wo + ox) * C)[cv] = m;
  }
}

// PyTorch-layout fp32 weight -> packed bf16 GEMM-B  [rows][Ktot]; every tap's Cin is zero-padded to Cp = kc*64
__device__ __forceinline__ void pack_weight_body(const float* __restrict__ w, int kind, int Cin, int Cp, int Cout, long long total,
                                                 __nv_bfloat16* __restrict__ out, long long ldo, int koff, long long i0, long long istep) {
  for (long long i = i0; i < total; i += istep) {
    float v = 0.f;
    if (kind == HPE_CONV_STEM7x7) {
      const int co = static_cast<int>(i / 256), k = static_cast<int>(i % 256);
      const int khp = k / 64, kwp = (k % 64) / 16, q = k % 16, py = q / 8, px = (q / 4) % 2, c = q % 4;
      const int ky = 2 * khp + py - 1, kx = 2 * kwp + px - 1;
      if (c < 3 && ky >= 0 && kx >= 0) v = w[((static_cast<size_t>(co) * 3 + c) * 7 + ky) * 7 + kx];   // [64][3][7][7]
    } else if (kind == HPE_CONV_DECONV4) {   // torch [Cin][Cout][4][4] -> [(cls*Cout + co)][(ty*2+tx)*Cp + ci]
      const int K = 4 * Cp;
      const int row = static_cast<int>(i / K), k = static_cast<int>(i % K);
      const int cls = row / Cout, co = row % Cout, t = k / Cp, ci = k % Cp;
      const int py = cls >> 1, px = cls & 1, ty = t >> 1, tx = t & 1;
      const int ky = py ? (ty ? 0 : 2) : (ty ? 3 : 1);
      const int kx = px ? (tx ? 0 : 2) : (tx ? 3 : 1);
      if (ci < Cin) v = w[((static_cast<size_t>(ci) * Cout + co) * 4 + ky) * 4 + kx];
    } else {                                  // Conv2d [Cout][Cin][kh][kw] -> [co][tap*Cp + ci], tap = ky*kw + kx
      const int taps = kind == HPE_CONV_3x3 ? 9 : (kind == HPE_CONV_2x2S2 ? 4 : 1);
      const int K = taps * Cp, co = static_cast<int>(i / K), k = static_cast<int>(i % K);
      const int tap = k / Cp, ci = k % Cp;
      if (ci < Cin) v = w[(static_cast<size_t>(co) * Cin + ci) * taps + tap];
      if (ldo) { out[static_cast<size_t>(co) * ldo + koff + k] = __float2bfloat16(v); continue; }   // K-slice of a wider row
    }
    out[i] = __float2bfloat16(v);
  }
}
__global__ void __launch_bounds__(256)
pack_weight_kernel(const float* __restrict__ w, int kind, int Cin, int Cp, int Cout, long long total,
                   __nv_bfloat16* __restrict__ out, long long ldo = 0, int koff = 0) {
  pack_weight_body(w, kind, Cin, Cp, Cout, total, out, ldo, koff, static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x,
                   static_cast<long long>(gridDim.x) * blockDim.x);
}

}  // namespace hpe

using namespace hpe;

struct hpe_conv_plan {
  hpe_conv_desc d;
  ConvParams p;
  int block_n, grid;
  int pair;              // cta_group::2 kernel: 2-CTA clusters, each CTA loads half of the weight tile
  int Ho, Wo;            // real output size
  size_t smem;
  CUtensorMap tmA, tmW, tmRes;
  OutMaps tmOut;
  const void *x, *x2, *w, *res, *w2;
  void *y, *y2;                       // y2: CHAIN, the chained conv's output
  CUtensorMap tmA2;
  const void* fn;
  int epi2;                           // two epilogue groups (608-thread kernel)
  int halo_taps_y;                    // != 0: the layer runs on the y-halo variant (conv_halo.cu)
  int ah;                             // != 0: halo'd A operand inside conv_gemm_kernel (3x3 / stride 1, 128 -> 128; AH template flag)
  HaloParams hp;
};

// CTAs a conv launch may use: all SMs, or fewer (HPE_B200_MAX_CTAS, read when the plan is created) so that two half-batch plans on
// two streams occupy disjoint SM halves and an HBM-bound layer of one overlaps a tensor-bound layer of the other.
static int cta_budget() {
  const char* e = getenv("HPE_B200_MAX_CTAS");
  const int n = e ? atoi(e) : 0;
  return (n > 0 && n < num_sms()) ? n : num_sms();
}

// Layers whose generic kernel is operand-delivery bound and that fit the y-halo variant: the stem, and 3x3/s1 64->64.
static bool halo_eligible(const hpe_conv_desc* d, int Ho, int Wo) {
  const char* e = getenv("HPE_B200_HALO");
  if (e && e[0] == '0') return false;
  if (d->has_residual || d->in_cstride || d->out_cstride || d->Cin2) return false;
  const bool shape_ok = d->Cout == 64 && (Wo + 7) / 8 * 8 * ((Ho + kHaloRows - 1) / kHaloRows * kHaloRows) <= (long long)Wo * Ho * 9 / 8;   // <= 12.5 % padded
  if (d->kind == HPE_CONV_STEM7x7) return shape_ok || d->pool;      // the pooled stem has its own (overlapping) tiling
  if (d->kind == HPE_CONV_3x3) return shape_ok && d->Cin == 64 && d->stride == 1;
  return false;
}

static int desc_check(const hpe_conv_desc* d, const char* fn) {
  HPE_CHECK_ARG(d, "%s: null desc", fn);
  HPE_CHECK_ARG(d->kind >= HPE_CONV_1x1 && d->kind <= HPE_CONV_2x2S2, "%s: bad kind %d", fn, d->kind);
  HPE_CHECK_ARG(d->B > 0 && d->H > 0 && d->W > 0 && d->Cin > 0 && d->Cout > 0, "%s: non-positive size", fn);
  HPE_CHECK_ARG(d->in_cstride >= 0 && d->out_cstride >= 0, "%s: negative channel stride", fn);
  if (d->pool && (d->kind != HPE_CONV_STEM7x7 || d->act != HPE_ACT_RELU || d->has_residual || d->in_cstride || d->out_cstride))
    return fail(HPE_EUNSUPPORTED, "%s: the fused max-pool needs the dense ReLU stem (kind %d, act %d)", fn, d->kind, d->act);
  if (d->kind == HPE_CONV_STEM7x7) {
    if (d->Cin != 3 || d->Cout != 64 || (d->H % 2) || (d->W % 2))
      return fail(HPE_EUNSUPPORTED, "%s: stem needs Cin=3, Cout=64, even H/W (got %d,%d,%dx%d)", fn, d->Cin, d->Cout, d->H, d->W);
    return HPE_OK;
  }
  if (d->Cin % 8) return fail(HPE_EUNSUPPORTED, "%s: Cin=%d must be a multiple of 8", fn, d->Cin);
  if (d->head2) {
    if (d->head2 < 0 || d->head2 > 32) return fail(HPE_EUNSUPPORTED, "%s: the fused head supports at most 32 joints (got %d)", fn, d->head2);
    const bool kind_ok = d->kind == HPE_CONV_DECONV4 || ((d->kind == HPE_CONV_1x1 || d->kind == HPE_CONV_3x3) && d->stride == 1);
    if (!kind_ok || d->Cout != 256 || d->has_residual || d->out_cstride || d->Cin2 || d->act != HPE_ACT_RELU)
      return fail(HPE_EUNSUPPORTED, "%s: the fused head needs a ReLU deconv / stride-1 conv with Cout = 256, no residual, no second input", fn);
  }
  if (d->chain) {
    const bool shape_ok = (d->Cout == 256 && (d->chain == 64 || d->chain == 128)) || (d->Cout == 512 && d->chain == 128);
    if (!shape_ok || d->kind != HPE_CONV_1x1 || d->stride != 1 || d->head2 || d->out_cstride || d->act != HPE_ACT_RELU)
      return fail(HPE_EUNSUPPORTED, "%s: a chained 1x1 conv needs a dense ReLU 1x1 stride-1 layer with Cout = 256 (64 | 128 chained outputs) or 512 (128)", fn);
  }
  if (d->Cin2) {
    if (d->kind != HPE_CONV_1x1 || d->stride != 1) return fail(HPE_EUNSUPPORTED, "%s: a second input needs a 1x1 stride-1 main conv", fn);
    if (d->Cin2 < 0 || d->Cin2 % 8) return fail(HPE_EUNSUPPORTED, "%s: Cin2=%d must be a positive multiple of 8", fn, d->Cin2);
    if (d->stride2 != 1 && d->stride2 != 2) return fail(HPE_EUNSUPPORTED, "%s: stride2 %d not in {1,2}", fn, d->stride2);
  }
  if (d->in_cstride && (d->in_cstride % 8 || d->in_cstride < d->Cin))
    return fail(HPE_EINVAL, "%s: in_cstride=%d must be a multiple of 8 and >= Cin", fn, d->in_cstride);
  if (d->kind == HPE_CONV_FINAL) {
    if (d->Cout > 256) return fail(HPE_EUNSUPPORTED, "%s: final conv supports at most 256 joints (got %d)", fn, d->Cout);
    if (d->act != HPE_ACT_NONE || d->has_residual) return fail(HPE_EUNSUPPORTED, "%s: final conv has no activation/residual", fn);
    if (d->out_cstride && d->out_cstride < d->Cout) return fail(HPE_EINVAL, "%s: out_cstride < Cout", fn);
    return HPE_OK;
  }
  if (d->Cout % 8) return fail(HPE_EUNSUPPORTED, "%s: Cout=%d must be a multiple of 8", fn, d->Cout);
  if (d->out_cstride && (d->out_cstride % 8 || d->out_cstride < d->Cout))
    return fail(HPE_EINVAL, "%s: out_cstride=%d must be a multiple of 8 and >= Cout", fn, d->out_cstride);
  if (d->Cout > kMaxCout) return fail(HPE_EUNSUPPORTED, "%s: Cout=%d exceeds %d", fn, d->Cout, kMaxCout);
  if (d->act < HPE_ACT_NONE || d->act > HPE_ACT_GELU_LN) return fail(HPE_EINVAL, "%s: bad act %d", fn, d->act);
  if (d->act == HPE_ACT_GELU_LN && (d->kind != HPE_CONV_1x1 || d->stride != 1 || d->Cin2 || d->chain || d->head2 || d->has_residual))
    return fail(HPE_EUNSUPPORTED, "%s: the LayerNorm-folded GELU epilogue is for plain 1x1 / stride 1 layers", fn);
  if (d->act == HPE_ACT_GELU && d->has_residual) return fail(HPE_EUNSUPPORTED, "%s: GELU epilogue has no residual variant", fn);
  if (d->kind == HPE_CONV_1x1 || d->kind == HPE_CONV_3x3) {
    if (d->stride != 1 && d->stride != 2) return fail(HPE_EUNSUPPORTED, "%s: stride %d not in {1,2}", fn, d->stride);
    if (d->stride == 2 && ((d->H % 2) || (d->W % 2))) return fail(HPE_EUNSUPPORTED, "%s: stride 2 needs even H/W", fn);
  }
  if ((d->kind == HPE_CONV_4x4S2 || d->kind == HPE_CONV_2x2S2) && ((d->H % 2) || (d->W % 2)))
    return fail(HPE_EUNSUPPORTED, "%s: stride-2 kinds need even H/W", fn);
  if (d->kind >= HPE_CONV_DGRAD3x3S2 && d->kind <= HPE_CONV_4x4S2 && ((d->Cin % 64) || (d->Cout % 64)))
    return fail(HPE_EUNSUPPORTED, "%s: backward-data kinds need Cin, Cout multiples of 64", fn);
  return HPE_OK;
}

// BLOCK_N: the widest tile dividing Cout.  (Measured on B200: narrower tiles that fill the last wave of the
// persistent grid better are still slower -- layer4 3x3 66 -> 78 us -- because every extra N tile re-reads
// the A operand through the L2->SM fabric, which is what bounds these kernels; see profiles/.)
// 3x3 / stride 1 / 128 -> 128 (ResNet layer2 conv2): the generic kernel re-reads the A brick for each of the 9 taps through the
// L2 -> SM fabric (576 KB of operands per 128-pixel tile, 125 B per cycle of MMA work: the layer ran at 0.53 of the tensor peak);
// with two halo'd boxes per tile it is 334 KB.  HPE_B200_HALO128=0 keeps the generic kernel (1: single CTAs, 2: CTA pairs, default).
static bool halo128_eligible(const hpe_conv_desc* d) {
  const char* e = getenv("HPE_B200_HALO128");
  if (!(e ? atoi(e) : 2)) return false;
  return d->kind == HPE_CONV_3x3 && d->stride == 1 && d->Cin == 128 && d->Cout == 128 && d->act == HPE_ACT_RELU && !d->has_residual &&
         !d->in_cstride && !d->out_cstride && !d->Cin2 && !d->chain && !d->head2 && !d->pool && d->W % 8 == 0;
}

static bool is_class_strided(int kind);
static int pick_block_n(const hpe_conv_desc* d, long long m_tiles) {
  if (d->kind == HPE_CONV_FINAL) return d->Cout <= 32 ? 32 : (d->Cout <= 64 ? 64 : (d->Cout <= 128 ? 128 : 256));
  {
    // Wave quantisation of the cta_group::2 kernels: layer4's conv1 / conv2 (Cout 512 over 96 M tiles) are 96 work items of 256
    // columns on 74 clusters = 2 waves at 65 %; as 144 items of 192 columns (the third N tile a third full) they are 2 waves of
    // 3/4 the length.  Measured (per-layer log, three boxes): layer4 3x3 176 / 169 / 175 -> 157 / 143 / 130 us, layer4 1x1
    // 246 / 235 / 245 -> 244 / 224 / 209 us.  HPE_B200_WAVE192=0 keeps the widest tile that divides Cout.
    const char* ew = getenv("HPE_B200_WAVE192");            // read per plan (plans are created once per layer and shape)
    const int wave = ew ? atoi(ew) : 1;
    // (not for the GELU layers: their 256-wide tiles run two epilogue groups, which the 192-wide kernel does not have)
    if (wave && d->Cout % 256 == 0 && d->Cout >= 512 && !d->chain && !d->has_residual && d->act != HPE_ACT_GELU && d->act != HPE_ACT_GELU_LN && !is_class_strided(d->kind) && m_tiles >= 2 &&
        (d->Cout + 191) / 192 * 192 <= kMaxCout) {
      const long long clusters = cta_budget() / 2, pairs = (m_tiles + 1) / 2;
      const long long w256 = (pairs * (d->Cout / 256) + clusters - 1) / clusters * 256;
      const long long w192 = (pairs * ((d->Cout + 191) / 192) + clusters - 1) / clusters * 192;
      if (w192 < w256) return 192;
    }
  }
  // the tile width wasting the fewest padded columns; ties -> the wider tile (fewer re-reads of A)
  const int cand[4] = {256, 192, 128, 64};
  int best = 64, best_waste = 1 << 30;
  for (int i = 0; i < 4; ++i) {
    const int bn = cand[i], waste = (d->Cout + bn - 1) / bn * bn - d->Cout;
    if (waste < best_waste) { best_waste = waste; best = bn; }
  }
  return best;
}

// Launch-order / L2 hints of a plan (performance only; results are bit-identical).  Consecutive layers walk their tiles in the same
// order, so a layer starts by reading the part of its input that its producer wrote FIRST -- with tensors near the size of the
// 126 MB L2 exactly the part that has been evicted.  reverse_tiles walks the work items last to first; l2_out / l2_res / l2_in =
// 0 (none), 1 (evict first), 2 (evict last) put an eviction priority on the output stores / the residual loads / the input
// (A operand) loads.  Plans on the y-halo kernel take the direction only.  Returns the reverse flag in effect, negative on error.
extern "C" int hpe_conv_set_schedule(hpe_conv_plan* pl, int reverse_tiles, int l2_out, int l2_res, int l2_in) {
  if (!pl) return fail(HPE_EINVAL, "hpe_conv_set_schedule: null plan");
  if (l2_out < 0 || l2_out > 2 || l2_res < 0 || l2_res > 2 || l2_in < 0 || l2_in > 2)
    return fail(HPE_EINVAL, "hpe_conv_set_schedule: hints are 0 (none), 1 (evict first), 2 (evict last)");
  if (pl->halo_taps_y) { pl->hp.reverse = reverse_tiles ? 1 : 0; return pl->hp.reverse; }
  static const unsigned long long pol[3] = {0ull, kL2EvictFirst, kL2EvictLast};
  pl->p.reverse = reverse_tiles ? 1 : 0;
  pl->p.l2_out = pol[l2_out];
  pl->p.l2_res = pol[l2_res];
  pl->p.l2_in = l2_in ? pol[l2_in] : kL2EvictNormal;
  return pl->p.reverse;
}

extern "C" int hpe_conv_out_hw(const hpe_conv_desc* d, int* Ho, int* Wo) {
  int rc = desc_check(d, "hpe_conv_out_hw");
  if (rc) return rc;
  HPE_CHECK_ARG(Ho && Wo, "hpe_conv_out_hw: null out");
  switch (d->kind) {
    case HPE_CONV_STEM7x7:                      // pool: MaxPool2d(3, 2, 1) of the H/2 x W/2 convolution output
      *Ho = d->pool ? (d->H / 2 - 1) / 2 + 1 : d->H / 2; *Wo = d->pool ? (d->W / 2 - 1) / 2 + 1 : d->W / 2; break;
    case HPE_CONV_DECONV4: case HPE_CONV_DGRAD3x3S2: case HPE_CONV_DGRAD1x1S2: *Ho = d->H * 2; *Wo = d->W * 2; break;
    case HPE_CONV_4x4S2: case HPE_CONV_2x2S2: *Ho = d->H / 2; *Wo = d->W / 2; break;
    case HPE_CONV_FINAL:   *Ho = d->H; *Wo = d->W; break;
    default:               *Ho = d->H / d->stride; *Wo = d->W / d->stride; break;
  }
  return HPE_OK;
}

static int cin_padded(const hpe_conv_desc* d) { return (d->Cin + 63) / 64 * 64; }
static size_t ktot(const hpe_conv_desc* d) {
  const size_t cp = cin_padded(d);
  switch (d->kind) {
    case HPE_CONV_3x3:     return 9 * cp;
    case HPE_CONV_STEM7x7: return 256;
    case HPE_CONV_DECONV4: case HPE_CONV_DGRAD3x3S2: case HPE_CONV_2x2S2: return 4 * cp;
    case HPE_CONV_4x4S2:   return 16 * cp;
    default:               return cp + (size_t)(d->Cin2 + 63) / 64 * 64;
  }
}
static bool is_class_strided(int kind) { return kind == HPE_CONV_DECONV4 || kind == HPE_CONV_DGRAD3x3S2; }
static size_t wrows(const hpe_conv_desc* d) { return is_class_strided(d->kind) ? (size_t)4 * d->Cout : d->Cout; }

extern "C" size_t hpe_conv_packed_weight_elems(const hpe_conv_desc* d) {
  if (desc_check(d, "hpe_conv_packed_weight_elems")) return 0;
  return wrows(d) * ktot(d);
}

extern "C" int hpe_conv_pack_weight(const hpe_conv_desc* d, const float* w_torch, void* w_packed, void* stream) {
  int rc = desc_check(d, "hpe_conv_pack_weight");
  if (rc) return rc;
  HPE_CHECK_ARG(w_torch && w_packed, "hpe_conv_pack_weight: null pointer");
  if (d->Cin2) return fail(HPE_EINVAL, "hpe_conv_pack_weight: the layer has a second input; use hpe_conv_pack_weight2");
  if (d->kind > HPE_CONV_FINAL && d->kind != HPE_CONV_2x2S2) return fail(HPE_EINVAL, "hpe_conv_pack_weight: kind %d is a backward-data kind; use hpe_conv_pack_weight_dgrad", d->kind);
  const long long total = (long long)(wrows(d) * ktot(d));
  const int grid = (int)((total + 255) / 256 > 4096 ? 4096 : (total + 255) / 256);
  pack_weight_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w_torch, d->kind, d->Cin, cin_padded(d), d->Cout, total,
                                                            (__nv_bfloat16*)w_packed);
  return launch_status("pack_weight_kernel");
}

extern "C" int hpe_conv_pack_weight2(const hpe_conv_desc* d, const float* w_torch, const float* w2_torch, void* w_packed,
                                     void* stream) {
  int rc = desc_check(d, "hpe_conv_pack_weight2");
  if (rc) return rc;
  HPE_CHECK_ARG(w_torch && w2_torch && w_packed, "hpe_conv_pack_weight2: null pointer");
  if (!d->Cin2) return fail(HPE_EINVAL, "hpe_conv_pack_weight2: the layer has no second input");
  const int cp1 = cin_padded(d), cp2 = (d->Cin2 + 63) / 64 * 64;
  const long long ldo = (long long)ktot(d);
  for (int src = 0; src < 2; ++src) {
    const int cp = src ? cp2 : cp1;
    const long long total = (long long)d->Cout * cp;
    const int grid = (int)((total + 255) / 256 > 4096 ? 4096 : (total + 255) / 256);
    pack_weight_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src ? w2_torch : w_torch, HPE_CONV_1x1, src ? d->Cin2 : d->Cin, cp,
                                                              d->Cout, total, (__nv_bfloat16*)w_packed, ldo, src ? cp1 : 0);
    rc = launch_status("pack_weight_kernel");
    if (rc) return rc;
  }
  return HPE_OK;
}

namespace hpe {
// PyTorch-layout fp32 FORWARD weight -> packed bf16 GEMM-B of the layer's backward-data convolution
__device__ __forceinline__ void pack_weight_dgrad_body(const float* __restrict__ w, int fkind, int fstride, int Cin, int Cout,
                                                       long long total, __nv_bfloat16* __restrict__ out, long long i0, long long istep) {
  // Cin/Cout are the FORWARD layer's; the dgrad GEMM has N = Cin (rows) and K over Cout.
  for (long long i = i0; i < total; i += istep) {
    float v = 0.f;
    if (fkind == HPE_CONV_1x1) {                                        // [ci][co] = w[co][ci]
      const int ci = static_cast<int>(i / Cout), co = static_cast<int>(i % Cout);
      v = w[static_cast<size_t>(co) * Cin + ci];
    } else if (fkind == HPE_CONV_FINAL) {                               // [ci][64]: joints zero-padded
      const int ci = static_cast<int>(i / 64), k = static_cast<int>(i % 64);
      if (k < Cout) v = w[static_cast<size_t>(k) * Cin + ci];
    } else if (fkind == HPE_CONV_3x3 && fstride == 1) {                 // [ci][t*Cout+co] = w[co][ci][2-ty][2-tx]
      const int K = 9 * Cout, ci = static_cast<int>(i / K), k = static_cast<int>(i % K);
      const int t = k / Cout, co = k % Cout, ty = t / 3, tx = t % 3;
      v = w[((static_cast<size_t>(co) * Cin + ci) * 3 + (2 - ty)) * 3 + (2 - tx)];
    } else if (fkind == HPE_CONV_3x3) {                                 // stride 2: [(cls*Cin+ci)][t*Cout+co]
      const int K = 4 * Cout;
      const int row = static_cast<int>(i / K), k = static_cast<int>(i % K);
      const int cls = row / Cin, ci = row % Cin, t = k / Cout, co = k % Cout;
      const int py = cls >> 1, px = cls & 1, nx = 1 + px, ny = 1 + py, ty = t / nx, tx = t % nx;
      if (ty < ny) {
        const int ky = py ? (ty ? 0 : 2) : 1, kx = px ? (tx ? 0 : 2) : 1;
        v = w[((static_cast<size_t>(co) * Cin + ci) * 3 + ky) * 3 + kx];
      }
    } else {                                                            // deconv: [ci][t*Cout+co] = w[ci][co][ky][kx]
      const int K = 16 * Cout, ci = static_cast<int>(i / K), k = static_cast<int>(i % K);
      const int t = k / Cout, co = k % Cout;
      v = w[(static_cast<size_t>(ci) * Cout + co) * 16 + t];
    }
    out[i] = __float2bfloat16(v);
  }
}
__global__ void __launch_bounds__(256)
pack_weight_dgrad_kernel(const float* __restrict__ w, int fkind, int fstride, int Cin, int Cout, long long total,
                         __nv_bfloat16* __restrict__ out) {
  pack_weight_dgrad_body(w, fkind, fstride, Cin, Cout, total, out, static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x,
                         static_cast<long long>(gridDim.x) * blockDim.x);
}

// Every layer's forward and backward-data pack in ONE launch (training re-packs all bf16 operands from the updated fp32
// parameters each step: ~110 separate launches of a few microseconds each were 1.4 ms of a 25 ms step).
struct PackJob {
  const float* src; __nv_bfloat16* dst; long long total;
  int dgrad, kind, stride, Cin, Cp, Cout, block_begin, n_blocks;
};
__global__ void __launch_bounds__(256)
pack_jobs_kernel(const PackJob* __restrict__ jobs, int n_jobs) {
  int lo = 0, hi = n_jobs - 1;                       // last job whose first block is <= this block
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (jobs[mid].block_begin <= static_cast<int>(blockIdx.x)) lo = mid; else hi = mid - 1;
  }
  const PackJob j = jobs[lo];
  const long long i0 = static_cast<long long>(blockIdx.x - j.block_begin) * blockDim.x + threadIdx.x;
  const long long istep = static_cast<long long>(j.n_blocks) * blockDim.x;
  if (j.dgrad) pack_weight_dgrad_body(j.src, j.kind, j.stride, j.Cin, j.Cout, j.total, j.dst, i0, istep);
  else         pack_weight_body(j.src, j.kind, j.Cin, j.Cp, j.Cout, j.total, j.dst, 0, 0, i0, istep);
}

}  // namespace hpe
using namespace hpe;

extern "C" size_t hpe_pack_job_bytes(void) { return sizeof(PackJob); }

extern "C" int hpe_conv_pack_job_fill(const hpe_conv_desc* fwd_desc, int dgrad, const float* w_torch, void* w_packed,
                                      int block_begin, void* job_host, int* n_blocks) {
  HPE_CHECK_ARG(fwd_desc && w_torch && w_packed && job_host && n_blocks && block_begin >= 0, "hpe_conv_pack_job_fill: bad argument");
  PackJob j;
  j.src = w_torch; j.dst = (__nv_bfloat16*)w_packed; j.dgrad = dgrad ? 1 : 0;
  j.kind = fwd_desc->kind; j.stride = fwd_desc->stride; j.Cin = fwd_desc->Cin; j.Cout = fwd_desc->Cout;
  if (dgrad) {
    hpe_conv_desc g;
    int rc = hpe_conv_dgrad_desc(fwd_desc, &g);
    if (rc) return rc;
    j.total = (long long)hpe_conv_packed_weight_elems(&g); j.Cp = 0;
  } else {
    int rc = desc_check(fwd_desc, "hpe_conv_pack_job_fill");
    if (rc) return rc;
    if (fwd_desc->Cin2 || (fwd_desc->kind > HPE_CONV_FINAL && fwd_desc->kind != HPE_CONV_2x2S2))
      return fail(HPE_EINVAL, "hpe_conv_pack_job_fill: forward jobs take single-input forward kinds");
    j.total = (long long)(wrows(fwd_desc) * ktot(fwd_desc)); j.Cp = cin_padded(fwd_desc);
  }
  const long long nb = (j.total + 1023) / 1024;              // ~4 elements per thread
  j.n_blocks = (int)(nb < 1 ? 1 : (nb > 4096 ? 4096 : nb));
  j.block_begin = block_begin;
  *n_blocks = j.n_blocks;
  memcpy(job_host, &j, sizeof(j));
  return HPE_OK;
}

extern "C" int hpe_conv_pack_jobs_run(const void* jobs_device, int n_jobs, int total_blocks, void* stream) {
  HPE_CHECK_ARG(jobs_device && n_jobs > 0 && total_blocks > 0, "hpe_conv_pack_jobs_run: bad argument");
  pack_jobs_kernel<<<total_blocks, 256, 0, (cudaStream_t)stream>>>((const PackJob*)jobs_device, n_jobs);
  return launch_status("pack_jobs_kernel");
}

extern "C" int hpe_conv_dgrad_desc(const hpe_conv_desc* f, hpe_conv_desc* g) {
  HPE_CHECK_ARG(f && g, "hpe_conv_dgrad_desc: null pointer");
  int Ho = 0, Wo = 0;
  int rc = hpe_conv_out_hw(f, &Ho, &Wo);
  if (rc) return rc;
  *g = *f;
  g->H = Ho; g->W = Wo; g->Cin = f->Cout; g->Cout = f->Cin; g->stride = 1; g->act = HPE_ACT_NONE; g->has_residual = 0;
  g->Cin2 = 0; g->stride2 = 0; g->head2 = 0; g->pool = 0; g->chain = 0;
  if (f->Cin2) return fail(HPE_EUNSUPPORTED, "hpe_conv_dgrad_desc: fused two-input layers are inference-only");
  switch (f->kind) {
    case HPE_CONV_1x1:     g->kind = f->stride == 2 ? HPE_CONV_DGRAD1x1S2 : HPE_CONV_1x1; break;
    case HPE_CONV_3x3:     g->kind = f->stride == 2 ? HPE_CONV_DGRAD3x3S2 : HPE_CONV_3x3; break;
    case HPE_CONV_DECONV4: g->kind = HPE_CONV_4x4S2; g->stride = 2; break;
    case HPE_CONV_FINAL:   g->kind = HPE_CONV_1x1; g->Cin = 64; break;
    default: return fail(HPE_EUNSUPPORTED, "hpe_conv_dgrad_desc: kind %d has no backward-data convolution (the stem's input needs no gradient)", f->kind);
  }
  return HPE_OK;
}

extern "C" int hpe_conv_pack_weight_dgrad(const hpe_conv_desc* f, const float* w_torch, void* w_packed, void* stream) {
  hpe_conv_desc g;
  int rc = hpe_conv_dgrad_desc(f, &g);
  if (rc) return rc;
  HPE_CHECK_ARG(w_torch && w_packed, "hpe_conv_pack_weight_dgrad: null pointer");
  const long long total = (long long)hpe_conv_packed_weight_elems(&g);
  const int grid = (int)((total + 255) / 256 > 4096 ? 4096 : (total + 255) / 256);
  pack_weight_dgrad_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w_torch, f->kind, f->stride, f->Cin, f->Cout, total,
                                                                  (__nv_bfloat16*)w_packed);
  return launch_status("pack_weight_dgrad_kernel");
}

typedef void (*ConvKernelFn)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const OutMaps, const CUtensorMap, const ConvParams);

// ST_N / ST_R / ST_F: operand-ring depth without / with a residual / for the fp32 heatmap epilogue
template <int BN, int ST_N, int ST_R, int ST_F>
static ConvKernelFn pick_kernel_bn(int act, bool res, bool fin, size_t* smem) {
  *smem = res ? SmemLayout<BN, ST_R, false, 4>::kDynamic : SmemLayout<BN, ST_N, false, 3>::kDynamic;
  if (fin) { *smem = SmemLayout<BN, ST_F, false, 0>::kDynamic; return conv_gemm_kernel<BN, ST_F, HPE_ACT_NONE, false, true>; }
  if constexpr (BN >= 64) {
    switch (act) {
      case HPE_ACT_RELU: return res ? conv_gemm_kernel<BN, ST_R, HPE_ACT_RELU, true, false> : conv_gemm_kernel<BN, ST_N, HPE_ACT_RELU, false, false>;
      case HPE_ACT_SILU: return res ? conv_gemm_kernel<BN, ST_R, HPE_ACT_SILU, true, false> : conv_gemm_kernel<BN, ST_N, HPE_ACT_SILU, false, false>;
      case HPE_ACT_GELU: return res ? nullptr : conv_gemm_kernel<BN, ST_N, HPE_ACT_GELU, false, false>;
      default:           return res ? conv_gemm_kernel<BN, ST_R, HPE_ACT_NONE, true, false> : conv_gemm_kernel<BN, ST_N, HPE_ACT_NONE, false, false>;
    }
  }
  return nullptr;
}

static ConvKernelFn pick_kernel_ring6(int act, size_t* smem) {
  *smem = SmemLayout<256, 6, true, 2>::kDynamic;
  switch (act) {
    case HPE_ACT_RELU: return conv_gemm_kernel<256, 6, HPE_ACT_RELU, false, false, true>;
    case HPE_ACT_SILU: return conv_gemm_kernel<256, 6, HPE_ACT_SILU, false, false, true>;
    case HPE_ACT_NONE: return conv_gemm_kernel<256, 6, HPE_ACT_NONE, false, false, true>;
    default:           return nullptr;
  }
}

static ConvKernelFn pick_kernel_ring6_128(int act, size_t* smem) {       // single-CTA 128-wide tiles (layer2's 3x3, 256 -> 128 1x1)
  *smem = SmemLayout<128, 6, false, 2>::kDynamic;
  switch (act) {
    case HPE_ACT_RELU: return conv_gemm_kernel<128, 6, HPE_ACT_RELU, false, false>;
    case HPE_ACT_SILU: return conv_gemm_kernel<128, 6, HPE_ACT_SILU, false, false>;
    case HPE_ACT_NONE: return conv_gemm_kernel<128, 6, HPE_ACT_NONE, false, false>;
    default:           return nullptr;
  }
}

template <int BN, int ST_N, int ST_R>
static ConvKernelFn pick_kernel_pair(int act, bool res, size_t* smem) {
  *smem = res ? SmemLayout<BN, ST_R, true, 4>::kDynamic : SmemLayout<BN, ST_N, true, 3>::kDynamic;
  switch (act) {
    case HPE_ACT_RELU: return res ? conv_gemm_kernel<BN, ST_R, HPE_ACT_RELU, true, false, true> : conv_gemm_kernel<BN, ST_N, HPE_ACT_RELU, false, false, true>;
    case HPE_ACT_SILU: return res ? conv_gemm_kernel<BN, ST_R, HPE_ACT_SILU, true, false, true> : conv_gemm_kernel<BN, ST_N, HPE_ACT_SILU, false, false, true>;
    case HPE_ACT_GELU: return res ? nullptr : conv_gemm_kernel<BN, ST_N, HPE_ACT_GELU, false, false, true>;
    default:           return res ? conv_gemm_kernel<BN, ST_R, HPE_ACT_NONE, true, false, true> : conv_gemm_kernel<BN, ST_N, HPE_ACT_NONE, false, false, true>;
  }
}

// two epilogue groups (608 threads) on the 256-wide pair kernel: with a residual 3 ring stages + 2 x 3 staging buffers; without
// one (the GELU epilogue of ConvNeXt's pwconv1: 5 FMAs + 2 MUFUs per element over 4C channels) 4 stages + 2 x 2 buffers
static ConvKernelFn pick_kernel_epi2(int act, bool res, size_t* smem) {
  if (!res) {
    *smem = SmemLayout<256, 4, true, 4, false, true>::kDynamic;
    if (act == HPE_ACT_GELU_LN) return conv_gemm_kernel<256, 4, HPE_ACT_GELU_LN, false, false, true, false, true>;
    return act == HPE_ACT_GELU ? conv_gemm_kernel<256, 4, HPE_ACT_GELU, false, false, true, false, true> : nullptr;
  }
  *smem = SmemLayout<256, 3, true, 6, false, true>::kDynamic;
  switch (act) {
    case HPE_ACT_RELU: return conv_gemm_kernel<256, 3, HPE_ACT_RELU, true, false, true, false, true>;
    case HPE_ACT_SILU: return conv_gemm_kernel<256, 3, HPE_ACT_SILU, true, false, true, false, true>;
    case HPE_ACT_GELU: return nullptr;
    default:           return conv_gemm_kernel<256, 3, HPE_ACT_NONE, true, false, true, false, true>;
  }
}

// chained 1x1 (N2 = 64 | 128) behind a 256-wide single-CTA ReLU tile: 2 ring stages of 48 KB, 4 (N2 = 128 with a residual: 3)
// staging buffers, the chained weights resident
static ConvKernelFn pick_kernel_chain(int cout, int n2, bool res, int num_kb, size_t* smem) {
  if (cout == 512) {       // layer2: four 128-wide quarters per pixel tile, streamed weights (ring 3 x 32 KB, 3 chained K slices)
    if (n2 != 128) return nullptr;
    *smem = SmemLayout<128, 3, false, 4, false, false, 128, 0, 4>::kDynamic;
    return res ? conv_gemm_kernel<128, 3, HPE_ACT_RELU, true, false, false, false, false, 128, 0, 4>
               : conv_gemm_kernel<128, 3, HPE_ACT_RELU, false, false, false, false, false, 128, 0, 4>;
  }
  // (N2, residual, k-blocks per tile) of the three layer1 cases: conv3 (K = 64) + identity shortcut -> conv1 64 | 128, and the
  // first block's two-source conv3 (K = 64 + 64, projection shortcut folded in) -> conv1 64
  if (n2 == 64 && res && num_kb == 1) {
    *smem = SmemLayout<256, 3, false, 6, false, false, 64, 1>::kDynamic;
    return conv_gemm_kernel<256, 3, HPE_ACT_RELU, true, false, false, false, false, 64, 1>;
  }
  if (n2 == 128 && res && num_kb == 1) {
    *smem = SmemLayout<256, 3, false, 4, false, false, 128, 1>::kDynamic;
    return conv_gemm_kernel<256, 3, HPE_ACT_RELU, true, false, false, false, false, 128, 1>;
  }
  if (n2 == 64 && !res && num_kb == 2) {
    *smem = SmemLayout<256, 3, false, 4, false, false, 64, 2>::kDynamic;
    return conv_gemm_kernel<256, 3, HPE_ACT_RELU, false, false, false, false, false, 64, 2>;
  }
  return nullptr;
}

// ring depths: as many stages as fit beside the staging buffers in the 227 KB a CTA may use
static ConvKernelFn pick_kernel(int block_n, int act, bool res, bool fin, bool pair, bool head2, bool epi2, size_t* smem) {
  if (head2) { *smem = SmemLayout<256, 4, true, 4, true>::kDynamic; return conv_gemm_kernel<256, 4, HPE_ACT_RELU, false, false, true, true>; }
  if (epi2) return pick_kernel_epi2(act, res, smem);
  if (pair && block_n == 256 && !res && act != HPE_ACT_GELU) {
    static int ring6 = -1;
    if (ring6 < 0) { const char* e = getenv("HPE_B200_RING6"); ring6 = (e && e[0] == '0') ? 0 : 1; }       // default on: forward 3.228 -> 3.214 ms (two A/B pairs)
    if (ring6) return pick_kernel_ring6(act, smem);
  }
  if (pair && block_n == 256) return pick_kernel_pair<256, 5, 5>(act, res, smem);       // 32 KB stages (A + half of B)
  if (pair && block_n == 192 && act == HPE_ACT_GELU_LN) {
    *smem = SmemLayout<192, 6, true, 3>::kDynamic;
    return conv_gemm_kernel<192, 6, HPE_ACT_GELU_LN, false, false, true>;
  }
  if (pair && block_n == 192) return pick_kernel_pair<192, 6, 5>(act, res, smem);       // 28 KB stages (ConvNeXt C = 192/384/768)
  if (pair) return pick_kernel_pair<128, 7, 6>(act, res, smem);                         // 24 KB stages
  if (!pair && block_n == 128 && !res && !fin && act != HPE_ACT_GELU) {
    static int r128 = -1;
    if (r128 < 0) { const char* e = getenv("HPE_B200_RING6_128"); r128 = (e && e[0] == '1') ? 1 : 0; }
    if (r128) return pick_kernel_ring6_128(act, smem);
  }
  switch (block_n) {
    case 32:  return pick_kernel_bn<32, 6, 6, 11>(act, res, fin, smem);     // 20 KB stages (FINAL only)
    case 64:  return pick_kernel_bn<64, 7, 6, 9>(act, res, fin, smem);      // 24 KB stages
    case 128: return pick_kernel_bn<128, 5, 5, 7>(act, res, fin, smem);     // 32 KB
    case 192: return pick_kernel_bn<192, 4, 4, 4>(act, res, fin, smem);     // 40 KB
    default:  return pick_kernel_bn<256, 3, 3, 4>(act, res, fin, smem);     // 48 KB
  }
}

extern "C" int hpe_conv_plan_create(const hpe_conv_desc* d, hpe_conv_plan** out) {
  int rc = desc_check(d, "hpe_conv_plan_create");
  if (rc) return rc;
  HPE_CHECK_ARG(out, "hpe_conv_plan_create: null out");
  rc = hpe_check_device();
  if (rc) return rc;
  hpe_conv_plan* pl = new hpe_conv_plan();
  pl->d = *d;
  hpe_conv_out_hw(d, &pl->Ho, &pl->Wo);
  ConvParams& p = pl->p;
  p.kind = d->kind;
  p.l2_in = kL2EvictNormal;
  p.in_stride = (d->kind == HPE_CONV_1x1 || d->kind == HPE_CONV_3x3) ? d->stride
                : ((d->kind == HPE_CONV_4x4S2 || d->kind == HPE_CONV_2x2S2) ? 2 : 1);
  switch (d->kind) {
    case HPE_CONV_3x3: p.taps = 9; break;
    case HPE_CONV_4x4S2: p.taps = 16; break;
    case HPE_CONV_1x1: case HPE_CONV_FINAL: case HPE_CONV_DGRAD1x1S2: p.taps = 1; break;
    default: p.taps = 4; break;                 // stem, deconv, (dgrad 3x3/s2: upper bound, see kb_count)
  }
  p.kc_per_tap = d->kind == HPE_CONV_STEM7x7 ? 1 : cin_padded(d) / 64;
  p.kc2 = (d->Cin2 + 63) / 64; p.in_stride2 = d->Cin2 ? d->stride2 : 1;
  p.head2_k = d->head2; p.bias2 = nullptr;
  p.n_classes = is_class_strided(d->kind) ? 4 : 1;
  // tile grid extent: deconv-like tiles walk the INPUT-resolution grid of one parity class
  const bool up2 = is_class_strided(d->kind) || d->kind == HPE_CONV_DGRAD1x1S2;
  p.Ho = up2 ? d->H : pl->Ho;
  p.Wo = up2 ? d->W : pl->Wo;
  p.B = d->B; p.Cout = d->Cout;
  p.act = d->act; p.has_res = d->has_residual ? 1 : 0;
  int tw = 1, th = 1, nb = 128;
  pick_brick(128, p.Wo, p.Ho, p.B, (p.in_stride == 2 || p.in_stride2 == 2) ? 128 : 256, &tw, &th, &nb);
  pl->ah = halo128_eligible(d) ? 1 : 0;
  if (pl->ah) { tw = 8; th = 16; nb = 1; }         // the brick the two halo'd boxes {64, 10, 18, 1} cover
  p.tw = tw; p.th = th; p.nb = nb;
  p.tiles_x = (p.Wo + tw - 1) / tw; p.tiles_y = (p.Ho + th - 1) / th; p.tiles_b = (p.B + nb - 1) / nb;
  pl->block_n = pick_block_n(d, (long long)p.tiles_x * p.tiles_y * p.tiles_b * p.n_classes);
  if (d->chain && d->Cout == 512) pl->block_n = 128;       // four quarters per pixel tile inside one CTA
  p.n_tiles_n = (d->kind == HPE_CONV_FINAL || d->chain) ? 1 : (d->Cout + pl->block_n - 1) / pl->block_n;   // chain: all of Cout in one CTA
  p.cout_total = d->out_cstride ? d->out_cstride : d->Cout;
  if (is_class_strided(d->kind) && d->Cout % pl->block_n) { delete pl; return fail(HPE_EUNSUPPORTED, "hpe_conv_plan_create: parity-class kinds need Cout %% %d == 0", pl->block_n); }
  if (p.n_tiles_n * pl->block_n > kMaxCout) { delete pl; return fail(HPE_EUNSUPPORTED, "hpe_conv_plan_create: padded Cout exceeds %d", kMaxCout); }
  const long long nt = (long long)p.tiles_x * p.tiles_y * p.tiles_b * p.n_tiles_n * p.n_classes;
  if (nt > 0x7fffffff) { delete pl; return fail(HPE_EUNSUPPORTED, "hpe_conv_plan_create: too many tiles"); }
  p.num_tiles = (int)nt;
  pl->grid = p.num_tiles < cta_budget() ? p.num_tiles : cta_budget();
  {
    // cta_group::2 for the 256-wide tiles (HPE_B200_PAIR=0 disables): pairs of M tiles share one MMA stream
    const char* e = getenv("HPE_B200_PAIR");
    // 128-wide tiles: opt-in only (HPE_B200_PAIR128=1) -- measured slower (layer2 3x3 128->128: 74 -> 88 us): their 24 KB
    // stages hold only 256 cycles of MMA work, so the longer cross-CTA hand-off outweighs the two extra ring stages
    const char* e128 = getenv("HPE_B200_PAIR128");
    const char* e192 = getenv("HPE_B200_PAIR192");
    const char* eah = getenv("HPE_B200_HALO128");
    // HPE_B200_HALO128 = 2 (default): the halo'd-A kernel as a cta_group::2 pair (8 KB weight half-tiles, ten stages: layer 55.4 ->
    // 53.2 us, forward 3.285 -> 3.275 ms); 1: single CTAs
    const bool ah_pair = (eah ? atoi(eah) : 2) == 2;
    const long long m_tiles = (long long)p.tiles_x * p.tiles_y * p.tiles_b;
    pl->pair = !d->chain && !(e && e[0] == '0') && (pl->block_n == 256 || (pl->block_n == 192 && !(e192 && e192[0] == '0')) || (pl->block_n == 128 && ((e128 && e128[0] == '1') || (pl->ah && ah_pair)))) && d->kind != HPE_CONV_FINAL && m_tiles >= 2 && !halo_eligible(d, pl->Ho, pl->Wo);
    if (d->chain && halo_eligible(d, pl->Ho, pl->Wo)) { delete pl; return fail(HPE_EUNSUPPORTED, "hpe_conv_plan_create: chained layer on the halo kernel"); }
  if (d->head2 && !pl->pair) { delete pl; return fail(HPE_EUNSUPPORTED, "hpe_conv_plan_create: the fused head runs on the cta_group::2 kernel only (needs >= 2 M tiles, HPE_B200_PAIR != 0)"); }
    if (pl->pair) {
      const long long np = (m_tiles + 1) / 2 * p.n_tiles_n * p.n_classes;
      p.num_tiles = (int)np;
      const int max_clusters = cta_budget() / 2;
      pl->grid = 2 * (int)(np < max_clusters ? np : max_clusters);
    }
  }
  {
    // two epilogue groups for the 256-wide pair kernel with a residual (HPE_B200_EPI2=0 disables, =2 also takes the deep-K
    // layers): these layers are epilogue-bound (8.2 k cycles per tile against 0.5-4 k of MMA).  The second group's staging
    // buffers cost two ring stages (3 instead of 5), so layers with more than 8 k-blocks per tile keep the single group.
    const char* e = getenv("HPE_B200_EPI2");
    const int mode = e ? atoi(e) : 1;
    const int num_kb = p.taps * p.kc_per_tap + p.kc2;
    const char* eg = getenv("HPE_B200_EPI2_GELU");
    const bool gelu2 = ((d->act == HPE_ACT_GELU && !(eg && eg[0] == '0')) || d->act == HPE_ACT_GELU_LN) && !p.has_res;   // epilogue-bound GELU layers (pwconv1)
    pl->epi2 = mode != 0 && pl->pair && pl->block_n == 256 && !d->head2 &&
               ((p.has_res && d->act != HPE_ACT_GELU && (mode == 2 || num_kb <= 8)) || gelu2);
  }
  if (d->act == HPE_ACT_GELU_LN && !pl->epi2 && !(pl->pair && pl->block_n == 192)) {
    delete pl;
    return fail(HPE_EUNSUPPORTED, "hpe_conv_plan_create: the LayerNorm-folded GELU layer runs on the cta_group::2 kernels only (>= 2 pixel tiles; 256-wide tiles with two epilogue groups or 192-wide tiles; HPE_B200_PAIR / HPE_B200_EPI2 != 0)");
  }
  if (pl->ah) {
    // 6 weight stages of 16 KB + 2 staging buffers (the `wait_group.read 0` protocol of the 6-stage kernels): 5 stages hold
    // 1800 cycles of MMA work, less than the L2 round trip
    pl->smem = SmemLayout<128, 6, false, 2, false, false, 0, 0, 1, true>::kDynamic;
    pl->fn = (const void*)conv_gemm_kernel<128, 6, HPE_ACT_RELU, false, false, false, false, false, 0, 0, 1, true>;
    if (pl->pair) {       // cta_group::2: 8 KB weight half-tiles, ten stages
      pl->smem = SmemLayout<128, 10, true, 3, false, false, 0, 0, 1, true>::kDynamic;
      pl->fn = (const void*)conv_gemm_kernel<128, 10, HPE_ACT_RELU, false, false, true, false, false, 0, 0, 1, true>;
    }
  } else
  pl->fn = d->chain ? (const void*)pick_kernel_chain(d->Cout, d->chain, p.has_res != 0, p.taps * p.kc_per_tap + p.kc2, &pl->smem)
                    : (const void*)pick_kernel(pl->block_n, d->act, p.has_res != 0, d->kind == HPE_CONV_FINAL, pl->pair != 0, d->head2 != 0, pl->epi2 != 0, &pl->smem);
  if (!pl->fn) { const int bn = pl->block_n; delete pl; return fail(HPE_EUNSUPPORTED, "hpe_conv_plan_create: no kernel for BLOCK_N=%d%s", bn, d->chain ? " (chained layers: Cin = 64 with a residual, or Cin = Cin2 = 64 without)" : ""); }
  {
    cudaError_t e = cudaFuncSetAttribute((const void*)pl->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl->smem);
    if (e != cudaSuccess) { delete pl; return fail(HPE_ECUDA, "cudaFuncSetAttribute(smem=%zu) failed: %s", pl->smem, cudaGetErrorString(e)); }
  }
  if (pl->pair) {
    // the persistent schedule strides by the number of clusters, so every cluster must be co-resident: ask the driver how
    // many 2-CTA clusters of this kernel fit (not every TPC of a B200 has both SMs enabled)
    // (queried per plan: the answer depends on the kernel variant, its shared memory and the current device)
    int max_clusters = -1;
    {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(num_sms() / 2 * 2); cfg.blockDim = dim3(pl->epi2 ? kConvThreads2 : kConvThreads); cfg.dynamicSmemBytes = pl->smem;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      int n = 0;
      cudaError_t e = cudaOccupancyMaxActiveClusters(&n, (const void*)pl->fn, &cfg);
      if (e != cudaSuccess || n <= 0) { delete pl; return fail(HPE_ECUDA, "cudaOccupancyMaxActiveClusters failed: %s", cudaGetErrorString(e)); }
      const char* ov = getenv("HPE_B200_PAIR_CLUSTERS");
      max_clusters = ov ? atoi(ov) : n;
      if (max_clusters > cta_budget() / 2) max_clusters = cta_budget() / 2;
      if (getenv("HPE_B200_DEBUG")) fprintf(stderr, "[hpe] max active 2-CTA clusters: %d (using %d)\n", n, max_clusters);
    }
    pl->grid = 2 * (p.num_tiles < max_clusters ? p.num_tiles : max_clusters);
  }
  pl->x = pl->x2 = pl->w = pl->res = pl->w2 = nullptr; pl->y = pl->y2 = nullptr;
  pl->halo_taps_y = 0;
  if (halo_eligible(d, pl->Ho, pl->Wo)) {
    // 8-wide x 32-tall bricks; the 3x3 reads ONE 10-pixel-wide halo'd box per brick whose column taps are 128-byte shifts of
    // the descriptor start.  Measured on B200: SWIZZLE_128B is a function of the absolute shared-memory address, so such starts
    // need base offset 0 (setting the descriptor's base-offset field to (start >> 7) & 7 instead FAILS parity).
    const bool stem = d->kind == HPE_CONV_STEM7x7;
    pl->halo_taps_y = stem ? 4 : 3;
    pl->hp.tiles_x = (pl->Wo + 7) / 8; pl->hp.tiles_y = (pl->Ho + kHaloRows - 1) / kHaloRows;
    pl->hp.x_lo = stem ? 0 : 1; pl->hp.y_lo = stem ? 2 : 1;
    pl->hp.pool = d->pool ? 1 : 0; pl->hp.Wc = d->W / 2; pl->hp.Hc = d->H / 2;
    pl->hp.Wo = pl->Wo; pl->hp.Ho = pl->Ho; pl->hp.out = nullptr;
    if (d->pool) {                              // 3 x 15 pooled pixels per (overlapping) brick; pl->Wo/Ho are the POOLED extent
      pl->hp.tiles_x = (pl->Wo + 2) / 3; pl->hp.tiles_y = (pl->Ho + 14) / 15;
    }
    pl->hp.num_tiles = pl->hp.tiles_x * pl->hp.tiles_y * d->B;
    pl->hp.nbatch = d->B; pl->hp.reverse = 0;
    pl->grid = pl->hp.num_tiles < cta_budget() ? pl->hp.num_tiles : cta_budget();
    p.tw = 8; p.th = kHaloRows; p.nb = 1;        // brick of the A map encoded below
  } else if (d->pool) {
    delete pl;
    return fail(HPE_EUNSUPPORTED, "hpe_conv_plan_create: the fused max-pool runs on the y-halo stem kernel only (HPE_B200_HALO=0 disables it)");
  }
  *out = pl;
  return HPE_OK;
}

extern "C" void hpe_conv_plan_destroy(hpe_conv_plan* plan) { delete plan; }

static int encode_plan_maps(hpe_conv_plan* pl, const void* x, const void* x2, const void* w, const void* res, void* y,
                            const void* w2 = nullptr, void* y2 = nullptr) {
  const hpe_conv_desc& d = pl->d;
  const ConvParams& p = pl->p;
  int rc;
  const uint32_t ones[4] = {1, 1, 1, 1};
  if (x != pl->x) {
    HPE_CHECK_ARG(((uintptr_t)x & 15) == 0, "hpe_conv_run: x must be 16-byte aligned");
    if (d.kind == HPE_CONV_STEM7x7) {
      const uint64_t Hs = d.H / 2, Ws = d.W / 2, Wp = Ws + kS2DPad;
      const uint64_t dims[4] = {64, Ws, Hs, (uint64_t)d.B};
      const uint64_t str[3] = {32, Wp * 32, Hs * Wp * 32};           // overlapping 64-element windows, 16 el apart
      const uint32_t box[4] = {64, (uint32_t)p.tw, (uint32_t)(pl->halo_taps_y ? kHaloRows + pl->halo_taps_y - 1 : p.th), (uint32_t)p.nb};
      rc = encode_map(&pl->tmA, x, 4, dims, str, box, ones, "stem A");
    } else {
      const uint64_t C = d.Cin, P = d.in_cstride ? d.in_cstride : d.Cin;      // P = channel pitch of a pixel
      const uint64_t dims[4] = {C, (uint64_t)d.W, (uint64_t)d.H, (uint64_t)d.B};
      const uint64_t str[3] = {P * 2, (uint64_t)d.W * P * 2, (uint64_t)d.H * d.W * P * 2};
      const uint32_t s = (uint32_t)p.in_stride;
      const uint32_t box[4] = {64, (uint32_t)((pl->halo_taps_y || pl->ah) ? 10 : p.tw * s),
                               (uint32_t)(pl->halo_taps_y ? kHaloRows + pl->halo_taps_y - 1 : (pl->ah ? 18 : p.th * s)), (uint32_t)p.nb};
      const uint32_t es[4] = {1, s, s, 1};
      rc = encode_map(&pl->tmA, x, 4, dims, str, box, es, "A");
    }
    if (rc) return rc;
    pl->x = x;
  }
  if (d.Cin2 && x2 != pl->x2) {
    HPE_CHECK_ARG(x2 && ((uintptr_t)x2 & 15) == 0, "hpe_conv_run2: x2 must be non-null and 16-byte aligned");
    const uint32_t s = (uint32_t)p.in_stride2;                // dense [B, Ho*s, Wo*s, Cin2]
    const uint64_t C = d.Cin2, W2 = (uint64_t)pl->Wo * s, H2 = (uint64_t)pl->Ho * s;
    const uint64_t dims[4] = {C, W2, H2, (uint64_t)d.B};
    const uint64_t str[3] = {C * 2, W2 * C * 2, H2 * W2 * C * 2};
    const uint32_t box[4] = {64, (uint32_t)p.tw * s, (uint32_t)p.th * s, (uint32_t)p.nb};
    const uint32_t es[4] = {1, s, s, 1};
    rc = encode_map(&pl->tmA2, x2, 4, dims, str, box, es, "A2");
    if (rc) return rc;
    pl->x2 = x2;
  }
  if (w != pl->w) {
    HPE_CHECK_ARG(((uintptr_t)w & 15) == 0, "hpe_conv_run: w must be 16-byte aligned");
    const uint64_t K = ktot(&d), R = wrows(&d);
    const uint64_t dims[2] = {K, R};
    const uint64_t str[1] = {K * 2};
    const uint32_t box[2] = {64, (uint32_t)(pl->halo_taps_y ? 64 : (pl->pair ? pl->block_n / 2 : pl->block_n))};
    rc = encode_map(&pl->tmW, w, 2, dims, str, box, ones, "W");
    if (rc) return rc;
    pl->w = w;
  }
  if (d.head2 && w2 != pl->w2) {               // fused head: joint weights [32][256] bf16 (rows >= K zero), 16 rows per CTA
    HPE_CHECK_ARG(w2 && ((uintptr_t)w2 & 15) == 0, "hpe_conv_run_head2: w2 must be non-null and 16-byte aligned");
    const uint64_t dims[2] = {256, 32};
    const uint64_t str[1] = {256 * 2};
    const uint32_t box[2] = {64, 16};
    rc = encode_map(&pl->tmRes, w2, 2, dims, str, box, ones, "head2 W");
    if (rc) return rc;
    pl->w2 = w2;
  }
  if (d.kind != HPE_CONV_FINAL && !d.head2 && !pl->halo_taps_y && y != pl->y) {      // (the y-halo kernel stores without TMA)
    HPE_CHECK_ARG(((uintptr_t)y & 15) == 0, "hpe_conv_run: y must be 16-byte aligned");
    const uint64_t C = d.Cout, P = d.out_cstride ? d.out_cstride : d.Cout;
    const uint32_t box[4] = {64, (uint32_t)p.tw, (uint32_t)p.th, (uint32_t)p.nb};
    if (is_class_strided(d.kind) || d.kind == HPE_CONV_DGRAD1x1S2) {
      const uint64_t Wo = pl->Wo, Ho = pl->Ho;
      const int ncls = d.kind == HPE_CONV_DGRAD1x1S2 ? 1 : 4;
      for (int cls = 0; cls < 4; ++cls) {
        if (cls >= ncls) { pl->tmOut.m[cls] = pl->tmOut.m[0]; continue; }
        const int py = cls >> 1, px = cls & 1;
        const char* base = (const char*)y + ((size_t)py * Wo + px) * P * 2;
        const uint64_t dims[4] = {C, (uint64_t)d.W, (uint64_t)d.H, (uint64_t)d.B};
        const uint64_t str[3] = {2 * P * 2, 2 * Wo * P * 2, Ho * Wo * P * 2};
        rc = encode_map(&pl->tmOut.m[cls], base, 4, dims, str, box, ones, "deconv out");
        if (rc) return rc;
      }
    } else {
      const uint64_t dims[4] = {C, (uint64_t)pl->Wo, (uint64_t)pl->Ho, (uint64_t)d.B};
      const uint64_t str[3] = {P * 2, (uint64_t)pl->Wo * P * 2, (uint64_t)pl->Ho * pl->Wo * P * 2};
      rc = encode_map(&pl->tmOut.m[0], y, 4, dims, str, box, ones, "out");
      if (rc) return rc;
      if (d.chain) pl->tmOut.m[3] = pl->tmOut.m[0];             // m[1], m[2]: the chained conv's output / weights
      else pl->tmOut.m[1] = pl->tmOut.m[2] = pl->tmOut.m[3] = pl->tmOut.m[0];
    }
    pl->y = y;
  }
  if (d.chain && (w2 != pl->w2 || y2 != pl->y2)) {       // chained 1x1: weights [N2][256] -> tmOut.m[2], dense output [B,Ho,Wo,N2] -> tmOut.m[1]
    HPE_CHECK_ARG(w2 && y2 && (((uintptr_t)w2 | (uintptr_t)y2) & 15) == 0, "hpe_conv_run_chain: w2 / y2 must be non-null and 16-byte aligned");
    {
      const uint64_t dims[2] = {(uint64_t)d.Cout, (uint64_t)d.chain};      // [N2 rows][K = Cout]
      const uint64_t str[1] = {(uint64_t)d.Cout * 2};
      const uint32_t box[2] = {64, (uint32_t)d.chain};
      rc = encode_map(&pl->tmOut.m[2], w2, 2, dims, str, box, ones, "chain W");
      if (rc) return rc;
    }
    {
      const uint64_t C = d.chain;
      const uint64_t dims[4] = {C, (uint64_t)pl->Wo, (uint64_t)pl->Ho, (uint64_t)d.B};
      const uint64_t str[3] = {C * 2, (uint64_t)pl->Wo * C * 2, (uint64_t)pl->Ho * pl->Wo * C * 2};
      const uint32_t box[4] = {64, (uint32_t)p.tw, (uint32_t)p.th, (uint32_t)p.nb};
      rc = encode_map(&pl->tmOut.m[1], y2, 4, dims, str, box, ones, "chain out");
      if (rc) return rc;
    }
    pl->w2 = w2; pl->y2 = y2;
  }
  if (p.has_res && res != pl->res) {
    HPE_CHECK_ARG(res && ((uintptr_t)res & 15) == 0, "hpe_conv_run: residual must be non-null and 16-byte aligned");
    const uint64_t C = d.Cout;                  // the residual is always a dense [B,Ho,Wo,Cout] tensor
    const uint64_t dims[4] = {C, (uint64_t)pl->Wo, (uint64_t)pl->Ho, (uint64_t)d.B};
    const uint64_t str[3] = {C * 2, (uint64_t)pl->Wo * C * 2, (uint64_t)pl->Ho * pl->Wo * C * 2};
    const uint32_t box[4] = {64, (uint32_t)p.tw, (uint32_t)p.th, (uint32_t)p.nb};
    rc = encode_map(&pl->tmRes, res, 4, dims, str, box, ones, "residual");
    if (rc) return rc;
    pl->res = res;
  }
  return HPE_OK;
}

extern "C" int hpe_conv_run(hpe_conv_plan* pl, const void* x, const void* w_packed, const float* scale,
                            const float* shift, const void* residual, void* y, void* stream) {
  return hpe_conv_run2(pl, x, nullptr, w_packed, scale, shift, residual, y, stream);
}

static int conv_launch(hpe_conv_plan* pl, const void* x, const void* x2, const void* w_packed, const float* scale,
                       const float* shift, const void* residual, void* y, const void* w2, const float* bias2, void* stream);

static int conv_launch(hpe_conv_plan* pl, const void* x, const void* x2, const void* w_packed, const float* scale,
                       const float* shift, const void* residual, void* y, const void* w2, const float* bias2, void* stream,
                       const float* scale2, const float* shift2, void* y2);

extern "C" int hpe_conv_run2(hpe_conv_plan* pl, const void* x, const void* x2, const void* w_packed, const float* scale,
                             const float* shift, const void* residual, void* y, void* stream) {
  HPE_CHECK_ARG(pl && !pl->d.head2, "hpe_conv_run: the plan has a fused head; use hpe_conv_run_head2");
  HPE_CHECK_ARG(!pl->d.chain, "hpe_conv_run: the plan has a chained 1x1 conv; use hpe_conv_run_chain");
  HPE_CHECK_ARG(pl->d.act != HPE_ACT_GELU_LN, "hpe_conv_run: the plan folds a LayerNorm; use hpe_conv_run_ln");
  return conv_launch(pl, x, x2, w_packed, scale, shift, residual, y, nullptr, nullptr, stream);
}

extern "C" int hpe_conv_run_ln(hpe_conv_plan* pl, const void* x, const void* w_packed, const float* colsum, const float* shift,
                               const float* rowstats, int n_slices, int ln_channels, float eps, void* y, void* stream) {
  HPE_CHECK_ARG(pl && pl->d.act == HPE_ACT_GELU_LN, "hpe_conv_run_ln: the plan's activation is not HPE_ACT_GELU_LN");
  HPE_CHECK_ARG(rowstats && ((uintptr_t)rowstats & 7) == 0 && n_slices > 0 && ln_channels > 0 && eps > 0.f, "hpe_conv_run_ln: bad LayerNorm arguments");
  pl->p.rowstats = reinterpret_cast<const float2*>(rowstats);
  pl->p.ln_slices = n_slices; pl->p.ln_inv_c = 1.f / (float)ln_channels; pl->p.ln_eps = eps;
  return conv_launch(pl, x, nullptr, w_packed, colsum, shift, nullptr, y, nullptr, nullptr, stream);
}

extern "C" int hpe_conv_run_chain(hpe_conv_plan* pl, const void* x, const void* x2, const void* w_packed, const float* scale,
                                  const float* shift, const void* residual, void* y, const void* w2_packed, const float* scale2,
                                  const float* shift2, void* y2, void* stream) {
  HPE_CHECK_ARG(pl && pl->d.chain, "hpe_conv_run_chain: the plan has no chained conv (desc.chain = 0)");
  HPE_CHECK_ARG(w2_packed && scale2 && shift2 && y2, "hpe_conv_run_chain: null pointer");
  return conv_launch(pl, x, x2, w_packed, scale, shift, residual, y, w2_packed, nullptr, stream, scale2, shift2, y2);
}

extern "C" int hpe_conv_run_head2(hpe_conv_plan* pl, const void* x, const void* w_packed, const float* scale, const float* shift,
                                  const void* w2_packed, const float* bias2, float* heatmaps, void* stream) {
  HPE_CHECK_ARG(pl && pl->d.head2, "hpe_conv_run_head2: the plan has no fused head (desc.head2 = 0)");
  HPE_CHECK_ARG(w2_packed && bias2 && heatmaps, "hpe_conv_run_head2: null pointer");
  return conv_launch(pl, x, nullptr, w_packed, scale, shift, nullptr, heatmaps, w2_packed, bias2, stream);
}

static int conv_launch(hpe_conv_plan* pl, const void* x, const void* x2, const void* w_packed, const float* scale,
                       const float* shift, const void* residual, void* y, const void* w2, const float* bias2, void* stream) {
  return conv_launch(pl, x, x2, w_packed, scale, shift, residual, y, w2, bias2, stream, nullptr, nullptr, nullptr);
}

static int conv_launch(hpe_conv_plan* pl, const void* x, const void* x2, const void* w_packed, const float* scale,
                       const float* shift, const void* residual, void* y, const void* w2, const float* bias2, void* stream,
                       const float* scale2, const float* shift2, void* y2) {
  HPE_CHECK_ARG(pl && x && w_packed && scale && shift && y, "hpe_conv_run: null pointer");
  HPE_CHECK_ARG((pl->d.Cin2 != 0) == (x2 != nullptr), "hpe_conv_run2: x2 must be given exactly when the plan has a second input (Cin2=%d)", pl->d.Cin2);
  HPE_CHECK_ARG((((uintptr_t)scale | (uintptr_t)shift) & 15) == 0, "hpe_conv_run: scale/shift must be 16-byte aligned");
  if ((is_class_strided(pl->d.kind) || pl->d.kind == HPE_CONV_DGRAD1x1S2) && pl->p.has_res)
    return fail(HPE_EUNSUPPORTED, "hpe_conv_run: the strided-store kinds have no residual input");
  if (pl->d.kind == HPE_CONV_DGRAD1x1S2)      // odd rows/columns of dx receive no gradient
    HPE_CUDA_OK(cudaMemsetAsync(y, 0, (size_t)pl->d.B * pl->Ho * pl->Wo * pl->d.Cout * 2, (cudaStream_t)stream));
  int rc = encode_plan_maps(pl, x, x2, w_packed, residual, y, w2, y2);
  if (rc) return rc;
  if (pl->halo_taps_y) {
    pl->hp.scale = scale; pl->hp.shift = shift; pl->hp.out = y;
    return halo_launch(pl->halo_taps_y, pl->d.act, &pl->tmA, &pl->tmW, pl->hp, pl->grid, (cudaStream_t)stream);
  }
  ConvParams p = pl->p;
  p.scale = scale; p.shift = shift;
  p.hm_out = (pl->d.kind == HPE_CONV_FINAL || pl->d.head2) ? (float*)y : nullptr;
  p.bias2 = bias2;
  p.scale2 = scale2; p.shift2 = shift2;
  {
    // residual L2 prefetch one tile ahead (experiment, OFF: measured slower -- forward 3.12 / 3.20 ms off, 3.20 / 3.23 ms on the
    // chained kernels, 3.23 ms on every residual layer; like the A-stream prefetch of round 1, the prefetches only compete with
    // the demand loads of an HBM-bound kernel): HPE_B200_RES_PREFETCH = 1 chained kernels, 2 every residual layer
    static int mode = -1;
    if (mode < 0) { const char* e = getenv("HPE_B200_RES_PREFETCH"); mode = e ? atoi(e) : 0; }
    p.res_prefetch = p.has_res && (mode == 2 || (mode == 1 && pl->d.chain)) ? 1 : 0;
    // chained kernels: drain tile i's chained-conv output two chunks into tile i+1 (chain_seq); HPE_B200_CHAIN_DEFER=0 = right behind
    // tile i, where the group that gets it waits for the second GEMM over the chunk the OTHER group has only just staged
    static int defer = -1;
    if (defer < 0) { const char* e = getenv("HPE_B200_CHAIN_DEFER"); defer = (e && atoi(e) == 0) ? 0 : 2; }
    p.chain_defer = pl->d.chain ? defer : 0;
  }
  if (!p.has_res && !pl->d.head2) pl->tmRes = pl->tmA;      // unused, but must be a valid object to copy
  if (pl->d.kind == HPE_CONV_FINAL || pl->d.head2) for (int c = 0; c < 4; ++c) pl->tmOut.m[c] = pl->tmA;
  if (!pl->d.Cin2) pl->tmA2 = pl->tmA;      // unused, but must be a valid object to copy
  void* args[6] = {&pl->tmA, &pl->tmA2, &pl->tmW, &pl->tmOut, &pl->tmRes, &p};
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(pl->grid); cfg.blockDim = dim3(pl->d.chain ? kConvThreadsC : (pl->epi2 ? kConvThreads2 : kConvThreads));
  cfg.dynamicSmemBytes = pl->smem; cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (pl->pair) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr; cfg.numAttrs = na;
  cudaError_t e = cudaLaunchKernelExC(&cfg, (const void*)pl->fn, args);
  if (e != cudaSuccess) return fail(HPE_ECUDA, "launch of conv_gemm_kernel failed: %s", cudaGetErrorString(e));
  return launch_status("conv_gemm_kernel");
}

extern "C" size_t hpe_preprocess_out_elems(int B, int H, int W) {
  if (B <= 0 || H <= 0 || W <= 0 || (H % 2) || (W % 2)) return 0;
  return (size_t)B * (H / 2) * (W / 2 + kS2DPad) * 16;
}

extern "C" int hpe_preprocess(const float* img, int B, int H, int W, void* out, void* stream) {
  HPE_CHECK_ARG(img && out, "hpe_preprocess: null pointer");
  HPE_CHECK_ARG(B > 0 && H > 0 && W > 0 && H % 2 == 0 && W % 2 == 0, "hpe_preprocess: H and W must be positive and even");
  HPE_CHECK_ARG((((uintptr_t)img & 7) | ((uintptr_t)out & 15)) == 0, "hpe_preprocess: misaligned pointer");
  const long long total = (long long)B * (H / 2) * (W / 2 + kS2DPad);
  const int grid = (int)((total + 255) / 256 > (long long)num_sms() * 16 ? (long long)num_sms() * 16 : (total + 255) / 256);
  preprocess_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(img, B, H, W, (__nv_bfloat16*)out);
  return launch_status("preprocess_kernel");
}

extern "C" int hpe_preprocess_u8(const unsigned char* img, int B, int H, int W, void* out, void* stream) {
  HPE_CHECK_ARG(img && out, "hpe_preprocess_u8: null pointer");
  HPE_CHECK_ARG(B > 0 && H > 0 && W > 0 && H % 2 == 0 && W % 2 == 0, "hpe_preprocess_u8: H and W must be positive and even");
  HPE_CHECK_ARG((((uintptr_t)img & 1) | ((uintptr_t)out & 15)) == 0, "hpe_preprocess_u8: misaligned pointer");
  if (W % 8 == 0 && ((uintptr_t)img & 3) == 0) {
    // four space-to-depth pixels per thread from 4-byte loads (the 2-byte loads of the generic kernel made the uint8 path
    // SLOWER than the fp32 one -- 68 vs 40 us at B = 256 -- although it reads a quarter of the bytes)
    const long long total = (long long)B * (H / 2) * ((W / 2 + kS2DPad) / 4);
    const int grid = (int)((total + 255) / 256 > (long long)num_sms() * 16 ? (long long)num_sms() * 16 : (total + 255) / 256);
    preprocess_u8x4_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(img, B, H, W, (__nv_bfloat16*)out);
    return launch_status("preprocess_u8x4_kernel");
  }
  const long long total = (long long)B * (H / 2) * (W / 2 + kS2DPad);
  const int grid = (int)((total + 255) / 256 > (long long)num_sms() * 16 ? (long long)num_sms() * 16 : (total + 255) / 256);
  preprocess_kernel<unsigned char><<<grid, 256, 0, (cudaStream_t)stream>>>(img, B, H, W, (__nv_bfloat16*)out);
  return launch_status("preprocess_kernel<u8>");
}

extern "C" int hpe_maxpool3x3s2(const void* x, int B, int H, int W, int C, void* y, void* stream) {
  HPE_CHECK_ARG(x && y, "hpe_maxpool3x3s2: null pointer");
  HPE_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0, "hpe_maxpool3x3s2: C must be a positive multiple of 8");
  HPE_CHECK_ARG((((uintptr_t)x | (uintptr_t)y) & 15) == 0, "hpe_maxpool3x3s2: misaligned pointer");
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  const long long total = (long long)B * Ho * Wo * (C / 8);
  const int grid = (int)((total + 255) / 256 > (long long)num_sms() * 16 ? (long long)num_sms() * 16 : (total + 255) / 256);
  maxpool_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, B, H, W, C, (__nv_bfloat16*)y);
  return launch_status("maxpool_kernel");
}
