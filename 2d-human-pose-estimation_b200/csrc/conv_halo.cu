// "y-halo" implicit-GEMM convolution for the layers whose generic kernel is bound by operand delivery from L2
// (profiles/: one TMA box per filter tap moves 7x-27x the algorithmic bytes through the L2->SM fabric):
//   * the 7x7/s2 stem on the space-to-depth image (4 row taps, K = 64 per tap, Cout = 64),
//   * 3x3 / stride 1 / 64 -> 64 (ResNet layer1 conv2; also its backward-data conv).
// An M tile is an 8-wide x 16-tall brick of ONE image, so a filter-row shift dy is a shift of 8 shared-memory rows =
// 1024 bytes = exactly one SWIZZLE_128B atom: the A operand of every row tap is the SAME halo'd TMA box
// {64 ch, 8 px, 16 + taps_y - 1 rows} read through a 1024-byte-aligned descriptor offset.  Column taps (3x3: dx = -1,0,1)
// use one such box each.  The whole packed weight (<= 72 KB) is loaded into shared memory once per CTA.
// L2->SM traffic per 128-pixel tile: stem 19 KB instead of 96 KB, 3x3 54 KB instead of 216 KB.
//
// Warp roles as in conv_gemm.cu: warp 0 TMA producer, warp 1 TMEM + tcgen05.mma issuer, warps 3-10 epilogue
// (tcgen05.ld -> scale/shift -> ReLU|SiLU -> bf16 -> swizzled smem -> TMA store).
#include "hpe_common.cuh"
#include "ptx_sm100.cuh"
#include "tma_host.cuh"
#include "conv_halo.cuh"

namespace hpe {
using namespace ptx;

constexpr int kHThreads = 352, kHEpiThreads = 256, kHEpiWarp0 = 3;
constexpr int kHN = 64;                       // Cout = BLOCK_N
constexpr int kHNumSB = 2;                    // epilogue staging buffers (16 KB each)
constexpr int kHSBBytes = 128 * 64 * 2;

__device__ __forceinline__ uint4 h_lds_v4(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ float4 h_lds_f4(uint32_t addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void h_sts_v4(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ uint32_t h_max2(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}

// K-major SWIZZLE_128B descriptor from a shared-memory byte address with an explicit 8-row-group stride.  Measured on B200:
// the 128-byte swizzle is a function of the ABSOLUTE shared-memory address bits, so a TMA-written tile can be read through
// any 128-byte-aligned start (row shift) with base offset 0 -- this is what makes the column taps of the wide box work.
__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr, uint32_t sbo_bytes) {
  return static_cast<uint64_t>((saddr & 0x3FFFFu) >> 4) | (static_cast<uint64_t>(1) << 16) |
         (static_cast<uint64_t>(sbo_bytes >> 4) << 32) | (static_cast<uint64_t>(1) << 46) | (static_cast<uint64_t>(2) << 61);
}

// TAPS_Y row taps per box, BOXES column taps (one halo'd box each)
// POOL (stem, ReLU): the epilogue takes MaxPool2d(3, 2, 1) of the tile while it is in shared memory and stores 3x7 pooled pixels;
// tiles overlap by one output row/column (brick origin (6*tx - 1, 14*ty - 1)) so every pooling window lies inside one brick.
// Convolution positions outside the image are staged as 0, which is the identity of max over ReLU outputs (each window holds
// at least one real position), so the result equals maxpool(relu(bn(conv))) of the unfused pair bit for bit.
template <int TAPS_Y, int BOXES, int ACT, int kHStages, int TX = 1, bool POOL = false>
__global__ void __launch_bounds__(kHThreads, 1)
conv_yhalo_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                  const __grid_constant__ CUtensorMap tmOut, const HaloParams p) {
  constexpr int kTaps = TAPS_Y * BOXES * TX;
  constexpr int kBW = 8 + TX - 1;                                  // box width in pixels (TX column taps inside one box)
  constexpr int kBoxRows = (16 + TAPS_Y - 1) * kBW;                // smem rows (128 B each) of one halo'd box
  constexpr int kBoxBytes = (kBoxRows * 128 + 1023) / 1024 * 1024;
  constexpr int kABytes = BOXES * kBoxBytes;
  constexpr int kWBytes = kTaps * kHN * 128;                       // [tap][64 rows][64 k] bf16
  constexpr int kOffW = kHStages * kABytes;
  constexpr int kOffSB = kOffW + kWBytes;
  constexpr int kOffPB = kOffSB + kHNumSB * kHSBBytes;            // POOL: 2 pooled staging buffers, 21 rows x 128 B in 3 KB each
  constexpr int kPBBytes = 3072;
  constexpr int kOffSS = kOffPB + (POOL ? 2 * kPBBytes : 0);
  constexpr int kOffBar = kOffSS + 2 * kHN * 4;
  constexpr uint32_t kIdesc = make_idesc_bf16_m128(kHN);
  constexpr int kTmemCols = 128;                                   // 2 accumulators x 64 columns

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
    prefetch_tmap(&tmA); prefetch_tmap(&tmW); prefetch_tmap(&tmOut);
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
      // written by the pack kernel earlier in the stream: keep it after griddep_wait)
      mbar_arrive_expect_tx(wfull_bar, kWBytes);
      for (int t = 0; t < kTaps; ++t) tma_load_2d(sW + t * kHN * 128, &tmW, wfull_bar, t * 64, 0);
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        const int tx = tile % p.tiles_x, ty = (tile / p.tiles_x) % p.tiles_y, b = tile / (p.tiles_x * p.tiles_y);
        const int x0 = POOL ? 6 * tx - 1 : tx * 8, y0 = POOL ? 14 * ty - 1 : ty * 16;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[stage], BOXES * kBoxRows * 128);      // bytes the boxes deliver (a stage may be padded to 1 KB)
#pragma unroll
        for (int bx = 0; bx < BOXES; ++bx)
          tma_load_4d(sA + stage * kABytes + bx * kBoxBytes, &tmA, &full_bar[stage], 0, x0 + bx - p.x_lo, y0 - p.y_lo, b);
        if (++stage == kHStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 || warp == 2) {
    // TWO issuing warps: warp 1 takes the even tiles (accumulator 0), warp 2 the odd ones (accumulator 1), so the issue paths
    // of consecutive tiles overlap; the tensor pipe serialises the execution.
    // The whole warp walks the tile loop (uniform control flow keeps the stage / descriptor arithmetic in uniform registers);
    // only lane 0 issues the tcgen05 instructions.  With N = 64 an MMA occupies the tensor pipe for just 32 cycles, so the
    // issue path must stay that short.
    mbar_wait(wfull_bar, 0);
    int stage = 0; uint32_t phase = 0; int it = 0;
    const uint32_t sA0 = smem_u32(sA), sW0 = smem_u32(sW);
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
      const int acc = it & 1;
      if (acc != warp - 1) { if (++stage == kHStages) { stage = 0; phase ^= 1; } continue; }
      mbar_wait(&tempty_bar[acc], ((it >> 1) & 1) ^ 1);
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * kHN;
      const uint32_t a_stage = sA0 + stage * kABytes;
#pragma unroll
      for (int bx = 0; bx < BOXES; ++bx)
#pragma unroll
        for (int ty = 0; ty < TAPS_Y; ++ty)
#pragma unroll
          for (int tx = 0; tx < TX; ++tx) {
            const int tap = ty * (BOXES * TX) + bx * TX + tx;       // packed weight order: tap = ky * kw + kx
            // low descriptor word = (address >> 4) | LBO field; high word = SBO | version | SWIZZLE_128B (constants)
            const uint32_t a_lo = (((a_stage + bx * kBoxBytes + (ty * kBW + tx) * 128) & 0x3FFFFu) >> 4) | (1u << 16);
            const uint32_t b_lo = (((sW0 + tap * kHN * 128) & 0x3FFFFu) >> 4) | (1u << 16);
            constexpr uint32_t a_hi = ((kBW * 128) >> 4) | (1u << 14) | (2u << 29);
            constexpr uint32_t b_hi = (1024 >> 4) | (1u << 14) | (2u << 29);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16_elect_lo(tmem_d, a_lo + 2 * k, a_hi, b_lo + 2 * k, b_hi, kIdesc, (bx | ty | tx | k) != 0);
          }
      umma_commit_elect(&empty_bar[stage]); umma_commit_elect(&tfull_bar[acc]);
      __syncwarp();
      if (++stage == kHStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp >= kHEpiWarp0) {
    // Two independent epilogue groups of 4 warps: group g drains accumulator g (tiles with it % 2 == g) through its own
    // staging buffer, so the tcgen05.ld -> math -> fence -> TMA-store chains of consecutive tiles overlap.
    const int ew = warp & 3, grp = (warp - kHEpiWarp0) >> 2;
    const int row = ew * 32 + lane;
    const bool leader = (warp - kHEpiWarp0) % 4 == 0 && lane == 0;
    const uint32_t rowp = smem_u32(sSB + grp * kHSBBytes) + row * 128;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + grp * kHN;
    int it = 0;
    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0}; long long t0 = 0, t1;
#define PROF(i) if (p.prof) { t1 = clock64(); pc[i] += t1 - t0; t0 = t1; }
    if (p.prof) t0 = clock64();
    // group g takes tiles blockIdx.x + (2k + g) * gridDim.x; (tx, ty, b) advance incrementally (no divisions on the per-tile path)
    const int step = 2 * gridDim.x, per_img = p.tiles_x * p.tiles_y;
    const int s_b = step / per_img, s_y = (step % per_img) / p.tiles_x, s_x = (step % per_img) % p.tiles_x;
    int tile = blockIdx.x + grp * gridDim.x;
    int b = tile / per_img, ty = (tile % per_img) / p.tiles_x, tx = (tile % per_img) % p.tiles_x;
    for (; tile < p.num_tiles; tile += step, ++it) {
      const int x0 = POOL ? 6 * tx - 1 : tx * 8, y0 = POOL ? 14 * ty - 1 : ty * 16;
      // POOL: this thread's convolution position (row = y * 8 + x inside the brick) may lie outside the image
      const bool inside = !POOL || (static_cast<unsigned>(x0 + (row & 7)) < static_cast<unsigned>(p.Wc) &&
                                    static_cast<unsigned>(y0 + (row >> 3)) < static_cast<unsigned>(p.Hc));
      PROF(0)
      mbar_wait(&tfull_bar[grp], it & 1);
      tc_fence_after();
      PROF(1)
      uint32_t v[64];
      tmem_ld_x32(taddr, v);
      tmem_ld_x32(taddr + 32, v + 32);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[grp]);                // accumulator drained: hand it back before anything else
      PROF(2)
      const uint32_t sb = smem_u32(sSB + grp * kHSBBytes);
      if (p.direct) {
        // Staging buffer -> global memory with plain 16-byte stores (8 lanes per 128-byte pixel): no proxy fence, no bulk-store
        // wait and, without the pool, no block barrier -- a warp re-reads only the 32 rows it staged itself.
        if (POOL) named_bar_sync(1 + grp, 128); else __syncwarp();   // the previous tile's reads of the buffer are done
      } else {
        if (leader) bulk_wait_read<0>();                           // this group's previous store has left its staging buffer
        named_bar_sync(1 + grp, 128);                              // leader's wait is visible before anyone writes the buffer
      }
      PROF(3)
      // scale/shift through explicit ld.shared, fetched one chunk ahead of the st.shared of the previous one
      const uint32_t ss = smem_u32(sScale), hs = smem_u32(sShift);
      float4 nsa = h_lds_f4(ss), nsb = h_lds_f4(ss + 16), nha = h_lds_f4(hs), nhb = h_lds_f4(hs + 16);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 sa = nsa, sb4 = nsb, ha = nha, hb = nhb;
        if (j < 7) { nsa = h_lds_f4(ss + (j + 1) * 32); nsb = h_lds_f4(ss + (j + 1) * 32 + 16); nha = h_lds_f4(hs + (j + 1) * 32); nhb = h_lds_f4(hs + (j + 1) * 32 + 16); }
        const float sc[8] = {sa.x, sa.y, sa.z, sa.w, sb4.x, sb4.y, sb4.z, sb4.w};
        const float sh[8] = {ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w};
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          o[e] = fmaf(__uint_as_float(v[j * 8 + e]), sc[e], sh[e]);
          if (ACT == HPE_ACT_RELU) o[e] = fmaxf(o[e], 0.f);
          if (ACT == HPE_ACT_SILU) o[e] = o[e] / (1.f + __expf(-o[e]));
        }
        uint4 pk = make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
        if (POOL && !inside) pk = make_uint4(0u, 0u, 0u, 0u);
        h_sts_v4(rowp + ((j ^ (row & 7)) << 4), pk);
      }
      PROF(4)
      if (POOL) {
        named_bar_sync(3 + grp, 128);                              // the whole brick is staged
        PROF(5)
        const uint32_t pb = smem_u32(smem + kOffPB + grp * kPBBytes);
        const int t = ew * 32 + lane;
#pragma unroll
        for (int o = t; o < 21 * 8; o += 128) {                    // (pooled pixel, 16-byte channel chunk)
          const int pp = o >> 3, j = o & 7, ppy = pp / 3, ppx = pp - 3 * ppy;
          uint4 m = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
          for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
              const int r = (2 * ppy + ky) * 8 + 2 * ppx + kx;
              const uint4 q = h_lds_v4(sb + r * 128 + ((j ^ (r & 7)) << 4));
              m.x = h_max2(m.x, q.x); m.y = h_max2(m.y, q.y); m.z = h_max2(m.z, q.z); m.w = h_max2(m.w, q.w);
            }
          if (p.direct) {
            const int oy = 7 * ty + ppy, ox = 3 * tx + ppx;
            if (oy < p.Ho && ox < p.Wo)
              reinterpret_cast<uint4*>(p.out)[((static_cast<size_t>(b) * p.Ho + oy) * p.Wo + ox) * 8 + j] = m;
          } else {
            h_sts_v4(pb + pp * 128 + ((j ^ (pp & 7)) << 4), m);
          }
        }
        if (!p.direct) {
          fence_proxy_async_smem();
          named_bar_sync(5 + grp, 128);
          if (leader) { tma_store_4d(&tmOut, smem + kOffPB + grp * kPBBytes, 0, 3 * tx, 7 * ty, b); bulk_commit(); }
        }
      } else if (p.direct) {
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; ++i) {                              // 4 pixels (512 contiguous bytes of one image row) per instruction
          const int r = ew * 32 + i * 4 + (lane >> 3), c = lane & 7;
          const uint4 q = h_lds_v4(sb + r * 128 + ((c ^ (r & 7)) << 4));
          const int oy = y0 + (r >> 3), ox = x0 + (r & 7);
          if (oy < p.Ho && ox < p.Wo)
            reinterpret_cast<uint4*>(p.out)[((static_cast<size_t>(b) * p.Ho + oy) * p.Wo + ox) * 8 + c] = q;
        }
      } else {
        fence_proxy_async_smem();
        named_bar_sync(3 + grp, 128);
        if (leader) { tma_store_4d(&tmOut, sSB + grp * kHSBBytes, 0, x0, y0, b); bulk_commit(); }
      }
      PROF(6)
      tx += s_x; if (tx >= p.tiles_x) { tx -= p.tiles_x; ++ty; }
      ty += s_y; if (ty >= p.tiles_y) { ty -= p.tiles_y; ++b; }
      b += s_b;
    }
    if (p.prof && blockIdx.x == 0 && lane == 0 && (warp == kHEpiWarp0 || warp == kHEpiWarp0 + 5))
      for (int i = 0; i < 8; ++i) p.prof[(warp == kHEpiWarp0 ? 0 : 8) + i] = pc[i];
    if (leader) bulk_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, kTmemCols); }
}

template <int TAPS_Y, int BOXES, int kHStages, int TX = 1, bool POOL = false>
static size_t halo_smem_bytes() {
  return (POOL ? 2 * 3072 : 0) + (size_t)kHStages * BOXES * (((16 + TAPS_Y - 1) * (8 + TX - 1) * 128 + 1023) / 1024 * 1024) + (size_t)TAPS_Y * BOXES * TX * kHN * 128 + kHNumSB * kHSBBytes + 2 * kHN * 4 +
         (2 * kHStages + 5) * 8 + 16 + 1024;
}

typedef void (*HaloFn)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const HaloParams);

int halo_launch(int taps_y, int boxes, int act, int pool, const CUtensorMap* tmA, const CUtensorMap* tmW, const CUtensorMap* tmOut,
                const HaloParams& p, int grid, cudaStream_t stream) {
  HaloFn fn = nullptr; size_t smem = 0;
  if (pool) {
    if (taps_y != 4 || boxes != 1 || act != HPE_ACT_RELU) return fail(HPE_EUNSUPPORTED, "halo_launch: the fused max-pool exists for the ReLU stem only");
    smem = halo_smem_bytes<4, 1, 6, 1, true>();
    fn = conv_yhalo_kernel<4, 1, HPE_ACT_RELU, 6, 1, true>;
  } else if (taps_y == 4 && boxes == 1) {
    smem = halo_smem_bytes<4, 1, 6>();            // 19 KB per tile: 6 tiles in flight hide the HBM latency of the image read
    fn = act == HPE_ACT_RELU ? conv_yhalo_kernel<4, 1, HPE_ACT_RELU, 6> : (act == HPE_ACT_SILU ? conv_yhalo_kernel<4, 1, HPE_ACT_SILU, 6> : conv_yhalo_kernel<4, 1, HPE_ACT_NONE, 6>);
  } else if (taps_y == 3 && boxes == 3) {
    smem = halo_smem_bytes<3, 3, 2>();            // 54 KB per tile + 72 KB of resident weights
    fn = act == HPE_ACT_RELU ? conv_yhalo_kernel<3, 3, HPE_ACT_RELU, 2> : (act == HPE_ACT_SILU ? conv_yhalo_kernel<3, 3, HPE_ACT_SILU, 2> : conv_yhalo_kernel<3, 3, HPE_ACT_NONE, 2>);
  } else if (taps_y == 3 && boxes == 0) {          // wide box: 23 KB per tile, 4 tiles in flight
    smem = halo_smem_bytes<3, 1, 4, 3>();
    fn = act == HPE_ACT_RELU ? conv_yhalo_kernel<3, 1, HPE_ACT_RELU, 4, 3> : (act == HPE_ACT_SILU ? conv_yhalo_kernel<3, 1, HPE_ACT_SILU, 4, 3> : conv_yhalo_kernel<3, 1, HPE_ACT_NONE, 4, 3>);
  } else {
    return fail(HPE_EUNSUPPORTED, "halo_launch: no kernel for %d x %d taps", taps_y, boxes);
  }
  cudaError_t e = cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return fail(HPE_ECUDA, "cudaFuncSetAttribute(halo smem=%zu) failed: %s", smem, cudaGetErrorString(e));
  HaloParams pp = p;
  static long long* prof_dev = nullptr;
  const bool prof = getenv("HPE_B200_HALO_PROF") != nullptr;
  if (prof && !prof_dev) cudaMalloc(&prof_dev, 16 * sizeof(long long));
  pp.prof = prof ? prof_dev : nullptr;
  void* args[4] = {(void*)tmA, (void*)tmW, (void*)tmOut, &pp};
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kHThreads); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  e = cudaLaunchKernelExC(&cfg, (const void*)fn, args);
  if (e != cudaSuccess) return fail(HPE_ECUDA, "launch of conv_yhalo_kernel failed: %s", cudaGetErrorString(e));
  if (prof) {
    long long h[16];
    cudaDeviceSynchronize(); cudaMemcpy(h, prof_dev, sizeof(h), cudaMemcpyDeviceToHost);
    const int tiles = (p.num_tiles + grid - 1) / grid / 2;
    fprintf(stderr, "[halo prof] tiles/group %d; cycles per tile: ", tiles);
    for (int i = 0; i < 16; ++i) fprintf(stderr, "%s%lld", i == 8 ? " | " : " ", h[i] / (tiles ? tiles : 1));
    fprintf(stderr, "\n");
  }
  return launch_status("conv_yhalo_kernel");
}

}  // namespace hpe
