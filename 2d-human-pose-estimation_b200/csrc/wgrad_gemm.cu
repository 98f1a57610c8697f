// Weight-gradient GEMM for the convolution stack on sm_100a (tcgen05 + TMEM + TMA), used by the training step.
//
//   dW[(tap, l-channel)][r-channel] = sum over output pixels p of  L_tap[p][l-channel] * R[p][r-channel]
//
// The reduction (GEMM-K) dimension is the PIXEL index, which is the strided dimension of NHWC tensors, so both
// operands are MN-major: a TMA box {64 channels, brick of 64 pixels} lands in shared memory as 64 rows (pixels)
// x 128 B (64 bf16 channels) with the 128-byte swizzle, which is exactly the canonical SWIZZLE_128B MN-major UMMA
// layout (atom = 64 MN elements x 8 K rows; K groups 1024 B apart = SBO; 64-channel blocks 8 KB apart = LBO).
//   * L ("left", the M side) is the tensor that carries the filter-tap shift / stride: the layer input x for
//     convolutions (tap offset dy,dx, element stride 1|2 via the TMA map, zero padding = TMA OOB fill), dz for the
//     transposed convolutions (dW[ci,co,ky,kx] = sum x[iy,ix,ci] dz[2iy-1+ky, 2ix-1+kx, co]).
//     An M tile = two 64-channel blocks, each with its own (tap, channel block).
//   * R ("right", the N side, BLOCK_N <= 256 channels) is read unshifted on the pixel grid.
//   * split-K over pixel bricks across the persistent grid; every (tile, split) item writes its fp32 partial tile with plain
//     stores into ITS OWN slice of the workspace [splits][n_lblocks*64][N]; a small kernel then adds the slices in a fixed
//     order while permuting into the PyTorch weight layout.  No atomics, no memset: the weight gradients -- like the forward --
//     are run-to-run bit-stable (round 1 accumulated with red.global.add.f32, whose order varies from run to run).
// Warp roles: warp 0 TMA producer, warp 1 TMEM alloc + MMA issuer, warps 2-9 epilogue (TMEM -> red.global).
#include "hpe_common.cuh"
#include "ptx_sm100.cuh"
#include "tma_host.cuh"

namespace hpe {
using namespace ptx;

constexpr int kWgThreads  = 320;
constexpr int kWgEpiWarp0 = 2;
constexpr int kWgEpiWarps = 8;
constexpr int kBrickPx    = 64;                     // pixels per pipeline stage (4 MMAs of K=16)
constexpr int kBlkBytes   = kBrickPx * 128;         // one 64-channel block of one brick: 8 KB

struct WgradParams {
  int kind, swapped;
  int n_lblocks, lblocks_per_tap;
  int n_mtiles, n_ntiles, splits, bricks, bricks_per_split, num_items;
  int kw, kh, knb, tiles_x, tiles_y;
  int lstride;
  int out_cols;
  long long ws_slice;          // elements of one split's slice of the workspace
  float* ws;
};

__device__ __forceinline__ void wg_tap_offset(int kind, int t, int& dx, int& dy) {
  if (kind == HPE_CONV_3x3)          { dy = t / 3 - 1; dx = t % 3 - 1; }
  else if (kind == HPE_CONV_STEM7x7) { dy = t - 2;     dx = 0; }
  else if (kind == HPE_CONV_DECONV4) { dy = (t >> 2) - 1; dx = (t & 3) - 1; }
  else { dx = 0; dy = 0; }
}

// MN-major SWIZZLE_128B operands: 64-element MN blocks kBlkBytes apart (LBO), 8-row K groups 1024 B apart (SBO); the 64-bit
// descriptors are assembled in the MMA warp from a per-stage low word and a constant high word
__host__ __device__ constexpr uint32_t make_idesc_bf16_m128_mnmajor(int n) {
  return make_idesc_bf16_m128(n) | (1u << 15) | (1u << 16);
}

__device__ __forceinline__ void st_v4f(float* p, float a, float b, float c, float d) {
  asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <int BLOCK_N, int STAGES>
struct WgSmem {
  static constexpr int kLBytes = 2 * kBlkBytes;
  static constexpr int kRBytes = (BLOCK_N / 64) * kBlkBytes;
  static constexpr int kStageBytes = kLBytes + kRBytes;
  static constexpr int kOffBar = STAGES * kStageBytes;
  static constexpr int kNumBar = 2 * STAGES + 4;
  static constexpr int kDynamic = kOffBar + kNumBar * 8 + 16 + 1024;
};

template <int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(kWgThreads, 1)
wgrad_gemm_kernel(const __grid_constant__ CUtensorMap tmL, const __grid_constant__ CUtensorMap tmR, const WgradParams p) {
  using S = WgSmem<BLOCK_N, STAGES>;
  constexpr int kTmemCols = 2 * BLOCK_N < 32 ? 32 : 2 * BLOCK_N;
  constexpr uint32_t kIdesc = make_idesc_bf16_m128_mnmajor(BLOCK_N);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars       = reinterpret_cast<uint64_t*>(smem + S::kOffBar);
  uint64_t* full_bar   = bars;
  uint64_t* empty_bar  = bars + STAGES;
  uint64_t* tfull_bar  = bars + 2 * STAGES;
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;
  uint32_t* tmem_slot  = reinterpret_cast<uint32_t*>(bars + S::kNumBar);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmL); prefetch_tmap(&tmR);
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], kWgEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 1) { tmem_alloc(tmem_slot, kTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch_dependents();
  griddep_wait();

  // work item -> (m tile, n tile, brick range)
  auto item_range = [&](int item, int& mt, int& nt, int& q0, int& q1, int* split = nullptr) {
    const int sp = item % p.splits; int r = item / p.splits;
    if (split) *split = sp;
    nt = r % p.n_ntiles; mt = r / p.n_ntiles;
    q0 = sp * p.bricks_per_split;
    q1 = q0 + p.bricks_per_split; if (q1 > p.bricks) q1 = p.bricks;
  };

  if (warp == 0) {
    // producer: the whole warp runs the loop converged, one elected lane issues (see conv_gemm.cu: issued from a divergent
    // `if (lane == 0)` every UTMALDG is wrapped in an R2UR.BROADCAST loop and the 2 + BLOCK_N/64 loads of a stage cost more
    // issue time than the stage's 512 cycles of MMA work)
    int stage = 0; uint32_t phase = 0;
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
      int mt, nt, q0, q1; item_range(item, mt, nt, q0, q1);
      int lch[2], ldx[2], ldy[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        int lb = 2 * mt + h; if (lb >= p.n_lblocks) lb = p.n_lblocks - 1;
        const int tap = lb / p.lblocks_per_tap;
        lch[h] = (lb - tap * p.lblocks_per_tap) * 64;
        wg_tap_offset(p.swapped ? HPE_CONV_1x1 : p.kind, tap, ldx[h], ldy[h]);
      }
      // brick coordinates advance incrementally (no per-stage division)
      int tx = q0 % p.tiles_x, r0 = q0 / p.tiles_x, ty = r0 % p.tiles_y, tb = r0 / p.tiles_y;
      for (int q = q0; q < q1; ++q) {
        const int x0 = tx * p.kw, y0 = ty * p.kh, b0 = tb * p.knb;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        const uint32_t fb = smem_u32(&full_bar[stage]);
        const uint32_t sL = smem_u32(smem + stage * S::kStageBytes), sR = sL + S::kLBytes;
        mbar_arrive_expect_tx_elect(fb, S::kStageBytes);
#pragma unroll
        for (int h = 0; h < 2; ++h)
          tma_load_4d_elect(sL + h * kBlkBytes, &tmL, fb, lch[h], x0 * p.lstride + ldx[h], y0 * p.lstride + ldy[h], b0);
#pragma unroll
        for (int j = 0; j < BLOCK_N / 64; ++j)
          tma_load_4d_elect(sR + j * kBlkBytes, &tmR, fb, nt * BLOCK_N + j * 64, x0, y0, b0);
        if (++tx == p.tiles_x) { tx = 0; if (++ty == p.tiles_y) { ty = 0; ++tb; } }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // MMA issuer: converged warp, elected issue, descriptors as 32-bit halves (constant high word)
    constexpr uint32_t kDescHi = (1024u >> 4) | (1u << 14) | (2u << 29);     // SBO = 1024 B, version 1, SWIZZLE_128B
    constexpr uint32_t kLboLo  = static_cast<uint32_t>(kBlkBytes >> 4) << 16;  // 64-element MN blocks 8 KB apart
    int stage = 0; uint32_t phase = 0; int it = 0;
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x, ++it) {
      int mt, nt, q0, q1; item_range(item, mt, nt, q0, q1);
      const int acc = it & 1;
      mbar_wait(&tempty_bar[acc], ((it >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
      for (int q = q0; q < q1; ++q) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sL = smem_u32(smem + stage * S::kStageBytes);
        const uint32_t a_lo = ((sL & 0x3FFFFu) >> 4) | kLboLo;
        const uint32_t b_lo = (((sL + S::kLBytes) & 0x3FFFFu) >> 4) | kLboLo;
#pragma unroll
        for (int k = 0; k < kBrickPx / 16; ++k)     // 16 pixel rows = 2048 B per MMA -> +128 in the (addr>>4) field
          umma_bf16_elect_lo(tmem_d, a_lo + 128 * k, kDescHi, b_lo + 128 * k, kDescHi, kIdesc, ((q - q0) | k) != 0);
        umma_commit_elect(&empty_bar[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      umma_commit_elect(&tfull_bar[acc]);
    }
  } else {
    // epilogue: TMEM lane quarter = warp & 3; the two warps of a quarter alternate 32-column groups
    const int ew = warp & 3;
    const int half = (warp - kWgEpiWarp0) >> 2;
    const int row = ew * 32 + lane;                 // tile row: block h = row / 64, channel-in-block = row % 64
    int it = 0;
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x, ++it) {
      int mt, nt, q0, q1, sp; item_range(item, mt, nt, q0, q1, &sp);
      const int acc = it & 1;
      mbar_wait(&tfull_bar[acc], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + acc * BLOCK_N;
      const int lb = 2 * mt + (row >> 6);
      const bool live = lb < p.n_lblocks;
      float* dst = p.ws + static_cast<size_t>(sp) * p.ws_slice + (static_cast<size_t>(lb) * 64 + (row & 63)) * p.out_cols + nt * BLOCK_N;
      for (int g = half; g * 32 < BLOCK_N; g += 2) {
        uint32_t v[32];
        tmem_ld_x32(taddr + g * 32, v);
        tmem_ld_wait();
        if (live) {
#pragma unroll
          for (int e = 0; e < 32; e += 4)
            st_v4f(dst + g * 32 + e, __uint_as_float(v[e]), __uint_as_float(v[e + 1]), __uint_as_float(v[e + 2]),
                       __uint_as_float(v[e + 3]));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, kTmemCols); }
}

// workspace [rows][cols] fp32 -> PyTorch weight layout (optionally accumulating)
__global__ void __launch_bounds__(256)
wgrad_unpack_kernel(const float* __restrict__ ws, int kind, int swapped, int Cin, int Cout, int cols, long long total,
                    float* __restrict__ dw, int accumulate, int splits, long long ws_slice) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    // The walk follows the WORKSPACE layout where that differs from the weight layout (consecutive threads = consecutive output
    // channels of one workspace row), so the `splits` loads per element are coalesced and the single store is the strided access:
    // walking the weight layout made every load of a warp touch 32 different lines (9 us per launch, 57-65 launches per step).
    size_t src, dst = static_cast<size_t>(i);
    if (kind == HPE_CONV_1x1 && swapped) {              // dw[co][ci] <- ws[co][ci]
      const int co = static_cast<int>(i / Cin), ci = static_cast<int>(i % Cin);
      src = static_cast<size_t>(co) * cols + ci;
    } else if (kind == HPE_CONV_1x1) {                  // dw[co][ci] <- ws[ci][co]
      const int co = static_cast<int>(i % Cout), ci = static_cast<int>(i / Cout);
      src = static_cast<size_t>(ci) * cols + co;
      dst = static_cast<size_t>(co) * Cin + ci;
    } else if (kind == HPE_CONV_FINAL) {                // dw[k][ci]  <- ws[ci][kpad]
      const int k = static_cast<int>(i / Cin), ci = static_cast<int>(i % Cin);
      src = static_cast<size_t>(ci) * cols + k;
    } else if (kind == HPE_CONV_3x3) {                  // dw[co][ci][tap] <- ws[tap*Cin + ci][co]
      const int co = static_cast<int>(i % Cout), ci = static_cast<int>((i / Cout) % Cin), tap = static_cast<int>(i / (static_cast<long long>(Cout) * Cin));
      src = (static_cast<size_t>(tap) * Cin + ci) * cols + co;
      dst = (static_cast<size_t>(co) * Cin + ci) * 9 + tap;
    } else if (kind == HPE_CONV_STEM7x7) {              // dw[co][c][ky][kx] <- ws[khp*64 + kwp*16 + (py*8+px*4+c)][co]
      const int kx = static_cast<int>(i % 7), ky = static_cast<int>((i / 7) % 7), c = static_cast<int>((i / 49) % 3),
                co = static_cast<int>(i / 147);
      const int khp = (ky + 1) >> 1, py = (ky + 1) & 1, kwp = (kx + 1) >> 1, px = (kx + 1) & 1;
      src = static_cast<size_t>(khp * 64 + kwp * 16 + py * 8 + px * 4 + c) * cols + co;
    } else {                                            // deconv: dw[ci][co][ky][kx] <- ws[(ky*4+kx)*Cout + co][ci]
      const int t = static_cast<int>(i % 16), co = static_cast<int>((i / 16) % Cout), ci = static_cast<int>(i / (16LL * Cout));
      src = (static_cast<size_t>(t) * Cout + co) * cols + ci;
    }
    float v = 0.f;
#pragma unroll 8
    for (int sp = 0; sp < splits; ++sp) v += __ldg(ws + static_cast<size_t>(sp) * ws_slice + src);      // fixed order: deterministic
    dw[dst] = accumulate ? dw[dst] + v : v;
  }
}

}  // namespace hpe

using namespace hpe;

struct hpe_wgrad_plan {
  hpe_conv_desc d;
  WgradParams p;
  int block_n, grid;
  int Ho, Wo;               // conv output size
  int gW, gH;               // pixel grid the bricks walk (the unstrided tensor's spatial size)
  int Lc, Rc;               // channel counts of the left / right tensors
  size_t smem, ws_elems;
  CUtensorMap tmL, tmR;
  const void *l, *r;
  const void* fn;
};

static size_t wg_weight_elems(const hpe_conv_desc* d) {
  switch (d->kind) {
    case HPE_CONV_3x3:     return (size_t)9 * d->Cin * d->Cout;
    case HPE_CONV_STEM7x7: return (size_t)64 * 147;
    case HPE_CONV_DECONV4: return (size_t)16 * d->Cin * d->Cout;
    default:               return (size_t)d->Cin * d->Cout;
  }
}

typedef void (*WgKernelFn)(const CUtensorMap, const CUtensorMap, const WgradParams);

extern "C" int hpe_wgrad_plan_create(const hpe_conv_desc* d, hpe_wgrad_plan** out) {
  HPE_CHECK_ARG(d && out, "hpe_wgrad_plan_create: null pointer");
  int Ho = 0, Wo = 0;
  int rc = hpe_conv_out_hw(d, &Ho, &Wo);      // validates the descriptor
  if (rc) return rc;
  if (d->pool || d->head2 || d->Cin2) return fail(HPE_EUNSUPPORTED, "hpe_wgrad_plan_create: describe the plain layer (no fused pool / head / second input)");
  rc = hpe_check_device();
  if (rc) return rc;
  hpe_wgrad_plan* pl = new hpe_wgrad_plan();
  pl->d = *d; pl->Ho = Ho; pl->Wo = Wo;
  WgradParams& p = pl->p;
  p.kind = d->kind; p.swapped = 0; p.lstride = 1;
  int taps = 1;
  switch (d->kind) {
    case HPE_CONV_1x1:
      p.lstride = d->stride;
      if (d->stride == 1 && d->Cout > d->Cin) { p.swapped = 1; pl->Lc = d->Cout; pl->Rc = d->Cin; }
      else { pl->Lc = d->Cin; pl->Rc = d->Cout; }
      pl->gW = Wo; pl->gH = Ho; break;
    case HPE_CONV_3x3:
      taps = 9; p.lstride = d->stride; pl->Lc = d->Cin; pl->Rc = d->Cout; pl->gW = Wo; pl->gH = Ho; break;
    case HPE_CONV_STEM7x7:
      taps = 4; pl->Lc = 64; pl->Rc = 64; pl->gW = Wo; pl->gH = Ho; break;
    case HPE_CONV_DECONV4:      // left = dz (2H x 2W, stride 2, 16 taps), right = x
      taps = 16; p.lstride = 2; pl->Lc = d->Cout; pl->Rc = d->Cin; pl->gW = d->W; pl->gH = d->H; break;
    default:                    // FINAL: left = x, right = dz padded to 64 joints
      if (d->Cout > 64) { delete pl; return fail(HPE_EUNSUPPORTED, "hpe_wgrad_plan_create: final conv wgrad supports at most 64 joints"); }
      pl->Lc = d->Cin; pl->Rc = 64; pl->gW = d->W; pl->gH = d->H; break;
  }
  p.lblocks_per_tap = pl->Lc / 64;
  p.n_lblocks = taps * p.lblocks_per_tap;
  p.n_mtiles = (p.n_lblocks + 1) / 2;
  pl->block_n = pl->Rc % 256 == 0 ? 256 : (pl->Rc % 128 == 0 ? 128 : 64);
  p.n_ntiles = pl->Rc / pl->block_n;
  p.out_cols = pl->Rc;
  int kw = 1, kh = 1, knb = kBrickPx;
  pick_brick(kBrickPx, pl->gW, pl->gH, d->B, p.lstride == 2 ? 64 : 64, &kw, &kh, &knb);
  p.kw = kw; p.kh = kh; p.knb = knb;
  p.tiles_x = (pl->gW + kw - 1) / kw; p.tiles_y = (pl->gH + kh - 1) / kh;
  const long long bricks = (long long)p.tiles_x * p.tiles_y * ((d->B + knb - 1) / knb);
  if (bricks > 0x7fffffff) { delete pl; return fail(HPE_EUNSUPPORTED, "hpe_wgrad_plan_create: too many pixel bricks"); }
  p.bricks = (int)bricks;
  const int tiles = p.n_mtiles * p.n_ntiles;
  int splits = num_sms() / tiles; if (splits < 1) splits = 1;
  if (splits > (p.bricks + 3) / 4) splits = (p.bricks + 3) / 4;
  p.bricks_per_split = (p.bricks + splits - 1) / splits;
  p.splits = (p.bricks + p.bricks_per_split - 1) / p.bricks_per_split;
  p.num_items = tiles * p.splits;
  pl->grid = p.num_items < num_sms() ? p.num_items : num_sms();
  p.ws_slice = (long long)p.n_lblocks * 64 * p.out_cols;
  pl->ws_elems = (size_t)p.ws_slice * p.splits;
  WgKernelFn fn;
  switch (pl->block_n) {
    case 64:  fn = wgrad_gemm_kernel<64, 8>;  pl->smem = WgSmem<64, 8>::kDynamic; break;
    case 128: fn = wgrad_gemm_kernel<128, 6>; pl->smem = WgSmem<128, 6>::kDynamic; break;
    default:  fn = wgrad_gemm_kernel<256, 4>; pl->smem = WgSmem<256, 4>::kDynamic; break;
  }
  pl->fn = (const void*)fn;
  cudaError_t e = cudaFuncSetAttribute(pl->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl->smem);
  if (e != cudaSuccess) { delete pl; return fail(HPE_ECUDA, "cudaFuncSetAttribute(wgrad smem=%zu) failed: %s", pl->smem, cudaGetErrorString(e)); }
  pl->l = pl->r = nullptr;
  *out = pl;
  return HPE_OK;
}

extern "C" void hpe_wgrad_plan_destroy(hpe_wgrad_plan* plan) { delete plan; }

extern "C" size_t hpe_wgrad_workspace_elems(const hpe_wgrad_plan* plan) { return plan ? plan->ws_elems : 0; }

static int wg_encode(hpe_wgrad_plan* pl, const void* x, const void* dz) {
  const hpe_conv_desc& d = pl->d;
  const WgradParams& p = pl->p;
  const bool left_is_dz = p.swapped || d.kind == HPE_CONV_DECONV4;
  const void* l = left_is_dz ? dz : x;
  const void* r = left_is_dz ? x : dz;
  const uint32_t ones[4] = {1, 1, 1, 1};
  int rc;
  if (l != pl->l) {
    HPE_CHECK_ARG(((uintptr_t)l & 15) == 0, "hpe_wgrad_run: tensors must be 16-byte aligned");
    const uint32_t s = (uint32_t)p.lstride;
    const uint32_t box[4] = {64, (uint32_t)p.kw * s, (uint32_t)p.kh * s, (uint32_t)p.knb};
    const uint32_t es[4] = {1, s, s, 1};
    if (d.kind == HPE_CONV_STEM7x7) {
      const uint64_t Hs = d.H / 2, Ws = d.W / 2, Wp = Ws + 4;
      const uint64_t dims[4] = {64, Ws, Hs, (uint64_t)d.B};
      const uint64_t str[3] = {32, Wp * 32, Hs * Wp * 32};
      rc = encode_map(&pl->tmL, l, 4, dims, str, box, ones, "wgrad stem L");
    } else {
      // spatial size of the left tensor: x (H x W) for convs, dz (2H x 2W) for the deconv, dz (Ho x Wo) if swapped
      const uint64_t LW = d.kind == HPE_CONV_DECONV4 ? 2 * d.W : (p.swapped ? pl->Wo : d.W);
      const uint64_t LH = d.kind == HPE_CONV_DECONV4 ? 2 * d.H : (p.swapped ? pl->Ho : d.H);
      const uint64_t C = pl->Lc;
      const uint64_t dims[4] = {C, LW, LH, (uint64_t)d.B};
      const uint64_t str[3] = {C * 2, LW * C * 2, LH * LW * C * 2};
      rc = encode_map(&pl->tmL, l, 4, dims, str, box, es, "wgrad L");
    }
    if (rc) return rc;
    pl->l = l;
  }
  if (r != pl->r) {
    HPE_CHECK_ARG(((uintptr_t)r & 15) == 0, "hpe_wgrad_run: tensors must be 16-byte aligned");
    const uint64_t C = pl->Rc;
    const uint64_t dims[4] = {C, (uint64_t)pl->gW, (uint64_t)pl->gH, (uint64_t)d.B};
    const uint64_t str[3] = {C * 2, (uint64_t)pl->gW * C * 2, (uint64_t)pl->gH * pl->gW * C * 2};
    const uint32_t box[4] = {64, (uint32_t)p.kw, (uint32_t)p.kh, (uint32_t)p.knb};
    rc = encode_map(&pl->tmR, r, 4, dims, str, box, ones, "wgrad R");
    if (rc) return rc;
    pl->r = r;
  }
  return HPE_OK;
}

extern "C" int hpe_wgrad_run(hpe_wgrad_plan* pl, const void* x, const void* dz, float* workspace, float* dw_torch,
                             int accumulate, void* stream) {
  HPE_CHECK_ARG(pl && x && dz && workspace && dw_torch, "hpe_wgrad_run: null pointer");
  HPE_CHECK_ARG(((uintptr_t)workspace & 15) == 0, "hpe_wgrad_run: workspace must be 16-byte aligned");
  int rc = wg_encode(pl, x, dz);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  WgradParams p = pl->p;                       // (no memset: every slice element that is read is written by exactly one item)
  p.ws = workspace;
  void* args[3] = {&pl->tmL, &pl->tmR, &p};
  cudaError_t e = cudaLaunchKernel(pl->fn, dim3(pl->grid), dim3(kWgThreads), args, pl->smem, st);
  if (e != cudaSuccess) return fail(HPE_ECUDA, "launch of wgrad_gemm_kernel failed: %s", cudaGetErrorString(e));
  const long long total = (long long)wg_weight_elems(&pl->d);
  const int grid = (int)((total + 255) / 256 > 2048 ? 2048 : (total + 255) / 256);
  wgrad_unpack_kernel<<<grid, 256, 0, st>>>(workspace, pl->d.kind, p.swapped, pl->d.Cin, pl->d.Cout, p.out_cols, total,
                                           dw_torch, accumulate, p.splits, p.ws_slice);
  return launch_status("wgrad_unpack_kernel");
}
