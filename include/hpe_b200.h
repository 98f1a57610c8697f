/*
 * hpe_b200.h -- C ABI of libhpe_b200.so: the B200 (sm_100a) device path for the top-down heatmap
 * pose hot path of AI-sangbeom/2D-Human-Pose-Estimation.
 *
 * The reference is pure Python/PyTorch and defines no FFI ("the plugin API is Python
 * duck-typing").  Each entry point below therefore cites the reference *call site* (file:line under
 * /root/reference) whose device work it replaces, or -- for the stages the reference does not
 * implement -- the paragraph of SURVEY.md Appendix A that specifies it.
 *
 * Conventions (all entry points):
 *   - return 0 on success, a negative HPE_E* code otherwise; hpe_last_error() returns a thread-local
 *     human-readable message for the last failure on the calling thread;
 *   - stream-ordered on the caller's cudaStream_t (passed as void*; NULL = legacy default stream);
 *     no allocation, no host synchronisation, no host<->device copy inside unless the name ends in
 *     `_host` (those take HOST pointers and do the copies on `stream`, then synchronise it);
 *   - every buffer is caller-owned device memory (plain pointers + sizes; no torch types);
 *   - one host thread per stream; different streams may be driven from different threads.
 */
#ifndef HPE_B200_H_
#define HPE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HPE_OK            0
#define HPE_EINVAL       -1   /* bad argument (null pointer, non-positive size, bad dtype, misalignment) */
#define HPE_ECUDA        -2   /* a CUDA runtime/driver call or kernel launch failed                   */
#define HPE_EUNSUPPORTED -3   /* shape outside what the sm_100a kernels cover                          */
#define HPE_EARCH        -4   /* device is not compute capability 10.x                                 */

#define HPE_F32  0
#define HPE_BF16 1

/* activation codes of the fused conv epilogue */
#define HPE_ACT_NONE 0
#define HPE_ACT_RELU 1        /* timm Bottleneck act (resnet.py:244-270)                               */
#define HPE_ACT_SILU 2        /* first-party ConvBlock default act (HPE/models/nn/modules/block.py:20-45) */
#define HPE_ACT_GELU 3        /* exact erf GELU of the ConvNeXt block MLP (HPE/thirdparty/dinov3/dinov3/models/convnext.py:60) */
#define HPE_ACT_GELU_LN 4     /* GELU of a 1x1 conv with the LayerNorm in front of it folded in (convnext.py:74-77: norm -> pwconv1 -> act); run with hpe_conv_run_ln */

int         hpe_abi_version(void);
const char* hpe_last_error(void);
/* 0 if the current device can run the sm_100a kernels, HPE_EARCH otherwise. */
int         hpe_check_device(void);

/* ------------------------------------------------------------------------------------------------
 * Heatmap decode.  Replaces the decode the reference's metric/test loop would run on the output of
 * `output = model(data)` (HPE/engine/trainer.py:90, :110-113; HPE/test.py:335-361 is the reference's
 * regression analogue).  Semantics: SURVEY.md App. A.2 (first-max argmax, coords zeroed when
 * maxval <= 0), A.3 (quarter-pixel, strictly interior peaks) and A.4 (per-sample 2x3 inverse affine),
 * one pass over the heatmaps.
 *   hm      [B,K,H,W] dtype HPE_F32|HPE_BF16, contiguous
 *   affine  [B,2,3] f32 or NULL (identity: heatmap coordinates are returned)
 *   refine  0 = A.2 only (get_max_preds), 1 = A.2+A.3
 *   coords  [B,K,2] f32 out; maxvals [B,K] f32 out; idx [B,K] i32 out or NULL
 * ---------------------------------------------------------------------------------------------- */
int hpe_decode_heatmaps(const void* hm, int dtype, int B, int K, int H, int W,
                        const float* affine, int refine,
                        float* coords, float* maxvals, int32_t* idx, void* stream);

/* Same decode writing packed rows xyv[B,K,3] = (x, y, maxval) f32: the layout the data-parallel result gather ships
 * (one all-gather of (B/W)*K*12 bytes per rank), so no concatenation kernel follows the decode. */
int hpe_decode_heatmaps_xyv(const void* hm, int dtype, int B, int K, int H, int W,
                            const float* affine, int refine, float* xyv, int32_t* idx, void* stream);

/* Same call with HOST buffers (pageable or pinned): H2D of hm/affine, kernel, D2H of the results on
 * `stream`, then a stream synchronise.  `d_scratch` is caller-owned DEVICE scratch of at least
 * hpe_decode_host_scratch_bytes(...) bytes. */
size_t hpe_decode_host_scratch_bytes(int dtype, int B, int K, int H, int W);
int hpe_decode_heatmaps_host(const void* h_hm, int dtype, int B, int K, int H, int W,
                             const float* h_affine, int refine,
                             float* h_coords, float* h_maxvals, int32_t* h_idx,
                             void* d_scratch, size_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Target generation + target-weighted joints MSE.  Replaces `loss = criterion(output, target,
 * target_weight[B,K,1])` (HPE/engine/trainer.py:92) and the dataloader-side target rasterisation.
 * Semantics: SURVEY.md App. A.5 (Gaussian targets) and A.6 (JointsMSELoss, use_target_weight=True).
 *
 * Fused mode (targets never materialised):
 *   pred     [B,K,H,W] f32|bf16          joints [B,K,3] f32 (x, y in input px, visibility)
 *   partials [B*K] f32 scratch (per-map sum of w^2 (p-g)^2, fixed reduction order)
 *   loss     [1] f32 out = 0.5/(K*B*H*W) * sum(partials)   (second, single-CTA, fixed-order stage)
 *   grad     [B,K,H,W] in pred's dtype or NULL; when non-NULL also writes
 *            dL/dpred = grad_scale * w^2 (p-g)/(K*B*H*W) in the same pass
 *   weight_out [B,K] f32 or NULL: the effective target weight (A.5, incl. the out-of-map rule)
 * ---------------------------------------------------------------------------------------------- */
int hpe_joints_mse_fused(const void* pred, int dtype, const float* joints,
                         int B, int K, int H, int W, float sigma, float stride_x, float stride_y,
                         float* partials, float* loss, void* grad, float grad_scale,
                         float* weight_out, void* stream);

/* Allocation-free variant of hpe_joints_mse_fused.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments and results (bit-identical) minus `partials`: the per-map partial sums live in a 512 KB array
 * built into the library (every GPU that loads it holds the array until exit), so B*K <= hpe_static_loss_max_maps() = 131072;
 * larger problems are rejected with HPE_EUNSUPPORTED.  The call allocates nothing (no cudaMalloc / cudaFree / allocating
 * library routine).  All calls share the array: calls that overlap on one GPU (different streams or threads) race on it --
 * serialise them or use hpe_joints_mse_fused with separate scratch.  Nothing needs clearing between calls (each call rewrites
 * the B*K entries it reads).  The device the pointers live on must be current. */
int hpe_static_loss_max_maps(void);
int hpe_joints_mse_fused_noalloc(const void* pred, int dtype, const float* joints,
                                 int B, int K, int H, int W, float sigma, float stride_x, float stride_y,
                                 float* loss, void* grad, float grad_scale, float* weight_out, void* stream);

/* Drop-in mode: materialised target [B,K,H,W] f32 and weight [B,K] f32 (the trainer's call shape). */
int hpe_joints_mse_dense(const void* pred, int dtype, const float* target, const float* weight,
                         int B, int K, int H, int W,
                         float* partials, float* loss, void* grad, float grad_scale, void* stream);

/* A.5 alone: rasterise target [B,K,H,W] f32 and weight [B,K] f32 from joints. */
int hpe_generate_target(const float* joints, int B, int K, int H, int W,
                        float sigma, float stride_x, float stride_y,
                        float* target, float* weight, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Convolution stack (NHWC bf16 activations, fp32 accumulation on tcgen05/TMEM, operands by TMA).
 * Replaces the cuDNN conv + BatchNorm2d + activation (+ residual add) launches behind
 * `output = model(data)` (HPE/engine/trainer.py:90) for the config-default backbone:
 * timm ResNet-50 stem/Bottleneck/forward_features (HPE/thirdparty/pytorch-image-models/timm/models/
 * resnet.py:554-576, 244-270, 742-756), and -- for the first-party conv unit -- ConvBlock.forward_fuse
 * (HPE/models/nn/modules/block.py:44-45).  The deconv head and final 1x1 follow SURVEY.md App. A.1.
 *
 * A plan owns the TMA descriptors and launch geometry for ONE layer at ONE shape; create once, run
 * every step.  Tensors (device, caller-owned):
 *   x        [B,H,W,Cin]     bf16 NHWC         w  packed bf16, see hpe_conv_pack_weight_elems()
 *   scale,shift [Cout] f32   folded BN (y = acc*scale + shift), shift doubles as conv bias
 *   residual [B,Ho,Wo,Cout]  bf16 or NULL      y  [B,Ho,Wo,Cout] bf16 NHWC
 *                                              (HPE_CONV_FINAL: y = heatmaps [B,Cout,Ho,Wo] f32 NCHW)
 * ---------------------------------------------------------------------------------------------- */
typedef struct hpe_conv_plan hpe_conv_plan;

#define HPE_CONV_1x1      0   /* 1x1, stride 1|2, pad 0                                         */
#define HPE_CONV_3x3      1   /* 3x3, stride 1|2, pad 1                                         */
#define HPE_CONV_STEM7x7  2   /* 7x7 s2 p3 on the space-to-depth image written by hpe_preprocess */
#define HPE_CONV_DECONV4  3   /* ConvTranspose2d k4 s2 p1 as four 2x2 sub-pixel convolutions    */
#define HPE_CONV_FINAL    4   /* 1x1 -> K joints (+bias), writes fp32 NCHW heatmaps             */
/* backward-data kinds (training): built with hpe_conv_dgrad_desc / hpe_conv_pack_weight_dgrad        */
#define HPE_CONV_DGRAD3x3S2 5 /* dgrad of 3x3/s2/p1: four output-parity classes with 1/2/2/4 taps  */
#define HPE_CONV_DGRAD1x1S2 6 /* dgrad of 1x1/s2: GEMM stored to the even pixels, rest zeroed      */
#define HPE_CONV_4x4S2      7 /* dgrad of the k4/s2/p1 transposed conv = 4x4 stride-2 conv, 16 taps */
#define HPE_CONV_2x2S2      8 /* 2x2 stride-2 pad-0 conv: ConvNeXt downsample layers (dinov3 convnext.py:163-167) */

typedef struct hpe_conv_desc {
  int kind;                 /* HPE_CONV_*                                         */
  int B, H, W, Cin, Cout;   /* input geometry (H,W = input height/width)          */
  int stride;               /* 1 or 2 (ignored for STEM/DECONV/FINAL)             */
  int act;                  /* HPE_ACT_*                                          */
  int has_residual;         /* epilogue adds `residual` before the activation     */
  int in_cstride;           /* channels between consecutive pixels of x (0 = Cin): x may be a channel slice of a wider NHWC tensor */
  int out_cstride;          /* same for y (the residual stays dense); 0 = Cout.  This is how torch.cat along channels disappears:
                               producers write straight into their slice (HPE/models/nn/modules/block.py:100,113,142-155).
                               HPE_CONV_FINAL: channel count of the NCHW tensor y is a slice of (head.py:176-179) */
  int Cin2;                 /* != 0 (HPE_CONV_1x1, stride 1 only): a SECOND 1x1 conv over x2 [B,Ho*stride2,Wo*stride2,Cin2] is
                               accumulated into the same output tile before the epilogue -- the projection shortcut of a timm
                               Bottleneck (`shortcut = self.downsample(x)`; `x += shortcut`, resnet.py:247-268) fused into conv3,
                               so the shortcut tensor is never written or re-read.  Both BN scales are folded into the weights
                               by the caller (scale = 1, shift = shift3 + shift_ds). */
  int stride2;              /* stride of the second conv: 1 or 2 */
  int head2;                /* K != 0 (K <= 32; ReLU HPE_CONV_DECONV4 or stride-1 1x1/3x3 with Cout = 256): the layer's bf16 output is
                               NOT written; a second GEMM against the K-joint 1x1 weights runs on the tile while it is still in shared
                               memory and fp32 NCHW heatmaps [B,K,Ho,Wo] are written instead -- the last deconv stage + final_layer of
                               the App. A.1 head in one kernel.  Run with hpe_conv_run_head2. */
  int pool;                 /* != 0 (HPE_CONV_STEM7x7 with HPE_ACT_RELU only): `x = self.maxpool(x)` of timm ResNet.forward_features
                               (MaxPool2d(3, 2, 1) after conv1/bn1/act1) is taken in the stem's epilogue; y is the POOLED
                               [B, Hp, Wp, 64] tensor and the H/2 x W/2 stem output is never written.  Bit-identical to
                               hpe_conv_run followed by hpe_maxpool3x3s2. */
  int chain;                /* N2 = 64 | 128 (dense ReLU HPE_CONV_1x1, stride 1, Cout = 256): the NEXT layer's 1x1 conv 256 -> N2 (+ folded
                               BN + ReLU) runs on every output tile while it is staged for its store -- `x = self.act3(x)` of one timm
                               Bottleneck and `x = self.conv1(x); x = self.bn1(x); x = self.act1(x)` of the next (resnet.py:247-270) in
                               one kernel.  y is written as usual (it is the next block's shortcut); the chained output y2
                               [B,Ho,Wo,N2] is bit-identical to running the second conv as its own layer.  Run with
                               hpe_conv_run_chain. */
} hpe_conv_desc;
/* Channel counts: Cin and Cout must be multiples of 8 (16-byte pixels); partial 64-channel blocks are zero-filled by TMA. */

int  hpe_conv_plan_create(const hpe_conv_desc* desc, hpe_conv_plan** out);
void hpe_conv_plan_destroy(hpe_conv_plan* plan);
/* number of bf16 elements of the packed weight buffer for this layer */
size_t hpe_conv_packed_weight_elems(const hpe_conv_desc* desc);
/* Pack a PyTorch-layout fp32 weight (Conv2d [Cout,Cin,kh,kw]; ConvTranspose2d [Cin,Cout,4,4]) that
 * lives on the DEVICE into the kernel's bf16 GEMM-B layout (device).  Stream-ordered. */
int  hpe_conv_pack_weight(const hpe_conv_desc* desc, const float* w_torch, void* w_packed, void* stream);
int  hpe_conv_run(hpe_conv_plan* plan, const void* x, const void* w_packed,
                  const float* scale, const float* shift, const void* residual, void* y, void* stream);
/* two-input variant (desc.Cin2 != 0): w_packed from hpe_conv_pack_weight2 = [Cout][ceil64(Cin) + ceil64(Cin2)];
 * x2 is the dense bf16 NHWC tensor the second 1x1 conv reads with stride2 */
int  hpe_conv_pack_weight2(const hpe_conv_desc* desc, const float* w_torch, const float* w2_torch, void* w_packed, void* stream);
int  hpe_conv_run2(hpe_conv_plan* plan, const void* x, const void* x2, const void* w_packed,
                   const float* scale, const float* shift, const void* residual, void* y, void* stream);
/* fused-head variant (desc.head2 = K): w2_packed = the final 1x1 weights packed with hpe_conv_pack_weight (kind HPE_CONV_FINAL,
 * Cin 256, Cout K) into a ZERO-initialised buffer of 32*256 bf16; bias2 [K] f32; heatmaps [B,K,Ho,Wo] f32 NCHW */
int  hpe_conv_run_head2(hpe_conv_plan* plan, const void* x, const void* w_packed, const float* scale, const float* shift,
                        const void* w2_packed, const float* bias2, float* heatmaps, void* stream);
/* LayerNorm-folded variant (desc.act = HPE_ACT_GELU_LN; plain 1x1 / stride 1 layers on the cta_group::2 kernels): computes
 * `x = self.act(self.pwconv1(self.norm(x)))` of the ConvNeXt block (HPE/thirdparty/dinov3/dinov3/models/convnext.py:74-77) from
 * the RAW x: w_packed = pwconv1's weights times the LayerNorm gamma (per input channel), colsum[n] = sum over the input
 * channels of the bf16-rounded folded weights, shift[n] = b[n] + sum_c W[n][c] beta[c]; rowstats / n_slices as written by
 * hpe_dwconv7_bias_stats, ln_channels = C, eps = the LayerNorm's.  y = gelu(rstd_r (x W'^T - mu_r colsum) + shift). */
int  hpe_conv_run_ln(hpe_conv_plan* plan, const void* x, const void* w_packed, const float* colsum, const float* shift,
                     const float* rowstats, int n_slices, int ln_channels, float eps, void* y, void* stream);
/* chained variant (desc.chain = N2): w2_packed = the next 1x1 conv's weights packed with hpe_conv_pack_weight (kind HPE_CONV_1x1,
 * Cin 256, Cout N2); scale2 / shift2 [N2] its folded BN; y2 [B,Ho,Wo,N2] bf16 NHWC.  x2 as in hpe_conv_run2 (NULL without Cin2). */
int  hpe_conv_run_chain(hpe_conv_plan* plan, const void* x, const void* x2, const void* w_packed, const float* scale,
                        const float* shift, const void* residual, void* y, const void* w2_packed, const float* scale2,
                        const float* shift2, void* y2, void* stream);
/* Batched packing (training re-packs every layer's forward and backward-data operands each step): fill one job record per
 * (layer, direction) on the HOST (hpe_pack_job_bytes() bytes each, consecutive block ranges: pass the running sum of
 * *n_blocks as block_begin), copy the records to the device once, then ONE launch re-packs everything. */
size_t hpe_pack_job_bytes(void);
int  hpe_conv_pack_job_fill(const hpe_conv_desc* fwd_desc, int dgrad, const float* w_torch, void* w_packed, int block_begin,
                            void* job_host, int* n_blocks);
int  hpe_conv_pack_jobs_run(const void* jobs_device, int n_jobs, int total_blocks, void* stream);
/* Launch order / L2 eviction hints of a plan (performance only, results bit-identical; no counterpart in the reference, whose
 * cuDNN calls schedule themselves).  reverse_tiles != 0 walks the tiles last to first: layer n+1 then starts on the part of its
 * input that layer n wrote last, which is still in L2.  l2_out / l2_res / l2_in: 0 none, 1 evict-first, 2 evict-last on the
 * output stores / the residual loads / the input (A operand) loads.  Plans on the y-halo kernel take the direction only.
 * Returns the reverse flag in effect, < 0 on error. */
int  hpe_conv_set_schedule(hpe_conv_plan* plan, int reverse_tiles, int l2_out, int l2_res, int l2_in);
/* output spatial size of the layer */
int  hpe_conv_out_hw(const hpe_conv_desc* desc, int* Ho, int* Wo);

/* fp32 NCHW image [B,3,H,W] -> bf16 space-to-depth NHWC staging the stem conv reads
 * (replaces `data.float()` + cuDNN's layout transform, HPE/engine/trainer.py:88-90). */
size_t hpe_preprocess_out_elems(int B, int H, int W);
int  hpe_preprocess(const float* img_nchw, int B, int H, int W, void* s2d_out, void* stream);
/* same from a uint8 NCHW image: the reference's `imgs.float() / 255.0` (HPE/train.py:117-120, HPE/test.py:16-29) fused in */
int  hpe_preprocess_u8(const unsigned char* img_nchw, int B, int H, int W, void* s2d_out, void* stream);
/* MaxPool 3x3 s2 p1 on NHWC bf16 (timm resnet.py:574-576). */
int  hpe_maxpool3x3s2(const void* x, int B, int H, int W, int C, void* y, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Training step (BASELINE configs[2]; call sites `loss.backward()` / `optimizer.step()`,
 * HPE/engine/trainer.py:96-100; in the reference these are cuDNN dgrad/wgrad + ATen BatchNorm backward
 * launched by autograd).  Activation gradients are bf16 NHWC, weight gradients fp32 in the PyTorch
 * weight layout.
 *
 * Weight gradient of one conv layer described by the FORWARD descriptor:
 *   x   the layer's forward input  (bf16 NHWC [B,H,W,Cin]; STEM: the hpe_preprocess staging buffer)
 *   dz  gradient w.r.t. the conv output (bf16 NHWC [B,Ho,Wo,Cout]; FINAL: [B,H,W,64], joints zero-padded)
 *   workspace  fp32 scratch of hpe_wgrad_workspace_elems(plan) elements
 *   dw_torch   fp32, same shape as the layer's PyTorch weight; overwritten, or added to if accumulate != 0
 * ---------------------------------------------------------------------------------------------- */
/* Backward-data convolution of a forward layer: `g` receives the descriptor of the conv that maps
 * dz [B,Ho,Wo,Cout] to dx [B,H,W,Cin] (run it with hpe_conv_plan_create/hpe_conv_run; the epilogue's
 * residual input doubles as gradient accumulation: dx = dgrad(dz) + residual).  FINAL: dz is the
 * [B,H,W,64] zero-padded NHWC bf16 heatmap gradient.  The stem has no dgrad (image needs no gradient). */
int hpe_conv_dgrad_desc(const hpe_conv_desc* fwd_desc, hpe_conv_desc* g);
/* pack the FORWARD PyTorch-layout fp32 weight for that backward-data conv
 * (hpe_conv_packed_weight_elems(g) bf16 elements) */
int hpe_conv_pack_weight_dgrad(const hpe_conv_desc* fwd_desc, const float* w_torch, void* w_packed, void* stream);

typedef struct hpe_wgrad_plan hpe_wgrad_plan;
int    hpe_wgrad_plan_create(const hpe_conv_desc* fwd_desc, hpe_wgrad_plan** out);
void   hpe_wgrad_plan_destroy(hpe_wgrad_plan* plan);
size_t hpe_wgrad_workspace_elems(const hpe_wgrad_plan* plan);
int    hpe_wgrad_run(hpe_wgrad_plan* plan, const void* x, const void* dz, float* workspace, float* dw_torch,
                     int accumulate, void* stream);


/* BatchNorm2d in training mode over an NHWC bf16 tensor z [M = B*H*W][C] (timm resnet.py:244-270 runs
 * nn.BatchNorm2d with batch statistics under model.train(), HPE/engine/trainer.py:77-80).
 * hpe_bn_stats: batch mean / biased variance -> scale = gamma*rstd, shift = beta - mean*scale, saved mean / rstd,
 *   running statistics updated with `momentum` (unbiased variance), as torch does.
 *   scratch = fp32 device scratch of hpe_bn_scratch_elems(C) elements (per-block partial sums, added in a fixed
 *   order: no atomics, the statistics are bit-reproducible run to run).
 * hpe_bn_apply: y = act(z*scale + shift (+ residual)), act = ReLU if relu != 0.
 * hpe_bn_backward: g = relu ? dy*(y>0) : dy;  dbeta = sum g;  dgamma = sum g*xhat;
 *   dz = gamma*rstd*(g - dbeta/M - xhat*dgamma/M);  g_out (optional) = g, the gradient of the residual input. */
size_t hpe_bn_scratch_elems(int C);
int hpe_bn_stats(const void* z, long long M, int C, const float* gamma, const float* beta, float eps, float momentum,
                 float* running_mean, float* running_var, float* scratch, float* scale, float* shift,
                 float* mean, float* rstd, void* stream);
int hpe_bn_apply(const void* z, long long M, int C, const float* scale, const float* shift, const void* residual,
                 int relu, void* y, void* stream);
int hpe_bn_backward(const void* dy, const void* y, const void* z, long long M, int C, const float* mean,
                    const float* rstd, const float* gamma, int relu, float* scratch, float* dbeta, float* dgamma,
                    void* dz, void* g_out, void* stream);
/* Same with y == NULL allowed for ReLU units WITHOUT a residual: the mask y > 0 is recomputed as z*scale + shift > 0 from the
 * forward's own scale/shift (hpe_bn_stats outputs), so the backward reads 4 instead of 6 bytes per element. */
int hpe_bn_backward_zmask(const void* dy, const void* y, const void* z, long long M, int C, const float* mean,
                          const float* rstd, const float* gamma, int relu, const float* scale, const float* shift,
                          float* scratch, float* dbeta, float* dgamma, void* dz, void* g_out, void* stream);
/* ReLU units WITH a residual: the forward writes one mask byte per 8 channels (bit e = pre-activation of channel 8j+e > 0,
 * bits[M*C/8]) next to y; the backward reads the mask instead of y (1/16 of the bytes). */
int hpe_bn_apply_bits(const void* z, long long M, int C, const float* scale, const float* shift, const void* residual,
                      int relu, void* y, uint8_t* bits, void* stream);
int hpe_bn_backward_bits(const void* dy, const uint8_t* bits, const void* z, long long M, int C, const float* mean,
                         const float* rstd, const float* gamma, float* scratch, float* dbeta, float* dgamma, void* dz,
                         void* g_out, void* stream);
/* MaxPool 3x3 s2 p1 that records the winning window position (first max, torch semantics) and its backward. */
int hpe_maxpool3x3s2_train(const void* x, int B, int H, int W, int C, void* y, uint8_t* idx, void* stream);
int hpe_maxpool3x3s2_backward(const void* dy, const uint8_t* idx, int B, int H, int W, int C, void* dx, void* stream);
/* dL/dheatmap [B,K,H,W] NCHW (f32|bf16) -> [B,H,W,64] bf16 NHWC, joints zero-padded; dbias[K] (optional) = sum. */
int hpe_heatmap_grad_nhwc(const void* dhm, int dtype, int B, int K, int H, int W, void* out, float* dbias, void* stream);
/* torch.optim.SGD step (momentum, weight decay, dampening 0) over a flat fp32 parameter buffer
 * (`optimizer.step()`, HPE/engine/trainer.py:98-100): g' = grad*grad_scale + wd*p; m = mu*m + g'; p -= lr*m. */
int hpe_sgd_step(float* param, const float* grad, float* momentum_buf, long long n, float lr, float momentum,
                 float weight_decay, float grad_scale, void* stream);
/* The optimizer of the reference's WORKING trainer: optim.AdamW (two learning-rate groups, HPE/train.py:52-55) after
 * torch.nn.utils.clip_grad_norm_(model.parameters(), 10.0) (HPE/train.py:136).
 * hpe_grad_sumsq: out[0] (+)= sum grad^2 over a flat fp32 buffer, deterministic two-stage reduction; `scratch` holds at least
 *   hpe_grad_sumsq_scratch_elems() floats.  Call once per parameter group with accumulate = 1 to get the global norm.
 * hpe_adamw_step: decoupled weight decay, bias correction for the 1-based `step`, no amsgrad (torch.optim.AdamW):
 *   clip = min(1, max_norm / (sqrt(*grad_sumsq) * |grad_scale| + 1e-6))  (grad_sumsq NULL or max_norm <= 0: no clipping)
 *   g = grad*grad_scale*clip;  p *= 1 - lr*wd;  m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;
 *   p -= lr/(1-b1^step) * m / (sqrt(v)/sqrt(1-b2^step) + eps).   The norm is read on the device: no host synchronisation. */
size_t hpe_grad_sumsq_scratch_elems(void);
int hpe_grad_sumsq(const float* grad, long long n, float* scratch, float* out, int accumulate, void* stream);
int hpe_adamw_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n, float lr, float beta1,
                   float beta2, float eps, float weight_decay, int step, float grad_scale, const float* grad_sumsq,
                   float max_norm, void* stream);

/* ------------------------------------------------------------------------------------------------
 * The reference's OWN pose-head post-processing, loss and metrics: these stages
 * exist in the reference, so their oracle is pinned against the reference's outputs (tests/golden/detpose_golden.npz).
 *
 * hpe_pose_decode = PoseHead._inference_pose + kpts_decode + make_anchors
 *   (HPE/models/nn/modules/head.py:185-211, HPE/models/utils.py:15-27).
 *   levels[l]  raw head map [B, nc + nkpts*ndim, Hs[l], Ws[l]] (dtype f32|bf16, NCHW), strides[l] = 8/16/32
 *   out        [B, nc + nkpts*ndim, A] f32, A = sum Hs*Ws: sigmoid(cls) | (raw-0.5+anchor)*stride | sigmoid(vis)
 * hpe_pose_nms = non_max_suppression_pose (HPE/test.py:69-185) incl. torchvision.ops.nms (greedy, stable order):
 *   pred [B,C,A] f32 (output of hpe_pose_decode); per image: score = max class prob > conf_thres, first-max label,
 *   box = min/max of the keypoint xy, class offset label*7680, IoU > iou_thres (compared as double) suppresses.
 *   Outputs are fixed-capacity: boxes [B,max_det,4], scores [B,max_det], labels [B,max_det] i32,
 *   keypoints [B,max_det,nkpts,3] (conf = 1 when ndim == 2), count [B] i32 = detections kept (descending score).
 *   scratch = f32 device scratch of hpe_pose_nms_scratch_elems(B, A) elements.  A <= 16384.
 * ---------------------------------------------------------------------------------------------- */
int hpe_pose_decode(const void* const* levels, const int* Hs, const int* Ws, const float* strides, int n_levels,
                    int dtype, int B, int nc, int nkpts, int ndim, float* out, void* stream);
size_t hpe_pose_nms_scratch_elems(int B, int A);
int hpe_pose_nms(const float* pred, int B, int C, int A, int nc, int nkpts, float conf_thres, double iou_thres,
                 int max_det, float* scratch, float* boxes, float* scores, int32_t* labels, float* keypoints,
                 int32_t* count, void* stream);

/* hpe_compute_loss = ComputeLoss.__call__ (HPE/loss.py:349-448; varifocal class term over all cells, wing-OKS +
 * 0.5 L1 + 0.1 smoothness keypoint term, BCE visibility term, robust keypoint box + CIoU soft label) and -- when
 * `grads` is non-NULL -- its analytic gradient (what `loss.backward()` produces, HPE/train.py training step).
 *   preds[l]  [B, nc + 3*nkpts, Hs[l], Ws[l]] f32 NCHW (the head's training output, head.py:176-181)
 *   targets   [n_targets, 6 + 3*nkpts] f32 rows (batch, class, cx, cy, w, h, (x, y, v)*K), normalised (HPE/dataset.py:76-86)
 *   sigmas    [nkpts] f32 (COCO sigmas / 10 for 17 keypoints, loss.py:250-257)
 *   loss_out  [4] f32 = {total, cls, kpt, vis};  grads[l] same shape as preds[l], overwritten with grad_scale * dL/dpred
 *   workspace device scratch of hpe_compute_loss_workspace_bytes(...) bytes (zeroed inside, stream-ordered). */
size_t hpe_compute_loss_workspace_bytes(int n_levels, int B, const int* Hs, const int* Ws, int nc, int n_targets);
int hpe_compute_loss(const float* const* preds, const int* Hs, const int* Ws, int n_levels, int B, int nc, int nkpts,
                     const float* targets, int n_targets, const float* sigmas, void* workspace, size_t workspace_bytes,
                     float* loss_out, float* const* grads, float grad_scale, void* stream);

/* Keypoint metrics / losses on decoded keypoints (HPE/core/metric/oksap.py:32-74, pck.py:14-104, HPE/core/loss/keypoint.py:13-21).
 *   hpe_oks: gt [N,K,3] (x,y,v), pred [N,K,2], areas [N], sigmas [K] -> out [N]
 *   hpe_pck: gt,pred [N,K,2], visible [N,K] u8 or NULL, normalizer NULL (gt bbox) | [1] | [N] -> counts[2] i64 = {correct, visible};
 *            scratch = i32 [2*N]
 *   hpe_keypoint_loss: pred/gt rows of pred_stride/gt_stride floats (x,y first), mask [N,K], area [N], sigmas [K]
 *            -> loss [1]; scratch = f32 [N] */
int hpe_oks(const float* gt, const float* pred, const float* areas, const float* sigmas, int N, int K,
            float in_vis_thres, float eps, float* out, void* stream);
int hpe_pck(const float* gt, const float* pred, const unsigned char* visible, const float* normalizer, int norm_n,
            int N, int K, float alpha, float eps, int32_t* scratch, long long* counts, void* stream);
int hpe_keypoint_loss(const float* pred, int pred_stride, const float* gt, int gt_stride, const float* mask,
                      const float* area, const float* sigmas, int N, int K, float* scratch, float* loss, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Bandwidth-bound NHWC bf16 ops of the reference's real pose model (DINOv3Pose, HPE/models/pose.py:9-65: ConvNeXt
 * backbone + FeatureAdaptor + SPPF + PAN + PoseHead).  Every tensor takes an explicit channel
 * pitch (elements between consecutive pixels) so producers can write into a slice of a concat buffer.
 *   hpe_dwconv            depthwise ksize x ksize (3|7), pad ksize/2, stride 1.  w_taps = fp32 [ksize*ksize][C].
 *                         mode 0: y = SiLU(acc*p0 + p1)            (ConvBlock(k=3, g=C), head.py:143; block.py:20-45)
 *                         mode 1: y = LayerNorm_C(acc + p0; gamma=p1, beta=p2, eps)   (ConvNeXt Block, convnext.py:73-76)
 *   hpe_layernorm_nhwc    per-pixel LayerNorm over C (channels_first LayerNorm, convnext.py:112-117)
 *   hpe_patchify4         fp32 NCHW [B,3,H,W] -> bf16 [B,H/4,W/4,64]: the 48 inputs (c,ky,kx) of the 4x4/s4 stem conv
 *                         (convnext.py:157-160) + 16 zeros, so the stem is a 1x1 GEMM with Cin = 64
 *   hpe_maxpool5          MaxPool2d(5,1,2) (SPPF, block.py:102-114)
 *   hpe_upsample_bilinear F.interpolate(size=(Ho,Wo), mode='bilinear', align_corners=False) (PAN, block.py:132-151)
 * ---------------------------------------------------------------------------------------------- */
int hpe_dwconv(const void* x, int in_pitch, int B, int H, int W, int C, int ksize, int mode, const float* w_taps,
               const float* p0, const float* p1, const float* p2, float eps, void* y, int out_pitch, void* stream);
/* depthwise 7x7 + bias only (FMA-bound tiled formulation; the LayerNorm that follows is hpe_layernorm_nhwc) */
int hpe_dwconv7_bias(const void* x, int in_pitch, int B, int H, int W, int C, const float* w_taps, const float* bias,
                     void* y, int out_pitch, void* stream);
/* ... that also writes, per output pixel and 64-channel slice, (sum, sum of squares) of the STORED bf16 values:
 * rowstats [B*H*W][ceil(C/64)][2] fp32 = hpe_dwconv7_stats_elems(B,H,W,C) floats -- the statistics of the block's LayerNorm
 * (convnext.py:73-74), consumed by hpe_conv_run_ln, so that the LayerNorm pass itself disappears. */
size_t hpe_dwconv7_stats_elems(int B, int H, int W, int C);
int hpe_dwconv7_bias_stats(const void* x, int in_pitch, int B, int H, int W, int C, const float* w_taps, const float* bias,
                           void* y, int out_pitch, float* rowstats, void* stream);
int hpe_layernorm_nhwc(const void* x, int in_pitch, long long M, int C, const float* gamma, const float* beta, float eps,
                       void* y, int out_pitch, void* stream);
int hpe_patchify4(const float* img_nchw, int B, int H, int W, void* out, void* stream);
int hpe_patchify4_u8(const unsigned char* img_nchw, int B, int H, int W, void* out, void* stream);   /* uint8 / 255 fused in */
int hpe_maxpool5(const void* x, int in_pitch, int B, int H, int W, int C, void* y, int out_pitch, void* stream);
int hpe_upsample_bilinear(const void* x, int in_pitch, int B, int Hi, int Wi, int C, int Ho, int Wo, void* y, int out_pitch,
                          void* stream);

/* ------------------------------------------------------------------------------------------------
 * Training-mode pieces of the reference's real pose model beyond its (frozen) backbone: the neck and head of DINOv3Pose are
 * trained by HPE/train.py:116-139 (`finetuning=True` freezes the ConvNeXt, HPE/models/pose.py:47-49); they are built from
 * ConvBlock = Conv2d + BatchNorm2d (batch statistics under model.train()) + SiLU (HPE/models/nn/modules/block.py:20-45).
 * The convolutions use hpe_conv_run / hpe_conv_dgrad_desc / hpe_wgrad_*; these are the bandwidth-bound rest, for any channel
 * count that is a multiple of 8.  act: 0 none, 1 ReLU, 2 SiLU.  All reductions are two-stage in a fixed order (no atomics).
 *   hpe_bnact_stats     as hpe_bn_stats (scale/shift/mean/rstd + running statistics); scratch = hpe_bnact_scratch_elems(C) floats
 *   hpe_bnact_apply     y = act(z*scale + shift)
 *   hpe_bnact_backward  g = dy * act'(z*scale + shift); dbeta = sum g; dgamma = sum g*xhat; dz = gamma*rstd*(g - dbeta/M - xhat*dgamma/M);
 *                       frozen_stats != 0 (the module is in eval mode, mean / rstd are its running statistics): dz = gamma*rstd*g
 *   hpe_dwconv3_wgrad   weight gradient [C][1][3][3] of a depthwise 3x3 conv (pad 1, stride 1); its forward and backward-data
 *                       passes are hpe_dwconv (mode 0 with scale 1, shift 0 and eps < 0 = no activation; flipped taps for dgrad)
 *   hpe_maxpool5_train / _backward  MaxPool2d(5, 1, 2) (SPPF, block.py:102-114) that also records the window position of the first
 *                                   maximum (idx: one byte per element) / its backward from that record
 *   hpe_upsample_bilinear_backward  backward of F.interpolate(mode='bilinear', align_corners=False) (PAN, block.py:132-151)
 * ---------------------------------------------------------------------------------------------- */
size_t hpe_bnact_scratch_elems(int C);
int hpe_bnact_stats(const void* z, long long M, int C, const float* gamma, const float* beta, float eps, float momentum,
                    float* running_mean, float* running_var, float* scratch, float* scale, float* shift, float* mean,
                    float* rstd, void* stream);
int hpe_bnact_apply(const void* z, long long M, int C, const float* scale, const float* shift, int act, void* y, void* stream);
int hpe_bnact_backward(const void* dy, const void* z, long long M, int C, const float* mean, const float* rstd,
                       const float* gamma, const float* scale, const float* shift, int act, int frozen_stats, float* scratch,
                       float* dbeta, float* dgamma, void* dz, void* stream);
size_t hpe_dwconv3_wgrad_scratch_elems(int C);
int hpe_dwconv3_wgrad(const void* x, const void* dz, int B, int H, int W, int C, float* scratch, float* dw_torch, void* stream);
int hpe_maxpool5_train(const void* x, int B, int H, int W, int C, void* y, unsigned char* idx, void* stream);
int hpe_maxpool5_backward(const unsigned char* idx, const void* dy, int B, int H, int W, int C, void* dx, void* stream);
int hpe_upsample_bilinear_backward(const void* dy, int B, int Hi, int Wi, int C, int Ho, int Wo, void* dx, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HPE_B200_H_ */
