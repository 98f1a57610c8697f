"""The reference's real pose model on the sm_100a kernels: ``DINOv3Pose`` = ConvNeXt backbone + FeatureAdaptor + SPPF +
PAN + PoseHead, behind the reference's constructor signature

    DINOv3Pose(backbone='dinov3_convnext_base', nkpts=(4, 3), ncls=10, backbone_ckps=None, finetuning=True, device='cuda')
    (/root/reference/HPE/models/pose.py:9-28)

``model.eval()(x)`` -> ``[B, nc+nk, A]`` decoded predictions (head.py:185-199); ``model.train()(x)`` -> the list of raw
per-level maps ``[B, nc+nk, H_l, W_l]`` (head.py:176-181).  The parameter tree has the reference's ``state_dict`` keys
(``backbone.stages.2.5.pwconv1.weight``, ``pan.fpn_conv2.cv3.weight``, ``head.cv4.0.2.bias`` ...), so reference checkpoints
load with ``load_state_dict``.  The ``torch.nn`` modules below are parameter containers only: the forward pass runs
  * every Conv2d / Linear as the tcgen05 implicit-GEMM kernel (csrc/conv_gemm.cu) with BN / bias / layer-scale / residual /
    SiLU / GELU fused in its epilogue,
  * depthwise 7x7 + LayerNorm, depthwise 3x3 + BN + SiLU, LayerNorm, 5x5 max-pool, bilinear upsampling as NHWC bf16
    bandwidth kernels (csrc/convnext_ops.cu),
  * ``torch.cat`` along channels as nothing at all: producers write into channel slices of one buffer.
There is no CPU path.  ``model.train()`` under autograd runs posetrain.py (batch-statistics BatchNorm in the neck / head and their
backward pass behind the frozen backbone, HPE/train.py:116-139).  Not built: the ViT backbones (pose.py:34-46).
"""
from __future__ import annotations

import math
import os

import torch
import torch.nn as nn

from . import _lib
from ._lib import ACT_GELU, ACT_GELU_LN, ACT_NONE, ACT_RELU, ACT_SILU, CONV_1x1, CONV_2x2S2, CONV_3x3, CONV_DECONV4, CONV_FINAL, ConvDesc
from .conv import ConvLayer, fold_bn
from .detpose import pose_decode

convnext_sizes = {          # /root/reference/HPE/models/backbones/dinov3convnext.py:93-110
    "tiny": dict(depths=[3, 3, 9, 3], dims=[96, 192, 384, 768]),
    "small": dict(depths=[3, 3, 27, 3], dims=[96, 192, 384, 768]),
    "base": dict(depths=[3, 3, 27, 3], dims=[128, 256, 512, 1024]),
    "large": dict(depths=[3, 3, 27, 3], dims=[192, 384, 768, 1536]),
}


# ---------------------------------------------------------------------------------- parameter containers
class _LN(nn.Module):
    def __init__(self, c, eps=1e-6):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(c)); self.bias = nn.Parameter(torch.zeros(c)); self.eps = eps


class _Block(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.dwconv = nn.Conv2d(dim, dim, 7, padding=3, groups=dim)
        self.norm = _LN(dim)
        self.pwconv1 = nn.Linear(dim, 4 * dim); self.pwconv2 = nn.Linear(4 * dim, dim)
        self.gamma = nn.Parameter(1e-6 * torch.ones(dim))


class _ConvNeXt(nn.Module):
    def __init__(self, depths, dims):
        super().__init__()
        self.downsample_layers = nn.ModuleList([nn.Sequential(nn.Conv2d(3, dims[0], 4, 4), _LN(dims[0]))])
        for i in range(3):
            self.downsample_layers.append(nn.Sequential(_LN(dims[i]), nn.Conv2d(dims[i], dims[i + 1], 2, 2)))
        self.stages = nn.ModuleList([nn.Sequential(*[_Block(dims[i]) for _ in range(depths[i])]) for i in range(4)])
        self.norm = nn.LayerNorm(dims[-1], eps=1e-6)
        self.norms = nn.ModuleList([nn.Identity() for _ in range(3)] + [self.norm])
        for m in self.modules():                                   # convnext.py:196-205
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                nn.init.trunc_normal_(m.weight, std=0.02); nn.init.zeros_(m.bias)


class _ConvBlock(nn.Module):
    def __init__(self, c1, c2, k=1, s=1, g=1):
        super().__init__()
        self.conv = nn.Conv2d(c1, c2, k, s, k // 2, groups=g, bias=False)
        self.bn = nn.BatchNorm2d(c2)
        self.k, self.s = k, s


class _Bottleneck(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.cv1 = _ConvBlock(c, c, 3); self.cv2 = _ConvBlock(c, c, 3)


class _CSP(nn.Module):
    def __init__(self, c1, c2, n=2):
        super().__init__()
        c_ = int(c2 * 0.5)
        self.cv1 = _ConvBlock(c1, c_, 1)
        self.cv2 = nn.Conv2d(c1, c_, 1, 1, bias=False); self.cv3 = nn.Conv2d(c_, c_, 1, 1, bias=False)
        self.cv4 = _ConvBlock(2 * c_, c2, 1)
        self.bn = nn.BatchNorm2d(2 * c_)
        self.m = nn.Sequential(*(_Bottleneck(c_) for _ in range(n)))
        self.c1, self.c2, self.c_ = c1, c2, c_


class _SPPF(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.conv1 = _ConvBlock(c, c // 2, 1); self.conv2 = _ConvBlock(c // 2 * 4, c, 1)


class _PAN(nn.Module):
    def __init__(self, ch):
        super().__init__()
        c1, c2, c3 = ch
        self.fpn_conv1 = _ConvBlock(c3, c2, 1); self.fpn_conv2 = _CSP(c2 + c2, c2)
        self.fpn_conv3 = _ConvBlock(c2, c1, 1); self.fpn_conv4 = _CSP(c1 + c1, c1)
        self.down1 = _ConvBlock(c1, c1, 3, 2); self.pan_conv1 = _CSP(c1 + c2, c2)
        self.down2 = _ConvBlock(c2, c2, 3, 2); self.pan_conv2 = _CSP(c2 + c3, c3)


class _Adaptor(nn.Module):
    def __init__(self, sc, tc):
        super().__init__()
        self.adapters = nn.ModuleList([nn.Sequential(_ConvBlock(a, b, 1), _ConvBlock(b, b, 3)) for a, b in zip(sc, tc)])


class _Head(nn.Module):
    def __init__(self, ncls, kpt_shape, in_ch):
        super().__init__()
        self.nc, self.kpt_shape, self.nk = ncls, tuple(kpt_shape), kpt_shape[0] * kpt_shape[1]
        self.nl = len(in_ch)
        self.register_buffer("stride", torch.tensor([8, 16, 32], dtype=torch.float32))
        c3 = max(in_ch[0], min(ncls, 100)); c4 = max(in_ch[0] // 4, self.nk)          # head.py:139-140
        self.c3, self.c4 = c3, c4
        self.cv3 = nn.ModuleList(nn.Sequential(nn.Sequential(_ConvBlock(x, x, 3, g=x), _ConvBlock(x, c3, 1)),
                                               nn.Sequential(_ConvBlock(c3, c3, 3), _ConvBlock(c3, c3, 1)),
                                               nn.Conv2d(c3, ncls, 1)) for x in in_ch)
        self.cv4 = nn.ModuleList(nn.Sequential(_ConvBlock(x, c4, 3), _ConvBlock(c4, c4, 3), nn.Conv2d(c4, self.nk, 1)) for x in in_ch)
        for m in self.cv3:                                                             # head.py:163-171
            nn.init.constant_(m[-1].bias, -math.log((1 - 0.01) / 0.01))


# ---------------------------------------------------------------------------------- elementwise op wrappers
def _pitch(t):
    return t.stride(-2)


def _dwconv(x, w_taps, p0, p1, p2, eps, ksize, mode, y):
    B, H, W, C = x.shape
    with torch.cuda.device(x.device):
        rc = _lib.lib().hpe_dwconv(_lib.ptr(x), _pitch(x), B, H, W, C, ksize, mode, _lib.ptr(w_taps), _lib.ptr(p0), _lib.ptr(p1),
                                   _lib.ptr(p2), float(eps), _lib.ptr(y), _pitch(y), _lib.stream_ptr())
    _lib.check(rc, "hpe_dwconv")


def _dwconv7_bias(x, w_taps, bias, y):
    B, H, W, C = x.shape
    with torch.cuda.device(x.device):
        rc = _lib.lib().hpe_dwconv7_bias(_lib.ptr(x), _pitch(x), B, H, W, C, _lib.ptr(w_taps), _lib.ptr(bias), _lib.ptr(y), _pitch(y),
                                         _lib.stream_ptr())
    _lib.check(rc, "hpe_dwconv7_bias")


def _dwconv7_bias_stats(x, w_taps, bias, y, rowstats):
    B, H, W, C = x.shape
    with torch.cuda.device(x.device):
        rc = _lib.lib().hpe_dwconv7_bias_stats(_lib.ptr(x), _pitch(x), B, H, W, C, _lib.ptr(w_taps), _lib.ptr(bias), _lib.ptr(y), _pitch(y),
                                               _lib.ptr(rowstats), _lib.stream_ptr())
    _lib.check(rc, "hpe_dwconv7_bias_stats")


def _layernorm(x, gamma, beta, eps, y):
    B, H, W, C = x.shape
    with torch.cuda.device(x.device):
        rc = _lib.lib().hpe_layernorm_nhwc(_lib.ptr(x), _pitch(x), B * H * W, C, _lib.ptr(gamma), _lib.ptr(beta), float(eps),
                                           _lib.ptr(y), _pitch(y), _lib.stream_ptr())
    _lib.check(rc, "hpe_layernorm_nhwc")


def _maxpool5(x, y):
    B, H, W, C = x.shape
    with torch.cuda.device(x.device):
        rc = _lib.lib().hpe_maxpool5(_lib.ptr(x), _pitch(x), B, H, W, C, _lib.ptr(y), _pitch(y), _lib.stream_ptr())
    _lib.check(rc, "hpe_maxpool5")


def _upsample(x, y):
    B, Hi, Wi, C = x.shape
    with torch.cuda.device(x.device):
        rc = _lib.lib().hpe_upsample_bilinear(_lib.ptr(x), _pitch(x), B, Hi, Wi, C, y.shape[1], y.shape[2], _lib.ptr(y), _pitch(y),
                                              _lib.stream_ptr())
    _lib.check(rc, "hpe_upsample_bilinear")


def _patchify4(img, out):
    B, _, H, W = img.shape
    with torch.cuda.device(img.device):
        fn = _lib.lib().hpe_patchify4_u8 if img.dtype == torch.uint8 else _lib.lib().hpe_patchify4
        rc = fn(_lib.ptr(img), B, H, W, _lib.ptr(out), _lib.stream_ptr())
    _lib.check(rc, "hpe_patchify4")


def _pad_to(t, dim, n):
    if t.shape[dim] == n:
        return t
    shape = list(t.shape); shape[dim] = n - t.shape[dim]
    return torch.cat([t, t.new_zeros(shape)], dim)


class _PosePlan:
    """Every kernel launch of one DINOv3Pose forward at one (B, H, W), with its buffers; replayed as a CUDA graph.
    ``heatmap_head=True``: ``model`` is a ``HeatmapPose`` with a ConvNeXt backbone -- the four ConvNeXt stages feed the App. A.1
    deconv heatmap head instead of adaptor + SPPF + PAN + PoseHead."""

    def __init__(self, model, B, H, W, device, heatmap_head=False, backbone_only=False):
        if H % 32 or W % 32:
            raise ValueError(f"input {H}x{W} must be a multiple of 32")
        self.B, self.H, self.W, self.dev = B, H, W, device
        self.steps = []                   # callables, in launch order
        self.keep = []                    # device tensors owned by the plan (packed weights live in the ConvLayers)
        self.n_launches = 0
        self.graph = None
        self.levels = None                # backbone_only plans have no head outputs
        self.ln_folded = 0                # ConvNeXt blocks whose LayerNorm runs inside pwconv1's epilogue
        dims, depths = model.dims, model.depths
        bf = lambda *shape: torch.zeros(shape, dtype=torch.bfloat16, device=device)      # noqa: E731
        f32 = lambda t: t.detach().to(device, torch.float32).contiguous()               # noqa: E731
        ones = lambda n: torch.ones(n, dtype=torch.float32, device=device)              # noqa: E731

        def conv(kind, src, dst, weight, scale, shift, act, stride=1, residual=None, cin=None, cout=None):
            """src/dst: NHWC views (possibly channel slices).  weight: torch-layout fp32 (already padded)."""
            Bx, Hx, Wx, Cx = src.shape
            cin = cin or Cx
            final = kind == CONV_FINAL
            n_out = cout or weight.shape[0]
            if not final and n_out > 2048:           # the kernel stages <= 2048 scale/shift pairs: split wide layers by output slice
                pieces = -(-n_out // 2048)
                step_c = -(-n_out // pieces // 64) * 64
                scale_f, shift_f = f32(scale), f32(shift)
                for c0 in range(0, n_out, step_c):
                    c1 = min(c0 + step_c, n_out)
                    conv(kind, src, dst[..., c0:c1], weight[c0:c1], scale_f[c0:c1], shift_f[c0:c1], act, stride,
                         None if residual is None else residual[..., c0:c1], cin, c1 - c0)
                return None
            layer = ConvLayer(kind, Bx, Hx, Wx, cin, cout or weight.shape[0], stride, act, residual is not None, device,
                              in_cstride=0 if src.is_contiguous() else _pitch(src),
                              out_cstride=(dst[1] if final else (0 if dst.is_contiguous() else _pitch(dst))))
            layer.load(f32(weight), f32(scale), f32(shift))
            out_t = dst[0] if final else dst
            self.steps.append(lambda: layer(src, residual, out_t))
            self.n_launches += 1
            return layer

        def pwconv1_ln(blk, src, dst, rowstats):
            """`x = self.act(self.pwconv1(self.norm(x)))` (convnext.py:74-77) with the LayerNorm folded into the GEMM (conv.run_ln);
            False when the layer has no kernel for it (tiny batch: no cta_group::2 tiles) -- the caller keeps the separate pass."""
            Bx, Hx, Wx, Cx = src.shape
            W1 = blk.pwconv1.weight.detach().to(device, torch.float32)                         # [4C, C]
            g, beta = blk.norm.weight.detach().to(device, torch.float32), blk.norm.bias.detach().to(device, torch.float32)
            Wf = W1 * g[None, :]
            colsum = Wf.to(torch.bfloat16).float().sum(1)                                       # of the weights the GEMM really uses
            shift = blk.pwconv1.bias.detach().to(device, torch.float32) + W1 @ beta
            n_out = W1.shape[0]
            pieces = -(-n_out // 2048)
            step_c = -(-n_out // pieces // 64) * 64
            layers = []
            try:
                for c0 in range(0, n_out, step_c):
                    c1 = min(c0 + step_c, n_out)
                    d = dst[..., c0:c1]
                    layer = ConvLayer(CONV_1x1, Bx, Hx, Wx, Cx, c1 - c0, 1, ACT_GELU_LN, False, device,
                                      out_cstride=0 if d.is_contiguous() else _pitch(d))
                    layer.load(Wf[c0:c1].reshape(c1 - c0, Cx, 1, 1).contiguous(), colsum[c0:c1].contiguous(), shift[c0:c1].contiguous())
                    layers.append((layer, d))
            except RuntimeError:
                return False
            ns = -(-Cx // 64)
            for layer, d in layers:
                self.steps.append(lambda layer=layer, d=d: layer.run_ln(src, rowstats, ns, Cx, blk.norm.eps, d))
                self.n_launches += 1
            return True

        def convblock(cb, src, dst, extra_bn=None, pad_out=None, pad_in=None):
            """ConvBlock (block.py:20-45) with its BatchNorm folded; 1x1 / 3x3, stride 1|2, SiLU."""
            w = cb.conv.weight.detach().float()
            scale, shift = fold_bn(cb.bn.weight, cb.bn.bias, cb.bn.running_mean, cb.bn.running_var, cb.bn.eps)
            if pad_in:
                w = _pad_to(w, 1, pad_in)
            if pad_out:
                w, scale, shift = _pad_to(w, 0, pad_out), _pad_to(scale, 0, pad_out), _pad_to(shift, 0, pad_out)
            return conv(CONV_3x3 if cb.k == 3 else CONV_1x1, src, dst, w, scale, shift, ACT_SILU, cb.s)

        def step(fn):
            self.steps.append(fn); self.n_launches += 1

        # ---------------- backbone (convnext.py:157-180, 71-84)
        bb = model.backbone
        h, w = H // 4, W // 4
        self.patches = bf(B, h, w, 64)
        stem_conv, stem_ln = bb.downsample_layers[0][0], bb.downsample_layers[0][1]
        t_a, t_b = bf(B, h, w, dims[0]), bf(B, h, w, dims[0])
        wst = _pad_to(stem_conv.weight.detach().float().reshape(dims[0], 48), 1, 64).reshape(dims[0], 64, 1, 1)
        conv(CONV_1x1, self.patches, t_a, wst, ones(dims[0]), stem_conv.bias, ACT_NONE)
        g, b_ = f32(stem_ln.weight), f32(stem_ln.bias); self.keep += [g, b_]
        step(lambda x=t_a, y=t_b, g=g, b=b_, e=stem_ln.eps: _layernorm(x, g, b, e, y))
        cur, other = t_b, t_a
        feats = []
        for i in range(4):
            C = dims[i]
            if i > 0:
                ln, cv = bb.downsample_layers[i][0], bb.downsample_layers[i][1]
                g, b_ = f32(ln.weight), f32(ln.bias); self.keep += [g, b_]
                step(lambda x=cur, y=other, g=g, b=b_, e=ln.eps: _layernorm(x, g, b, e, y))
                h, w = h // 2, w // 2
                nxt, nxt2 = bf(B, h, w, C), bf(B, h, w, C)
                conv(CONV_2x2S2, other, nxt, cv.weight, ones(C), cv.bias, ACT_NONE, 2)
                cur, other = nxt, nxt2
            t1, t2 = bf(B, h, w, C), bf(B, h, w, 4 * C)
            # LayerNorm folded into pwconv1 (HPE_B200_LN_FOLD=0: separate LayerNorm pass): the depthwise conv emits the per-pixel
            # statistics, pwconv1 normalises in its epilogue
            ln_fold = os.environ.get("HPE_B200_LN_FOLD", "1") != "0" and os.environ.get("HPE_B200_DWCONV", "tiled") != "fused" \
                and os.environ.get("HPE_B200_DWCONV_TMA", "1") != "0"
            rowstats = torch.zeros(B * h * w * (-(-C // 64)) * 2, dtype=torch.float32, device=device) if ln_fold else None
            for blk in bb.stages[i]:
                wt = f32(blk.dwconv.weight.detach().reshape(C, 49).t())             # [49][C]
                p0, p1, p2 = f32(blk.dwconv.bias), f32(blk.norm.weight), f32(blk.norm.bias); self.keep += [wt, p0, p1, p2]
                folded = False
                if os.environ.get("HPE_B200_DWCONV", "tiled") == "fused":      # one kernel: dwconv + bias + LayerNorm
                    step(lambda x=cur, y=t1, wt=wt, p0=p0, p1=p1, p2=p2, e=blk.norm.eps: _dwconv(x, wt, p0, p1, p2, e, 7, 1, y))
                elif ln_fold:                                                  # dwconv (+ LayerNorm statistics), LayerNorm inside pwconv1
                    n_before = len(self.steps)
                    step(lambda x=cur, y=t1, wt=wt, p0=p0, rs=rowstats: _dwconv7_bias_stats(x, wt, p0, y, rs))
                    folded = pwconv1_ln(blk, t1, t2, rowstats)
                    if not folded:
                        del self.steps[n_before:]
                        self.n_launches -= 1
                if not folded and os.environ.get("HPE_B200_DWCONV", "tiled") != "fused":
                    step(lambda x=cur, y=t1, wt=wt, p0=p0: _dwconv7_bias(x, wt, p0, y))      # FMA-bound tiled dwconv, then LayerNorm in place
                    step(lambda y=t1, p1=p1, p2=p2, e=blk.norm.eps: _layernorm(y, p1, p2, e, y))
                if not folded:
                    conv(CONV_1x1, t1, t2, blk.pwconv1.weight.detach().reshape(4 * C, C, 1, 1), ones(4 * C), blk.pwconv1.bias, ACT_GELU)
                self.ln_folded += int(folded)
                gam = blk.gamma.detach().float()
                conv(CONV_1x1, t2, other, blk.pwconv2.weight.detach().reshape(C, 4 * C, 1, 1), gam, gam * blk.pwconv2.bias.detach().float(),
                     ACT_NONE, residual=cur)
                cur, other = other, cur
            if i > 0:                                # later stages own fresh buffers, so `cur` survives as a pyramid level
                feats.append((cur, h, w, C))
        self.feats = [f for f, _, _, _ in feats]      # stage 1-3 outputs (NHWC bf16): the pyramid the neck reads
        if backbone_only:                                # training: the frozen backbone only; the neck and head run in posetrain.py
            return
        if heatmap_head:
            self._build_heatmap_head(model, feats[-1], B, device)
            return
        # ---------------- neck (pose.py:58-65; block.py:236-254, 102-114, 117-157)
        c1n, c2n, c3n = model.tc
        (f1, h1, w1, _), (f2, h2, w2, _), (f3, h3, w3, _) = feats
        cat1 = bf(B, h1, w1, 2 * c1n)             # [c1 | up(fpn_conv3(c2f))]
        cat2 = bf(B, h2, w2, 2 * c2n)             # [c2 | up(fpn_conv1(c3))]
        cat3 = bf(B, h2, w2, c1n + c2n)           # [down1(c1f) | c2f]
        cat4 = bf(B, h3, w3, c2n + c3n)           # [down2(c2p) | c3]
        ad = model.adaptor.adapters
        ta = bf(B, h1, w1, c1n); convblock(ad[0][0], f1, ta); convblock(ad[0][1], ta, cat1[..., :c1n])
        tb = bf(B, h2, w2, c2n); convblock(ad[1][0], f2, tb); convblock(ad[1][1], tb, cat2[..., :c2n])
        tc_, a3 = bf(B, h3, w3, c3n), bf(B, h3, w3, c3n); convblock(ad[2][0], f3, tc_); convblock(ad[2][1], tc_, a3)
        hc = c3n // 2
        spp = bf(B, h3, w3, 4 * hc)
        convblock(model.sppf.conv1, a3, spp[..., :hc])
        for j in range(3):
            step(lambda x=spp[..., j * hc:(j + 1) * hc], y=spp[..., (j + 1) * hc:(j + 2) * hc]: _maxpool5(x, y))
        c3 = cat4[..., c2n:]
        convblock(model.sppf.conv2, spp, c3)

        def csp(m, src, dst):
            c_ = m.c_
            a, b2 = bf(*src.shape[:3], c_), bf(*src.shape[:3], c_)
            cc = bf(*src.shape[:3], 2 * c_)
            convblock(m.cv1, src, a)
            x_ = a
            for bn_ in m.m:
                convblock(bn_.cv1, x_, b2)
                nx = bf(*src.shape[:3], c_)
                convblock(bn_.cv2, b2, nx)
                x_ = nx
            sc, sh = fold_bn(m.bn.weight, m.bn.bias, m.bn.running_mean, m.bn.running_var, m.bn.eps)
            conv(CONV_1x1, x_, cc[..., :c_], m.cv3.weight, sc[:c_], sh[:c_], ACT_SILU)
            conv(CONV_1x1, src, cc[..., c_:], m.cv2.weight, sc[c_:], sh[c_:], ACT_SILU)
            convblock(m.cv4, cc, dst)

        pan = model.pan
        u1 = bf(B, h3, w3, c2n); convblock(pan.fpn_conv1, c3, u1)
        step(lambda x=u1, y=cat2[..., c2n:]: _upsample(x, y))
        c2f = cat3[..., c1n:]
        csp(pan.fpn_conv2, cat2, c2f)
        u2 = bf(B, h2, w2, c1n); convblock(pan.fpn_conv3, c2f, u2)
        step(lambda x=u2, y=cat1[..., c1n:]: _upsample(x, y))
        c1f = bf(B, h1, w1, c1n); csp(pan.fpn_conv4, cat1, c1f)
        convblock(pan.down1, c1f, cat3[..., :c1n])               # F.interpolate to the same size is the identity
        c2p = bf(B, h2, w2, c2n); csp(pan.pan_conv1, cat3, c2p)
        convblock(pan.down2, c2p, cat4[..., :c2n])
        c3p = bf(B, h3, w3, c3n); csp(pan.pan_conv2, cat4, c3p)
        self.features = [c1f, c2p, c3p]
        # ---------------- head (head.py:131-181)
        hd = model.head
        nc, nk, c3h = hd.nc, hd.nk, hd.c3
        c4p = (hd.c4 + 63) // 64 * 64                # 51 -> 64 zero-padded channels (SiLU(0) = 0 keeps them zero)
        self.levels = []
        for i, f in enumerate(self.features):
            Bf, hf, wf, cf = f.shape
            out = torch.zeros((B, nc + nk, hf, wf), dtype=torch.float32, device=device)
            self.levels.append(out)
            dw = hd.cv3[i][0][0]
            wt = f32(dw.conv.weight.detach().reshape(cf, 9).t())
            sc, sh = fold_bn(dw.bn.weight, dw.bn.bias, dw.bn.running_mean, dw.bn.running_var, dw.bn.eps)
            sc, sh = f32(sc), f32(sh); self.keep += [wt, sc, sh]
            d1, d2, d3, d4 = bf(B, hf, wf, cf), bf(B, hf, wf, c3h), bf(B, hf, wf, c3h), bf(B, hf, wf, c3h)
            step(lambda x=f, y=d1, wt=wt, sc=sc, sh=sh: _dwconv(x, wt, sc, sh, None, 0.0, 3, 0, y))
            convblock(hd.cv3[i][0][1], d1, d2); convblock(hd.cv3[i][1][0], d2, d3); convblock(hd.cv3[i][1][1], d3, d4)
            fc = hd.cv3[i][2]
            conv(CONV_FINAL, d4, (out, nc + nk), fc.weight, ones(nc), fc.bias, ACT_NONE)
            k1, k2 = bf(B, hf, wf, c4p), bf(B, hf, wf, c4p)
            convblock(hd.cv4[i][0], f, k1, pad_out=c4p); convblock(hd.cv4[i][1], k1, k2, pad_out=c4p, pad_in=c4p)
            fk = hd.cv4[i][2]
            conv(CONV_FINAL, k2, (out[:, nc:], nc + nk), _pad_to(fk.weight.detach().float(), 1, c4p), ones(nk), fk.bias, ACT_NONE)

    def _build_heatmap_head(self, model, feat, B, device):
        """SURVEY App. A.1 on the last ConvNeXt stage: 3 x [ConvTranspose2d(k4,s2,p1) + BN + ReLU] + final 1x1; the last deconv
        stage and the final 1x1 run as ONE kernel (fused head) when K <= 32, exactly as behind the ResNet-50 trunk."""
        cur, h, w, cin = feat
        K = model.njoints
        self.features = [cur]
        self.fused_head = False
        for d, idx in enumerate((0, 3, 6)):
            dc, bn = model.deconv_layers[idx], model.deconv_layers[idx + 1]
            scale, shift = fold_bn(bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps)
            layer, out = None, None
            if d == 2 and K <= 32 and os.environ.get("HPE_B200_FUSE_HEAD", "1") != "0":
                try:
                    layer = ConvLayer(CONV_DECONV4, B, h, w, cin, 256, 2, ACT_RELU, False, device, head2=K)
                    layer.load(dc.weight, scale, shift)
                    layer.load_head2(model.final_layer.weight, model.final_layer.bias)
                    out = torch.zeros((B, K, 2 * h, 2 * w), dtype=torch.float32, device=device)
                    self.fused_head = True
                except RuntimeError:
                    layer = None                    # too few tiles for a CTA pair (tiny batch): separate layers
            if layer is None:
                layer = ConvLayer(CONV_DECONV4, B, h, w, cin, 256, 2, ACT_RELU, False, device)
                layer.load(dc.weight, scale, shift)
                out = torch.zeros((B, 2 * h, 2 * w, 256), dtype=torch.bfloat16, device=device)
            self.steps.append(lambda layer=layer, src=cur, out=out: layer(src, None, out))
            self.n_launches += 1
            cur, cin, h, w = out, 256, 2 * h, 2 * w
        if cur.dtype != torch.float32:              # separate final 1x1 (K > 32 or tiny batch)
            fl = model.final_layer
            layer = ConvLayer(CONV_FINAL, B, h, w, 256, K, 1, ACT_NONE, False, device)
            layer.load(fl.weight, torch.ones(K), fl.bias)
            out = torch.zeros((B, K, h, w), dtype=torch.float32, device=device)
            self.steps.append(lambda layer=layer, src=cur, out=out: layer(src, None, out))
            self.n_launches += 1
            cur = out
        self.heatmaps = cur
        self.levels = [cur]

    def _launch_all(self):
        for s in self.steps:
            s()

    def run(self, x):
        _patchify4(x, self.patches)
        # (inside an outer capture -- PoseTrainStep(cuda_graph=True) -- the launches are recorded into THAT graph)
        if os.environ.get("HPE_B200_CUDA_GRAPH", "1") == "0" or torch.cuda.is_current_stream_capturing():
            self._launch_all()
            return self.levels
        if self.graph is None:
            self._launch_all()                       # warm-up: encodes the TMA descriptors once
            torch.cuda.synchronize(self.dev)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._launch_all()
            self.graph = g
        self.graph.replay()
        return self.levels


class DINOv3Pose(nn.Module):
    """Reference constructor (pose.py:10-15).  ``backbone_ckps``: optional local ConvNeXt checkpoint (the reference's
    default paths under ./checkpoints need downloaded weights; without a file the backbone keeps its seeded init)."""

    def __init__(self, backbone="dinov3_convnext_base", nkpts=(4, 3), ncls=10, backbone_ckps=None, finetuning=True, device="cuda"):
        super().__init__()
        info = backbone.split("_")
        if len(info) != 3 or "convnext" not in info[1]:
            raise TypeError(f"not supported backbone type : {backbone} (the ViT backbones are not built)")
        size = info[2]
        self.tc = [192, 384, 768]
        self.dims, self.depths = convnext_sizes[size]["dims"], convnext_sizes[size]["depths"]
        self.backbone = _ConvNeXt(self.depths, self.dims)
        if backbone_ckps:
            self.backbone.load_state_dict(torch.load(backbone_ckps, map_location="cpu"), strict=True)
        self.adaptor = _Adaptor(self.dims[1:], self.tc)
        if finetuning:
            for p in self.backbone.parameters():
                p.requires_grad = False
        self.sppf = _SPPF(self.tc[-1])
        self.pan = _PAN(self.tc)
        self.head = _Head(ncls, nkpts, self.tc)
        self._plans = {}
        self.to(device)

    def _apply(self, fn, *a, **kw):
        # a no-op `.to(device)` (the reference trainer does one per batch, HPE/engine/trainer.py:87) keeps the plans
        p0 = next(self.parameters(), None)
        before = None if p0 is None else (p0.device, p0.dtype, p0.data_ptr())
        out = super()._apply(fn, *a, **kw)
        p1 = next(self.parameters(), None)
        if before != (None if p1 is None else (p1.device, p1.dtype, p1.data_ptr())):
            self._plans = {}
        return out

    def load_state_dict(self, state_dict, strict=True, **kw):
        out = super().load_state_dict(state_dict, strict=strict, **kw)
        self._plans = {}
        return out

    def refresh(self):
        """Re-pack weights after parameters were modified in place."""
        self._plans = {}

    def _plan(self, x):
        if x.dim() != 4 or x.shape[1] != 3:
            raise ValueError(f"expected [B,3,H,W] input, got {tuple(x.shape)}")
        if not x.is_cuda:
            raise RuntimeError("libhpe_b200 has no CPU path: input must live on a CUDA (sm_100a) device")
        key = (tuple(x.shape), x.device.index)
        plan = self._plans.get(key)
        if plan is None:
            with torch.no_grad():
                plan = _PosePlan(self, x.shape[0], x.shape[2], x.shape[3], x.device)
            self._plans[key] = plan
        return plan

    @torch.no_grad()
    def forward_features(self, x):
        """PAN outputs as fp32 NCHW (pose.py:60-65)."""
        plan = self._plan(x)
        plan.run(x.contiguous() if x.dtype == torch.uint8 else x.float().contiguous())
        return [f.float().permute(0, 3, 1, 2) for f in plan.features]

    def forward(self, x):
        if self.training and torch.is_grad_enabled():
            # model.train() under autograd (HPE/train.py:116-139): frozen backbone on the inference kernels, the neck and head with
            # BatchNorm batch statistics and a backward pass through the dgrad / wgrad kernels (posetrain.py)
            from .posetrain import train_forward
            return train_forward(self, x)
        with torch.no_grad():
            return self._forward_nograd(x)

    def _forward_nograd(self, x):
        plan = self._plan(x)
        levels = plan.run(x.contiguous() if x.dtype == torch.uint8 else x.float().contiguous())   # uint8 = x.float()/255 (train.py:120)
        if self.training:
            # head.py:176-181 (forward only: no autograd graph is built).  Fresh tensors: the plan's level buffers are
            # overwritten by the next call.
            return [lv.clone() for lv in levels]
        return pose_decode(levels, self.head.nc, self.head.kpt_shape, tuple(self.head.stride.tolist()))
