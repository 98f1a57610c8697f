"""Host-side handles for the tcgen05 implicit-GEMM convolution layers (csrc/conv_gemm.cu).

One ``ConvLayer`` = one ``hpe_conv_plan`` (TMA descriptors + launch geometry) + packed bf16 weights +
folded-BN ``scale``/``shift``.  It replaces ``conv -> bn -> act (-> += shortcut)`` of the reference's
blocks: timm ``Bottleneck.forward`` (/root/reference/HPE/thirdparty/pytorch-image-models/timm/models/
resnet.py:244-270) and the first-party ``ConvBlock.forward_fuse`` (HPE/models/nn/modules/block.py:44-45).
Activations are NHWC bf16 tensors; PyTorch only owns the memory.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import (ACT_GELU, ACT_GELU_LN, ACT_NONE, ACT_RELU, ACT_SILU, CONV_1x1, CONV_2x2S2, CONV_3x3, CONV_4x4S2, CONV_DECONV4,
                   CONV_DGRAD1x1S2, CONV_DGRAD3x3S2, CONV_FINAL, CONV_STEM7x7, ConvDesc)


def fold_bn(weight, bias, running_mean, running_var, eps):
    """BN(eval) as y = x*scale + shift (same algebra as ConvBlock's fuse, block.py:44-45 / SURVEY §8 a3)."""
    scale = weight.float() / torch.sqrt(running_var.float() + eps)
    shift = bias.float() - running_mean.float() * scale
    return scale.contiguous(), shift.contiguous()


class ConvLayer:
    def __init__(self, kind, B, H, W, Cin, Cout, stride=1, act=ACT_RELU, has_residual=False, device="cuda", desc=None,
                 in_cstride=0, out_cstride=0, Cin2=0, stride2=1, head2=0, pool=0, chain=0):
        self.device = torch.device(device)
        self.desc = desc if desc is not None else ConvDesc(kind, B, H, W, Cin, Cout, stride, act, int(has_residual),
                                                           in_cstride, out_cstride, Cin2, stride2 if Cin2 else 0, head2, int(pool),
                                                           int(chain))
        kind, B, Cout = self.desc.kind, self.desc.B, self.desc.Cout
        lib = _lib.lib()
        ho, wo = C.c_int(), C.c_int()
        with torch.cuda.device(self.device):
            _lib.check(lib.hpe_conv_out_hw(C.byref(self.desc), C.byref(ho), C.byref(wo)), "hpe_conv_out_hw")
            plan = C.c_void_p()
            _lib.check(lib.hpe_conv_plan_create(C.byref(self.desc), C.byref(plan)), "hpe_conv_plan_create")
        self.plan = plan
        self.Ho, self.Wo = ho.value, wo.value
        self.kind, self.B, self.Cout = kind, B, Cout
        n = lib.hpe_conv_packed_weight_elems(C.byref(self.desc))
        self.w_packed = torch.zeros(n, dtype=torch.bfloat16, device=self.device)
        self.head2 = int(self.desc.head2)
        if self.head2:       # fused head: the final 1x1's weights (32 zero-padded rows x 256) and bias
            self.w2_packed = torch.zeros(32 * 256, dtype=torch.bfloat16, device=self.device)
            self.bias2 = torch.zeros(self.head2, dtype=torch.float32, device=self.device)
        self.chain = int(self.desc.chain)
        if self.chain:       # chained 1x1 conv 256 -> chain (the next layer): its packed weights and folded BN
            self.w2_packed = torch.zeros(self.chain * Cout, dtype=torch.bfloat16, device=self.device)
            self.scale2 = torch.ones(self.chain, dtype=torch.float32, device=self.device)
            self.shift2 = torch.zeros(self.chain, dtype=torch.float32, device=self.device)
        self.scale = torch.ones(Cout, dtype=torch.float32, device=self.device)
        self.shift = torch.zeros(Cout, dtype=torch.float32, device=self.device)

    def __del__(self):
        try:
            if getattr(self, "plan", None):
                _lib.lib().hpe_conv_plan_destroy(self.plan)
                self.plan = None
        except Exception:
            pass

    def run_ln(self, x: torch.Tensor, rowstats: torch.Tensor, n_slices: int, channels: int, eps: float, out: torch.Tensor):
        """``act = ACT_GELU_LN``: ``gelu(pwconv1(norm(x)))`` of the ConvNeXt block (dinov3 convnext.py:74-77) on the RAW ``x`` --
        the LayerNorm is folded into the weights (``load`` them as ``gamma * W``), the column table (``scale`` = column sums of
        the bf16 folded weights, ``shift`` = ``b + W beta``) and a per-row (mean, rstd) the epilogue derives from ``rowstats``
        ([pixels][n_slices][2] fp32 partial sums written by ``hpe_dwconv7_bias_stats``)."""
        _lib.require_cuda(x, rowstats)
        if self.desc.out_cstride:          # a channel slice of a wider NHWC buffer
            if not out.is_cuda or out.stride(-1) != 1 or out.stride(-2) != self.desc.out_cstride:
                raise RuntimeError("libhpe_b200: channel-slice tensor does not match the plan's channel pitch")
        else:
            _lib.require_cuda(out)
        with torch.cuda.device(self.device):
            rc = _lib.lib().hpe_conv_run_ln(self.plan, _lib.ptr(x), _lib.ptr(self.w_packed), _lib.ptr(self.scale), _lib.ptr(self.shift),
                                            _lib.ptr(rowstats), int(n_slices), int(channels), float(eps), _lib.ptr(out), _lib.stream_ptr())
        _lib.check(rc, "hpe_conv_run_ln")
        return out

    def set_schedule(self, reverse=False, l2_out=0, l2_res=0, l2_in=0) -> bool:
        """Tile order / L2 eviction hints of this layer's launches (hpe_conv_set_schedule; performance only).  Returns the tile
        direction in effect (True = reversed)."""
        rc = _lib.lib().hpe_conv_set_schedule(self.plan, int(bool(reverse)), int(l2_out), int(l2_res), int(l2_in))
        if rc < 0:
            _lib.check(rc, "hpe_conv_set_schedule")
        return bool(rc)

    def out_shape(self):
        if self.head2:
            return (self.B, self.head2, self.Ho, self.Wo)
        if self.kind == CONV_FINAL:
            return (self.B, self.Cout, self.Ho, self.Wo)
        return (self.B, self.Ho, self.Wo, self.Cout)

    def alloc_out(self):
        dt = torch.float32 if (self.kind == CONV_FINAL or self.head2) else torch.bfloat16
        return torch.empty(self.out_shape(), dtype=dt, device=self.device)

    def load(self, weight: torch.Tensor, scale: torch.Tensor = None, shift: torch.Tensor = None):
        """weight: PyTorch layout fp32 (Conv2d [Cout,Cin,kh,kw] / ConvTranspose2d [Cin,Cout,4,4])."""
        w = weight.detach().to(self.device, torch.float32).contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().hpe_conv_pack_weight(C.byref(self.desc), _lib.ptr(w), _lib.ptr(self.w_packed),
                                                       _lib.stream_ptr()), "hpe_conv_pack_weight")
        if scale is not None:
            self.scale.copy_(scale.detach().to(self.device, torch.float32))
        if shift is not None:
            self.shift.copy_(shift.detach().to(self.device, torch.float32))
        return self

    def load_head2(self, weight: torch.Tensor, bias: torch.Tensor):
        """Fused head (``head2 = K``): ``weight`` [K,256,1,1] / ``bias`` [K] of the final 1x1 conv (SURVEY App. A.1
        ``final_layer``), applied to this layer's output tile while it is still in shared memory."""
        w = weight.detach().to(self.device, torch.float32).contiguous()
        fd = ConvDesc(CONV_FINAL, self.B, self.Ho, self.Wo, 256, self.head2, 1, ACT_NONE, 0, 0, 0, 0, 0, 0)
        self.w2_packed.zero_()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().hpe_conv_pack_weight(C.byref(fd), _lib.ptr(w), _lib.ptr(self.w2_packed), _lib.stream_ptr()),
                       "hpe_conv_pack_weight")
        self.bias2.copy_(bias.detach().to(self.device, torch.float32))
        return self

    def load_chain(self, weight: torch.Tensor, scale: torch.Tensor, shift: torch.Tensor):
        """Chained layer (``chain = N2``): ``weight`` [N2,Cout,1,1] and folded BN ``scale``/``shift`` [N2] of the NEXT 1x1 conv
        (timm ``Bottleneck.conv1``/``bn1``, resnet.py:247-251), applied to this layer's output tile while it is staged."""
        w = weight.detach().to(self.device, torch.float32).contiguous()
        cd = ConvDesc(CONV_1x1, self.B, self.Ho, self.Wo, self.Cout, self.chain, 1, ACT_RELU, 0)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().hpe_conv_pack_weight(C.byref(cd), _lib.ptr(w), _lib.ptr(self.w2_packed), _lib.stream_ptr()),
                       "hpe_conv_pack_weight")
        self.scale2.copy_(scale.detach().to(self.device, torch.float32))
        self.shift2.copy_(shift.detach().to(self.device, torch.float32))
        return self

    def load2(self, weight: torch.Tensor, scale: torch.Tensor, shift: torch.Tensor, weight2: torch.Tensor, scale2: torch.Tensor,
              shift2: torch.Tensor):
        """Two-input layer (``Cin2``): ``y = act(bn(conv(x)) + bn2(conv2(x2)))`` -- conv3 + bn3 + the projection shortcut
        ``downsample(x)`` + add + ReLU of timm ``Bottleneck.forward`` (resnet.py:247-268) as ONE GEMM over the concatenated
        K.  The two accumulations share one TMEM tile, so both BN scales are folded into the weights (in fp32, before the
        bf16 rounding) and the epilogue applies ``shift + shift2`` only."""
        f32 = dict(device=self.device, dtype=torch.float32)
        w = (weight.detach().to(**f32) * scale.detach().to(**f32).view(-1, 1, 1, 1)).contiguous()
        w2 = (weight2.detach().to(**f32) * scale2.detach().to(**f32).view(-1, 1, 1, 1)).contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().hpe_conv_pack_weight2(C.byref(self.desc), _lib.ptr(w), _lib.ptr(w2), _lib.ptr(self.w_packed),
                                                        _lib.stream_ptr()), "hpe_conv_pack_weight2")
        self.scale.fill_(1.0)
        self.shift.copy_(shift.detach().to(**f32) + shift2.detach().to(**f32))
        return self

    def __call__(self, x: torch.Tensor, residual: torch.Tensor = None, out: torch.Tensor = None, x2: torch.Tensor = None,
                 out2: torch.Tensor = None):
        if out is None:
            out = self.alloc_out()
        if self.desc.in_cstride or self.desc.out_cstride:
            # channel slices of wider NHWC buffers: the innermost stride must be 1 and the pixel pitch the declared one
            _lib.require_cuda(residual)
            for t, pitch in ((x, self.desc.in_cstride), (out, self.desc.out_cstride)):
                if not t.is_cuda or (pitch and self.kind != CONV_FINAL and (t.stride(-1) != 1 or t.stride(-2) != pitch)):
                    raise RuntimeError("libhpe_b200: channel-slice tensor does not match the plan's channel pitch")
        else:
            _lib.require_cuda(x, residual, out)
        _lib.require_cuda(x2)
        if self.head2:
            if out.dtype != torch.float32:
                raise TypeError("the fused head writes float32 heatmaps")
            with torch.cuda.device(self.device):
                rc = _lib.lib().hpe_conv_run_head2(self.plan, _lib.ptr(x), _lib.ptr(self.w_packed), _lib.ptr(self.scale),
                                                   _lib.ptr(self.shift), _lib.ptr(self.w2_packed), _lib.ptr(self.bias2),
                                                   _lib.ptr(out), _lib.stream_ptr())
            _lib.check(rc, "hpe_conv_run_head2")
            return out
        if self.chain:
            if out2 is None:
                out2 = torch.empty((self.B, self.Ho, self.Wo, self.chain), dtype=torch.bfloat16, device=self.device)
            _lib.require_cuda(out2)
            with torch.cuda.device(self.device):
                rc = _lib.lib().hpe_conv_run_chain(self.plan, _lib.ptr(x), _lib.ptr(x2), _lib.ptr(self.w_packed), _lib.ptr(self.scale),
                                                   _lib.ptr(self.shift), _lib.ptr(residual), _lib.ptr(out), _lib.ptr(self.w2_packed),
                                                   _lib.ptr(self.scale2), _lib.ptr(self.shift2), _lib.ptr(out2), _lib.stream_ptr())
            _lib.check(rc, "hpe_conv_run_chain")
            return out, out2
        with torch.cuda.device(self.device):
            rc = _lib.lib().hpe_conv_run2(self.plan, _lib.ptr(x), _lib.ptr(x2), _lib.ptr(self.w_packed), _lib.ptr(self.scale),
                                          _lib.ptr(self.shift), _lib.ptr(residual), _lib.ptr(out), _lib.stream_ptr())
        _lib.check(rc, "hpe_conv_run")
        return out


def preprocess(img: torch.Tensor, out: torch.Tensor = None):
    """fp32 (or uint8, = ``img.float() / 255``: HPE/train.py:117-120) NCHW [B,3,H,W] -> the bf16 space-to-depth staging
    tensor the stem conv reads."""
    _lib.require_cuda(img)
    B, Cc, H, W = img.shape
    if Cc != 3 or img.dtype not in (torch.float32, torch.uint8):
        raise ValueError("preprocess expects float32 or uint8 [B,3,H,W]")
    n = _lib.lib().hpe_preprocess_out_elems(B, H, W)
    if n == 0:
        raise ValueError("H and W must be even")
    if out is None:
        out = torch.empty(n, dtype=torch.bfloat16, device=img.device)
    fn = _lib.lib().hpe_preprocess_u8 if img.dtype == torch.uint8 else _lib.lib().hpe_preprocess
    with torch.cuda.device(img.device):
        _lib.check(fn(_lib.ptr(img), B, H, W, _lib.ptr(out), _lib.stream_ptr()), "hpe_preprocess")
    return out


def maxpool3x3s2(x: torch.Tensor, out: torch.Tensor = None):
    """MaxPool2d(3, stride 2, padding 1) on NHWC bf16 (timm resnet.py:574-576)."""
    _lib.require_cuda(x, out)
    B, H, W, Cc = x.shape
    if out is None:
        out = torch.empty((B, (H - 1) // 2 + 1, (W - 1) // 2 + 1, Cc), dtype=torch.bfloat16, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().hpe_maxpool3x3s2(_lib.ptr(x), B, H, W, Cc, _lib.ptr(out), _lib.stream_ptr()), "hpe_maxpool3x3s2")
    return out


class DgradLayer(ConvLayer):
    """Backward-data convolution of a forward layer: ``dx = dgrad(dz) (+ residual)`` on the same tcgen05 kernel
    (weights re-packed: transposed / 180-degree rotated / per-parity-class).  ``fwd`` = the forward layer's
    (kind, B, H, W, Cin, Cout, stride)."""

    def __init__(self, kind, B, H, W, Cin, Cout, stride=1, accumulate=False, device="cuda"):
        self.fwd_desc = ConvDesc(kind, B, H, W, Cin, Cout, stride, ACT_NONE, 0)
        g = ConvDesc()
        _lib.check(_lib.lib().hpe_conv_dgrad_desc(C.byref(self.fwd_desc), C.byref(g)), "hpe_conv_dgrad_desc")
        g.has_residual = int(bool(accumulate))
        super().__init__(None, None, None, None, None, None, device=device, desc=g)

    def load(self, weight: torch.Tensor, scale=None, shift=None):
        """weight: the FORWARD layer's PyTorch-layout fp32 weight."""
        w = weight.detach().to(self.device, torch.float32).contiguous()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().hpe_conv_pack_weight_dgrad(C.byref(self.fwd_desc), _lib.ptr(w), _lib.ptr(self.w_packed),
                                                             _lib.stream_ptr()), "hpe_conv_pack_weight_dgrad")
        return self


class WgradLayer:
    """Weight gradient of one conv layer (csrc/wgrad_gemm.cu): ``dW = wgrad(x, dz)`` in the PyTorch weight
    layout, fp32.  Replaces the cuDNN wgrad autograd launches behind ``loss.backward()``
    (/root/reference/HPE/engine/trainer.py:96-97)."""

    def __init__(self, kind, B, H, W, Cin, Cout, stride=1, device="cuda"):
        self.device = torch.device(device)
        self.desc = ConvDesc(kind, B, H, W, Cin, Cout, stride, ACT_NONE, 0)
        plan = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().hpe_wgrad_plan_create(C.byref(self.desc), C.byref(plan)), "hpe_wgrad_plan_create")
        self.plan = plan
        n = _lib.lib().hpe_wgrad_workspace_elems(plan)
        self.workspace = torch.empty(n, dtype=torch.float32, device=self.device)

    def __del__(self):
        try:
            if getattr(self, "plan", None):
                _lib.lib().hpe_wgrad_plan_destroy(self.plan)
                self.plan = None
        except Exception:
            pass

    def __call__(self, x: torch.Tensor, dz: torch.Tensor, dw: torch.Tensor, accumulate=False):
        _lib.require_cuda(x, dz, dw)
        if dw.dtype != torch.float32:
            raise TypeError("weight gradients are float32")
        with torch.cuda.device(self.device):
            rc = _lib.lib().hpe_wgrad_run(self.plan, _lib.ptr(x), _lib.ptr(dz), _lib.ptr(self.workspace), _lib.ptr(dw),
                                          int(bool(accumulate)), _lib.stream_ptr())
        _lib.check(rc, "hpe_wgrad_run")
        return dw
