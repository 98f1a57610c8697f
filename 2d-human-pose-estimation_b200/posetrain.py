"""Training-mode forward + backward of the reference's real pose model on the sm_100a kernels.

The reference trains ``DINOv3Pose`` with its backbone frozen (``finetuning=True`` -> ``requires_grad = False`` on every ConvNeXt
parameter, /root/reference/HPE/models/pose.py:47-49; training loop HPE/train.py:116-139: ``preds = model(imgs)`` under
``model.train()``, ``loss = criterion(preds, targets)``, ``loss.backward()``, ``clip_grad_norm_``, ``AdamW.step()``).  What needs
gradients is the FeatureAdaptor, SPPF, PAN and PoseHead: 70-odd ``ConvBlock`` units (Conv2d without bias + BatchNorm2d with batch
statistics + SiLU, block.py:20-45), two plain 1x1 convs and a BatchNorm over a concat per BottleneckCSP (block.py:74-100), a
depthwise 3x3 ConvBlock per level, three 5x5 max-pools, two bilinear up-samplings, six biased 1x1 output convs.

Here every one of those ops is a ``torch.autograd.Function`` whose forward AND backward are launches of this library:

* convolutions: the tcgen05 implicit-GEMM kernel for the forward (raw output, ``ACT_NONE``), for the backward-data pass
  (``DgradLayer``: transposed / rotated / parity-class weights) and the tcgen05 wgrad GEMM (``WgradLayer``) -- the kernels the
  ResNet training step uses (train.py);
* BatchNorm + SiLU, depthwise 3x3, max-pool, up-sampling and their backward passes: csrc/neck_train_ops.cu, convnext_ops.cu;
* the frozen backbone: the inference launch plan (posemodel._PosePlan, CUDA-graph replay), no autograd.

PyTorch's autograd only does the bookkeeping of the graph (which Function feeds which, summing the gradients of a tensor with
several consumers, splitting the gradient of ``torch.cat``); activations are NHWC bf16, parameters and their gradients fp32, exactly
the tensors ``model.parameters()`` exposes, so ``torch.optim.AdamW`` / ``clip_grad_norm_`` -- or ``hpe_adamw_step`` -- work unchanged.
Channel counts that are not multiples of 64 (96 inside ``fpn_conv4``, 51 in the keypoint branch) run zero-padded to the next
multiple of 64; the padding stays exactly zero through BN (beta_pad = 0) and SiLU and is sliced off before a concat.
There is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.nn.functional as F

from . import _lib
from ._lib import ACT_NONE, CONV_1x1, CONV_3x3, CONV_FINAL
from .conv import ConvLayer, DgradLayer, WgradLayer
from .posemodel import _PosePlan, _dwconv, _upsample

ACT_SILU_BN = 2          # act code of the hpe_bnact_* kernels: 0 none, 1 ReLU, 2 SiLU


def _cp(c: int) -> int:
    return (c + 63) // 64 * 64


def _plans(model):
    if not hasattr(model, "_tplans"):
        model._tplans = {}
    return model._tplans


def _f32(n, dev, fill=None):
    t = torch.empty(n, dtype=torch.float32, device=dev)
    if fill is not None:
        t.fill_(fill)
    return t


# ------------------------------------------------------------------------------------------------- convolution
class _ConvSpec:
    """Forward / backward-data / weight-gradient plans of one conv at one input shape (channels padded to multiples of 64)."""

    def __init__(self, kind, B, H, W, cin, cout, stride, dev, need_dx):
        self.cin, self.cout, self.cinp, self.coutp = cin, cout, _cp(cin), _cp(cout)
        self.k = 3 if kind == CONV_3x3 else 1
        self.fwd = ConvLayer(kind, B, H, W, self.cinp, self.coutp, stride, ACT_NONE, False, dev)
        self.dgrad = DgradLayer(kind, B, H, W, self.cinp, self.coutp, stride, device=dev) if need_dx else None
        self.wgrad = WgradLayer(kind, B, H, W, self.cinp, self.coutp, stride, dev)
        self.mod = None              # the nn.Conv2d whose weight this spec packs (set by conv())
        self.wsrc = None             # persistent zero-padded fp32 copy of the weight (layers with padded channels), see prepack()
        self.prepacked = False       # True while a PoseTrainStep has packed both operands for the step in flight

    def pad_weight(self, w):
        if self.cinp == self.cin and self.coutp == self.cout:
            return w.detach().float().contiguous()
        wp = torch.zeros((self.coutp, self.cinp, self.k, self.k), dtype=torch.float32, device=w.device)
        wp[:self.cout, :self.cin] = w.detach()
        return wp


class ConvFn(torch.autograd.Function):
    """``z = conv(x, w)`` (no bias): Conv2d of a ConvBlock / BottleneckCSP.cv2 / cv3 (block.py:24, 82-83)."""

    @staticmethod
    def forward(ctx, x, w, spec):
        if spec.prepacked:           # operands packed by prepack() in one launch for the whole step
            wp = spec.wsrc if spec.wsrc is not None else w.detach()
        else:
            wp = spec.pad_weight(w)
            spec.fwd.load(wp)
        z = spec.fwd(x)
        ctx.save_for_backward(x, wp)
        ctx.spec = spec
        return z

    @staticmethod
    def backward(ctx, dz):
        x, wp = ctx.saved_tensors
        spec = ctx.spec
        dz = dz.contiguous()
        dx = dw = None
        if ctx.needs_input_grad[0]:
            if not spec.prepacked:
                spec.dgrad.load(wp)
            dx = spec.dgrad(dz)
        if ctx.needs_input_grad[1]:
            tgt = _direct_target(spec.mod, "weight") if spec.mod is not None else None
            if tgt is not None and tuple(wp.shape) == tuple(tgt.shape):
                spec.wgrad(x, dz, tgt)                       # straight into the flat gradient buffer
            else:
                dwp = torch.empty_like(wp)
                spec.wgrad(x, dz, dwp)
                if tgt is not None:
                    tgt.copy_(dwp[:spec.cout, :spec.cin])
                else:
                    dw = dwp[:spec.cout, :spec.cin].contiguous() if dwp.shape[:2] != (spec.cout, spec.cin) else dwp
        return dx, dw, None


def conv(model, mod, x, kind, stride=1, need_dx=True):
    """``mod``: the nn.Conv2d parameter container; x NHWC bf16 with ceil64(Cin) channels."""
    B, H, W, _ = x.shape
    key = ("conv", id(mod), B, H, W, need_dx)
    spec = _plans(model).get(key)
    if spec is None:
        spec = _plans(model)[key] = _ConvSpec(kind, B, H, W, mod.in_channels, mod.out_channels, stride, x.device, need_dx)
        spec.mod = mod
    return ConvFn.apply(x.contiguous(), mod.weight, spec)


def prepack(model):
    """Pack the forward and backward-data operands of EVERY ConvFn layer of the model's training plans in ONE launch
    (hpe_conv_pack_jobs_run, as ``TrainPlan.repack`` does for the ResNet step) instead of two launches per layer per step, and
    mark the specs so that ConvFn skips its own packing.  Layers with padded channels pack from a persistent zero-padded fp32
    copy of the weight (one strided copy per step).  Returns the specs it marked (``release`` un-marks them); [] before the
    first forward has created the plans."""
    specs = [sp for k, sp in _plans(model).items() if k[0] == "conv" and sp.mod is not None]
    if not specs:
        return []
    items = []
    for sp in specs:
        w = sp.mod.weight
        if sp.cinp != sp.cin or sp.coutp != sp.cout:
            if sp.wsrc is None:
                sp.wsrc = torch.zeros((sp.coutp, sp.cinp, sp.k, sp.k), dtype=torch.float32, device=w.device)
            sp.wsrc[:sp.cout, :sp.cin].copy_(w.detach())
            src = sp.wsrc
        else:
            src = w.detach()
            if not (src.dtype == torch.float32 and src.is_contiguous()):
                return []
        items.append((sp.fwd.desc, 0, src, sp.fwd.w_packed))
        if sp.dgrad is not None:
            items.append((sp.dgrad.fwd_desc, 1, src, sp.dgrad.w_packed))
    cache = _plans(model).setdefault(("prepack",), {})
    key = tuple((w.data_ptr(), wp.data_ptr()) for _, _, w, wp in items)
    lib = _lib.lib()
    if cache.get("key") != key:
        jb = lib.hpe_pack_job_bytes()
        host = (C.c_char * (jb * len(items)))()
        begin = 0
        for k, (desc, dg, w, wp) in enumerate(items):
            nb = C.c_int()
            _lib.check(lib.hpe_conv_pack_job_fill(C.byref(desc), dg, w.data_ptr(), wp.data_ptr(), begin,
                                                  C.addressof(host) + k * jb, C.byref(nb)), "hpe_conv_pack_job_fill")
            begin += nb.value
        cache.update(key=key, jobs=torch.frombuffer(bytearray(host.raw), dtype=torch.uint8).to(items[0][2].device), n=len(items), blocks=begin)
    with torch.cuda.device(items[0][2].device):
        _lib.check(lib.hpe_conv_pack_jobs_run(_lib.ptr(cache["jobs"]), cache["n"], cache["blocks"], _lib.stream_ptr()),
                   "hpe_conv_pack_jobs_run")
    for sp in specs:
        sp.prepacked = True
    return specs


def release(specs):
    for sp in specs:
        sp.prepacked = False


# Direct gradient targets (PoseTrainStep): a parameter whose gradient is produced by exactly one kernel per step (conv weight:
# wgrad; BatchNorm gamma / beta: hpe_bnact_backward) can have that kernel write straight into the parameter's slice of the flat
# gradient buffer -- no temporary, no slice copy, no autograd accumulation kernel (181 elementwise adds per step otherwise).
# `_DIRECT[0]` is the current step's token (0 = off); a module remembers the token of its last direct write in `_hpe_grad_step`,
# so a module that runs twice in one step (shared weights) falls back to the accumulating autograd path on its second use.
_DIRECT = [0]
_NBT = [None]        # PoseTrainStep: set of id(bn) of the training-mode BatchNorm modules the forward in flight ran (else None)


def _direct_target(mod, name):
    """The flat-buffer view registered for `mod.<name>` if this use may write it directly, else None."""
    tok = _DIRECT[0]
    tgt = getattr(mod, "_hpe_grad", None) if tok else None
    if tgt is None or name not in tgt:
        return None
    seen = mod.__dict__.setdefault("_hpe_grad_step", {})
    if seen.get(name) == tok:
        return None
    seen[name] = tok
    return tgt[name]


# ------------------------------------------------------------------------------------------------- BatchNorm + activation
class BNActFn(torch.autograd.Function):
    """``y = act(BatchNorm2d(z))``: BATCH statistics when the module is in training mode (running statistics updated as torch
    does), its running statistics as constants when it is in eval mode (``bn.eval()`` inside a training model)."""

    @staticmethod
    def forward(ctx, z, gamma, beta, bn, act):
        Cp, Cr = z.shape[-1], gamma.numel()
        M = z.numel() // Cp
        dev = z.device
        g, b = gamma.detach().float(), beta.detach().float()
        rm, rv = bn.running_mean, bn.running_var
        if Cp != Cr:                                  # zero-padded channels: gamma 1, beta 0 keep them exactly zero
            g = torch.cat([g, torch.ones(Cp - Cr, device=dev)]); b = torch.cat([b, torch.zeros(Cp - Cr, device=dev)])
            rm = torch.cat([rm, torch.zeros(Cp - Cr, device=dev)]); rv = torch.cat([rv, torch.ones(Cp - Cr, device=dev)])
        g, b = g.contiguous(), b.contiguous()
        frozen = not bn.training
        if frozen:                                    # eval-mode BatchNorm: y = z * gamma / sqrt(running_var + eps) + (beta - ...)
            mean, rstd = rm.detach().float().contiguous(), torch.rsqrt(rv.detach().float() + bn.eps)
            scale = (g * rstd).contiguous(); shift = (b - mean * scale).contiguous()
        else:
            scale, shift, mean, rstd = (_f32(Cp, dev) for _ in range(4))
            scratch = _f32(_lib.lib().hpe_bnact_scratch_elems(Cp), dev)
            mom = bn.momentum if bn.momentum is not None else 0.1
            with torch.cuda.device(dev):
                _lib.check(_lib.lib().hpe_bnact_stats(_lib.ptr(z), M, Cp, _lib.ptr(g), _lib.ptr(b), float(bn.eps), float(mom), _lib.ptr(rm),
                                                      _lib.ptr(rv), _lib.ptr(scratch), _lib.ptr(scale), _lib.ptr(shift), _lib.ptr(mean),
                                                      _lib.ptr(rstd), _lib.stream_ptr()), "hpe_bnact_stats")
        with torch.cuda.device(dev):
            y = torch.empty_like(z)
            _lib.check(_lib.lib().hpe_bnact_apply(_lib.ptr(z), M, Cp, _lib.ptr(scale), _lib.ptr(shift), int(act), _lib.ptr(y),
                                                  _lib.stream_ptr()), "hpe_bnact_apply")
        if not frozen:
            if Cp != Cr:
                bn.running_mean.copy_(rm[:Cr]); bn.running_var.copy_(rv[:Cr])
            if bn.num_batches_tracked is not None:
                if _NBT[0] is not None:
                    _NBT[0].add(id(bn))              # PoseTrainStep adds 1 to all of them with ONE launch after the forward
                else:
                    bn.num_batches_tracked += 1
        ctx.save_for_backward(z, g, scale, shift, mean, rstd)
        ctx.act, ctx.Cr, ctx.frozen, ctx.bn = int(act), Cr, frozen, bn
        return y

    @staticmethod
    def backward(ctx, dy):
        z, g, scale, shift, mean, rstd = ctx.saved_tensors
        Cp = z.shape[-1]
        M = z.numel() // Cp
        dev = z.device
        dy = dy.contiguous()
        dz = torch.empty_like(z)
        direct = False
        if Cp == ctx.Cr and ctx.needs_input_grad[1] and ctx.needs_input_grad[2]:
            tg, tb = _direct_target(ctx.bn, "weight"), _direct_target(ctx.bn, "bias")
            direct = tg is not None and tb is not None
        dbeta, dgamma = (tb, tg) if direct else (_f32(Cp, dev), _f32(Cp, dev))
        scratch = _f32(_lib.lib().hpe_bnact_scratch_elems(Cp), dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().hpe_bnact_backward(_lib.ptr(dy), _lib.ptr(z), M, Cp, _lib.ptr(mean), _lib.ptr(rstd), _lib.ptr(g),
                                                     _lib.ptr(scale), _lib.ptr(shift), ctx.act, int(ctx.frozen), _lib.ptr(scratch), _lib.ptr(dbeta),
                                                     _lib.ptr(dgamma), _lib.ptr(dz), _lib.stream_ptr()), "hpe_bnact_backward")
        if direct:
            return dz, None, None, None, None
        return dz, dgamma[:ctx.Cr].contiguous(), dbeta[:ctx.Cr].contiguous(), None, None


def bn_act(bn, z, act=ACT_SILU_BN):
    return BNActFn.apply(z.contiguous(), bn.weight, bn.bias, bn, act)


def conv_block(model, cb, x, need_dx=True):
    """ConvBlock.forward (block.py:41-42): act(bn(conv(x))), 1x1 / 3x3, stride 1 | 2, SiLU."""
    z = conv(model, cb.conv, x, CONV_3x3 if cb.k == 3 else CONV_1x1, cb.s, need_dx)
    return bn_act(cb.bn, z)


# ------------------------------------------------------------------------------------------------- depthwise 3x3, pool, up-sampling
class DwConv3Fn(torch.autograd.Function):
    """Depthwise 3x3 conv (groups = C, pad 1): the head's class branch opens with ConvBlock(x, x, 3, g=x) (head.py:143-144)."""

    @staticmethod
    def forward(ctx, x, w):
        Cc = x.shape[-1]
        wt = w.detach().float().reshape(Cc, 9).t().contiguous()                 # [9][C]
        one, zero = _f32(Cc, x.device, 1.0), _f32(Cc, x.device, 0.0)
        y = torch.empty_like(x)
        _dwconv(x, wt, one, zero, None, -1.0, 3, 0, y)                            # eps < 0: raw output (BN + SiLU follow as their own op)
        ctx.save_for_backward(x, wt, one, zero)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, wt, one, zero = ctx.saved_tensors
        B, H, W, Cc = x.shape
        dy = dy.contiguous()
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            _dwconv(dy, wt.flip(0).contiguous(), one, zero, None, -1.0, 3, 0, dx)  # correlation with the 180-degree rotated taps
        if ctx.needs_input_grad[1]:
            dw = torch.empty((Cc, 1, 3, 3), dtype=torch.float32, device=x.device)
            scratch = _f32(_lib.lib().hpe_dwconv3_wgrad_scratch_elems(Cc), x.device)
            with torch.cuda.device(x.device):
                _lib.check(_lib.lib().hpe_dwconv3_wgrad(_lib.ptr(x), _lib.ptr(dy), B, H, W, Cc, _lib.ptr(scratch), _lib.ptr(dw),
                                                        _lib.stream_ptr()), "hpe_dwconv3_wgrad")
        return dx, dw


class MaxPool5Fn(torch.autograd.Function):
    """MaxPool2d(5, 1, 2) of SPPF (block.py:108, 111-113)."""

    @staticmethod
    def forward(ctx, x):
        B, H, W, Cc = x.shape
        y = torch.empty_like(x)
        idx = torch.empty(x.shape, dtype=torch.uint8, device=x.device)          # window position of the first maximum
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().hpe_maxpool5_train(_lib.ptr(x), B, H, W, Cc, _lib.ptr(y), _lib.ptr(idx), _lib.stream_ptr()),
                       "hpe_maxpool5_train")
        ctx.save_for_backward(idx)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        B, H, W, Cc = idx.shape
        dx = torch.empty(idx.shape, dtype=dy.dtype, device=dy.device)
        with torch.cuda.device(dy.device):
            _lib.check(_lib.lib().hpe_maxpool5_backward(_lib.ptr(idx), _lib.ptr(dy.contiguous()), B, H, W, Cc, _lib.ptr(dx),
                                                        _lib.stream_ptr()), "hpe_maxpool5_backward")
        return dx


class UpsampleFn(torch.autograd.Function):
    """F.interpolate(size=(Ho, Wo), mode='bilinear', align_corners=False) of PAN (block.py:138, 141)."""

    @staticmethod
    def forward(ctx, x, Ho, Wo):
        B, Hi, Wi, Cc = x.shape
        y = torch.empty((B, Ho, Wo, Cc), dtype=x.dtype, device=x.device)
        _upsample(x, y)
        ctx.shape = (B, Hi, Wi, Cc, Ho, Wo)
        return y

    @staticmethod
    def backward(ctx, dy):
        B, Hi, Wi, Cc, Ho, Wo = ctx.shape
        dx = torch.empty((B, Hi, Wi, Cc), dtype=dy.dtype, device=dy.device)
        with torch.cuda.device(dy.device):
            _lib.check(_lib.lib().hpe_upsample_bilinear_backward(_lib.ptr(dy.contiguous()), B, Hi, Wi, Cc, Ho, Wo, _lib.ptr(dx),
                                                                 _lib.stream_ptr()), "hpe_upsample_bilinear_backward")
        return dx, None, None


# ------------------------------------------------------------------------------------------------- biased 1x1 output conv
class _FinalSpec:
    def __init__(self, B, H, W, cin, K, dev):
        self.cin, self.cinp, self.K = cin, _cp(cin), K
        self.fwd = ConvLayer(CONV_FINAL, B, H, W, self.cinp, K, 1, ACT_NONE, False, dev)
        self.dgrad = DgradLayer(CONV_FINAL, B, H, W, self.cinp, K, 1, device=dev)
        self.wgrad = WgradLayer(CONV_FINAL, B, H, W, self.cinp, K, 1, dev)
        self.ones = torch.ones(K, dtype=torch.float32, device=dev)


class FinalConvFn(torch.autograd.Function):
    """``nn.Conv2d(c, K, 1)`` with bias -> fp32 NCHW maps: the last layer of the head's class / keypoint branches
    (head.py:154-160)."""

    @staticmethod
    def forward(ctx, x, w, b, spec):
        wp = w.detach().float()
        if spec.cinp != spec.cin:
            wp = F.pad(wp, (0, 0, 0, 0, 0, spec.cinp - spec.cin))
        wp = wp.contiguous()
        spec.fwd.load(wp, spec.ones, b.detach())
        out = spec.fwd(x)                                   # [B, K, H, W] fp32
        ctx.save_for_backward(x, wp)
        ctx.spec = spec
        return out

    @staticmethod
    def backward(ctx, dout):
        from .train import heatmap_grad_nhwc
        x, wp = ctx.saved_tensors
        spec = ctx.spec
        B, H, W, _ = x.shape
        dout = dout.contiguous().float()
        g = torch.empty((B, H, W, 64), dtype=torch.bfloat16, device=x.device)       # joints zero-padded to 64
        db = _f32(spec.K, x.device)
        heatmap_grad_nhwc(dout, g, db)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            spec.dgrad.load(wp)
            dx = spec.dgrad(g)
        if ctx.needs_input_grad[1]:
            dwp = torch.empty_like(wp)
            spec.wgrad(x, g, dwp)
            dw = dwp[:, :spec.cin].contiguous()
        return dx, dw, db, None


def final_conv(model, mod, x):
    B, H, W, _ = x.shape
    key = ("final", id(mod), B, H, W)
    spec = _plans(model).get(key)
    if spec is None:
        spec = _plans(model)[key] = _FinalSpec(B, H, W, mod.in_channels, mod.out_channels, x.device)
    return FinalConvFn.apply(x.contiguous(), mod.weight, mod.bias, spec)


# ------------------------------------------------------------------------------------------------- the model
def _csp(model, m, x):
    """BottleneckCSP.forward (block.py:96-100), shortcut=False."""
    c_ = m.c_
    y1 = conv_block(model, m.cv1, x)
    for b in m.m:
        y1 = conv_block(model, b.cv2, conv_block(model, b.cv1, y1))
    y1 = conv(model, m.cv3, y1, CONV_1x1)
    y2 = conv(model, m.cv2, x, CONV_1x1)
    cat = torch.cat([y1[..., :c_], y2[..., :c_]], -1)            # the zero padding of a 96-channel half is dropped here
    return conv_block(model, m.cv4, bn_act(m.bn, cat))


def train_forward(model, x):
    """``DINOv3Pose.forward`` under ``model.train()`` with autograd on: the list of raw per-level maps
    ``[B, nc + nk, H_l, W_l]`` fp32 (head.py:176-181), differentiable w.r.t. every adaptor / SPPF / PAN / head parameter."""
    if x.dim() != 4 or x.shape[1] != 3:
        raise ValueError(f"expected [B,3,H,W] input, got {tuple(x.shape)}")
    if not x.is_cuda:
        raise RuntimeError("libhpe_b200 has no CPU path: input must live on a CUDA (sm_100a) device")
    if any(p.requires_grad for p in model.backbone.parameters()):
        raise RuntimeError("training is built for the reference's configuration: frozen backbone (finetuning=True, pose.py:47-49)")
    B, _, H, W = x.shape
    key = ("backbone", B, H, W, x.device.index)
    plans = _plans(model)
    plan = plans.get(key)
    if plan is None:
        with torch.no_grad():
            plan = plans[key] = _PosePlan(model, B, H, W, x.device, backbone_only=True)
    with torch.no_grad():
        plan.run(x.contiguous() if x.dtype == torch.uint8 else x.float().contiguous())
        f1, f2, f3 = (f.clone() for f in plan.feats)             # the plan's buffers are overwritten by the next step
    # ---- FeatureAdaptor (block.py:236-254), SPPF (block.py:102-114)
    ad = model.adaptor.adapters
    c1 = conv_block(model, ad[0][1], conv_block(model, ad[0][0], f1, need_dx=False))
    c2 = conv_block(model, ad[1][1], conv_block(model, ad[1][0], f2, need_dx=False))
    c3 = conv_block(model, ad[2][1], conv_block(model, ad[2][0], f3, need_dx=False))
    y = [conv_block(model, model.sppf.conv1, c3)]
    for _ in range(3):
        y.append(MaxPool5Fn.apply(y[-1]))
    c3 = conv_block(model, model.sppf.conv2, torch.cat(y, -1))
    # ---- PAN (block.py:132-157); F.interpolate to an unchanged size (after down1 / down2) is the identity
    pan = model.pan
    up = lambda t, ref: UpsampleFn.apply(t, ref.shape[1], ref.shape[2])          # noqa: E731
    c2f = _csp(model, pan.fpn_conv2, torch.cat([c2, up(conv_block(model, pan.fpn_conv1, c3), c2)], -1))
    c1f = _csp(model, pan.fpn_conv4, torch.cat([c1, up(conv_block(model, pan.fpn_conv3, c2f), c1)], -1))
    c2p = _csp(model, pan.pan_conv1, torch.cat([conv_block(model, pan.down1, c1f), c2f], -1))
    c3p = _csp(model, pan.pan_conv2, torch.cat([conv_block(model, pan.down2, c2p), c3], -1))
    # ---- PoseHead (head.py:173-181)
    hd = model.head
    outs = []
    for i, f in enumerate((c1f, c2p, c3p)):
        dw = hd.cv3[i][0][0]
        t = bn_act(dw.bn, DwConv3Fn.apply(f.contiguous(), dw.conv.weight))
        t = conv_block(model, hd.cv3[i][0][1], t)
        t = conv_block(model, hd.cv3[i][1][1], conv_block(model, hd.cv3[i][1][0], t))
        cls = final_conv(model, hd.cv3[i][2], t)
        k = conv_block(model, hd.cv4[i][1], conv_block(model, hd.cv4[i][0], f))
        kpt = final_conv(model, hd.cv4[i][2], k)
        outs.append(torch.cat([cls, kpt], 1))
    return outs


# ------------------------------------------------------------------------------------------------- the training step
class PoseTrainStep:
    """The reference's training step for its real model (HPE/train.py:116-139) as one call, data-parallel:

        step = PoseTrainStep(model, lr=1e-4, weight_decay=1e-4, max_grad_norm=10.0)     # AdamW + clip_grad_norm_(10), train.py:52-55, 136
        loss, (cls, kpt, vis) = step(images, targets)

    ``images`` [B,3,H,W] uint8 or fp32 on the device (uint8 = ``imgs.float() / 255``, train.py:117-120, fused into the stem);
    ``targets`` rows ``[b, cls, cx, cy, w, h, (x, y, v) * K]`` normalised (HPE/dataset.py:23-37).  Forward (frozen backbone +
    batch-statistics neck / head), ``ComputeLoss`` with its analytic gradient, backward through the dgrad / wgrad kernels into ONE
    flat fp32 gradient buffer (the parameters' ``.grad`` are views of it), ONE NCCL all-reduce of that buffer under
    ``torch.distributed`` (per-rank BatchNorm statistics, averaged gradients: what the reference's DDP wrap does,
    pose/pose/core/builder.py:92-123), then the gradient-norm kernel and ONE AdamW kernel over the flat parameter buffer."""

    def __init__(self, model, lr=1e-4, weight_decay=1e-4, betas=(0.9, 0.999), eps=1e-8, max_grad_norm=10.0, group=None, cuda_graph=False):
        import torch.distributed as dist
        from .detpose import ComputeLoss
        from .train import ParamLayout
        self.model = model.train()
        self.criterion = ComputeLoss(model, items=False)
        self.lr, self.weight_decay, self.betas, self.eps, self.max_grad_norm = lr, weight_decay, tuple(betas), eps, float(max_grad_norm)
        self.group, self._t = group, 0
        # cuda_graph: capture forward + loss + backward (~2700 launches) once per (image shape, target shape) and replay it; the
        # all-reduce and the optimizer kernels follow the replay.  Needs a fixed number of target rows per step.
        self.use_graph, self.graph, self._key, self._warm = bool(cuda_graph), None, None, 0
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.layout = ParamLayout(model)                       # trainable parameters first
        dev = next(model.parameters()).device
        self.device = dev
        self.flat_param = torch.zeros(self.layout.total, dtype=torch.float32, device=dev)
        nt = self.layout.trainable_total
        self.flat_grad = torch.zeros(nt, dtype=torch.float32, device=dev)
        self.exp_avg, self.exp_avg_sq = torch.zeros_like(self.flat_grad), torch.zeros_like(self.flat_grad)
        self.grad_sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        pv = self.layout.views(self.flat_param)
        with torch.no_grad():
            for n, p in model.named_parameters():              # re-point the parameters at the flat buffer
                pv[n].copy_(p.data)
                p.data = pv[n]
        self._trainable = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        self._grad_views = {n: self.flat_grad.narrow(0, self.layout.offsets[n], p.numel()).view(p.shape) for n, p in self._trainable}
        # direct gradient targets of the conv weights and BatchNorm affine parameters (see _direct_target)
        mods = dict(model.named_modules())
        for n, p in self._trainable:
            mname, _, pname = n.rpartition(".")
            m = mods.get(mname)
            if isinstance(m, (torch.nn.Conv2d, torch.nn.BatchNorm2d)) and pname in ("weight", "bias") and \
                    not (isinstance(m, torch.nn.Conv2d) and pname == "bias"):
                m.__dict__.setdefault("_hpe_grad", {})[pname] = self._grad_views[n]
        self._step_token = 0
        # `num_batches_tracked` of every BatchNorm re-pointed at one flat int64 buffer: one masked add per step instead of one
        # tiny kernel per module (the mask = the modules that ran in training mode; torch increments only those)
        self._bns = [m for m in model.modules() if isinstance(m, torch.nn.BatchNorm2d) and m.num_batches_tracked is not None]
        self.flat_nbt = torch.zeros(max(len(self._bns), 1), dtype=torch.int64, device=dev)
        with torch.no_grad():
            for i, m in enumerate(self._bns):
                self.flat_nbt[i] = m.num_batches_tracked.to(dev)
                m.num_batches_tracked.data = self.flat_nbt[i]
        self._nbt_mask, self._nbt_key = None, None

    def broadcast_parameters(self, src=0):
        import torch.distributed as dist
        if self.world > 1:
            dist.broadcast(self.flat_param, src, group=self.group)
            for b in self.model.buffers():
                if b.dtype.is_floating_point:
                    dist.broadcast(b, src, group=self.group)

    def _compute(self, images, targets):
        """forward -> ComputeLoss -> backward into the flat gradient buffer (this rank's shard)."""
        self.flat_grad.zero_()
        for n, p in self._trainable:                           # autograd accumulates in place into the flat buffer
            p.grad = self._grad_views[n]
        packed = prepack(self.model) if os.environ.get("HPE_B200_POSE_PREPACK", "1") != "0" else []
        self._step_token += 1
        _DIRECT[0] = (id(self), self._step_token) if os.environ.get("HPE_B200_POSE_DIRECT_GRAD", "1") != "0" else 0
        _NBT[0] = set()
        try:
            levels = train_forward(self.model, images)
            ran = _NBT[0]
            _NBT[0] = None
            key = tuple(id(m) in ran for m in self._bns)
            if key != self._nbt_key:                      # (re)built outside any graph capture: the eager warm-up steps come first
                self._nbt_key = key
                self._nbt_mask = torch.tensor([int(k) for k in key] or [0], dtype=torch.int64).to(self.device)
            self.flat_nbt.add_(self._nbt_mask)
            loss, items = self.criterion(levels, targets)
            loss.backward()
        finally:
            release(packed)
            _DIRECT[0] = 0
            _NBT[0] = None
        self.loss, self.items = loss.detach(), items

    def _update(self):
        import torch.distributed as dist
        from .train import adamw_step, grad_sumsq
        if self.world > 1:
            dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM, group=self.group)
        self._t += 1
        clip = self.max_grad_norm > 0
        if clip:                                               # norm of the summed gradient; the kernel applies 1 / world to both
            grad_sumsq(self.flat_grad, self.grad_sumsq)
        nt = self.layout.trainable_total
        adamw_step(self.flat_param[:nt], self.flat_grad, self.exp_avg, self.exp_avg_sq, self._t, self.lr, self.betas, self.eps,
                   self.weight_decay, 1.0 / self.world, self.grad_sumsq if clip else None, self.max_grad_norm)

    def __call__(self, images, targets):
        if not self.use_graph:
            self._compute(images, targets)
            self._update()
            return self.loss, self.items
        key = (tuple(images.shape), images.dtype, tuple(targets.shape))
        if key != self._key:
            self._key, self.graph, self._warm = key, None, 0
            self.x_static, self.t_static = torch.empty_like(images), torch.empty_like(targets, dtype=torch.float32)
        self.x_static.copy_(images)
        self.t_static.copy_(targets)
        if self.graph is None and self._warm < 2:              # eager warm-up: plans, TMA descriptors, workspaces
            self._compute(self.x_static, self.t_static)
            self._warm += 1
        else:
            if self.graph is None:
                torch.cuda.synchronize(self.device)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._compute(self.x_static, self.t_static)
                self.graph = g
            self.graph.replay()
        self._update()
        return self.loss, self.items
