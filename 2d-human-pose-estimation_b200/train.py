"""Training step of the heatmap model on the sm_100a kernels (BASELINE configs[2]).

What the reference's trainer does per batch (/root/reference/HPE/engine/trainer.py:86-100):
``output = model(data)`` in ``model.train()`` mode (BatchNorm batch statistics), ``loss = criterion(...)``,
``loss.backward()``, ``optimizer.step()``; under DDP (pose/pose/core/builder.py:92-123) the gradients are
all-reduced.  Here:

* ``TrainPlan``  -- every layer of the net as forward conv (tcgen05) + BN-stats + BN-apply launches, and the
  matching backward: BN backward -> wgrad GEMM (tcgen05, MN-major) -> dgrad conv (tcgen05), maxpool backward,
  heatmap-gradient transposition.  Activations and activation gradients are bf16 NHWC, statistics / parameters /
  weight gradients fp32.  No PyTorch compute op is on this path; torch owns memory and streams.
* ``_TrainForward`` -- ``torch.autograd.Function`` so ``loss.backward()`` / ``torch.optim`` work as in the
  reference trainer (drop-in mode).
* ``TrainStep`` -- the fused data-parallel step: forward, fused Gaussian-target MSE loss + dL/dheatmap, backward
  into ONE flat fp32 gradient buffer, ONE NCCL all-reduce of that buffer, ONE SGD kernel over the
  flat parameter buffer.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.distributed as dist
import torch.nn as nn

from . import _lib
from .conv import (ACT_NONE, CONV_1x1, CONV_3x3, CONV_DECONV4, CONV_FINAL, CONV_STEM7x7, ConvLayer, DgradLayer, WgradLayer,
                   preprocess)

_LAYERS = (3, 4, 6, 3)
_PLANES = (64, 128, 256, 512)


# ---- thin wrappers over the C ABI -----------------------------------------------------------------------------
def bn_stats(z, bn: nn.BatchNorm2d, u):
    M, Cc = z.numel() // z.shape[-1], z.shape[-1]
    with torch.cuda.device(z.device):
        rc = _lib.lib().hpe_bn_stats(_lib.ptr(z), M, Cc, _lib.ptr(bn.weight), _lib.ptr(bn.bias), float(bn.eps),
                                     float(bn.momentum if bn.momentum is not None else 0.1),
                                     _lib.ptr(bn.running_mean), _lib.ptr(bn.running_var), _lib.ptr(bn_scratch(z.device)),
                                     _lib.ptr(u.scale),
                                     _lib.ptr(u.shift), _lib.ptr(u.mean), _lib.ptr(u.rstd), _lib.stream_ptr())
    _lib.check(rc, "hpe_bn_stats")


def bn_apply(z, scale, shift, residual, relu, y, bits=None):
    """``bits`` (uint8 [M*C/8]): also write the ReLU mask, one byte per 8 channels, for ``bn_backward(..., bits=)``."""
    M, Cc = z.numel() // z.shape[-1], z.shape[-1]
    with torch.cuda.device(z.device):
        rc = _lib.lib().hpe_bn_apply_bits(_lib.ptr(z), M, Cc, _lib.ptr(scale), _lib.ptr(shift), _lib.ptr(residual), int(relu),
                                          _lib.ptr(y), _lib.ptr(bits), _lib.stream_ptr())
    _lib.check(rc, "hpe_bn_apply")


_bn_scratch = {}


def bn_scratch(device):
    """One fp32 scratch per (device, stream) for the per-block partial sums of the BatchNorm reductions (stream-ordered
    reuse; a CUDA-graph capture stream gets its own, which the graph keeps using on replay)."""
    dev = torch.device(device)
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)     # per stream: two plans on different streams must not share it
    if key not in _bn_scratch:
        _bn_scratch[key] = torch.empty(_lib.lib().hpe_bn_scratch_elems(2048), dtype=torch.float32, device=device)
    return _bn_scratch[key]


def bn_backward(dy, y, z, mean, rstd, gamma, relu, dbeta, dgamma, dz, g_out=None, scale=None, shift=None, bits=None):
    """``y=None`` with ``relu`` and the forward's ``scale``/``shift``: the ReLU mask is recomputed from z (units without a
    residual), so y is not read.  ``bits``: the mask ``bn_apply`` wrote (units with a residual), read instead of y."""
    M, Cc = z.numel() // z.shape[-1], z.shape[-1]
    if bits is not None:
        with torch.cuda.device(z.device):
            rc = _lib.lib().hpe_bn_backward_bits(_lib.ptr(dy), _lib.ptr(bits), _lib.ptr(z), M, Cc, _lib.ptr(mean), _lib.ptr(rstd),
                                                 _lib.ptr(gamma), _lib.ptr(bn_scratch(z.device)), _lib.ptr(dbeta), _lib.ptr(dgamma),
                                                 _lib.ptr(dz), _lib.ptr(g_out), _lib.stream_ptr())
        _lib.check(rc, "hpe_bn_backward_bits")
        return
    with torch.cuda.device(z.device):
        rc = _lib.lib().hpe_bn_backward_zmask(_lib.ptr(dy), _lib.ptr(y), _lib.ptr(z), M, Cc, _lib.ptr(mean), _lib.ptr(rstd),
                                              _lib.ptr(gamma), int(relu), _lib.ptr(scale), _lib.ptr(shift),
                                              _lib.ptr(bn_scratch(z.device)), _lib.ptr(dbeta),
                                              _lib.ptr(dgamma), _lib.ptr(dz), _lib.ptr(g_out), _lib.stream_ptr())
    _lib.check(rc, "hpe_bn_backward")


def maxpool_train(x, y, idx):
    B, H, W, Cc = x.shape
    with torch.cuda.device(x.device):
        rc = _lib.lib().hpe_maxpool3x3s2_train(_lib.ptr(x), B, H, W, Cc, _lib.ptr(y), _lib.ptr(idx), _lib.stream_ptr())
    _lib.check(rc, "hpe_maxpool3x3s2_train")


def maxpool_backward(dy, idx, dx):
    B, H, W, Cc = dx.shape
    with torch.cuda.device(dx.device):
        rc = _lib.lib().hpe_maxpool3x3s2_backward(_lib.ptr(dy), _lib.ptr(idx), B, H, W, Cc, _lib.ptr(dx), _lib.stream_ptr())
    _lib.check(rc, "hpe_maxpool3x3s2_backward")


def heatmap_grad_nhwc(dhm, out, dbias):
    B, K, H, W = dhm.shape
    with torch.cuda.device(dhm.device):
        rc = _lib.lib().hpe_heatmap_grad_nhwc(_lib.ptr(dhm), _lib.dtype_code(dhm), B, K, H, W, _lib.ptr(out), _lib.ptr(dbias),
                                              _lib.stream_ptr())
    _lib.check(rc, "hpe_heatmap_grad_nhwc")


def sgd_step(param, grad, mom, lr, momentum, weight_decay, grad_scale=1.0):
    with torch.cuda.device(param.device):
        rc = _lib.lib().hpe_sgd_step(_lib.ptr(param), _lib.ptr(grad), _lib.ptr(mom), param.numel(), float(lr), float(momentum),
                                     float(weight_decay), float(grad_scale), _lib.stream_ptr())
    _lib.check(rc, "hpe_sgd_step")


_sumsq_scratch = {}


def grad_sumsq(grad, out, accumulate=False):
    """``out[0] (+)= sum(grad**2)`` over a flat fp32 buffer, deterministic (``clip_grad_norm_``'s norm without a host sync)."""
    key = (grad.device.index, torch.cuda.current_stream(grad.device).cuda_stream)
    if key not in _sumsq_scratch:
        _sumsq_scratch[key] = torch.empty(_lib.lib().hpe_grad_sumsq_scratch_elems(), dtype=torch.float32, device=grad.device)
    with torch.cuda.device(grad.device):
        rc = _lib.lib().hpe_grad_sumsq(_lib.ptr(grad), grad.numel(), _lib.ptr(_sumsq_scratch[key]), _lib.ptr(out), int(bool(accumulate)),
                                       _lib.stream_ptr())
    _lib.check(rc, "hpe_grad_sumsq")
    return out


def adamw_step(param, grad, exp_avg, exp_avg_sq, step, lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, grad_scale=1.0,
               sumsq=None, max_norm=0.0):
    """``torch.optim.AdamW`` step over a flat fp32 buffer with ``clip_grad_norm_(max_norm)`` folded in (HPE/train.py:52-55, 136);
    ``sumsq`` = device scalar from ``grad_sumsq`` (the squared global gradient norm)."""
    with torch.cuda.device(param.device):
        rc = _lib.lib().hpe_adamw_step(_lib.ptr(param), _lib.ptr(grad), _lib.ptr(exp_avg), _lib.ptr(exp_avg_sq), param.numel(),
                                       float(lr), float(betas[0]), float(betas[1]), float(eps), float(weight_decay), int(step),
                                       float(grad_scale), _lib.ptr(sumsq), float(max_norm), _lib.stream_ptr())
    _lib.check(rc, "hpe_adamw_step")


_BN_BITS = os.environ.get("HPE_B200_BN_BITS", "1") != "0"


# ---- one conv + BatchNorm (+ReLU) (+residual) unit in training mode ------------------------------------------------
class _Unit:
    def __init__(self, conv_name, bn_name, kind, B, H, W, Cin, Cout, stride, relu, dev, dgrad=True, dgrad_acc=False):
        self.conv_name, self.bn_name, self.relu = conv_name, bn_name, relu
        self.fwd = ConvLayer(kind, B, H, W, Cin, Cout, stride, ACT_NONE, False, dev)
        self.wg = WgradLayer(kind, B, H, W, Cin, Cout, stride, dev)
        self.dg = DgradLayer(kind, B, H, W, Cin, Cout, stride, accumulate=dgrad_acc, device=dev) if dgrad else None
        shape = self.fwd.out_shape()
        self.z = torch.empty(shape, dtype=torch.bfloat16, device=dev)
        self.y = torch.empty(shape, dtype=torch.bfloat16, device=dev)
        f = lambda n: torch.empty(n, dtype=torch.float32, device=dev)  # noqa: E731
        self.scale, self.shift, self.mean, self.rstd = f(Cout), f(Cout), f(Cout), f(Cout)
        self.bits = None         # ReLU mask of a residual unit (allocated on first use)
        self.x = None            # forward input (set by forward)
        self.bn = None           # nn.BatchNorm2d (set by load)
        self.in_shape = (B, H, W, Cin)

    def load(self, conv: nn.Module, bn: nn.BatchNorm2d):
        self.fwd.load(conv.weight)                 # scale = 1, shift = 0: the conv kernel writes the raw output z
        if self.dg is not None:
            self.dg.load(conv.weight)
        self.bn, self.conv_mod = bn, conv

    def forward(self, x, residual=None):
        self.x = x
        self.has_res = residual is not None
        self.fwd(x, None, self.z)
        bn_stats(self.z, self.bn, self)
        if self.has_res and self.relu and self.bits is None and _BN_BITS:   # residual units: keep the ReLU mask, 1 bit per element
            self.bits = torch.empty(self.z.numel() // 8, dtype=torch.uint8, device=self.z.device)
        bn_apply(self.z, self.scale, self.shift, residual, self.relu, self.y, self.bits if self.has_res else None)
        return self.y

    def backward(self, dy, grads, dz, g_out=None, dx_out=None, dx_res=None):
        """dy: gradient w.r.t. this unit's output y.  Writes the conv-weight / BN-affine gradients into ``grads``;
        returns dx (or None when the unit has no dgrad)."""
        dzv = dz.narrow(0, 0, self.z.numel()).view(self.z.shape)
        zmask = self.relu and not getattr(self, "has_res", True)      # y = relu(z*scale + shift): mask recomputed from z
        bits = self.bits if (self.relu and not zmask) else None       # residual unit: the mask bn_apply wrote
        bn_backward(dy, self.y if (self.relu and not zmask and bits is None) else None, self.z, self.mean, self.rstd,
                    self.bn.weight, self.relu, grads[self.bn_name + ".bias"], grads[self.bn_name + ".weight"], dzv, g_out,
                    self.scale if zmask else None, self.shift if zmask else None, bits)
        self.wg(self.x, dzv, grads[self.conv_name + ".weight"])
        if self.dg is None:
            return None
        return self.dg(dzv, dx_res, dx_out)


class TrainPlan:
    """All training-mode launches for one (B, H, W)."""

    def __init__(self, model, B, H, W, device):
        if H % 32 or W % 32:
            raise ValueError(f"input {H}x{W} must be a multiple of 32")
        self.B, self.H, self.W, self.device = B, H, W, device
        dev = device
        K = model.njoints
        self.K = K
        self.s2d = preprocess(torch.zeros(B, 3, H, W, device=dev))
        self.stem = _Unit("backbone.conv1", "backbone.bn1", CONV_STEM7x7, B, H, W, 3, 64, 2, True, dev, dgrad=False)
        h, w = H // 4, W // 4
        self.pool = torch.empty((B, h, w, 64), dtype=torch.bfloat16, device=dev)
        self.pool_idx = torch.empty((B, h, w, 64), dtype=torch.uint8, device=dev)
        self.blocks = []
        cin = 64
        for i, (n, p) in enumerate(zip(_LAYERS, _PLANES)):
            for j in range(n):
                s = 2 if (i > 0 and j == 0) else 1
                pre = f"backbone.layer{i + 1}.{j}"
                u1 = _Unit(pre + ".conv1", pre + ".bn1", CONV_1x1, B, h, w, cin, p, 1, True, dev, dgrad_acc=True)
                u2 = _Unit(pre + ".conv2", pre + ".bn2", CONV_3x3, B, h, w, p, p, s, True, dev)
                ds = None
                if j == 0:
                    ds = _Unit(pre + ".downsample.0", pre + ".downsample.1", CONV_1x1, B, h, w, cin, 4 * p, s, False, dev)
                h, w = h // s, w // s
                u3 = _Unit(pre + ".conv3", pre + ".bn3", CONV_1x1, B, h, w, p, 4 * p, 1, True, dev)
                self.blocks.append((u1, u2, u3, ds))
                cin = 4 * p
        self.deconvs = []
        for d, idx in enumerate((0, 3, 6)):
            self.deconvs.append(_Unit(f"deconv_layers.{idx}", f"deconv_layers.{idx + 1}", CONV_DECONV4, B, h, w, cin, 256, 2,
                                      True, dev))
            cin, h, w = 256, 2 * h, 2 * w
        self.hh, self.hw = h, w
        self.final = ConvLayer(CONV_FINAL, B, h, w, 256, K, 1, ACT_NONE, False, dev)
        self.final_wg = WgradLayer(CONV_FINAL, B, h, w, 256, K, 1, dev)
        self.final_dg = DgradLayer(CONV_FINAL, B, h, w, 256, K, 1, device=dev)
        self.heatmaps = torch.empty((B, K, h, w), dtype=torch.float32, device=dev)
        self.dhm_nhwc = torch.empty((B, h, w, 64), dtype=torch.bfloat16, device=dev)
        # gradient scratch (bf16): sized for the largest activation of the net
        units = self.units()
        big = max(max(u.z.numel() for u in units), max(int(torch.tensor(u.in_shape).prod()) for u in units))
        self._scratch = {k: torch.empty(big, dtype=torch.bfloat16, device=dev) for k in ("dz", "p", "q", "g", "ta", "tb")}
        self.weights_version = -1
        self.n_launches_fwd = 1 + 1 + len(units) * 4 + 1          # preprocess, maxpool, (conv, memset+reduce, finalize, apply), final
        self.n_launches_bwd = 2 + len(units) * 5 + 2 + 1          # approx: hm grad, final wgrad/dgrad, per unit (2 bn + 2 wgrad + dgrad)

    def units(self):
        out = [self.stem]
        for u1, u2, u3, ds in self.blocks:
            out += [u1, u2, u3] + ([ds] if ds is not None else [])
        return out + self.deconvs

    def _buf(self, key, shape):
        n = 1
        for s in shape:
            n *= s
        return self._scratch[key].narrow(0, 0, n).view(shape)

    @torch.no_grad()
    def load_weights(self, model):
        mods = dict(model.named_modules())
        for u in self.units():
            u.load(mods[u.conv_name], mods[u.bn_name])
        fl = mods["final_layer"]
        self._final_mod = fl
        self.final.scale.fill_(1.0)
        self.final.load(fl.weight, None, fl.bias)
        self.final_dg.load(fl.weight)
        self.weights_version = model._weights_version

    @torch.no_grad()
    def repack(self):
        """Re-pack the bf16 GEMM operands from the (updated) fp32 parameters: device-side kernels only, no host
        tensors, so it can live inside a captured CUDA graph.  Needs one prior load_weights()."""
        if os.environ.get("HPE_B200_BATCHED_PACK", "1") == "0":
            for u in self.units():
                u.load(u.conv_mod, u.bn)
            self.final.load(self._final_mod.weight, None, self._final_mod.bias)
            self.final_dg.load(self._final_mod.weight)
            return
        # ONE launch for every layer's forward + backward-data operand (job table built once; the parameters are updated in place)
        items = []
        for u in self.units():
            items.append((u.fwd.desc, 0, u.conv_mod.weight, u.fwd.w_packed))
            if u.dg is not None:
                items.append((u.dg.fwd_desc, 1, u.conv_mod.weight, u.dg.w_packed))
        items.append((self.final.desc, 0, self._final_mod.weight, self.final.w_packed))
        items.append((self.final_dg.fwd_desc, 1, self._final_mod.weight, self.final_dg.w_packed))
        key = tuple((w.data_ptr(), wp.data_ptr()) for _, _, w, wp in items)
        if getattr(self, "_pack_key", None) != key:
            lib = _lib.lib()
            jb = lib.hpe_pack_job_bytes()
            host = (C.c_char * (jb * len(items)))()
            begin = 0
            for k, (desc, dg, w, wp) in enumerate(items):
                if not (w.is_cuda and w.dtype == torch.float32 and w.is_contiguous()):
                    raise RuntimeError("libhpe_b200: parameters must be contiguous float32 CUDA tensors")
                nb = C.c_int()
                _lib.check(lib.hpe_conv_pack_job_fill(C.byref(desc), dg, w.data_ptr(), wp.data_ptr(), begin,
                                                      C.addressof(host) + k * jb, C.byref(nb)), "hpe_conv_pack_job_fill")
                begin += nb.value
            self._pack_jobs = torch.frombuffer(bytearray(host.raw), dtype=torch.uint8).to(self.device)
            self._pack_n, self._pack_blocks, self._pack_key = len(items), begin, key
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().hpe_conv_pack_jobs_run(_lib.ptr(self._pack_jobs), self._pack_n, self._pack_blocks,
                                                         _lib.stream_ptr()), "hpe_conv_pack_jobs_run")
        self.final.shift.copy_(self._final_mod.bias.detach())

    # ---- forward (training mode: batch statistics, running stats updated) --------------------------------------
    def forward(self, x):
        preprocess(x, self.s2d)
        y = self.stem.forward(self.s2d)
        maxpool_train(y, self.pool, self.pool_idx)
        cur = self.pool
        for u1, u2, u3, ds in self.blocks:
            t = u2.forward(u1.forward(cur))
            sc = ds.forward(cur) if ds is not None else cur
            cur = u3.forward(t, residual=sc)
        for u in self.deconvs:
            cur = u.forward(cur)
        self.final_in = cur
        self.final(cur, None, self.heatmaps)
        return self.heatmaps

    def export_acts(self):
        """NCHW fp32 CPU copies of every unit's stored input / raw conv output / output (parity tests, debugging)."""
        def nchw(t):
            return t.float().cpu().permute(0, 3, 1, 2).contiguous()
        acts = {}
        for u in self.units():
            acts[u.conv_name] = {"z": nchw(u.z), "y": nchw(u.y)}
            if u is not self.stem:
                acts[u.conv_name]["x"] = nchw(u.x)
        acts["final_layer"] = {"x": nchw(self.final_in)}
        return acts

    # ---- backward: dL/dheatmap -> every parameter gradient (views of one flat buffer) ----------------------------
    def backward(self, dhm, grads):
        dz = self._scratch["dz"]
        heatmap_grad_nhwc(dhm, self.dhm_nhwc, grads["final_layer.bias"])
        self.final_wg(self.final_in, self.dhm_nhwc, grads["final_layer.weight"])
        cur = self.final_dg(self.dhm_nhwc, None, self._buf("p", self.final_in.shape))
        pk, qk = "q", "p"                           # `cur` lives in "p"; next output goes to "q"
        for u in reversed(self.deconvs):
            cur = u.backward(cur, grads, dz, dx_out=self._buf(pk, u.in_shape))
            pk, qk = qk, pk
        for u1, u2, u3, ds in reversed(self.blocks):
            g3 = self._buf("g", u3.y.shape)
            dt2 = u3.backward(cur, grads, dz, g_out=g3, dx_out=self._buf("ta", u3.in_shape))
            dt1 = u2.backward(dt2, grads, dz, dx_out=self._buf("tb", u2.in_shape))
            if ds is not None:
                short = ds.backward(g3, grads, dz, dx_out=self._buf("ta", ds.in_shape))
            else:
                short = g3
            cur = u1.backward(dt1, grads, dz, dx_out=self._buf(pk, u1.in_shape), dx_res=short)
            pk, qk = qk, pk
        dstem = self._buf(pk, self.stem.y.shape)
        maxpool_backward(cur, self.pool_idx, dstem)
        self.stem.backward(dstem, grads, dz)


class ParamLayout:
    """Offsets of every parameter of the model inside one flat fp32 buffer (named_parameters order)."""

    def __init__(self, model):
        self.names, self.offsets, self.shapes = [], {}, {}
        off = 0
        params = list(model.named_parameters())
        self.trainable_total = 0
        # trainable parameters first: the optimizer kernel and the gradient all-reduce run over that prefix only, so
        # `requires_grad=False` parameters (a frozen backbone) are neither updated nor weight-decayed, like torch.optim.SGD
        for want in (True, False):
            for n, p in params:
                if p.requires_grad != want:
                    continue
                self.names.append(n)
                self.offsets[n] = off
                self.shapes[n] = tuple(p.shape)
                off += (p.numel() + 3) // 4 * 4            # keep every slice 16-byte aligned
            if want:
                self.trainable_total = off
        self.total = off

    def views(self, flat):
        out = {}
        for n in self.names:
            k = 1
            for s in self.shapes[n]:
                k *= s
            out[n] = flat.narrow(0, self.offsets[n], k).view(self.shapes[n])
        return out


class _TrainForward(torch.autograd.Function):
    """``heatmaps = model(x)`` in training mode with a backward that fills every parameter's gradient."""

    @staticmethod
    def forward(ctx, model, x, *params):
        plan = model._train_plan(x)
        hm = plan.forward(x.contiguous() if x.dtype == torch.uint8 else x.float().contiguous())   # uint8 = x.float()/255 fused in
        nbt = [m.num_batches_tracked for m in model.modules()
               if isinstance(m, nn.BatchNorm2d) and m.num_batches_tracked is not None]
        if nbt:
            torch._foreach_add_(nbt, 1)              # what every torch BatchNorm does in a training forward
        ctx.model, ctx.plan = model, plan
        return hm.clone()

    @staticmethod
    def backward(ctx, dhm):
        model, plan = ctx.model, ctx.plan
        layout = model._param_layout()
        flat = torch.zeros(layout.total, dtype=torch.float32, device=dhm.device)
        grads = layout.views(flat)
        plan.backward(dhm.contiguous(), grads)
        return (None, None) + tuple(grads[n] if p.requires_grad else None for n, p in model.named_parameters())


class TrainStep:
    """Fused data-parallel training step: ``loss = step(images, joints)``.

    images ``[B,3,H,W]`` fp32 (device), joints ``[B,K,3]`` (x, y in input pixels, visibility).  Gradients of the local
    shard are summed over ranks with one NCCL all-reduce of the flat buffer and divided by the world size in the
    SGD kernel (equal shards => the global-batch mean)."""

    def __init__(self, model, lr=1e-3, momentum=0.9, weight_decay=0.0, sigma=2.0, group=None, cuda_graph=True,
                 optimizer="sgd", betas=(0.9, 0.999), eps=1e-8, max_grad_norm=0.0):
        """``optimizer``: "sgd" (config default, HPE/configs/default.py:34: momentum SGD) or "adamw" (the reference's working
        trainer, HPE/train.py:52-55) with ``clip_grad_norm_(max_grad_norm)`` (HPE/train.py:136; 0 = no clipping)."""
        if optimizer not in ("sgd", "adamw"):
            raise ValueError(f"optimizer '{optimizer}' not built (sgd | adamw)")
        self.optimizer, self.betas, self.eps, self.max_grad_norm = optimizer, tuple(betas), eps, float(max_grad_norm)
        self._t = 0
        self.model = model.train()
        self.use_graph = bool(cuda_graph)
        self._key, self.graph, self.plan = None, None, None
        self.lr, self.momentum, self.weight_decay, self.sigma = lr, momentum, weight_decay, sigma
        self.group = group
        self.layout = ParamLayout(model)
        dev = next(model.parameters()).device
        self.device = dev
        self.flat_param = torch.zeros(self.layout.total, dtype=torch.float32, device=dev)
        self.flat_grad = torch.zeros_like(self.flat_param)
        self.flat_mom = torch.zeros_like(self.flat_param)              # SGD momentum / AdamW exp_avg
        self.flat_sq = torch.zeros_like(self.flat_param) if optimizer == "adamw" else None      # AdamW exp_avg_sq
        self.grad_sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        pv = self.layout.views(self.flat_param)
        with torch.no_grad():
            for n, p in model.named_parameters():       # re-point the parameters at the flat buffer
                pv[n].copy_(p.data)
                p.data = pv[n]
        self.grads = self.layout.views(self.flat_grad)
        # BatchNorm `num_batches_tracked` counters re-pointed at one flat int64 buffer: ONE add per step keeps saved
        # state_dicts identical to the reference's (torch BatchNorm increments it in every training forward)
        bns = [m for m in model.modules() if isinstance(m, nn.BatchNorm2d) and m.num_batches_tracked is not None]
        self.flat_nbt = torch.zeros(max(len(bns), 1), dtype=torch.int64, device=dev)
        with torch.no_grad():
            for i, m in enumerate(bns):
                self.flat_nbt[i] = m.num_batches_tracked.to(dev)
                m.num_batches_tracked.data = self.flat_nbt[i]
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.partials = None
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        self.dhm = None

    def broadcast_parameters(self, src=0):
        if self.world > 1:
            dist.broadcast(self.flat_param, src, group=self.group)
            for b in self.model.buffers():
                if b.dtype.is_floating_point:
                    dist.broadcast(b, src, group=self.group)

    def _compute(self, images, joints):
        """repack -> forward -> fused loss + dL/dheatmap -> backward into the flat gradient buffer (this rank's shard)."""
        plan = self.plan
        plan.repack()                                 # parameters changed last step: re-pack bf16 GEMM operands
        hm = plan.forward(images)
        B, K, H, W = hm.shape
        sx, sy = images.shape[3] / W, images.shape[2] / H
        with torch.cuda.device(hm.device):
            rc = _lib.lib().hpe_joints_mse_fused(_lib.ptr(hm), _lib.HPE_F32, _lib.ptr(joints), B, K, H, W, float(self.sigma),
                                                 float(sx), float(sy), _lib.ptr(self.partials), _lib.ptr(self.loss),
                                                 _lib.ptr(self.dhm), 1.0, 0, _lib.stream_ptr())
        _lib.check(rc, "hpe_joints_mse_fused")
        plan.backward(self.dhm, self.grads)

    def _update(self):
        """ONE NCCL all-reduce of the trainable prefix of the flat gradient buffer, then ONE SGD kernel over it."""
        nt = self.layout.trainable_total
        if self.world > 1:
            dist.all_reduce(self.flat_grad[:nt], op=dist.ReduceOp.SUM, group=self.group)
        if self.optimizer == "adamw":
            self._t += 1
            clip = self.max_grad_norm > 0
            if clip:                                  # global norm of the (summed) gradient; the kernel scales it by 1/world
                grad_sumsq(self.flat_grad[:nt], self.grad_sumsq)
            adamw_step(self.flat_param[:nt], self.flat_grad[:nt], self.flat_mom[:nt], self.flat_sq[:nt], self._t, self.lr, self.betas,
                       self.eps, self.weight_decay, 1.0 / self.world, self.grad_sumsq if clip else None, self.max_grad_norm)
        else:
            sgd_step(self.flat_param[:nt], self.flat_grad[:nt], self.flat_mom[:nt], self.lr, self.momentum, self.weight_decay,
                     1.0 / self.world)
        self.flat_nbt.add_(1)

    def _step(self, images, joints):
        self._compute(images, joints)
        self._update()

    def __call__(self, images, joints):
        model = self.model
        if images.dtype not in (torch.float32, torch.uint8):
            raise TypeError(f"TrainStep takes float32 or uint8 images, got {images.dtype}")
        key = (tuple(images.shape), tuple(joints.shape), images.dtype)
        if self._key != key:
            self._key, self.graph = key, None
            self.plan = model._train_plan(images)
            self.plan.load_weights(model)
            hm = self.plan.heatmaps
            self.dhm = torch.empty_like(hm)
            self.partials = torch.empty(hm.shape[0] * hm.shape[1], dtype=torch.float32, device=hm.device)
            # uint8 images stay uint8: `imgs.float() / 255` (HPE/train.py:117-120) is fused into the stem's layout kernel,
            # exactly as in the eval forward and PoseEstimator
            self.x_static = torch.empty_like(images)
            self.j_static = torch.empty_like(joints, dtype=torch.float32)
            self._warm = 0
        self.x_static.copy_(images)
        self.j_static.copy_(joints)
        if not self.use_graph:
            self._step(self.x_static, self.j_static)
        elif self.graph is None and self._warm < 1:
            self._step(self.x_static, self.j_static)          # eager warm-up: encodes the TMA descriptors, NCCL init
            self._warm += 1
        else:
            # The CUDA graph holds the ~430 compute launches of the step.  The NCCL all-reduce stays OUTSIDE it (one eager
            # collective + the optimizer kernel per step, ~2 launches): a captured collective keeps the communicator alive
            # inside the graph, which blocked `destroy_process_group()` at exit in round 1.
            if self.graph is None:
                torch.cuda.synchronize(self.device)
                g = torch.cuda.CUDAGraph()
                self._update_outside = self.world > 1 or self.optimizer == "adamw"    # AdamW's bias correction changes per step
                with torch.cuda.graph(g):
                    if self._update_outside:
                        self._compute(self.x_static, self.j_static)
                    else:
                        self._step(self.x_static, self.j_static)
                self.graph = g
            self.graph.replay()
            if self._update_outside:
                self._update()
        model._weights_version += 1
        return self.loss

    def close(self):
        """Drop the captured graph and the plan (call before tearing the process group down)."""
        self.graph = None
        self._key = None
