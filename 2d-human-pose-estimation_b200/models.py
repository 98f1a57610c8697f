"""Model registry + the top-down heatmap model behind the reference's constructor signature.

Boundary being mirrored:
  * ``models.__all__`` name table + ``getattr(models, cfg.model.name)(nkpts, backbone, pretrained)``
    (/root/reference/HPE/models/__init__.py:4-12, HPE/core/builder.py:15-19);
  * ``DeepPose(num_joints, backbone='resnet50', pretrained=True)``
    (HPE/models/reference/bottom_up/regression/models/DeepPose.py:4-10);
  * ``output = model(data)`` with ``data`` fp32 NCHW on the device (HPE/engine/trainer.py:87-90).

The parameter tree uses the timm key names (``backbone.conv1.weight``, ``backbone.layer1.0.bn1.*``,
``backbone.layer2.0.downsample.{0,1}.*`` ...) so reference checkpoints load with ``load_state_dict``;
the ``fc`` regression head of DeepPose is replaced by the App. A.1 deconv heatmap head
(``deconv_layers.{0,3,6}`` / ``.{1,4,7}`` / ``final_layer``).  The ``torch.nn`` modules below are
parameter containers only: the forward pass never calls them -- it runs the sm_100a kernels of
libhpe_b200.so through ``ConvLayer`` plans and raises if that library is unavailable.
"""
from __future__ import annotations

import os

import torch
import torch.nn as nn

from . import _lib
from .conv import (ACT_NONE, ACT_RELU, CONV_1x1, CONV_3x3, CONV_DECONV4, CONV_FINAL, CONV_STEM7x7, ConvLayer, fold_bn,
                   maxpool3x3s2, preprocess)

__all__ = ("DeepPose", "HeatmapPose")      # names Builder.model() may construct positionally (nkpts, backbone, pretrained)

_LAYERS = (3, 4, 6, 3)
_PLANES = (64, 128, 256, 512)


class _BottleneckParams(nn.Module):
    def __init__(self, inplanes, planes, stride, downsample):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride=stride, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, planes * 4, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(planes * 4)
        self.downsample = downsample
        self.stride = stride


class _ResNet50Params(nn.Module):
    """Parameter container with timm ResNet-50 names/shapes/init (resnet.py:554-621, 1548-1552)."""

    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        inplanes = 64
        for i, (n, p) in enumerate(zip(_LAYERS, _PLANES)):
            blocks = []
            for j in range(n):
                s = 2 if (i > 0 and j == 0) else 1
                ds = None
                if j == 0:
                    ds = nn.Sequential(nn.Conv2d(inplanes, p * 4, 1, stride=s, bias=False), nn.BatchNorm2d(p * 4))
                blocks.append(_BottleneckParams(inplanes, p, s, ds))
                inplanes = p * 4
            setattr(self, f"layer{i + 1}", nn.Sequential(*blocks))
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
        for m in self.modules():
            if isinstance(m, _BottleneckParams):
                nn.init.zeros_(m.bn3.weight)           # timm zero_init_last


class HeatmapPose(nn.Module):
    """``Model(nkpts, backbone, pretrained)`` -> heatmaps ``[B, K, H/4, W/4]`` (fp32) for fp32 NCHW input."""

    def __init__(self, num_joints=17, backbone="resnet50", pretrained=True):
        super().__init__()
        self.njoints = int(num_joints)
        self.backbone_name = backbone
        if backbone == "resnet50":
            self.backbone = _ResNet50Params()
            c = 2048
        elif backbone.startswith("dinov3_convnext_") and backbone.split("_")[2] in ("tiny", "small", "base", "large"):
            # the reference's ConvNeXt/DINOv3 backbone (HPE/models/backbones/dinov3convnext.py:76-110,
            # dinov3/models/convnext.py:121-180) as an alternate backbone behind the SAME heatmap head; parameter names are
            # the dinov3 checkpoint's (`backbone.stages.2.5.pwconv1.weight` ...)
            from .posemodel import _ConvNeXt, convnext_sizes
            cfg = convnext_sizes[backbone.split("_")[2]]
            self.dims, self.depths = cfg["dims"], cfg["depths"]
            self.backbone = _ConvNeXt(self.depths, self.dims)
            c = self.dims[-1]
        else:
            raise ValueError(f"backbone '{backbone}' not built: this path covers 'resnet50' (config default) and "
                             "'dinov3_convnext_{tiny,small,base,large}'")
        layers = []
        for _ in range(3):
            layers += [nn.ConvTranspose2d(c, 256, 4, stride=2, padding=1, output_padding=0, bias=False),
                       nn.BatchNorm2d(256), nn.ReLU(inplace=True)]
            c = 256
        self.deconv_layers = nn.Sequential(*layers)
        self.final_layer = nn.Conv2d(256, self.njoints, 1, bias=True)
        for m in self.deconv_layers:
            if isinstance(m, nn.ConvTranspose2d):
                nn.init.normal_(m.weight, std=0.001)
        nn.init.normal_(self.final_layer.weight, std=0.001)
        nn.init.zeros_(self.final_layer.bias)
        self._plans = {}
        self._layout = None
        self._weights_version = 0
        if pretrained:
            path = pretrained if isinstance(pretrained, str) else os.environ.get("HPE_B200_PRETRAINED", "")
            if not path:
                raise RuntimeError("pretrained=True needs a local timm-format ResNet-50 checkpoint: pass its path as "
                                   "`pretrained` or set HPE_B200_PRETRAINED (there is no network download path)")
            self.load_backbone(path)

    # ---- checkpoints -------------------------------------------------------------------------------
    def load_backbone(self, path):
        sd = torch.load(path, map_location="cpu")
        sd = sd.get("model_state", sd.get("state_dict", sd))
        sd = {k[len("backbone."):] if k.startswith("backbone.") else k: v for k, v in sd.items()}
        sd = {k: v for k, v in sd.items() if not k.startswith("fc.")}
        self.backbone.load_state_dict(sd, strict=True)
        self._weights_version += 1

    def load_state_dict(self, state_dict, strict=True, **kw):
        """Accepts reference DeepPose checkpoints: ``backbone.fc.*`` (the regression head) is ignored."""
        sd = {k: v for k, v in state_dict.items() if not k.startswith("backbone.fc.")}
        has_head = any(k.startswith("deconv_layers.") for k in sd)
        out = super().load_state_dict(sd, strict=strict and has_head, **kw)
        self._weights_version += 1
        return out

    def refresh(self):
        """Re-pack weights / re-fold BN after parameters were modified in place."""
        self._weights_version += 1

    def _apply(self, fn, *a, **kw):
        # The reference trainer calls ``model.to(device)`` on EVERY batch (HPE/engine/trainer.py:87): a no-op move must keep the
        # plans (TMA maps, activation buffers, captured graph).  Plans are dropped only when a parameter really changed device,
        # dtype or storage; the packed weights are refreshed either way (cheap, and ``fn`` may have changed values).
        p0 = next(self.parameters(), None)
        before = None if p0 is None else (p0.device, p0.dtype, p0.data_ptr())
        out = super()._apply(fn, *a, **kw)
        p1 = next(self.parameters(), None)
        after = None if p1 is None else (p1.device, p1.dtype, p1.data_ptr())
        if before != after:
            self._plans = {}
            self._layout = None
            self._weights_version += 1
        return out

    def train(self, mode: bool = True):
        # parameters / running statistics may have been updated in place by an optimizer or the training kernels while in
        # training mode: every eval plan re-packs its weights and re-folds BN on the next forward (stale eval plan)
        if mode != self.training:
            self._weights_version += 1
        return super().train(mode)

    # ---- forward -----------------------------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.dim() != 4 or x.shape[1] != 3:
            raise ValueError(f"expected [B,3,H,W] input, got {tuple(x.shape)}")
        if not x.is_cuda:
            raise RuntimeError("libhpe_b200 has no CPU path: input must live on a CUDA (sm_100a) device")
        if self.training and self.backbone_name != "resnet50":
            raise RuntimeError("training mode is built for the resnet50 backbone only (ConvNeXt backbones: inference)")
        if self.training:
            # model.train() (HPE/engine/trainer.py:77-80): BatchNorm batch statistics, and a backward pass through
            # the dgrad / wgrad kernels so `loss.backward()` fills every parameter's .grad (trainer.py:96-97)
            from .train import _TrainForward
            plan = self._train_plan(x)
            plan.load_weights(self)               # the optimizer updates parameters in place between steps
            self._weights_version += 1            # ... and this step's backward/optimizer will: eval plans are stale from here on
            return _TrainForward.apply(self, x, *self.parameters())
        # a fresh tensor, like the reference's `output = model(data)` (HPE/engine/trainer.py:90): the plan's persistent
        # heatmap buffer is overwritten by the next call (CUDA-graph replay), so flip-test averaging / collecting outputs
        # in a list must not alias it.  53 MB at B=256 = ~16 us; PoseEstimator and the benchmark use forward_static.
        return self.forward_static(x).clone()

    def forward_static(self, x: torch.Tensor) -> torch.Tensor:
        """Eval forward returning the plan's PERSISTENT heatmap buffer (zero-copy): valid until the next call with
        the same input shape.  The public ``forward`` returns a clone of it."""
        if x.dim() != 4 or x.shape[1] != 3:
            raise ValueError(f"expected [B,3,H,W] input, got {tuple(x.shape)}")
        if not x.is_cuda:
            raise RuntimeError("libhpe_b200 has no CPU path: input must live on a CUDA (sm_100a) device")
        if self.training:
            raise RuntimeError("forward_static is the inference path: call model.eval() first")
        key = (tuple(x.shape), x.device.index)
        plan = self._plans.get(key)
        xin = x.contiguous() if x.dtype == torch.uint8 else x.float().contiguous()      # uint8 = x.float()/255 fused in
        if self.backbone_name != "resnet50":
            # ConvNeXt backbone: the launch plan of posemodel.py with the deconv heatmap head; its weights are packed when the
            # plan is built, so a weight change rebuilds it
            if plan is None or plan.weights_version != self._weights_version:
                from .posemodel import _PosePlan
                with torch.no_grad():
                    plan = _PosePlan(self, x.shape[0], x.shape[2], x.shape[3], x.device, heatmap_head=True)
                plan.weights_version = self._weights_version
                self._plans[key] = plan
            plan.run(xin)
            return plan.heatmaps
        if plan is None:
            plan = _ForwardPlan(self, x.shape[0], x.shape[2], x.shape[3], x.device)
            self._plans[key] = plan
        if plan.weights_version != self._weights_version:
            plan.load_weights(self)
        return plan.run(xin)

    def _train_plan(self, x):
        from .train import TrainPlan
        key = ("train", tuple(x.shape), x.device.index)
        plan = self._plans.get(key)
        if plan is None:
            plan = TrainPlan(self, x.shape[0], x.shape[2], x.shape[3], x.device)
            self._plans[key] = plan
        return plan

    def _param_layout(self):
        from .train import ParamLayout
        if getattr(self, "_layout", None) is None:
            self._layout = ParamLayout(self)
        return self._layout

    def forward_features(self, x):
        """``[B, 2048, H/32, W/32]`` fp32 NCHW view of the trunk output (timm ``forward_features`` analogue)."""
        self.forward_static(x)
        plan = self._plans[(tuple(x.shape), x.device.index)]
        feat = plan.features if self.backbone_name == "resnet50" else plan.features[-1]
        return feat.float().permute(0, 3, 1, 2)              # .float() makes a fresh tensor


from .posemodel import DINOv3Pose  # noqa: E402  (the reference's real pose model, HPE/models/pose.py)

DeepPose = HeatmapPose     # config-default name (HPE/configs/default.py:13): same constructor, heatmap output


class _ForwardPlan:
    """All ConvLayer plans + activation buffers for one (B, H, W); optionally replayed as a CUDA graph."""

    def __init__(self, model: HeatmapPose, B, H, W, device):
        if H % 32 or W % 32:
            raise ValueError(f"input {H}x{W} must be a multiple of 32")
        self.B, self.H, self.W, self.device = B, H, W, device
        self.ops = []            # (layer, src_key, res_key, dst_key, src2_key, dst2_key), one per conv launch, in launch order
        # DEPTH-FIRST SECTION (experiment, OFF by default: HPE_B200_DF_CHUNK images per pass, HPE_B200_DF_STAGES ResNet stages).
        # Idea: the stem and the first stages move 403 MB tensors at B=256, but a chunk of 32 images keeps a bottleneck's working
        # set inside the 126 MB L2, so the section's layers are planned for `chunk` images with chunk-sized buffers that every pass
        # reuses; only the section's input (s2d image) and output are full-batch tensors, addressed per pass as batch slices.
        # Measured on B200 (profiles/r02_experiments.md): SLOWER at every setting (3.21 ms -> 3.40 ms at 64 x 2 stages, 3.65 ms at
        # 32 x 2): a conv launch has ~10-14 us of fixed cost (ramp, last-tile epilogue, wave tail), which the 66-150 extra launches
        # pay many times over.  Per-image results do not depend on the chunking (same kernels, same bricks within an image).
        chunk = int(os.environ.get("HPE_B200_DF_CHUNK", "0"))
        self.df_stages = min(3, int(os.environ.get("HPE_B200_DF_STAGES", "2"))) if (chunk > 0 and B > chunk and B % chunk == 0) else 0
        self.chunk = chunk if self.df_stages > 0 else B
        self.n_chunks = B // self.chunk
        self.n_section_ops = 0   # the first n_section_ops entries of self.ops run once per chunk
        self.full_keys = set()   # full-batch buffers the section addresses by batch slice
        self.fused_shortcut = {}  # conv3 path -> downsample path folded into it
        # projection shortcut (downsample conv + BN) accumulated into conv3's TMEM tile: the shortcut tensor never exists
        fuse_sc = os.environ.get("HPE_B200_FUSE_SHORTCUT", "1") != "0"
        # identity shortcut as extra k-blocks against an identity weight block (experiment, HPE_B200_RES_MMA=1, off by default): the
        # residual add moves from the epilogue to the tensor pipe.  Measured (profiles/r02_experiments.md): layer1 conv3 -13 us
        # each (its residual then streams through the 5-deep operand ring), every deeper layer slower (full-N identity k-blocks
        # cost 4 x 512 MMA cycles and 64 KB of weight traffic per tile): forward 3.34 -> 3.77 ms.
        res_mma = os.environ.get("HPE_B200_RES_MMA", "0") != "0"
        self.identity_shortcut = set()
        # conv3 (+ shortcut + ReLU) of a layer1 Bottleneck and conv1 (+ BN + ReLU) of the NEXT block in one kernel (desc.chain): the
        # 256-channel block output is read once (as the next shortcut) instead of twice, and one launch per block disappears
        fuse_chain = os.environ.get("HPE_B200_FUSE_CHAIN", "1") != "0"
        fuse_chain2 = int(os.environ.get("HPE_B200_FUSE_CHAIN2", "1"))         # ... and the 512 -> 128 chains inside layer2 (2: also behind its first block)
        self.chained = {}        # conv1 path -> the conv3 path whose kernel computes it
        pending_chain = None     # (conv3 name, t1 buffer key) while the next block's conv1 is still to be declared
        self.bufs = {}
        self.named = {}          # module path -> ConvLayer (for load_weights)
        K = model.njoints
        dev = device

        def buf(key, shape, dtype=torch.bfloat16):
            self.bufs[key] = torch.empty(shape, dtype=dtype, device=dev)
            return key

        def add(name, layer, src, dst, res=None, src2=None, full_out=False, dst2=None):
            self.named[name] = layer
            shape = layer.out_shape()
            if full_out:                    # last layer of the depth-first section: writes its batch slice of a full tensor
                shape = (Bf,) + tuple(shape[1:])
                self.full_keys.add(dst)
            buf(dst, shape, torch.float32 if (layer.kind == CONV_FINAL or layer.head2) else torch.bfloat16)
            if dst2:
                buf(dst2, (layer.B, layer.Ho, layer.Wo, layer.chain))
            self.ops.append((layer, src, res, dst, src2, dst2))

        self.s2d = preprocess(torch.zeros(B, 3, H, W, device=dev))
        Bf = B                              # full batch
        if self.df_stages > 0:
            B = self.chunk                  # the section's layers are planned for one chunk
            self.full_keys.add("s2d")
        # `x = self.maxpool(x)` (timm ResNet.forward_features) taken in the stem's epilogue: the H/2 x W/2 tensor never exists
        self.fused_pool = False
        h, w = H // 4, W // 4
        if os.environ.get("HPE_B200_FUSE_POOL", "1") != "0":
            try:
                add("backbone.conv1", ConvLayer(CONV_STEM7x7, B, H, W, 3, 64, 2, ACT_RELU, False, dev, pool=1), "s2d", "pool")
                self.fused_pool = True
            except Exception:      # e.g. HPE_B200_HALO=0: the pooled stem exists on the y-halo kernel only
                pass
        if not self.fused_pool:
            add("backbone.conv1", ConvLayer(CONV_STEM7x7, B, H, W, 3, 64, 2, ACT_RELU, False, dev), "s2d", "stem")
            buf("pool", (B, h, w, 64))
        cur, cin = "pool", 64
        for i, (n, p) in enumerate(zip(_LAYERS, _PLANES)):
            if self.df_stages > 0 and i == self.df_stages:
                self.n_section_ops = len(self.ops)
                B = Bf
            for j in range(n):
                last_of_section = self.df_stages > 0 and i == self.df_stages - 1 and j == n - 1
                s = 2 if (i > 0 and j == 0) else 1
                pre = f"backbone.layer{i + 1}.{j}"
                if pending_chain is not None:         # conv1 runs inside the previous block's conv3 kernel, which writes .t1
                    self.chained[pre + ".conv1"] = pending_chain
                    pending_chain = None
                else:
                    add(pre + ".conv1", ConvLayer(CONV_1x1, B, h, w, cin, p, 1, ACT_RELU, False, dev), cur, pre + ".t1")
                add(pre + ".conv2", ConvLayer(CONV_3x3, B, h, w, p, p, s, ACT_RELU, False, dev), pre + ".t1", pre + ".t2")
                sc = cur
                # chain the next block's conv1 (256 -> 64 | 128) behind this conv3 when the block output has 256 channels (layer1)
                nxt = (f"backbone.layer{i + 1}.{j + 1}", p) if j + 1 < n else ((f"backbone.layer{i + 2}.0", _PLANES[i + 1]) if i + 1 < 4 else None)
                # (layer2: identity blocks only -- the first block's two-source conv3 runs 6 k-blocks per 256-wide half through a
                #  2-stage ring and a single accumulator: measured 165 us chained against 81 + 45 us as two launches)
                ok = nxt and ((4 * p == 256 and nxt[1] in (64, 128)) or (4 * p == 512 and nxt[1] == 128 and fuse_chain2 and (j > 0 or fuse_chain2 == 2)))
                n2 = nxt[1] if (fuse_chain and ok and not res_mma and self.df_stages == 0) else 0
                d2 = nxt[0] + ".t1" if n2 else None
                if j == 0 and not fuse_sc:
                    add(pre + ".downsample.0", ConvLayer(CONV_1x1, B, h, w, cin, 4 * p, s, ACT_NONE, False, dev), cur, pre + ".sc")
                    sc = pre + ".sc"
                h, w = h // s, w // s
                if j == 0 and fuse_sc:
                    self.fused_shortcut[pre + ".conv3"] = pre + ".downsample"
                    add(pre + ".conv3", ConvLayer(CONV_1x1, B, h, w, p, 4 * p, 1, ACT_RELU, False, dev, Cin2=cin, stride2=s, chain=n2),
                        pre + ".t2", pre + ".out", None, cur, full_out=last_of_section, dst2=d2)
                elif res_mma:
                    self.identity_shortcut.add(pre + ".conv3")
                    add(pre + ".conv3", ConvLayer(CONV_1x1, B, h, w, p, 4 * p, 1, ACT_RELU, False, dev, Cin2=4 * p, stride2=1),
                        pre + ".t2", pre + ".out", None, sc, full_out=last_of_section)
                else:
                    add(pre + ".conv3", ConvLayer(CONV_1x1, B, h, w, p, 4 * p, 1, ACT_RELU, True, dev, chain=n2), pre + ".t2", pre + ".out", sc,
                        full_out=last_of_section, dst2=d2)
                if n2:
                    pending_chain = pre + ".conv3"
                cur, cin = pre + ".out", 4 * p
        self.features_key = cur
        # last deconv stage + final_layer in one kernel (the 256-channel deconv3 output never leaves shared memory) when the
        # joints fit one 32-column second GEMM and the layer runs on the CTA-pair kernel
        self.fused_head = None
        for d, idx in enumerate((0, 3, 6)):
            layer = None
            if d == 2 and K <= 32 and os.environ.get("HPE_B200_FUSE_HEAD", "1") != "0":
                try:
                    layer = ConvLayer(CONV_DECONV4, B, h, w, cin, 256, 2, ACT_RELU, False, dev, head2=K)
                    self.fused_head = f"deconv_layers.{idx}"
                except RuntimeError:
                    layer = None          # too few tiles for a CTA pair (tiny batch): separate layers
            if layer is not None:
                add(f"deconv_layers.{idx}", layer, cur, "heatmaps")
                h, w = 2 * h, 2 * w
                break
            add(f"deconv_layers.{idx}", ConvLayer(CONV_DECONV4, B, h, w, cin, 256, 2, ACT_RELU, False, dev), cur, f"deconv{d}")
            cur, cin, h, w = f"deconv{d}", 256, 2 * h, 2 * w
        if self.fused_head is None:
            add("final_layer", ConvLayer(CONV_FINAL, B, h, w, 256, K, 1, ACT_NONE, False, dev), cur, "heatmaps")
        self._schedule(int(os.environ.get("HPE_B200_L2_SCHED", "2")))
        self.weights_version = -1
        self.graph = None
        self.bufs["s2d"] = self.s2d.view(Bf, -1)
        self.B = Bf
        extra = 0 if self.fused_pool else 1                                       # separate max-pool launch
        self.n_launches = 1 + len(self.ops) + extra + (self.n_chunks - 1) * (self.n_section_ops + extra)      # 1 = preprocess

    @property
    def features(self):
        return self.bufs[self.features_key]

    @property
    def heatmaps(self):
        return self.bufs["heatmaps"]

    @torch.no_grad()
    def load_weights(self, model: HeatmapPose):
        mods = dict(model.named_modules())
        for name, layer in self.named.items():
            m = mods[name]
            if name == "final_layer":
                layer.load(m.weight, torch.ones(m.out_channels), m.bias)
                continue
            if name.startswith("deconv_layers."):
                bn = mods[f"deconv_layers.{int(name.split('.')[1]) + 1}"]
            elif name.endswith("downsample.0"):
                bn = mods[name[:-1] + "1"]
            else:
                bn = mods[name.replace("conv", "bn")]
            scale, shift = fold_bn(bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps)
            if layer.chain:
                c1 = next(k for k, v in self.chained.items() if v == name)
                bn1 = mods[c1.replace("conv", "bn")]
                layer.load_chain(mods[c1].weight, *fold_bn(bn1.weight, bn1.bias, bn1.running_mean, bn1.running_var, bn1.eps))
            if name == self.fused_head:
                layer.load_head2(mods["final_layer"].weight, mods["final_layer"].bias)
            if name in self.identity_shortcut:
                c = m.out_channels
                layer.load2(m.weight, scale, shift, torch.eye(c).view(c, c, 1, 1), torch.ones(c), torch.zeros(c))
                continue
            if name in self.fused_shortcut:
                ds = self.fused_shortcut[name]
                bn2 = mods[ds + ".1"]
                scale2, shift2 = fold_bn(bn2.weight, bn2.bias, bn2.running_mean, bn2.running_var, bn2.eps)
                layer.load2(m.weight, scale, shift, mods[ds + ".0"].weight, scale2, shift2)
                continue
            layer.load(m.weight, scale, shift)
        self.weights_version = model._weights_version
        self.graph = None         # packed buffers are updated in place, but be conservative

    def _schedule(self, mode):
        """Tile order and L2 eviction hints per launch (hpe_conv_set_schedule; results are bit-identical).  Every launch walks its
        tiles batch-major, so by default a layer starts on the part of its input that was written FIRST -- for tensors around
        the size of the 126 MB L2 (layer3: 100 MB) exactly the part already evicted.  mode & 2: a launch runs in the direction
        opposite to the one in which its largest input was last touched (it starts where that pass ended).  mode & 1: outputs
        up to HPE_B200_L2_KEEP_MB (104) are stored evict-last, the shortcut (read for the last time) is loaded evict-first.
        mode & 4: an input that no later launch reads is loaded evict-first.  HPE_B200_L2_SCHED, default 2.  Measured forward,
        five runs on two boxes: mode 0 3.151 ms, mode 2 3.107, mode 3 3.123, mode 7 3.175 (two runs); keep = 60 / 512 MB no better:
        the direction is what helps, the hints do not."""
        self.schedule = {}
        if not mode:
            return
        keep = int(os.environ.get("HPE_B200_L2_KEEP_MB", "104")) << 20
        last_dir = {}                       # buffer key -> direction (True = reversed) of the last launch that touched it
        ops = list(zip(self.named, self.ops))
        for k, (name, (layer, src, res, dst, src2, dst2)) in enumerate(ops):
            ins = [b for b in (src, res, src2) if b is not None and b in self.bufs]
            big = max(ins, key=lambda b: self.bufs[b].numel() * self.bufs[b].element_size(), default=None)
            rev = (not last_dir.get(big, False)) if (mode & 2 and big is not None and big in last_dir) else False
            out_bytes = self.bufs[dst].numel() * self.bufs[dst].element_size()
            l2_out = 2 if (mode & 1 and out_bytes <= keep and self.bufs[dst].dtype == torch.bfloat16) else 0
            l2_res = 1 if (mode & 1 and res is not None) else 0
            read_later = any(src in (o[1], o[2], o[4]) for _, o in ops[k + 1:])
            l2_in = 1 if (mode & 4 and not read_later and src2 is None) else 0
            rev = layer.set_schedule(rev, l2_out, l2_res, l2_in)
            self.schedule[name] = (rev, l2_out, l2_res, l2_in)
            for b in (src, res, src2, dst, dst2):
                if b is not None:
                    last_dir[b] = rev

    def _view(self, key, c):
        """Buffer `key` as pass `c` of the depth-first section sees it: full-batch tensors by batch slice, chunk buffers whole."""
        if key is None:
            return None
        t = self.bufs[key]
        return t[c * self.chunk:(c + 1) * self.chunk] if key in self.full_keys else t

    def _launch_op(self, op, c=None):
        layer, src, res, dst, src2, dst2 = op
        if c is None:
            b = self.bufs
            layer(b[src], b[res] if res else None, b[dst], b[src2] if src2 else None, *((b[dst2],) if dst2 else ()))
        else:
            layer(self._view(src, c), self._view(res, c), self._view(dst, c), self._view(src2, c))
        if dst == "stem":
            maxpool3x3s2(self.bufs["stem"], self.bufs["pool"])

    def _launch_all(self):
        ns = self.n_section_ops
        for c in range(self.n_chunks if ns else 0):
            for op in self.ops[:ns]:
                self._launch_op(op, c)
        for op in self.ops[ns:]:
            self._launch_op(op)

    def layer_table(self):
        """(module path, family, algorithmic FLOPs, ConvLayer, buffers) per conv launch, in launch order.  FLOPs = 2 x MAC
        over the REAL filter taps / channels (no padding, no zero insertion); a fused projection shortcut
        or fused final 1x1 counts the work of both convolutions."""
        rows = []
        for k, (name, layer, src, res, dst, src2, dst2) in enumerate((n, *op) for n, op in zip(self.named, self.ops)):
            d = layer.desc
            m = d.B * layer.Ho * layer.Wo * (self.n_chunks if k < self.n_section_ops else 1)      # per step
            taps = {CONV_1x1: 1, CONV_3x3: 9, CONV_STEM7x7: 49, CONV_DECONV4: 4, CONV_FINAL: 1}[d.kind]
            flops = 2.0 * m * d.Cin * d.Cout * taps + 2.0 * m * d.Cin2 * d.Cout + 2.0 * m * 256 * layer.head2 + 2.0 * m * d.Cout * layer.chain
            if name.startswith("deconv") or name == "final_layer":
                fam = "deconv_head"
            elif name == "backbone.conv1":
                fam = "stem"
            else:
                fam = name.split(".")[1] + (".3x3" if d.kind == CONV_3x3 else ".1x1")
            rows.append((name, fam, flops, layer, (src, res, dst, src2, dst2)))
        return rows

    def layer_bytes(self):
        """Algorithmic HBM bytes of every conv launch of a step, keyed by module path: the tensors it must read (input -- for a
        stride-2 1x1 conv only the pixels it samples --, shortcut, second source) and write (output, chained output), plus its
        packed weights once.  The other half of the per-layer roofline next to ``layer_table``'s FLOPs."""
        out = {}
        for k, (name, layer, src, res, dst, src2, dst2) in enumerate((n, *op) for n, op in zip(self.named, self.ops)):
            d = layer.desc
            passes = self.n_chunks if k < self.n_section_ops else 1
            nb = lambda key: self.bufs[key].numel() * self.bufs[key].element_size() if key in self.bufs else 0      # noqa: E731
            src_b = nb(src) if k >= self.n_section_ops or src not in self.full_keys else nb(src) / self.n_chunks
            if d.kind == CONV_1x1 and d.stride == 2:
                src_b /= 4
            src2_b = (nb(src2) / (4 if d.stride2 == 2 else 1)) if src2 else 0
            total = src_b + src2_b + (nb(res) if res else 0) + nb(dst) + (nb(dst2) if dst2 else 0)
            out[name] = total * passes + layer.w_packed.numel() * layer.w_packed.element_size()
        return out

    def export_acts(self):
        """NCHW fp32 CPU copies of every launch's stored inputs and output, keyed by module path: ``{"x", "res", "x2", "y"}``
        (the stem's ``x`` is omitted: its input is the image).  Teacher-forced parity tests feed these to the fp32 oracle
        layer by layer (tests/test_model_gpu.py)."""
        if self.n_section_ops and self.n_chunks > 1:
            raise RuntimeError("export_acts: the depth-first section reuses its chunk buffers (only the last pass is stored); "
                               "use a batch <= HPE_B200_DF_CHUNK or HPE_B200_DF_CHUNK=0")

        def nchw(t):
            return t.float().cpu() if t.dtype == torch.float32 else t.float().cpu().permute(0, 3, 1, 2).contiguous()
        acts = {}
        for c1, c3 in self.chained.items():              # a chained conv1: input = the block output its conv3 kernel staged
            op = self.ops[list(self.named).index(c3)]
            acts[c1] = {"x": nchw(self.bufs[op[3]]), "y": nchw(self.bufs[op[5]])}
        for name, (layer, src, res, dst, src2, dst2) in zip(self.named, self.ops):
            a = {"y": nchw(self.bufs[dst])}
            if src != "s2d":
                a["x"] = nchw(self.bufs[src])
            if res:
                a["res"] = nchw(self.bufs[res])
            if src2:
                a["x2"] = nchw(self.bufs[src2])
            acts[name] = a
        return acts

    def time_layers(self, reps=3):
        """CUDA-event time of every conv launch of the step, eagerly (no graph): per launch, one untimed call then ``reps``
        back-to-back calls between two events (the launch gap of all but the first is hidden behind the previous call).
        Returns [(name, family, flops, microseconds)].  Diagnostic / benchmark use."""
        out = []
        for k, (name, fam, flops, layer, (src, res, dst, src2, dst2)) in enumerate(self.layer_table()):
            passes = self.n_chunks if k < self.n_section_ops else 1            # a section layer is timed on pass 0, reported per step
            c = 0 if k < self.n_section_ops else None
            call = lambda: layer(self._view(src, c) if c is not None else self.bufs[src],          # noqa: E731
                                 (self._view(res, c) if c is not None else self.bufs[res]) if res else None,
                                 self._view(dst, c) if c is not None else self.bufs[dst],
                                 (self._view(src2, c) if c is not None else self.bufs[src2]) if src2 else None,
                                 *((self.bufs[dst2],) if dst2 else ()))
            call()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                call()
            e1.record()
            e1.synchronize()
            out.append((name, fam, flops, e0.elapsed_time(e1) / reps * 1e3 * passes))
        return out

    def run(self, x: torch.Tensor) -> torch.Tensor:
        preprocess(x, self.s2d)
        use_graph = os.environ.get("HPE_B200_CUDA_GRAPH", "1") != "0"
        if not use_graph:
            self._launch_all()
            return self.heatmaps
        if self.graph is None:
            self._launch_all()                       # warm-up: encodes the TMA descriptors once
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._launch_all()
            self.graph = g
        self.graph.replay()
        return self.heatmaps


def build_model(cfg):
    """``Builder.model()`` (HPE/core/builder.py:15-24): registry lookup, positional constructor call, optional
    checkpoint.  ``''``/``None`` checkpoint = none (the reference would try ``torch.load('')``)."""
    import sys
    registry = sys.modules[__name__]
    name = cfg.model.name
    assert name in __all__, f'Unknown model : "{name}"\n        Available models : {__all__}'
    model = getattr(registry, name)(cfg.model.nkpts, cfg.model.backbone, cfg.model.pretrained)
    if cfg.model.checkpoint:
        sd = torch.load(cfg.model.checkpoint, map_location="cpu")
        model.load_state_dict(sd.get("model_state", sd))
    return model
