"""CPU oracle (plain torch.nn, fp32) for the reference's real pose model ``DINOv3Pose``.

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.  Parity **pinned**: ``tests/golden/make_golden_dinov3pose.py``
loads this module's seeded ``state_dict`` into the REFERENCE's own ``DINOv3Pose`` (key-compatible, strict) and stores the
reference's train-mode and eval-mode outputs in ``tests/golden/dinov3pose_golden.npz``;
``tests/test_dinov3pose_oracle_cpu.py`` checks this restatement against them.

Follows (file:line under /root/reference/HPE):
  models/pose.py:9-65                          DINOv3Pose (backbone -> adaptor -> sppf -> pan -> head)
  thirdparty/dinov3/dinov3/models/convnext.py:42-84, 87-117, 121-180   Block, LayerNorm, ConvNeXt stem/downsample/stages
  models/backbones/dinov3convnext.py:76-84, 93-110                      forward_features_list, size table
  models/nn/modules/block.py:20-45, 47-71, 74-100, 102-114, 117-157, 236-254   ConvBlock, Bottleneck, BottleneckCSP, SPPF, PAN, FeatureAdaptor
  models/nn/modules/head.py:131-211            PoseHead
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

CONVNEXT_SIZES = {                                              # dinov3convnext.py:93-110
    "tiny": dict(depths=[3, 3, 9, 3], dims=[96, 192, 384, 768]),
    "small": dict(depths=[3, 3, 27, 3], dims=[96, 192, 384, 768]),
    "base": dict(depths=[3, 3, 27, 3], dims=[128, 256, 512, 1024]),
    "large": dict(depths=[3, 3, 27, 3], dims=[192, 384, 768, 1536]),
}


class LayerNorm(nn.Module):
    """convnext.py:87-117: channels_last = F.layer_norm; channels_first = the same statistics over dim 1."""

    def __init__(self, c, eps=1e-6, data_format="channels_last"):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(c)); self.bias = nn.Parameter(torch.zeros(c))
        self.eps, self.data_format, self.c = eps, data_format, c

    def forward(self, x):
        if self.data_format == "channels_last":
            return F.layer_norm(x, (self.c,), self.weight, self.bias, self.eps)
        u = x.mean(1, keepdim=True)
        s = (x - u).pow(2).mean(1, keepdim=True)
        return self.weight[:, None, None] * ((x - u) / torch.sqrt(s + self.eps)) + self.bias[:, None, None]


class Block(nn.Module):
    """convnext.py:42-84."""

    def __init__(self, dim):
        super().__init__()
        self.dwconv = nn.Conv2d(dim, dim, 7, padding=3, groups=dim)
        self.norm = LayerNorm(dim)
        self.pwconv1 = nn.Linear(dim, 4 * dim)
        self.act = nn.GELU()
        self.pwconv2 = nn.Linear(4 * dim, dim)
        self.gamma = nn.Parameter(1e-6 * torch.ones(dim))

    def forward(self, x):
        y = self.dwconv(x).permute(0, 2, 3, 1)
        y = self.pwconv2(self.act(self.pwconv1(self.norm(y))))
        return x + (self.gamma * y).permute(0, 3, 1, 2)


class ConvNeXt(nn.Module):
    """convnext.py:121-180 (+ the final ``norm`` the checkpoint carries; unused by the pose path)."""

    def __init__(self, depths, dims):
        super().__init__()
        self.downsample_layers = nn.ModuleList([nn.Sequential(nn.Conv2d(3, dims[0], 4, 4), LayerNorm(dims[0], data_format="channels_first"))])
        for i in range(3):
            self.downsample_layers.append(nn.Sequential(LayerNorm(dims[i], data_format="channels_first"), nn.Conv2d(dims[i], dims[i + 1], 2, 2)))
        self.stages = nn.ModuleList([nn.Sequential(*[Block(dims[i]) for _ in range(depths[i])]) for i in range(4)])
        self.norm = nn.LayerNorm(dims[-1], eps=1e-6)
        self.norms = nn.ModuleList([nn.Identity() for _ in range(3)] + [self.norm])      # convnext.py:175-176 (shared)

    def forward_features_list(self, x):            # dinov3convnext.py:76-84 (one image tensor)
        out = []
        for i in range(4):
            x = self.stages[i](self.downsample_layers[i](x))
            out.append(x)
        return out


class ConvBlock(nn.Module):
    """block.py:20-45."""

    def __init__(self, c1, c2, k=1, s=1, g=1, act=True):
        super().__init__()
        self.conv = nn.Conv2d(c1, c2, k, s, k // 2, groups=g, bias=False)
        self.bn = nn.BatchNorm2d(c2)
        self.act = nn.SiLU() if act else nn.Identity()

    def forward(self, x):
        return self.act(self.bn(self.conv(x)))


class Bottleneck(nn.Module):
    """block.py:47-71."""

    def __init__(self, c1, c2, shortcut=True, e=0.5):
        super().__init__()
        c_ = int(c2 * e)
        self.cv1 = ConvBlock(c1, c_, 3, 1); self.cv2 = ConvBlock(c_, c2, 3, 1)
        self.add = shortcut and c1 == c2

    def forward(self, x):
        y = self.cv2(self.cv1(x))
        return x + y if self.add else y


class BottleneckCSP(nn.Module):
    """block.py:74-100."""

    def __init__(self, c1, c2, n=1, shortcut=True, e=0.5):
        super().__init__()
        c_ = int(c2 * e)
        self.cv1 = ConvBlock(c1, c_, 1, 1)
        self.cv2 = nn.Conv2d(c1, c_, 1, 1, bias=False)
        self.cv3 = nn.Conv2d(c_, c_, 1, 1, bias=False)
        self.cv4 = ConvBlock(2 * c_, c2, 1, 1)
        self.bn = nn.BatchNorm2d(2 * c_)
        self.act = nn.SiLU()
        self.m = nn.Sequential(*(Bottleneck(c_, c_, shortcut, e=1.0) for _ in range(n)))

    def forward(self, x):
        y1 = self.cv3(self.m(self.cv1(x)))
        y2 = self.cv2(x)
        return self.cv4(self.act(self.bn(torch.cat((y1, y2), 1))))


class SPPF(nn.Module):
    """block.py:102-114."""

    def __init__(self, c1, c2, k=5):
        super().__init__()
        self.conv1 = ConvBlock(c1, c1 // 2, 1, 1); self.conv2 = ConvBlock(c1 // 2 * 4, c2, 1, 1)
        self.pool = nn.MaxPool2d(k, 1, k // 2)

    def forward(self, x):
        y = [self.conv1(x)]
        for _ in range(3):
            y.append(self.pool(y[-1]))
        return self.conv2(torch.cat(y, 1))


class PAN(nn.Module):
    """block.py:117-157."""

    def __init__(self, ch):
        super().__init__()
        c1, c2, c3 = ch
        self.fpn_conv1 = ConvBlock(c3, c2, 1); self.fpn_conv2 = BottleneckCSP(c2 + c2, c2, n=2, shortcut=False)
        self.fpn_conv3 = ConvBlock(c2, c1, 1); self.fpn_conv4 = BottleneckCSP(c1 + c1, c1, n=2, shortcut=False)
        self.down1 = ConvBlock(c1, c1, 3, 2); self.pan_conv1 = BottleneckCSP(c1 + c2, c2, n=2, shortcut=False)
        self.down2 = ConvBlock(c2, c2, 3, 2); self.pan_conv2 = BottleneckCSP(c2 + c3, c3, n=2, shortcut=False)

    def forward(self, feats):
        c1, c2, c3 = feats
        up = lambda t, ref: F.interpolate(t, size=ref.shape[2:], mode="bilinear", align_corners=False)   # noqa: E731
        c2f = self.fpn_conv2(torch.cat([c2, up(self.fpn_conv1(c3), c2)], 1))
        c1f = self.fpn_conv4(torch.cat([c1, up(self.fpn_conv3(c2f), c1)], 1))
        c2p = self.pan_conv1(torch.cat([up(self.down1(c1f), c2f), c2f], 1))
        c3p = self.pan_conv2(torch.cat([up(self.down2(c2p), c3), c3], 1))
        return [c1f, c2p, c3p]


class FeatureAdaptor(nn.Module):
    """block.py:236-254."""

    def __init__(self, sc, tc):
        super().__init__()
        self.adapters = nn.ModuleList([nn.Sequential(ConvBlock(a, b, 1), ConvBlock(b, b, 3)) for a, b in zip(sc, tc)])

    def forward(self, feats):
        return [m(f) for m, f in zip(self.adapters, feats)]


class PoseHead(nn.Module):
    """head.py:131-211 (``stride`` buffer included for state_dict compatibility)."""

    def __init__(self, ncls, kpt_shape, in_ch):
        super().__init__()
        self.nc, self.kpt_shape, self.nk = ncls, kpt_shape, kpt_shape[0] * kpt_shape[1]
        self.register_buffer("stride", torch.tensor([8, 16, 32], dtype=torch.float32))
        c3 = max(in_ch[0], min(ncls, 100)); c4 = max(in_ch[0] // 4, self.nk)
        self.cv3 = nn.ModuleList(nn.Sequential(nn.Sequential(ConvBlock(x, x, 3, g=x), ConvBlock(x, c3, 1)),
                                               nn.Sequential(ConvBlock(c3, c3, 3), ConvBlock(c3, c3, 1)),
                                               nn.Conv2d(c3, ncls, 1)) for x in in_ch)
        self.cv4 = nn.ModuleList(nn.Sequential(ConvBlock(x, c4, 3), ConvBlock(c4, c4, 3), nn.Conv2d(c4, self.nk, 1)) for x in in_ch)
        for m in self.cv3:
            nn.init.constant_(m[-1].bias, -math.log((1 - 0.01) / 0.01))

    def forward(self, feats):
        return [torch.cat([self.cv3[i](f), self.cv4[i](f)], 1) for i, f in enumerate(feats)]


class DINOv3PoseRef(nn.Module):
    """pose.py:9-65; ``forward`` returns the per-level raw maps (the reference's train-mode output, head.py:176-181)."""

    def __init__(self, size="tiny", nkpts=(17, 3), ncls=1):
        super().__init__()
        cfg = CONVNEXT_SIZES[size]
        self.tc = [192, 384, 768]
        self.backbone = ConvNeXt(cfg["depths"], cfg["dims"])
        self.adaptor = FeatureAdaptor(cfg["dims"][1:], self.tc)
        self.sppf = SPPF(self.tc[-1], self.tc[-1])
        self.pan = PAN(self.tc)
        self.head = PoseHead(ncls, nkpts, self.tc)

    def forward_features(self, x):
        f = self.adaptor(self.backbone.forward_features_list(x)[1:])
        f[-1] = self.sppf(f[-1])
        return self.pan(f)

    def forward(self, x):
        return self.head(self.forward_features(x))


def build_synthetic(size="tiny", nkpts=(17, 3), ncls=1, seed=0):
    """Seeded weights that make every op matter: trunc-normal convs/linears as the reference's init (convnext.py:196-205),
    but layer-scale gamma ~ U(0.2, 1) instead of 1e-6, biases N(0, 0.02), LayerNorm/BatchNorm affine and BN running
    statistics randomised (the defaults make LN/BN near-identities and the residual branches vanish)."""
    torch.manual_seed(seed)
    m = DINOv3PoseRef(size, nkpts, ncls)
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for name, mod in m.named_modules():
            if isinstance(mod, (nn.Conv2d, nn.Linear)):
                fan_in = mod.weight[0].numel()
                mod.weight.copy_(torch.randn(mod.weight.shape, generator=g) * (1.0 / math.sqrt(fan_in)))
                if mod.bias is not None and not name.startswith("head.cv3"):
                    mod.bias.copy_(torch.randn(mod.bias.shape, generator=g) * 0.02)
            elif isinstance(mod, (LayerNorm, nn.LayerNorm)):
                mod.weight.copy_(torch.rand(mod.weight.shape, generator=g) + 0.5)
                mod.bias.copy_(torch.randn(mod.bias.shape, generator=g) * 0.1)
            elif isinstance(mod, nn.BatchNorm2d):
                mod.weight.copy_(torch.rand(mod.weight.shape, generator=g) + 0.5)
                mod.bias.copy_(torch.randn(mod.bias.shape, generator=g) * 0.1)
                mod.running_mean.copy_(torch.randn(mod.running_mean.shape, generator=g) * 0.1)
                mod.running_var.copy_(torch.rand(mod.running_var.shape, generator=g) + 0.5)
            elif isinstance(mod, Block):
                mod.gamma.copy_(torch.rand(mod.gamma.shape, generator=g) * 0.8 + 0.2)
    return m.eval()


class HeatmapConvNeXtRef(nn.Module):
    """the reference's ConvNeXt/DINOv3 backbone (all four stages, ``forward_features_list``,
    dinov3convnext.py:76-84) behind the SAME App. A.1 deconv heatmap head as the ResNet-50 model
    (``oracle.resnet_ref.HeatmapHeadRef``).  The backbone half is pinned with the rest of this file (its weights load
    into the reference's own ConvNeXt); the head half is spec-pinned (App. A.1, parity unpinned)."""

    def __init__(self, size="tiny", num_joints=17):
        super().__init__()
        from .resnet_ref import HeatmapHeadRef
        cfg = CONVNEXT_SIZES[size]
        self.njoints = num_joints
        self.backbone = ConvNeXt(cfg["depths"], cfg["dims"])
        head = HeatmapHeadRef(cfg["dims"][-1], num_joints)
        self.deconv_layers, self.final_layer = head.deconv_layers, head.final_layer

    def forward_features(self, x):
        return self.backbone.forward_features_list(x)[-1]

    def forward(self, x):
        return self.final_layer(self.deconv_layers(self.forward_features(x)))


def build_synthetic_heatmap(size="tiny", num_joints=17, seed=0):
    """Seeded ConvNeXt + heatmap head: backbone randomised like ``build_synthetic``; head deconv / final weights
    N(0, 1/sqrt(fan_in)), BN affine and running statistics randomised; the final 1x1 is rescaled so that max|heatmap| = 1 on a
    seeded calibration image (absolute tolerances are then on the scale of a trained net's heatmaps)."""
    torch.manual_seed(seed)
    m = HeatmapConvNeXtRef(size, num_joints)
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for name, mod in m.named_modules():
            if isinstance(mod, nn.ConvTranspose2d):
                fan_in = mod.weight.shape[0] * 4                      # 2x2 taps reach every output pixel
                mod.weight.copy_(torch.randn(mod.weight.shape, generator=g) * (1.0 / math.sqrt(fan_in)))
            elif isinstance(mod, (nn.Conv2d, nn.Linear)):
                fan_in = mod.weight[0].numel()
                mod.weight.copy_(torch.randn(mod.weight.shape, generator=g) * (1.0 / math.sqrt(fan_in)))
                if mod.bias is not None:
                    mod.bias.copy_(torch.randn(mod.bias.shape, generator=g) * 0.02)
            elif isinstance(mod, (LayerNorm, nn.LayerNorm)):
                mod.weight.copy_(torch.rand(mod.weight.shape, generator=g) + 0.5)
                mod.bias.copy_(torch.randn(mod.bias.shape, generator=g) * 0.1)
            elif isinstance(mod, nn.BatchNorm2d):
                mod.weight.copy_(torch.rand(mod.weight.shape, generator=g) + 0.5)
                mod.bias.copy_(torch.randn(mod.bias.shape, generator=g) * 0.1)
                mod.running_mean.copy_(torch.randn(mod.running_mean.shape, generator=g) * 0.1)
                mod.running_var.copy_(torch.rand(mod.running_var.shape, generator=g) + 0.5)
            elif isinstance(mod, Block):
                mod.gamma.copy_(torch.rand(mod.gamma.shape, generator=g) * 0.8 + 0.2)
        m.eval()
        # head BatchNorm running statistics CALIBRATED on a seeded batch (like a trained net's; random statistics make the three
        # un-normalised deconv stages amplify every upstream rounding error), then the final 1x1 rescaled to max|heatmap| = 1
        x = torch.rand(2, 3, 128, 96, generator=torch.Generator().manual_seed(1234))
        f = m.forward_features(x)
        for i in (0, 3, 6):
            z = m.deconv_layers[i](f)
            bn = m.deconv_layers[i + 1]
            bn.running_mean.copy_(z.mean((0, 2, 3))); bn.running_var.copy_(z.var((0, 2, 3), unbiased=False))
            f = torch.relu(bn(z))
        m.final_layer.bias.zero_()
        m.final_layer.weight.div_(m.final_layer(f).abs().max())
    return m
