"""CPU oracle (plain ``torch.nn``, fp32) for the backbone + heatmap-head forward.

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.

``ResNet50Ref`` restates the arithmetic of the vendored timm ResNet-50 that the reference's
config-default model builds (``DeepPose.__init__``: HPE/models/reference/bottom_up/regression/models/
DeepPose.py:4-10 -> ``timm.create_model('resnet50')``):

* stem  conv7x7 s2 p3 (3->64, bias=False) + BN + ReLU + MaxPool 3x3 s2 p1
        (HPE/thirdparty/pytorch-image-models/timm/models/resnet.py:554-576)
* stages ``[3,4,6,3]`` of ``Bottleneck`` (resnet.py:154-270): 1x1 -> BN -> ReLU -> 3x3 (stride on the
  3x3) -> BN -> ReLU -> 1x1 -> BN -> (+shortcut; downsample = 1x1 stride-s conv + BN) -> ReLU
* ``forward_features`` order (resnet.py:742-756); ``resnet50`` cfg (resnet.py:1548-1552)
* init: kaiming_normal fan_out on every conv, zero last-BN gamma per block (resnet.py:610-621, 239-242)

Module / state_dict names equal timm's (``conv1, bn1, layer{1-4}.{i}.conv{1-3}/bn{1-3}/downsample.{0,1}``)
so a reference checkpoint loads into it unchanged.  PINNED against the reference's own code by
``tests/golden/make_golden.py`` (run in the build container, where ``/root/reference`` is mounted).

``HeatmapHeadRef`` / ``HeatmapPoseRef`` restate SURVEY.md App. A.1 (**parity unpinned**: the
reference has no deconv head).
"""
from __future__ import annotations

import torch
import torch.nn as nn


class BottleneckRef(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        out = planes * self.expansion
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride=stride, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, out, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(out)
        self.downsample = downsample

    def forward(self, x):
        s = x
        y = torch.relu(self.bn1(self.conv1(x)))
        y = torch.relu(self.bn2(self.conv2(y)))
        y = self.bn3(self.conv3(y))
        if self.downsample is not None:
            s = self.downsample(s)
        return torch.relu(y + s)


class ResNet50Ref(nn.Module):
    layers = (3, 4, 6, 3)
    planes = (64, 128, 256, 512)

    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.maxpool = nn.MaxPool2d(3, stride=2, padding=1)
        inplanes = 64
        for i, (n, p) in enumerate(zip(self.layers, self.planes)):
            stride = 1 if i == 0 else 2
            blocks = []
            for j in range(n):
                s = stride if j == 0 else 1
                ds = None
                if j == 0 and (s != 1 or inplanes != p * 4):
                    ds = nn.Sequential(nn.Conv2d(inplanes, p * 4, 1, stride=s, bias=False),
                                       nn.BatchNorm2d(p * 4))
                blocks.append(BottleneckRef(inplanes, p, s, ds))
                inplanes = p * 4
            setattr(self, f"layer{i + 1}", nn.Sequential(*blocks))
        self.num_features = inplanes
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
        for m in self.modules():
            if isinstance(m, BottleneckRef):
                nn.init.zeros_(m.bn3.weight)

    def forward_features(self, x):
        x = self.maxpool(torch.relu(self.bn1(self.conv1(x))))
        return self.layer4(self.layer3(self.layer2(self.layer1(x))))

    forward = forward_features


class HeatmapHeadRef(nn.Module):
    """3x [ConvTranspose2d(k4,s2,p1,op0,bias=False) -> BN -> ReLU] + Conv2d 1x1 (bias)."""

    def __init__(self, in_ch=2048, num_joints=17, mid=256):
        super().__init__()
        layers = []
        c = in_ch
        for _ in range(3):
            layers += [nn.ConvTranspose2d(c, mid, 4, stride=2, padding=1, output_padding=0, bias=False),
                       nn.BatchNorm2d(mid), nn.ReLU(inplace=True)]
            c = mid
        self.deconv_layers = nn.Sequential(*layers)
        self.final_layer = nn.Conv2d(mid, num_joints, 1, bias=True)

    def forward(self, x):
        return self.final_layer(self.deconv_layers(x))


class HeatmapPoseRef(nn.Module):
    """Same constructor shape as the reference's ``DeepPose(num_joints, backbone, pretrained)``
    (DeepPose.py:5) with the fc regression head replaced by the A.1 heatmap head."""

    def __init__(self, num_joints=17, backbone="resnet50", pretrained=False):
        super().__init__()
        assert backbone == "resnet50"
        self.njoints = num_joints
        self.backbone = ResNet50Ref()
        head = HeatmapHeadRef(2048, num_joints)
        self.deconv_layers = head.deconv_layers
        self.final_layer = head.final_layer

    def forward(self, x):
        return self.final_layer(self.deconv_layers(self.backbone.forward_features(x)))


def randomize_synthetic(model: nn.Module, seed_bn=1, seed_head=2):
    """SURVEY.md §8(d) synthetic weights: keep the default conv init, re-randomise every BN
    (gamma~U(0.5,1.5), beta~N(0,0.1), mean~N(0,0.1), var~U(0.5,1.5)) because timm's zero_init_last
    + (0,1) running stats make a random-init net degenerate; deconv/final weights N(0,0.02), bias 0."""
    g = torch.Generator().manual_seed(seed_bn)
    for m in model.modules():
        if isinstance(m, nn.BatchNorm2d):
            with torch.no_grad():
                m.weight.copy_(torch.rand(m.weight.shape, generator=g) + 0.5)
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
                m.running_mean.copy_(torch.randn(m.running_mean.shape, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(m.running_var.shape, generator=g) + 0.5)
    g = torch.Generator().manual_seed(seed_head)
    for m in model.modules():
        if isinstance(m, nn.ConvTranspose2d) or (isinstance(m, nn.Conv2d) and m.bias is not None):
            with torch.no_grad():
                m.weight.copy_(torch.randn(m.weight.shape, generator=g) * 0.02)
                if m.bias is not None:
                    m.bias.zero_()
    return model


def build_synthetic(num_joints=17, seed=0, normalize=False):
    """Seeded synthetic model (§8d).  ``normalize=True`` additionally rescales the final 1x1 conv so that
    max|heatmap| == 1 on a seeded 1x3x256x192 calibration image -- random-init ResNets produce heatmaps of
    magnitude ~75, while the north_star tolerances (max-abs 1e-2) are stated for trained-net heatmaps in
    [0,1]; normalising makes the absolute tolerance meaningful without changing any kernel input range."""
    torch.manual_seed(seed)
    model = randomize_synthetic(HeatmapPoseRef(num_joints)).eval()
    if normalize:
        x = torch.rand(1, 3, 256, 192, generator=torch.Generator().manual_seed(1234))
        with torch.no_grad():
            peak = model(x).abs().max()
            model.final_layer.weight.div_(peak)
    return model


@torch.no_grad()
def forward_bf16_emulated(model: HeatmapPoseRef, x: torch.Tensor, return_features=False):
    """The same network evaluated the way the bf16 CUDA path evaluates it: BN folded into a
    per-channel scale/shift applied to the fp32 conv accumulator, conv weights and every stored
    activation rounded to bf16, fp32 accumulation, heatmaps left in fp32.  Used for the tight
    (kernel-bug-catching) parity check; the loose check is against the plain fp32 forward."""
    import torch.nn.functional as F

    def r(t):
        return t.to(torch.bfloat16).to(torch.float32)

    def fold(bn):
        s = bn.weight / torch.sqrt(bn.running_var + bn.eps)
        return s, bn.bias - bn.running_mean * s

    def cbr(x, conv, bn, relu=True, res=None):
        y = F.conv2d(x, r(conv.weight), None, conv.stride, conv.padding)
        s, b = fold(bn)
        y = y * s[None, :, None, None] + b[None, :, None, None]
        if res is not None:
            y = y + res
        if relu:
            y = torch.relu(y)
        return r(y)

    bb = model.backbone
    x = cbr(r(x), bb.conv1, bb.bn1)
    x = bb.maxpool(x)
    for name in ("layer1", "layer2", "layer3", "layer4"):
        for blk in getattr(bb, name):
            s = x
            y = cbr(x, blk.conv1, blk.bn1)
            y = cbr(y, blk.conv2, blk.bn2)
            if blk.downsample is not None:
                s = cbr(x, blk.downsample[0], blk.downsample[1], relu=False)
            x = cbr(y, blk.conv3, blk.bn3, relu=True, res=s)
    feat = x
    for i in (0, 3, 6):
        dc, bn = model.deconv_layers[i], model.deconv_layers[i + 1]
        y = F.conv_transpose2d(x, r(dc.weight), None, 2, 1)
        s, b = fold(bn)
        x = r(torch.relu(y * s[None, :, None, None] + b[None, :, None, None]))
    hm = F.conv2d(x, r(model.final_layer.weight), model.final_layer.bias)
    return (hm, feat) if return_features else hm
