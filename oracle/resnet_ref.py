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


def forward_train_bf16_emulated(model: HeatmapPoseRef, x: torch.Tensor, taps: dict = None):
    """Training-mode forward (BatchNorm batch statistics) with the bf16 CUDA path's rounding points and a
    straight-through gradient, so autograd gives the fp32 gradient of the network *as the kernels evaluate it*:
    conv weights, the raw conv output z and the unit output y = relu(bn(z) + residual) are rounded to bf16,
    statistics/affine are fp32 and computed from the rounded z.  Needed because the §8(d) random-init net in
    train() mode amplifies any perturbation ~1.2x per block (BatchNorm removes a mean that is large against the
    spread), so the plain fp32 forward and ANY bf16 evaluation of it legitimately drift apart by tens of percent
    at the head; parity of the kernels is checked against this restatement, conditioning against the fp32 one.
    Does not update running statistics.  ``taps`` (optional dict) receives every unit's z and y by conv name."""
    import torch.nn.functional as F

    def r(t):                                  # round to bf16, identity gradient
        return t + (t.to(torch.bfloat16).to(torch.float32) - t).detach()

    def unit(name, x, conv, bn, relu=True, res=None, transposed=False):
        w = r(conv.weight)
        z = F.conv_transpose2d(x, w, None, 2, 1) if transposed else F.conv2d(x, w, None, conv.stride, conv.padding)
        z = r(z)
        mean = z.mean((0, 2, 3))
        var = ((z - mean[None, :, None, None]) ** 2).mean((0, 2, 3))
        scale = bn.weight * torch.rsqrt(var + bn.eps)
        shift = bn.bias - mean * scale
        y = z * scale[None, :, None, None] + shift[None, :, None, None]
        if res is not None:
            y = y + res
        if relu:
            y = torch.relu(y)
        y = r(y)
        if taps is not None:
            taps[name] = (z.detach(), y.detach())
        return y

    bb = model.backbone
    x = unit("backbone.conv1", r(x), bb.conv1, bb.bn1)
    x = bb.maxpool(x)
    for lname in ("layer1", "layer2", "layer3", "layer4"):
        for j, blk in enumerate(getattr(bb, lname)):
            pre = f"backbone.{lname}.{j}"
            s = x
            y = unit(pre + ".conv1", x, blk.conv1, blk.bn1)
            y = unit(pre + ".conv2", y, blk.conv2, blk.bn2)
            if blk.downsample is not None:
                s = unit(pre + ".downsample.0", x, blk.downsample[0], blk.downsample[1], relu=False)
            x = unit(pre + ".conv3", y, blk.conv3, blk.bn3, relu=True, res=s)
    for i in (0, 3, 6):
        x = unit(f"deconv_layers.{i}", x, model.deconv_layers[i], model.deconv_layers[i + 1], transposed=True)
    return F.conv2d(x, r(model.final_layer.weight), model.final_layer.bias)


def teacher_forced_backward(model: HeatmapPoseRef, acts: dict, dhm: torch.Tensor):
    """fp32 backward pass of the training step evaluated on SAVED forward state (``acts[conv_name] = {"x", "z"}``,
    NCHW fp32 copies of what the CUDA path stored: each unit's input and raw conv output), walking the same
    topology as ``HeatmapPoseRef.forward``.  Because the ReLU masks and BatchNorm statistics come from the same
    stored tensors the kernels used, this isolates the backward kernels (BN backward, dgrad, wgrad, maxpool
    backward) from the forward's chaotic sensitivity (see ``forward_train_bf16_emulated``).
    Returns ``(grads, fwd_err)``: parameter-name -> gradient, and conv-name -> (max-abs error of the stored z
    against conv(x) in fp32, relative to max|z|)."""
    import torch.nn.functional as F

    def r16(t):
        return t.to(torch.bfloat16).to(torch.float32)

    grads, fwd_err = {}, {}

    def unit_bwd(cname, bname, conv, bn, dy, relu=True, res=None, transposed=False, need_dx=True):
        x, z = acts[cname]["x"], acts[cname]["z"]
        zz = z.clone().requires_grad_(True)
        gam = bn.weight.detach().clone().requires_grad_(True)
        bet = bn.bias.detach().clone().requires_grad_(True)
        mean = zz.mean((0, 2, 3))
        var = ((zz - mean[None, :, None, None]) ** 2).mean((0, 2, 3))
        y = (zz - mean[None, :, None, None]) * torch.rsqrt(var + bn.eps)[None, :, None, None] * gam[None, :, None, None] \
            + bet[None, :, None, None]
        rr = None
        if res is not None:
            rr = res.clone().requires_grad_(True)
            y = y + rr
        if relu:
            y = torch.relu(y)
        y.backward(dy)
        grads[bname + ".weight"], grads[bname + ".bias"] = gam.grad, bet.grad
        dz = zz.grad
        xx = x.clone().requires_grad_(need_dx)
        ww = r16(conv.weight.detach()).requires_grad_(True)
        zc = F.conv_transpose2d(xx, ww, None, 2, 1) if transposed else F.conv2d(xx, ww, None, conv.stride, conv.padding)
        fwd_err[cname] = ((zc.detach() - z).abs().max() / z.abs().max().clamp_min(1e-30)).item()
        zc.backward(dz)
        grads[cname + ".weight"] = ww.grad
        return (xx.grad if need_dx else None), (rr.grad if rr is not None else None)

    # final 1x1 conv (+bias)
    xf = acts["final_layer"]["x"].clone().requires_grad_(True)
    wf = r16(model.final_layer.weight.detach()).requires_grad_(True)
    bf = model.final_layer.bias.detach().clone().requires_grad_(True)
    F.conv2d(xf, wf, bf).backward(dhm)
    grads["final_layer.weight"], grads["final_layer.bias"] = wf.grad, bf.grad
    d = xf.grad
    for i in (6, 3, 0):
        d, _ = unit_bwd(f"deconv_layers.{i}", f"deconv_layers.{i + 1}", model.deconv_layers[i], model.deconv_layers[i + 1], d,
                        transposed=True)
    bb = model.backbone
    for lname in ("layer4", "layer3", "layer2", "layer1"):
        blocks = list(getattr(bb, lname))
        for j in range(len(blocks) - 1, -1, -1):
            blk, pre = blocks[j], f"backbone.{lname}.{j}"
            xin = acts[pre + ".conv1"]["x"]
            if blk.downsample is not None:
                short = acts[pre + ".downsample.0"]["y"]
            else:
                short = xin
            dt2, g3 = unit_bwd(pre + ".conv3", pre + ".bn3", blk.conv3, blk.bn3, d, res=short)
            dt1, _ = unit_bwd(pre + ".conv2", pre + ".bn2", blk.conv2, blk.bn2, dt2)
            dx, _ = unit_bwd(pre + ".conv1", pre + ".bn1", blk.conv1, blk.bn1, dt1)
            if blk.downsample is not None:
                ds, _ = unit_bwd(pre + ".downsample.0", pre + ".downsample.1", blk.downsample[0], blk.downsample[1], g3,
                                 relu=False)
                d = dx + ds
            else:
                d = dx + g3
    ys = acts["backbone.conv1"]["y"].clone().requires_grad_(True)
    F.max_pool2d(ys, 3, 2, 1).backward(d)
    unit_bwd("backbone.conv1", "backbone.bn1", bb.conv1, bb.bn1, ys.grad, need_dx=False)
    return grads, fwd_err


@torch.no_grad()
def build_calibrated(num_joints=17, seed=0, gamma3=0.25, calib_batch=4, size=(256, 192)):
    """A WELL-CONDITIONED synthetic weight set: same architecture and seeded conv init as ``build_synthetic``, but every
    BatchNorm's running statistics are CALIBRATED layer by layer on a seeded synthetic batch (running_mean / running_var :=
    that layer's batch statistics, like a trained net's), gamma = 1 (``gamma3`` on each block's last BN, in the spirit of
    timm's zero_init_last: the residual stream is dominated by the identity path), beta = 0, and the final 1x1 conv rescaled
    so that max|heatmap| = 1 on the calibration batch.  Activations stay O(1) through all 53 layers, so an error is not
    amplified and north_star's absolute bar (max-abs <= 1e-2 on [0,1]-scale heatmaps in bf16) is meaningful and attainable;
    ``build_synthetic``'s randomised statistics amplify every perturbation (the bf16 noise floor there is 1.5e-2)."""
    import torch.nn.functional as F
    torch.manual_seed(seed)
    model = HeatmapPoseRef(num_joints).eval()
    g = torch.Generator().manual_seed(seed + 2)
    for m in model.modules():
        if isinstance(m, nn.ConvTranspose2d) or (isinstance(m, nn.Conv2d) and m.bias is not None):
            m.weight.copy_(torch.randn(m.weight.shape, generator=g) * 0.02)
            if m.bias is not None:
                m.bias.zero_()
    x = torch.rand(calib_batch, 3, size[0], size[1], generator=torch.Generator().manual_seed(4321))

    def calib(z, bn, gamma=1.0):
        bn.running_mean.copy_(z.mean((0, 2, 3)))
        bn.running_var.copy_(z.var((0, 2, 3), unbiased=False))
        bn.weight.fill_(gamma)
        bn.bias.zero_()
        return bn(z)

    bb = model.backbone
    x = bb.maxpool(torch.relu(calib(bb.conv1(x), bb.bn1)))
    for name in ("layer1", "layer2", "layer3", "layer4"):
        for blk in getattr(bb, name):
            s = x
            y = torch.relu(calib(blk.conv1(x), blk.bn1))
            y = torch.relu(calib(blk.conv2(y), blk.bn2))
            y = calib(blk.conv3(y), blk.bn3, gamma3)
            if blk.downsample is not None:
                s = calib(blk.downsample[0](x), blk.downsample[1])
            x = torch.relu(y + s)
    for i in (0, 3, 6):
        x = torch.relu(calib(model.deconv_layers[i](x), model.deconv_layers[i + 1]))
    hm = model.final_layer(x)
    model.final_layer.weight.div_(hm.abs().max())
    return model


@torch.no_grad()
def layer_forward_fp32(model: HeatmapPoseRef, name: str, x: torch.Tensor, res: torch.Tensor = None, x2: torch.Tensor = None,
                       pool: bool = False, head: bool = False):
    """ONE layer of the eval forward in fp32 on GIVEN inputs (teacher forcing): ``x`` (and ``res`` / ``x2``) are NCHW fp32
    copies of the bf16 tensors the CUDA path itself stored as this layer's inputs, so the comparison isolates this layer's
    kernel from everything upstream (any error is at most this layer's own: one bf16 rounding of the output).
    ``name`` = module path of the conv (``backbone.layer2.0.conv3``, ``deconv_layers.6``, ``final_layer``, ``backbone.conv1``).
      * conv weights rounded to bf16 (the kernels' operand type), BN(eval) as fp32 scale/shift on the fp32 accumulator;
      * ``x2``: the block input of a first block -- the projection shortcut ``downsample(x2)`` is added before the ReLU
        (timm Bottleneck.forward, resnet.py:262-268); the fused kernel folds both BN scales into the weights BEFORE the
        bf16 rounding, which is restated here;
      * ``pool``: stem followed by MaxPool2d(3, 2, 1) of the bf16-rounded stem output (resnet.py:574-576);
      * ``head``: last deconv stage (bf16-rounded) followed by ``final_layer`` in fp32 (App. A.1).
    Returns the layer output in fp32 BEFORE its final bf16 rounding (heatmaps for ``head`` / ``final_layer``)."""
    import torch.nn.functional as F

    def r(t):
        return t.to(torch.bfloat16).to(torch.float32)

    def fold(bn):
        s = bn.weight / torch.sqrt(bn.running_var + bn.eps)
        return s, bn.bias - bn.running_mean * s

    mods = dict(model.named_modules())
    conv = mods[name]
    if name == "final_layer":
        return F.conv2d(x, r(conv.weight), conv.bias)
    if name.startswith("deconv_layers."):
        bn = mods[f"deconv_layers.{int(name.split('.')[1]) + 1}"]
        s, b = fold(bn)
        y = torch.relu(F.conv_transpose2d(x, r(conv.weight), None, 2, 1) * s[None, :, None, None] + b[None, :, None, None])
        if head:
            return F.conv2d(r(y), r(model.final_layer.weight), model.final_layer.bias)
        return y
    bn = mods[name[:-1] + "1"] if name.endswith("downsample.0") else mods[name.replace("conv", "bn")]
    s, b = fold(bn)
    if x2 is not None:                         # conv3 + bn3 + downsample(x2) + add + ReLU, scales folded before rounding
        ds_conv, ds_bn = mods[name.rsplit(".", 1)[0] + ".downsample.0"], mods[name.rsplit(".", 1)[0] + ".downsample.1"]
        s2, b2 = fold(ds_bn)
        y = F.conv2d(x, r(conv.weight * s[:, None, None, None]), None, conv.stride, conv.padding)
        y = y + F.conv2d(x2, r(ds_conv.weight * s2[:, None, None, None]), None, ds_conv.stride, ds_conv.padding)
        return torch.relu(y + (b + b2)[None, :, None, None])
    y = F.conv2d(x, r(conv.weight), None, conv.stride, conv.padding) * s[None, :, None, None] + b[None, :, None, None]
    if name.endswith("downsample.0"):
        return y
    if res is not None:
        y = y + res
    y = torch.relu(y)
    if pool:
        return model.backbone.maxpool(r(y))
    return y


def teacher_forced_report(model: HeatmapPoseRef, x: torch.Tensor, acts: dict, fused_head=None, fused_pool=False):
    """Compare every launch of the CUDA eval forward with ``layer_forward_fp32`` fed that launch's OWN stored inputs.
    ``acts`` = ``_ForwardPlan.export_acts()`` (NCHW fp32 CPU copies keyed by module path), ``x`` the fp32 image batch.
    Returns [(name, max|err| / max|ref|, fraction of elements outside the bar)] with the bars: bf16 outputs within ONE bf16
    rounding (2^-8 |ref| + 2e-3 max|ref| for fp32 accumulation order); fp32 heatmaps within 1e-4 of max (1e-3 through the
    fused deconv+final kernel, whose bf16 mid-tensor is rounded inside the kernel and may flip single ulps)."""
    out = []
    for name, a in acts.items():
        head = name == fused_head
        xin = a.get("x")
        if xin is None:                                      # stem: input = the bf16-rounded image
            xin = x.to(torch.bfloat16).float()
        want = layer_forward_fp32(model, name, xin, res=a.get("res"), x2=a.get("x2"),
                                  pool=(name == "backbone.conv1" and fused_pool), head=head)
        got = a["y"]
        assert got.shape == want.shape, (name, tuple(got.shape), tuple(want.shape))
        err = (got - want).abs()
        scale = want.abs().max().item()
        if head or name == "final_layer":
            bound = (1e-3 if head else 1e-4) * scale * torch.ones_like(want)
        else:
            bound = 2.0 ** -8 * want.abs() + 2e-3 * scale
        out.append((name, err.max().item() / max(scale, 1e-30), (err > bound).float().mean().item()))
    return out
