"""GPU parity of every conv-kernel flavour (through the C ABI) against a plain PyTorch fp32 reference of the same
op evaluated on the same bf16-rounded inputs/weights.  Tolerance: one bf16 rounding of the output
(|err| <= 2^-8 |ref| + 2e-3 * max|ref| covers the fp32 accumulation-order difference)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def r16(t):
    return t.to(torch.bfloat16).to(torch.float32)


def check(out, ref, what):
    out = out.float().cpu()
    tol = ref.abs() * 2 ** -8 + 2e-3 * ref.abs().max()
    bad = (out - ref).abs() > tol
    assert not bad.any(), f"{what}: {int(bad.sum())}/{bad.numel()} outside tolerance, max err {(out - ref).abs().max():.4g} (ref max {ref.abs().max():.4g})"


def nhwc(t):      # NCHW fp32 cpu -> NHWC bf16 cuda
    return t.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()


def nchw(t):      # NHWC cuda -> NCHW fp32 cpu
    return t.float().cpu().permute(0, 3, 1, 2)


def make(g, *shape, s=1.0):
    return r16(torch.randn(*shape, generator=g) * s)


@pytest.mark.parametrize("B,H,W,Cin,Cout,stride,act,res", [
    (8, 16, 12, 64, 256, 1, 1, True),      # BLOCK_N=256, residual + ReLU (bottleneck conv3)
    (8, 16, 12, 256, 64, 1, 1, False),     # BLOCK_N=64  (bottleneck conv1)
    (4, 16, 12, 512, 128, 1, 1, False),    # BLOCK_N=128, 8 k-blocks > ring depth
    (4, 32, 24, 256, 512, 2, 0, False),    # stride-2 downsample, no activation
    (3, 10, 7, 64, 64, 1, 2, False),       # ragged tiles + SiLU (ConvBlock act)
    (2, 8, 6, 1024, 2048, 1, 1, True),     # many N tiles, many k-blocks
    (3, 16, 8, 128, 256, 1, 1, True),      # cta_group::2 pair kernel with an ODD number of M tiles (3): the last peer tile is out of range
    (1, 16, 8, 64, 512, 1, 2, False),      # a single M tile: too few for a pair -> 1-CTA 256-wide kernel, SiLU
    (5, 10, 6, 256, 256, 2, 1, False),     # pair kernel, stride 2, ragged bricks
])
def test_conv1x1(hpe, cuda_dev, B, H, W, Cin, Cout, stride, act, res):
    from importlib import import_module
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(1)
    x = make(g, B, Cin, H, W); w = make(g, Cout, Cin, 1, 1, s=Cin ** -0.5)
    scale = torch.rand(Cout, generator=g) + 0.5; shift = torch.randn(Cout, generator=g) * 0.1
    Ho, Wo = H // stride, W // stride
    rs = make(g, B, Cout, Ho, Wo) if res else None
    ref = F.conv2d(x, w, stride=stride) * scale[None, :, None, None] + shift[None, :, None, None]
    if res:
        ref = ref + rs
    ref = torch.relu(ref) if act == 1 else (F.silu(ref) if act == 2 else ref)
    layer = conv.ConvLayer(conv.CONV_1x1, B, H, W, Cin, Cout, stride, act, res).load(w, scale, shift)
    y = layer(nhwc(x), nhwc(rs) if res else None)
    torch.cuda.synchronize()
    assert tuple(y.shape) == (B, Ho, Wo, Cout)
    check(nchw(y), ref, "conv1x1")


@pytest.mark.parametrize("B,H,W,Cin,Cin2,Cout,stride2", [
    (8, 16, 12, 64, 64, 256, 1),        # layer1.0: conv3 + stride-1 projection shortcut
    (4, 16, 12, 128, 256, 512, 2),      # layer2.0: shortcut reads the block input with stride 2
    (2, 8, 6, 512, 1024, 2048, 2),      # layer4.0: many k-blocks from both sources, 8 N tiles
    (3, 10, 6, 64, 72, 128, 2),         # ragged bricks, second source with a partial 64-channel block
])
def test_conv1x1_fused_projection_shortcut(hpe, cuda_dev, B, H, W, Cin, Cin2, Cout, stride2):
    """conv3 + bn3 + downsample(x) + add + ReLU of timm Bottleneck.forward (resnet.py:247-268) as ONE two-source GEMM;
    both BN scales are folded into the weights in fp32 before the bf16 rounding."""
    from importlib import import_module
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(11)
    x = make(g, B, Cin, H, W); w = torch.randn(Cout, Cin, 1, 1, generator=g) * Cin ** -0.5
    x2 = make(g, B, Cin2, H * stride2, W * stride2); w2 = torch.randn(Cout, Cin2, 1, 1, generator=g) * Cin2 ** -0.5
    scale = torch.rand(Cout, generator=g) + 0.5; shift = torch.randn(Cout, generator=g) * 0.1
    scale2 = torch.rand(Cout, generator=g) + 0.5; shift2 = torch.randn(Cout, generator=g) * 0.1
    ref = (F.conv2d(x, r16(w * scale.view(-1, 1, 1, 1))) + F.conv2d(x2, r16(w2 * scale2.view(-1, 1, 1, 1)), stride=stride2)
           + (shift + shift2)[None, :, None, None])
    ref = torch.relu(ref)
    layer = conv.ConvLayer(conv.CONV_1x1, B, H, W, Cin, Cout, 1, 1, False, Cin2=Cin2, stride2=stride2)
    layer.load2(w, scale, shift, w2, scale2, shift2)
    y = layer(nhwc(x), None, None, nhwc(x2))
    torch.cuda.synchronize()
    assert tuple(y.shape) == (B, H, W, Cout)
    check(nchw(y), ref, "conv1x1 + projection shortcut")
    with pytest.raises(RuntimeError, match="x2"):
        layer(nhwc(x))


@pytest.mark.parametrize("B,H,W,C,Cout,stride", [
    (4, 16, 12, 64, 64, 1), (4, 16, 12, 128, 128, 1), (4, 32, 24, 128, 128, 2), (2, 16, 12, 256, 256, 2),
    (3, 10, 6, 64, 64, 1), (2, 8, 6, 512, 512, 1),
    (3, 32, 24, 64, 64, 1), (2, 16, 8, 64, 64, 1), (2, 64, 48, 64, 64, 1), (5, 48, 40, 64, 64, 1),   # y-halo variant (conv_halo.cu)
    (1, 120, 56, 64, 64, 1), (2, 124, 44, 64, 64, 1), (160, 32, 8, 64, 64, 1),                        # ... bricks hanging over the bottom / right edge
    (4, 32, 24, 128, 128, 1), (3, 20, 16, 128, 128, 1), (2, 8, 8, 128, 128, 1), (37, 32, 24, 128, 128, 1), (1, 50, 40, 128, 128, 1),   # halo'd A operand (AH)
])
def test_conv3x3(hpe, cuda_dev, monkeypatch, B, H, W, C, Cout, stride):
    from importlib import import_module
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(2)
    x = make(g, B, C, H, W); w = make(g, Cout, C, 3, 3, s=(9 * C) ** -0.5)
    scale = torch.rand(Cout, generator=g) + 0.5; shift = torch.randn(Cout, generator=g) * 0.1
    ref = torch.relu(F.conv2d(x, w, stride=stride, padding=1) * scale[None, :, None, None] + shift[None, :, None, None])
    layer = conv.ConvLayer(conv.CONV_3x3, B, H, W, C, Cout, stride, 1, False).load(w, scale, shift)
    y = layer(nhwc(x))
    torch.cuda.synchronize()
    check(nchw(y), ref, "conv3x3")
    if C == 128 and Cout == 128 and stride == 1 and W % 8 == 0:
        # the halo'd-A variant (two {64, 10, 18} boxes per 8 x 16 brick, taps through shifted descriptors) adds the same products
        # in the same order as the generic kernel (one A brick per tap through the ring): bit-identical
        for mode in ("0", "1"):           # generic kernel / single-CTA form of the halo'd-A kernel (the default is its cta_group::2 form)
            monkeypatch.setenv("HPE_B200_HALO128", mode)
            plain = conv.ConvLayer(conv.CONV_3x3, B, H, W, C, Cout, stride, 1, False).load(w, scale, shift)
            assert torch.equal(y, plain(nhwc(x))), mode
        assert layer.set_schedule(True, 0, 0, 0) is True and torch.equal(y, layer(nhwc(x)))


@pytest.mark.parametrize("kind,B,H,W,Cin,Cout", [("3x3", 256, 8, 6, 512, 512), ("1x1", 256, 8, 6, 2048, 512), ("3x3", 9, 8, 6, 512, 512)])
def test_wave_quantisation_tiles_and_schedule_are_bit_identical(hpe, cuda_dev, monkeypatch, kind, B, H, W, Cin, Cout):
    """Layer4's conv1 / conv2 at the benchmark batch: 96 cta_group::2 work items of 256 columns on 74 clusters, run as 144 items of
    192 columns instead (HPE_B200_WAVE192, the third N tile a third full) -- and any layer walked last tile to first with L2
    hints (hpe_conv_set_schedule): the same MMAs per output element, so the same bits as the 256-wide forward walk; parity
    against fp32 PyTorch on top."""
    from importlib import import_module
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(11)
    k = 3 if kind == "3x3" else 1
    x = make(g, B, Cin, H, W); w = make(g, Cout, Cin, k, k, s=(k * k * Cin) ** -0.5)
    scale = torch.rand(Cout, generator=g) + 0.5; shift = torch.randn(Cout, generator=g) * 0.1
    xin = nhwc(x)
    outs = []
    for wave in ("1", "0"):
        monkeypatch.setenv("HPE_B200_WAVE192", wave)
        layer = conv.ConvLayer(conv.CONV_3x3 if k == 3 else conv.CONV_1x1, B, H, W, Cin, Cout, 1, 1, False).load(w, scale, shift)
        outs.append(layer(xin).clone())
        assert layer.set_schedule(True, 2, 0, 1) is True                 # reversed walk, evict-last stores, evict-first input loads
        outs.append(layer(xin).clone())
    torch.cuda.synchronize()
    for o in outs[1:]:
        assert torch.equal(outs[0], o)
    if B <= 16:
        ref = torch.relu(F.conv2d(x, w, padding=k // 2) * scale[None, :, None, None] + shift[None, :, None, None])
        check(nchw(outs[0]), ref, "wave192 conv")
    with pytest.raises(RuntimeError):
        layer.set_schedule(False, 3, 0, 0)                               # hints are 0 | 1 | 2


@pytest.mark.parametrize("B,H,W", [(4, 64, 48), (2, 256, 192), (3, 36, 20), (3, 96, 80), (2, 240, 112), (1, 248, 90), (200, 64, 32)])
def test_stem_and_maxpool(hpe, cuda_dev, B, H, W):
    from importlib import import_module
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(3)
    img = torch.rand(B, 3, H, W, generator=g)
    w = make(g, 64, 3, 7, 7, s=0.1)
    scale = torch.rand(64, generator=g) + 0.5; shift = torch.randn(64, generator=g) * 0.1
    ref = torch.relu(F.conv2d(r16(img), w, stride=2, padding=3) * scale[None, :, None, None] + shift[None, :, None, None])
    s2d = conv.preprocess(img.cuda())
    layer = conv.ConvLayer(conv.CONV_STEM7x7, B, H, W, 3, 64, 2, 1, False).load(w, scale, shift)
    y = layer(s2d)
    torch.cuda.synchronize()
    check(nchw(y), ref, "stem")
    p = conv.maxpool3x3s2(y)
    pref = F.max_pool2d(nchw(y), 3, 2, 1)
    assert torch.equal(nchw(p), pref)


@pytest.mark.parametrize("B,H,W", [(4, 64, 48), (2, 256, 192), (3, 36, 20), (3, 96, 80), (1, 38, 44), (5, 128, 96), (2, 26, 14),
                                   (150, 64, 64)])
def test_stem_with_fused_maxpool(hpe, cuda_dev, B, H, W):
    """hpe_conv_desc.pool: MaxPool2d(3, 2, 1) taken in the stem's epilogue == stem launch followed by hpe_maxpool3x3s2, bit for bit
    (odd/even conv extents, bricks hanging over every image edge, more tiles than CTAs)."""
    from importlib import import_module
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(31)
    img = torch.rand(B, 3, H, W, generator=g) - 0.3
    w = make(g, 64, 3, 7, 7, s=0.1)
    scale = torch.rand(64, generator=g) + 0.5; shift = torch.randn(64, generator=g) * 0.1
    s2d = conv.preprocess(img.cuda())
    plain = conv.ConvLayer(conv.CONV_STEM7x7, B, H, W, 3, 64, 2, 1, False).load(w, scale, shift)
    want = conv.maxpool3x3s2(plain(s2d))
    fused = conv.ConvLayer(conv.CONV_STEM7x7, B, H, W, 3, 64, 2, 1, False, pool=1).load(w, scale, shift)
    assert fused.out_shape() == tuple(want.shape)
    out = torch.full(want.shape, float("nan"), dtype=torch.bfloat16, device="cuda")
    fused(s2d, None, out)
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    ref = F.max_pool2d(torch.relu(F.conv2d(r16(img), w, stride=2, padding=3) * scale[None, :, None, None] + shift[None, :, None, None]), 3, 2, 1)
    check(nchw(out), ref, "stem+pool")


def test_fused_maxpool_rejected_off_the_relu_stem(hpe, cuda_dev):
    from importlib import import_module
    conv = import_module("2d-human-pose-estimation_b200.conv")
    with pytest.raises(Exception):
        conv.ConvLayer(conv.CONV_3x3, 2, 16, 16, 64, 64, 1, 1, False, pool=1)
    with pytest.raises(Exception):
        conv.ConvLayer(conv.CONV_STEM7x7, 2, 64, 64, 3, 64, 2, 0, False, pool=1)      # no activation


@pytest.mark.parametrize("B,H,W,Cin", [(4, 8, 6, 256), (2, 8, 6, 2048), (3, 5, 7, 256), (2, 32, 24, 256), (3, 8, 8, 128)])
def test_deconv4x4s2(hpe, cuda_dev, B, H, W, Cin):
    from importlib import import_module
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(4)
    x = make(g, B, Cin, H, W); w = make(g, Cin, 256, 4, 4, s=(4 * Cin) ** -0.5)
    scale = torch.rand(256, generator=g) + 0.5; shift = torch.randn(256, generator=g) * 0.1
    ref = torch.relu(F.conv_transpose2d(x, w, stride=2, padding=1) * scale[None, :, None, None] + shift[None, :, None, None])
    layer = conv.ConvLayer(conv.CONV_DECONV4, B, H, W, Cin, 256, 2, 1, False).load(w, scale, shift)
    y = layer(nhwc(x))
    torch.cuda.synchronize()
    assert tuple(y.shape) == (B, 2 * H, 2 * W, 256)
    check(nchw(y), ref, "deconv")


@pytest.mark.parametrize("B,H,W,Cin,K", [(4, 32, 24, 256, 17), (3, 10, 6, 256, 17), (4, 8, 6, 2048, 32), (5, 16, 12, 64, 1)])
def test_deconv_with_fused_final_head(hpe, cuda_dev, B, H, W, Cin, K):
    """deconv (+BN+ReLU) and the K-joint final 1x1 in ONE kernel (desc.head2): the bf16 deconv output never leaves shared memory.
    Reference = the two separate layers' math: fp32 deconv -> BN -> ReLU -> round to bf16 -> 1x1 conv + bias."""
    from importlib import import_module
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(12)
    x = make(g, B, Cin, H, W); w = make(g, Cin, 256, 4, 4, s=(4 * Cin) ** -0.5)
    scale = torch.rand(256, generator=g) + 0.5; shift = torch.randn(256, generator=g) * 0.1
    wf = make(g, K, 256, 1, 1, s=1 / 16); bias = torch.randn(K, generator=g) * 0.1
    mid = torch.relu(F.conv_transpose2d(x, w, stride=2, padding=1) * scale[None, :, None, None] + shift[None, :, None, None])
    ref = F.conv2d(r16(mid), wf, bias)
    fused = conv.ConvLayer(conv.CONV_DECONV4, B, H, W, Cin, 256, 2, 1, False, head2=K).load(w, scale, shift).load_head2(wf, bias)
    hm = fused(nhwc(x))
    torch.cuda.synchronize()
    assert tuple(hm.shape) == (B, K, 2 * H, 2 * W) and hm.dtype == torch.float32
    # the separate layers through the same library: identical bf16 intermediate, so only fp32 accumulation order differs
    d3 = conv.ConvLayer(conv.CONV_DECONV4, B, H, W, Cin, 256, 2, 1, False).load(w, scale, shift)
    fin = conv.ConvLayer(conv.CONV_FINAL, B, 2 * H, 2 * W, 256, K, 1, 0, False).load(wf, torch.ones(K), bias)
    hm2 = fin(d3(nhwc(x)))
    torch.cuda.synchronize()
    assert (hm - hm2).abs().max() <= 1e-5 * hm2.abs().max() + 1e-6
    # vs PyTorch: the bf16 rounding of the intermediate may differ by one ulp where the fp32 accumulation order differs
    assert (hm.cpu() - ref).abs().max() <= 2e-3 * ref.abs().max() + 1e-4


@pytest.mark.parametrize("B,H,W,K", [(4, 64, 48, 17), (2, 64, 48, 133), (3, 10, 7, 17), (2, 96, 72, 17)])
def test_final_1x1_heatmaps(hpe, cuda_dev, B, H, W, K):
    from importlib import import_module
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(5)
    x = make(g, B, 256, H, W); w = make(g, K, 256, 1, 1, s=1 / 16)
    bias = torch.randn(K, generator=g) * 0.1
    ref = F.conv2d(x, w, bias)
    layer = conv.ConvLayer(conv.CONV_FINAL, B, H, W, 256, K, 1, 0, False).load(w, torch.ones(K), bias)
    y = layer(nhwc(x))
    torch.cuda.synchronize()
    assert tuple(y.shape) == (B, K, H, W) and y.dtype == torch.float32
    assert (y.cpu() - ref).abs().max() <= 1e-4 * ref.abs().max() + 1e-5     # fp32 out: only accumulation order differs


def test_conv_rejects_unsupported(hpe, cuda_dev):
    from importlib import import_module
    conv = import_module("2d-human-pose-estimation_b200.conv")
    with pytest.raises(RuntimeError, match="fused head"):      # a single M tile cannot form a CTA pair
        conv.ConvLayer(conv.CONV_DECONV4, 1, 8, 6, 256, 256, 2, 1, False, head2=17)
    with pytest.raises(RuntimeError, match="32 joints"):
        conv.ConvLayer(conv.CONV_DECONV4, 4, 8, 6, 256, 256, 2, 1, False, head2=133)
    with pytest.raises(RuntimeError, match="multiple of 8"):
        conv.ConvLayer(conv.CONV_1x1, 1, 8, 8, 44, 64)
    with pytest.raises(RuntimeError, match="stride"):
        conv.ConvLayer(conv.CONV_3x3, 1, 8, 8, 64, 64, stride=3)
