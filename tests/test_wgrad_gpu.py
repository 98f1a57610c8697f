"""GPU parity of the tcgen05 weight-gradient GEMM (through the C ABI) against PyTorch autograd in fp32 on the same
bf16-rounded activations / output gradients.  The kernel accumulates in fp32 (TMEM + fp32 atomics across the
split-K partitions), so the tolerance is an fp32 summation-order tolerance: |err| <= 2e-4 * max|ref|."""
from importlib import import_module

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def r16(t):
    return t.to(torch.bfloat16).to(torch.float32)


def nhwc16(t):
    return t.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()


def check(dw, ref, what):
    dw = dw.float().cpu()
    err = (dw - ref).abs().max().item()
    tol = 2e-4 * ref.abs().max().item() + 1e-6
    assert err <= tol, f"{what}: max err {err:.4g} > {tol:.4g} (ref max {ref.abs().max():.4g})"


def autograd_dw(fn, w, dz):
    w = w.clone().requires_grad_(True)
    fn(w).backward(dz)
    return w.grad


@pytest.mark.parametrize("B,H,W,Cin,Cout,stride", [
    (8, 16, 12, 256, 64, 1),      # left = x (4 blocks), N = 64
    (8, 16, 12, 64, 256, 1),      # swapped: left = dz
    (4, 16, 12, 64, 64, 1),       # single left block (duplicate half tile)
    (4, 32, 24, 256, 512, 2),     # stride-2 downsample
    (3, 10, 7, 128, 128, 1),      # ragged bricks
    (2, 8, 6, 2048, 512, 1),      # many M tiles, 2 N tiles
    (64, 16, 12, 256, 1024, 1),   # split-K over many bricks, swapped
])
def test_wgrad_1x1(hpe, cuda_dev, B, H, W, Cin, Cout, stride):
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(11)
    x = r16(torch.randn(B, Cin, H, W, generator=g))
    dz = r16(torch.randn(B, Cout, H // stride, W // stride, generator=g))
    w = torch.zeros(Cout, Cin, 1, 1)
    ref = autograd_dw(lambda w_: F.conv2d(x, w_, stride=stride), w, dz)
    layer = conv.WgradLayer(conv.CONV_1x1, B, H, W, Cin, Cout, stride)
    dw = torch.empty(Cout, Cin, 1, 1, device="cuda")
    layer(nhwc16(x), nhwc16(dz), dw)
    torch.cuda.synchronize()
    check(dw, ref, "wgrad1x1")
    layer(nhwc16(x), nhwc16(dz), dw, accumulate=True)
    torch.cuda.synchronize()
    check(dw, 2 * ref, "wgrad1x1 accumulate")


@pytest.mark.parametrize("B,H,W,Cin,Cout,stride", [
    (4, 16, 12, 64, 64, 1), (4, 16, 12, 128, 128, 1), (4, 32, 24, 128, 128, 2), (2, 16, 12, 256, 256, 2),
    (3, 10, 6, 64, 64, 1), (2, 8, 6, 512, 512, 1),
])
def test_wgrad_3x3(hpe, cuda_dev, B, H, W, Cin, Cout, stride):
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(12)
    x = r16(torch.randn(B, Cin, H, W, generator=g))
    dz = r16(torch.randn(B, Cout, H // stride, W // stride, generator=g))
    ref = autograd_dw(lambda w_: F.conv2d(x, w_, stride=stride, padding=1), torch.zeros(Cout, Cin, 3, 3), dz)
    layer = conv.WgradLayer(conv.CONV_3x3, B, H, W, Cin, Cout, stride)
    dw = torch.empty(Cout, Cin, 3, 3, device="cuda")
    layer(nhwc16(x), nhwc16(dz), dw)
    torch.cuda.synchronize()
    check(dw, ref, "wgrad3x3")


@pytest.mark.parametrize("B,H,W", [(4, 64, 48), (3, 36, 20)])
def test_wgrad_stem(hpe, cuda_dev, B, H, W):
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(13)
    x = torch.rand(B, 3, H, W, generator=g)
    dz = r16(torch.randn(B, 64, H // 2, W // 2, generator=g))
    ref = autograd_dw(lambda w_: F.conv2d(r16(x), w_, stride=2, padding=3), torch.zeros(64, 3, 7, 7), dz)
    s2d = conv.preprocess(x.cuda())
    layer = conv.WgradLayer(conv.CONV_STEM7x7, B, H, W, 3, 64, 2)
    dw = torch.empty(64, 3, 7, 7, device="cuda")
    layer(s2d, nhwc16(dz), dw)
    torch.cuda.synchronize()
    check(dw, ref, "wgrad stem")


@pytest.mark.parametrize("B,H,W,Cin,Cout", [(4, 8, 6, 512, 256), (2, 16, 12, 256, 256), (3, 5, 3, 128, 64)])
def test_wgrad_deconv(hpe, cuda_dev, B, H, W, Cin, Cout):
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(14)
    x = r16(torch.randn(B, Cin, H, W, generator=g))
    dz = r16(torch.randn(B, Cout, 2 * H, 2 * W, generator=g))
    ref = autograd_dw(lambda w_: F.conv_transpose2d(x, w_, stride=2, padding=1), torch.zeros(Cin, Cout, 4, 4), dz)
    layer = conv.WgradLayer(conv.CONV_DECONV4, B, H, W, Cin, Cout, 2)
    dw = torch.empty(Cin, Cout, 4, 4, device="cuda")
    layer(nhwc16(x), nhwc16(dz), dw)
    torch.cuda.synchronize()
    check(dw, ref, "wgrad deconv")


@pytest.mark.parametrize("B,H,W,K", [(4, 16, 12, 17), (2, 64, 48, 17)])
def test_wgrad_final(hpe, cuda_dev, B, H, W, K):
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(15)
    x = r16(torch.randn(B, 256, H, W, generator=g))
    dz = r16(torch.randn(B, K, H, W, generator=g))
    ref = autograd_dw(lambda w_: F.conv2d(x, w_), torch.zeros(K, 256, 1, 1), dz)
    dzp = torch.zeros(B, H, W, 64, dtype=torch.bfloat16, device="cuda")
    dzp[..., :K] = dz.permute(0, 2, 3, 1).to(torch.bfloat16).cuda()
    layer = conv.WgradLayer(conv.CONV_FINAL, B, H, W, 256, K, 1)
    dw = torch.empty(K, 256, 1, 1, device="cuda")
    layer(nhwc16(x), dzp, dw)
    torch.cuda.synchronize()
    check(dw, ref, "wgrad final")
