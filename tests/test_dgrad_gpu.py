"""GPU parity of the backward-data convolutions (same tcgen05 kernel, re-packed weights, through the C ABI) against
PyTorch autograd in fp32 on the same bf16-rounded dz / weights.  Tolerance: one bf16 rounding of the output."""
from importlib import import_module

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def r16(t):
    return t.to(torch.bfloat16).to(torch.float32)


def nhwc16(t):
    return t.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()


def nchw(t):
    return t.float().cpu().permute(0, 3, 1, 2)


def check(out, ref, what):
    tol = ref.abs() * 2 ** -8 + 2e-3 * ref.abs().max()
    bad = (out - ref).abs() > tol
    assert not bad.any(), f"{what}: {int(bad.sum())}/{bad.numel()} outside tolerance, max err {(out - ref).abs().max():.4g} (ref max {ref.abs().max():.4g})"


def autograd_dx(fn, x_shape, dz):
    x = torch.zeros(x_shape, requires_grad=True)
    fn(x).backward(dz)
    return x.grad


@pytest.mark.parametrize("kind,B,H,W,Cin,Cout,stride,acc", [
    ("1x1", 8, 16, 12, 64, 256, 1, True),
    ("1x1", 4, 16, 12, 512, 128, 1, False),
    ("1x1", 4, 32, 24, 256, 512, 2, False),     # downsample: even pixels only
    ("1x1", 3, 10, 8, 64, 64, 2, False),
    ("3x3", 4, 16, 12, 64, 64, 1, False),
    ("3x3", 2, 8, 6, 512, 512, 1, False),
    ("3x3", 4, 32, 24, 128, 128, 2, False),     # four parity classes, 1/2/2/4 taps
    ("3x3", 2, 16, 12, 256, 256, 2, False),
    ("3x3", 3, 12, 10, 64, 128, 2, False),
    ("deconv", 4, 8, 6, 512, 256, 2, False),
    ("deconv", 2, 16, 12, 256, 256, 2, False),
    ("deconv", 3, 5, 3, 128, 64, 2, False),
])
def test_dgrad(hpe, cuda_dev, kind, B, H, W, Cin, Cout, stride, acc):
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(21)
    if kind == "deconv":
        w = r16(torch.randn(Cin, Cout, 4, 4, generator=g) * (4 * Cout) ** -0.5)
        Ho, Wo = 2 * H, 2 * W
        fn = lambda x: F.conv_transpose2d(x, w, stride=2, padding=1)
        k = conv.CONV_DECONV4
    elif kind == "3x3":
        w = r16(torch.randn(Cout, Cin, 3, 3, generator=g) * (9 * Cout) ** -0.5)
        Ho, Wo = H // stride, W // stride
        fn = lambda x: F.conv2d(x, w, stride=stride, padding=1)
        k = conv.CONV_3x3
    else:
        w = r16(torch.randn(Cout, Cin, 1, 1, generator=g) * Cout ** -0.5)
        Ho, Wo = H // stride, W // stride
        fn = lambda x: F.conv2d(x, w, stride=stride)
        k = conv.CONV_1x1
    dz = r16(torch.randn(B, Cout, Ho, Wo, generator=g))
    ref = autograd_dx(fn, (B, Cin, H, W), dz)
    res = r16(torch.randn(B, Cin, H, W, generator=g)) if acc else None
    if acc:
        ref = ref + res
    layer = conv.DgradLayer(k, B, H, W, Cin, Cout, stride, accumulate=acc).load(w)
    dx = layer(nhwc16(dz), nhwc16(res) if acc else None)
    torch.cuda.synchronize()
    assert tuple(dx.shape) == (B, H, W, Cin)
    check(nchw(dx), ref, f"dgrad {kind} s{stride}")


def test_dgrad_final(hpe, cuda_dev):
    conv = import_module("2d-human-pose-estimation_b200.conv")
    g = torch.Generator().manual_seed(22)
    B, H, W, K = 4, 16, 12, 17
    w = r16(torch.randn(K, 256, 1, 1, generator=g) * 0.1)
    dz = r16(torch.randn(B, K, H, W, generator=g))
    ref = autograd_dx(lambda x: F.conv2d(x, w), (B, 256, H, W), dz)
    dzp = torch.zeros(B, H, W, 64, dtype=torch.bfloat16, device="cuda")
    dzp[..., :K] = dz.permute(0, 2, 3, 1).to(torch.bfloat16).cuda()
    layer = conv.DgradLayer(conv.CONV_FINAL, B, H, W, 256, K, 1).load(w)
    dx = layer(dzp)
    torch.cuda.synchronize()
    check(nchw(dx), ref, "dgrad final")
