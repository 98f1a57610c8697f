"""One depthwise 7x7 + LayerNorm launch at a ConvNeXt stage shape (for ncu): python tools/run_dwconv.py C H W [B=256] [iters=5]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib  # noqa: E402

pm = importlib.import_module("2d-human-pose-estimation_b200.posemodel")
C, H, W = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
B = int(sys.argv[4]) if len(sys.argv) > 4 else 256
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 5
dev = torch.device("cuda:0")
x = torch.randn(B, H, W, C, device=dev).bfloat16()
y = torch.empty_like(x)
wt = torch.randn(49, C, device=dev) * 0.1
p = [torch.randn(C, device=dev) for _ in range(3)]
ts = []
for i in range(iters):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); pm._dwconv(x, wt, p[0], p[1], p[2], 1e-6, 7, 1, y); b.record(); torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
ts2, ts3 = [], []
for i in range(iters):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); pm._dwconv7_bias(x, wt, p[0], y); b.record(); torch.cuda.synchronize(); ts2.append(a.elapsed_time(b))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); pm._layernorm(y, p[1], p[2], 1e-6, y); b.record(); torch.cuda.synchronize(); ts3.append(a.elapsed_time(b))
fl = 2.0 * B * H * W * C * 49
print(f"  tiled dwconv7+bias: best {min(ts2) * 1e3:.1f} us ({fl / min(ts2) / 1e9:.2f} TFLOP/s fp32) + layernorm {min(ts3) * 1e3:.1f} us")
print(f"dwconv7+LN C={C} {H}x{W} B={B}: best {min(ts) * 1e3:.1f} us, {fl / min(ts) / 1e9:.2f} TFLOP/s fp32, "
      f"{2 * x.numel() * 2 / min(ts) / 1e6:.0f} GB/s (r+w)")
