"""CUDA-event timing of the image layout kernels (fp32 and uint8 input), L2 flushed between launches -- development aid."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpe_b200 as hpe  # noqa: E402
from importlib import import_module  # noqa: E402
pre = import_module("2d-human-pose-estimation_b200.conv").preprocess

dev = torch.device("cuda:0")
B, H, W = 256, 256, 192
xf = torch.rand(B, 3, H, W, device=dev)
xu = torch.randint(0, 256, (B, 3, H, W), dtype=torch.uint8, device=dev)
out = pre(xf)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for name, x in (("fp32", xf), ("uint8", xu)):
    ts = []
    for i in range(12):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); pre(x, out); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = sorted(ts[2:])
    print(f"preprocess {name}: median {ts[len(ts) // 2]:.1f} us, min {ts[0]:.1f} us")
ref = pre(xu.float() / 255)
assert torch.equal(pre(xu), ref), "uint8 path differs from float(x) / 255"
print("uint8 == float / 255: bit-identical")
