"""time the stem (plain / pooled) and layer1 conv2 on the y-halo kernel"""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpe_b200  # noqa
conv = importlib.import_module("2d-human-pose-estimation_b200.conv")
B = 256
img = torch.rand(B, 3, 256, 192, device="cuda")
s2d = conv.preprocess(img)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timeit(layer, x, tag):
    y = layer.alloc_out(); ts = []
    for i in range(7):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); layer(x, None, y); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    print(f"{tag}: min {min(ts)*1e3:.1f} us median {sorted(ts)[3]*1e3:.1f} us", flush=True)
for pool in (0, 1):
    l = conv.ConvLayer(conv.CONV_STEM7x7, B, 256, 192, 3, 64, 2, 1, False, pool=pool).load(torch.randn(64, 3, 7, 7) * 0.1, torch.rand(64) + 0.5, torch.randn(64) * 0.1)
    timeit(l, s2d, f"stem pool={pool}")
x = torch.randn(B, 64, 48, 64, device="cuda").to(torch.bfloat16)
l = conv.ConvLayer(conv.CONV_3x3, B, 64, 48, 64, 64, 1, 1, False).load(torch.randn(64, 64, 3, 3) * 0.05, torch.rand(64) + 0.5, torch.randn(64) * 0.1)
timeit(l, x, f"l1conv2")
