"""Experiment: two half-batch forward plans on two streams, each capped to half of the SMs, against one full-batch plan.
    python tools/exp_dual.py [B] [iters]
HBM-bound layers of one half-batch can overlap tensor-bound layers of the other only if the two run on disjoint SMs
(every conv kernel is persistent with one 200 KB CTA per SM)."""
import importlib
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpe_b200  # noqa: F401,E402

models = importlib.import_module("2d-human-pose-estimation_b200.models")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 30
dev = torch.device("cuda:0")
torch.manual_seed(0)


def make():
    m = models.HeatmapPose(17, "resnet50", pretrained=False)
    return m.to(dev).eval()


def timeit(fn, n):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


x = torch.rand(B, 3, 256, 192, device=dev)
full = make()
t_full = timeit(lambda: full.forward_static(x), iters)
print(f"one plan, B={B}, all SMs: {t_full:.3f} ms/step  ({B / t_full:.1f} k img/s)")

for ways, cap in ((2, 74), (2, 148), (2, 96), (4, 37), (2, 64), (3, 50)):
    if B % ways:
        continue
    os.environ["HPE_B200_MAX_CTAS"] = str(cap)
    ms, xs, ss = [], [], []
    for i in range(ways):
        m = make()
        m.load_state_dict(full.state_dict())
        ms.append(m)
        xs.append(x[i * (B // ways):(i + 1) * (B // ways)].contiguous())
        ss.append(torch.cuda.Stream())
    cur = torch.cuda.current_stream()

    def step():
        for m, xi, s in zip(ms, xs, ss):
            s.wait_stream(cur)
            with torch.cuda.stream(s):
                m.forward_static(xi)
        for s in ss:
            cur.wait_stream(s)
    t = timeit(step, iters)
    print(f"{ways} plans of B={B // ways} on {ways} streams, {cap} CTAs each: {t:.3f} ms/step  ({B / t:.1f} k img/s)  x{t_full / t:.3f}")
    os.environ.pop("HPE_B200_MAX_CTAS")
    del ms
