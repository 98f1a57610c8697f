"""DINOv3Pose forward + decode + NMS timing (CUDA events) and a per-kernel breakdown -- development aid.
    python tools/bench_pose.py [size=small] [B=256] [H=256] [W=192]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpe_b200 as hpe  # noqa: E402

size = sys.argv[1] if len(sys.argv) > 1 else "small"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
H = int(sys.argv[3]) if len(sys.argv) > 3 else 256
W = int(sys.argv[4]) if len(sys.argv) > 4 else 192
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = hpe.DINOv3Pose(f"dinov3_convnext_{size}", nkpts=(17, 3), ncls=1, device=dev)
with torch.no_grad():
    for n, p in m.named_parameters():
        if n.endswith("gamma"):
            p.fill_(0.5)
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm2d):
            mod.running_var.uniform_(0.5, 1.5); mod.weight.uniform_(0.5, 1.5)
m.refresh()
m.eval()
x = [torch.rand(B, 3, H, W, device=dev) for _ in range(2)]


def timed(fn, n=10, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    ts = []
    for i in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(i); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


# 23.02 GFLOP/img for 'small' @256x192 (backbone 17.01); scale others by measured MACs later
gflop = {"tiny": None, "small": 23.02, "base": None}.get(size)
t_fwd = timed(lambda i: m(x[i & 1]))
print(f"{size} B={B} {H}x{W}: forward+decode {t_fwd:.3f} ms -> {B / t_fwd * 1e3:.0f} img/s" + (f", {gflop * B / t_fwd:.1f} TFLOP/s" if gflop else ""))
dec = m(x[0])
t_nms = timed(lambda i: hpe.pose_nms_raw(dec, 0.25, 0.45, nc=1, nkpts=17))
print(f"  NMS (conf 0.25): {t_nms:.3f} ms")
plan = next(iter(m._plans.values()))
print(f"  launches per forward: {plan.n_launches + 2}")
from torch.profiler import ProfilerActivity, profile  # noqa: E402
os.environ["HPE_B200_CUDA_GRAPH"] = "0"
m(x[0]); torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    m(x[0]); torch.cuda.synchronize()
rows = sorted(prof.key_averages(), key=lambda r: -r.device_time_total)[:16]
tot = sum(r.device_time_total for r in prof.key_averages())
print(f"  eager kernel time {tot / 1e3:.3f} ms")
for r in rows:
    print(f"  {r.device_time_total / 1e3:9.3f} ms  x{r.count:4d}  {r.key[:110]}")
