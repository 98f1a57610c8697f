"""DINOv3Pose training step (frozen backbone; ComputeLoss; flat-buffer AdamW + clip) timing + kernel breakdown -- development aid.
    python tools/bench_posetrain.py [size=small] [B=64] [H=256] [W=192]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpe_b200 as hpe  # noqa: E402

size = sys.argv[1] if len(sys.argv) > 1 else "small"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
H = int(sys.argv[3]) if len(sys.argv) > 3 else 256
W = int(sys.argv[4]) if len(sys.argv) > 4 else 192
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = hpe.DINOv3Pose(f"dinov3_convnext_{size}", nkpts=(17, 3), ncls=1, device=dev)
gen = torch.Generator().manual_seed(1)
x = [torch.randint(0, 256, (B, 3, H, W), generator=gen, dtype=torch.uint8).to(dev) for _ in range(2)]
tg = torch.zeros(B, 6 + 17 * 3)
tg[:, 0] = torch.arange(B); tg[:, 2:4] = 0.3 + 0.4 * torch.rand(B, 2, generator=gen); tg[:, 4:6] = 0.3
kp = torch.rand(B, 17, 3, generator=gen); kp[..., :2] = 0.2 + 0.6 * kp[..., :2]; kp[..., 2] = 2.0
tg[:, 6:] = kp.reshape(B, -1)
tg = tg.to(dev)
step = hpe.PoseTrainStep(m, lr=1e-4, weight_decay=1e-4, max_grad_norm=10.0, cuda_graph=os.environ.get("POSE_GRAPH", "1") == "1")
for i in range(3):
    loss, _ = step(x[i & 1], tg)
torch.cuda.synchronize()
ts = []
for i in range(10):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); loss, _ = step(x[i & 1], tg); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
t = sorted(ts)[len(ts) // 2]
n_tr = sum(p.numel() for p in m.parameters() if p.requires_grad)
print(f"DINOv3Pose({size}) training step B={B} {H}x{W}: {t:.2f} ms -> {B / t * 1e3:.0f} img/s; trainable parameters {n_tr / 1e6:.1f} M; loss {float(loss):.4f}")
from torch.profiler import ProfilerActivity, profile  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step(x[0], tg); torch.cuda.synchronize()
rows = sorted(prof.key_averages(), key=lambda r: -r.device_time_total)[:22]
tot = sum(r.device_time_total for r in prof.key_averages())
print(f"  device kernel time {tot / 1e3:.3f} ms over {sum(r.count for r in prof.key_averages())} launches")
for r in rows:
    print(f"  {r.device_time_total / 1e3:9.3f} ms  x{r.count:4d}  {r.key[:120]}")
