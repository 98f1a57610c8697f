"""Phase timing of one training step (CUDA events): forward / loss / backward / sgd, plus per-kernel-class sums."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpe_b200 as hpe
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
H, W, K = 256, 192, 17
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = hpe.HeatmapPose(K, "resnet50", pretrained=False).to(dev).train()
step = hpe.TrainStep(model, lr=1e-3, cuda_graph=False)
x = torch.rand(B, 3, H, W, device=dev)
joints = (torch.rand(B, K, 3, device=dev) * torch.tensor([W, H, 1.0], device=dev)).contiguous()
for _ in range(3):
    step(x, joints)
torch.cuda.synchronize()
plan = model._train_plan(x)
ev = lambda: torch.cuda.Event(enable_timing=True)
def timed(fn, n=5):
    ts = []
    for _ in range(n):
        a, b = ev(), ev(); a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]
print(f"B={B}: full step {timed(lambda: step(x, joints)):.3f} ms")
print(f"  repack (pack fwd+dgrad) {timed(lambda: plan.repack()):.3f} ms")
print(f"  forward {timed(lambda: plan.forward(x)):.3f} ms")
print(f"  backward {timed(lambda: plan.backward(step.dhm, step.grads)):.3f} ms")
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step(x, joints); torch.cuda.synchronize()
rows = sorted(prof.key_averages(), key=lambda r: -r.device_time_total)[:32]
for r in rows:
    print(f"  {r.device_time_total/1e3:9.3f} ms  x{r.count:4d}  {r.key[:90]}")
