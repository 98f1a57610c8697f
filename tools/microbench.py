"""Kernel-level micro-benchmarks (CUDA events, inputs larger than L2 or L2 flushed) -- development aid.
    python tools/microbench.py decode|loss
"""
import json
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpe_b200 as hpe  # noqa: E402

PEAK = 6468.9
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def timeit(fn, iters=20, warm=3, flush=None):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def decode():
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for dtype in (torch.float32, torch.bfloat16):
        for (B, K, H, W) in [(256, 17, 64, 48), (1024, 17, 64, 48), (16384, 17, 64, 48), (65536, 17, 64, 48),
                             (4096, 17, 96, 72), (1024, 133, 64, 48), (1024, 133, 128, 96)]:
            hm = torch.rand(B, K, H, W, device=dev).to(dtype)
            M = torch.rand(B, 2, 3, device=dev)
            med, best = timeit(lambda: hpe.get_final_preds_affine(hm, M, return_idx=True), flush=flush)
            nbytes = hm.numel() * hm.element_size() + B * 24 + B * K * 16
            print(f"decode {str(dtype)[6:]:9s} B={B:6d} K={K:3d} {H}x{W}: median {med*1e3:8.1f} us  best {best*1e3:8.1f} us "
                  f"-> {nbytes/med/1e6:7.1f} GB/s ({nbytes/med/1e6/PEAK*100:5.1f}% of measured {PEAK})", flush=True)
            del hm


def loss():
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    crit = hpe.JointsMSELoss()
    for dtype in (torch.float32, torch.bfloat16):
        for (B, K, H, W) in [(256, 17, 64, 48), (4096, 17, 64, 48), (16384, 17, 64, 48)]:
            pred = torch.rand(B, K, H, W, device=dev).to(dtype)
            joints = torch.rand(B, K, 3, device=dev) * torch.tensor([192., 256., 1.], device=dev) + torch.tensor([0., 0., .5], device=dev)
            med, best = timeit(lambda: crit(pred, joints, None), flush=flush)
            nbytes = pred.numel() * pred.element_size() + joints.numel() * 4
            print(f"loss-fwd {str(dtype)[6:]:9s} B={B:6d}: median {med*1e3:8.1f} us -> {nbytes/med/1e6:7.1f} GB/s ({nbytes/med/1e6/PEAK*100:5.1f}%)", flush=True)
            pg = pred.clone().requires_grad_(True)
            med, best = timeit(lambda: crit(pg, joints, None), flush=flush)
            print(f"loss-fwd+bwd {str(dtype)[6:]:9s} B={B:6d}: median {med*1e3:8.1f} us -> {2*nbytes/med/1e6:7.1f} GB/s ({2*nbytes/med/1e6/PEAK*100:5.1f}%)", flush=True)


def forward():
    import importlib
    models = importlib.import_module("2d-human-pose-estimation_b200.models")
    dev = torch.device("cuda:0")
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    H, W = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (256, 192)
    torch.manual_seed(0)
    m = models.HeatmapPose(17, "resnet50", False)
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm2d):
            mod.weight.data.uniform_(0.5, 1.5); mod.running_var.uniform_(0.5, 1.5)
    m = m.to(dev).eval()
    x = torch.rand(B, 3, H, W, device=dev)
    med, best = timeit(lambda: m(x), iters=10)
    flops = (8.008 + 2.846) * 1e9 * B * (H * W) / (256 * 192)
    print(f"forward B={B} {H}x{W}: median {med:.3f} ms best {best:.3f} ms -> {B/med*1e3:.0f} img/s, {flops/med/1e9:.1f} TFLOP/s", flush=True)
    plan = next(iter(m._plans.values()))
    # per-layer eager timing (L2 not flushed between layers: this is how they run in the step)
    rows = []
    for layer, src, res, dst, src2, _dst2 in plan.ops:
        name = [k for k, v in plan.named.items() if v is layer][0]
        fn = lambda: layer(plan.bufs[src], plan.bufs[res] if res else None, plan.bufs[dst], plan.bufs[src2] if src2 else None)
        med, best = timeit(fn, iters=5, warm=1)
        d = layer.desc
        taps = {0: 1, 1: 9, 2: 49 / 3 * 3, 3: 4, 4: 1}[d.kind]
        macs = layer.B * layer.Ho * layer.Wo * d.Cout * (d.Cin * taps + d.Cin2)
        byts = (plan.bufs[src2].numel() * 2 // (d.stride2 ** 2) if src2 else 0) + plan.bufs[src].numel() * 2 + plan.bufs[dst].numel() * plan.bufs[dst].element_size() + (plan.bufs[res].numel() * 2 if res else 0)
        rows.append((name, med, 2 * macs / med / 1e9, byts / med / 1e6))
        print(f"  {name:32s} k{d.kind} {d.Cin:4d}{'+%d' % d.Cin2 if d.Cin2 else ''}->{d.Cout:4d} s{d.stride} out {layer.Ho}x{layer.Wo}: {med*1e3:8.1f} us  {2*macs/med/1e9:7.1f} TFLOP/s  {byts/med/1e6:7.0f} GB/s", flush=True)
    print(f"  sum of layers {sum(r[1] for r in rows):.3f} ms")


if __name__ == "__main__":
    {"decode": decode, "loss": loss, "forward": forward}[sys.argv[1]]()
