"""Localise forward deviations of the training-mode DINOv3Pose path: per-module relative error vs the oracle (train mode)."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpe_b200 as hpe
from oracle import dinov3pose_ref as dr
pt = importlib.import_module("2d-human-pose-estimation_b200.posetrain")
dev = torch.device("cuda:0")
ref = dr.build_synthetic("tiny", nkpts=(17, 3), ncls=1, seed=3); ref.train()
m = hpe.DINOv3Pose("dinov3_convnext_tiny", nkpts=(17, 3), ncls=1, device=dev); m.load_state_dict(ref.state_dict()); m.train()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
x = torch.rand(B, 3, 128, 96, generator=torch.Generator().manual_seed(4))
acts = {}
def hook(name):
    def f(mod, inp, out): acts[name] = out.detach() if torch.is_tensor(out) else [o.detach() for o in out]
    return f
for name, mod in ref.named_modules():
    if name and (name.count(".") <= 2) and not name.startswith("backbone"):
        mod.register_forward_hook(hook(name))
feats_r = ref.backbone.forward_features_list(x)[1:]
# ours, instrumented: wrap conv_block / _csp to record outputs by module identity
rec = {}
names = {id(mod): n for n, mod in m.named_modules()}
orig_cb, orig_csp = pt.conv_block, pt._csp
def cb(model, mod, x_, need_dx=True):
    y = orig_cb(model, mod, x_, need_dx); rec[names[id(mod)]] = y.detach(); return y
def csp(model, mod, x_):
    y = orig_csp(model, mod, x_); rec[names[id(mod)]] = y.detach(); return y
pt.conv_block, pt._csp = cb, csp
outs = m(x.to(dev))
plan = [v for k, v in m._tplans.items() if k[0] == "backbone"][0]
if os.environ.get("EMULATE", "1") == "1":
    ste = lambda t: t + (t.to(torch.bfloat16).float() - t).detach()
    with torch.no_grad():
        for name, mod in ref.named_modules():
            if isinstance(mod, torch.nn.Conv2d) and name.startswith(("adaptor", "sppf", "pan", "head")):
                mod.weight.copy_(mod.weight.to(torch.bfloat16).float())
    for name, mod in ref.named_modules():
        if not name.startswith(("adaptor", "sppf", "pan", "head")): continue
        if isinstance(mod, torch.nn.Conv2d) and mod.bias is None: mod.register_forward_hook(lambda m_, i, o: ste(o))
        if isinstance(mod, dr.ConvBlock): mod.register_forward_hook(lambda m_, i, o: ste(o))
        if isinstance(mod, dr.BottleneckCSP): mod.act.register_forward_hook(lambda m_, i, o: ste(o))
    orig = dr.F.interpolate
    dr.F.interpolate = lambda *a, **k: ste(orig(*a, **k))
f = ref.adaptor([t.float().permute(0, 3, 1, 2).cpu() for t in plan.feats]); f[-1] = ref.sppf(f[-1]); outs_r = ref.head(ref.pan(f))
rel = lambda a, b: ((a - b).norm() / b.norm()).item()
for i, (f, fr) in enumerate(zip(plan.feats, feats_r)):
    print(f"backbone feat {i}: rel {rel(f.float().permute(0,3,1,2).cpu(), fr):.2e}")
for n, y in rec.items():
    if n in acts:
        r = acts[n]
        yy = y.float().permute(0, 3, 1, 2).cpu()[:, :r.shape[1]]
        print(f"{n:32s} C={r.shape[1]:4d} {tuple(r.shape[2:])}: rel {rel(yy, r):.2e}")
for i, (o, r) in enumerate(zip(outs, outs_r)):
    print(f"level {i}: rel {rel(o.detach().cpu(), r.detach()):.2e}  cls {rel(o.detach().cpu()[:, :1], r.detach()[:, :1]):.2e}  kpt {rel(o.detach().cpu()[:, 1:], r.detach()[:, 1:]):.2e}")
gen = torch.Generator().manual_seed(5)
loss = loss_r = 0.0
for o, r in zip(outs, outs_r):
    wgt = torch.randn(r.shape, generator=gen) / r.numel() ** 0.5
    loss = loss + (o * wgt.to(dev)).sum(); loss_r = loss_r + (r * wgt).sum()
loss.backward(); loss_r.backward()
rp = dict(ref.named_parameters())
rows = []
for name, p in m.named_parameters():
    if name.startswith("backbone."): continue
    gr = rp[name].grad
    rows.append((rel(p.grad.cpu(), gr), torch.nn.functional.cosine_similarity(p.grad.cpu().flatten(), gr.flatten(), dim=0).item(), name))
rows.sort(reverse=True)
print("params", len(rows), "worst:")
for r in rows[:12]: print("  rel %.3e cos %.5f %s" % r)
import statistics
print("median rel %.3e" % statistics.median(r[0] for r in rows), "min cos %.5f" % min(r[1] for r in rows))
