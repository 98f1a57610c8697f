"""Per-layer comparison of the training-mode forward (and backward) against the fp32 oracle -- development aid."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpe_b200 as hpe
from importlib import import_module
from oracle import heatmap_ref as hr, resnet_ref as rr
models = import_module("2d-human-pose-estimation_b200.models")

B, H, W = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
ref = rr.build_synthetic(17, seed=0, normalize=True).train()
damp = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
with torch.no_grad():
    for n, mod in ref.named_modules():
        if n.endswith("bn3"):
            mod.weight.mul_(damp)
m = models.HeatmapPose(17, "resnet50", pretrained=False); m.load_state_dict(ref.state_dict()); m = m.cuda().train()
x = torch.rand(B, 3, H, W, generator=torch.Generator().manual_seed(0))
acts = {}
def hook(name):
    def f(mod, inp, out): acts[name] = out.detach()
    return f
for n, mod in ref.named_modules():
    if isinstance(mod, (torch.nn.BatchNorm2d, torch.nn.Conv2d, torch.nn.ConvTranspose2d)):
        mod.register_forward_hook(hook(n))
hm_ref = ref(x)
taps = {}
hm_emu = rr.forward_train_bf16_emulated(ref, x, taps)
from oracle import heatmap_ref as hr
import numpy as np
joints = torch.from_numpy(hr.synth_joints(B, 17, (H, W)))
target, weight = hr.generate_target(joints.numpy(), hm_emu.shape[2:], 2.0)
t = torch.from_numpy(target); w = torch.from_numpy(weight).reshape(B, 17, 1, 1)
loss_emu = 0.5 * ((w * hm_emu - w * t) ** 2).sum() / hm_emu.numel()
ref.zero_grad(); loss_emu.backward()
g_emu = {n: p.grad.clone() for n, p in ref.named_parameters()}
plan = m._train_plan(x.cuda()); plan.load_weights(m)
hm = plan.forward(x.cuda()); torch.cuda.synchronize()
def rel(a, b): return ((a - b).norm() / (b.norm() + 1e-30)).item()
for u in plan.units():
    z = u.z.float().cpu().permute(0, 3, 1, 2)
    zr = acts[u.conv_name]
    bnr = acts[u.bn_name]
    # reconstruct bn output (pre residual / relu) from kernel scale/shift
    bn_out = (u.z.float() * u.scale + u.shift).cpu().permute(0, 3, 1, 2)
    ze, ye = taps[u.conv_name]
    y = u.y.float().cpu().permute(0, 3, 1, 2)
    print(f"{u.conv_name:34s} z rel {rel(z, zr):.4f}  bn rel {rel(bn_out, bnr):.4f}  |z| {zr.abs().mean():.3g}   vs emu: z {rel(z, ze):.5f} y {rel(y, ye):.5f}")
print("hm rel", rel(hm.cpu(), hm_ref.detach()), "max", hm_ref.abs().max().item(), "maxdiff", (hm.cpu() - hm_ref.detach()).abs().max().item())

print("hm vs emu rel", rel(hm.cpu(), hm_emu.detach()))
crit = hpe.JointsMSELoss()
m.zero_grad()
hm2 = m(x.cuda()); loss = crit(hm2, joints.cuda(), None); loss.backward(); torch.cuda.synchronize()
print("loss", loss.item(), "emu", loss_emu.item())
for n, p in m.named_parameters():
    a, b = p.grad.cpu(), g_emu[n]
    cos = (torch.dot(a.flatten(), b.flatten()) / (a.norm() * b.norm() + 1e-30)).item()
    print(f"grad {n:44s} rel {rel(a, b):.4f} cos {cos:.5f} |g| {b.norm():.3g}")

# ---- teacher-forced backward: fp32 backward on the kernel's own stored forward state
plan = m._train_plan(x.cuda())
acts = plan.export_acts()
acts["backbone.conv1"]["x"] = x.to(torch.bfloat16).float()
dhm = None
hm3 = m(x.cuda()); hm3.retain_grad(); loss3 = crit(hm3, joints.cuda(), None); m.zero_grad(); loss3.backward(); torch.cuda.synchronize()
dhm = hm3.grad.cpu()
acts = plan.export_acts(); acts["backbone.conv1"]["x"] = x.to(torch.bfloat16).float()
g_tf, fwd_err = rr.teacher_forced_backward(ref, acts, dhm)
print("max per-unit forward z err (rel to max|z|):", max(fwd_err.values()))
for n, p in m.named_parameters():
    a, b = p.grad.cpu(), g_tf[n]
    cos = (torch.dot(a.flatten(), b.flatten()) / (a.norm() * b.norm() + 1e-30)).item()
    print(f"tfgrad {n:44s} rel {rel(a, b):.4f} cos {cos:.5f} |g| {b.norm():.3g}")
