import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpe_b200 as hpe
from importlib import import_module
from oracle import heatmap_ref as hr, resnet_ref as rr
models = import_module("2d-human-pose-estimation_b200.models")
B, H, W = 8, 128, 96
ref = rr.build_synthetic(17, seed=0, normalize=True)
with torch.no_grad():
    for n, mod in ref.named_modules():
        if n.endswith("bn3"): mod.weight.mul_(0.1)
x = torch.rand(B, 3, H, W, generator=torch.Generator().manual_seed(0)).cuda()
joints = torch.from_numpy(hr.synth_joints(B, 17, (H, W))).cuda().float().contiguous()
def rel(a, b): return ((a - b).norm() / (b.norm() + 1e-30)).item()
def run():
    m = models.HeatmapPose(17, "resnet50", pretrained=False); m.load_state_dict(ref.state_dict()); m = m.cuda().train()
    hm = m(x); loss = hpe.JointsMSELoss()(hm, joints, None); loss.backward(); torch.cuda.synchronize()
    return hm.detach().clone(), {n: p.grad.clone() for n, p in m.named_parameters()}, m
h1, g1, m1 = run(); h2, g2, m2 = run()
print("hm run-to-run rel", rel(h1, h2))
for n in ["backbone.conv1.weight", "backbone.layer1.0.conv1.weight", "backbone.layer3.0.conv2.weight", "deconv_layers.0.weight", "final_layer.weight"]:
    print(n, "run-to-run rel", rel(g1[n], g2[n]))
m3 = models.HeatmapPose(17, "resnet50", pretrained=False); m3.load_state_dict(ref.state_dict()); m3 = m3.cuda().train()
step = hpe.TrainStep(m3, lr=0.05, momentum=0.9, weight_decay=0.0)
l = step(x, joints); torch.cuda.synchronize()
gv = step.layout.views(step.flat_grad)
for n in ["backbone.conv1.weight", "backbone.layer1.0.conv1.weight", "backbone.layer3.0.conv2.weight", "deconv_layers.0.weight", "final_layer.weight", "backbone.bn1.weight"]:
    print(n, "TrainStep vs drop-in rel", rel(gv[n], g1[n]))
