"""Run ONE conv layer configuration a few times (target for `ncu --set full` captures).
    python tools/run_layer.py <kind> B H W Cin Cout stride act res [iters] [head2_K]
"""
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hpe_b200  # noqa: F401,E402

conv = importlib.import_module("2d-human-pose-estimation_b200.conv")
kind, B, H, W, Cin, Cout, stride, act, res = [int(v) for v in sys.argv[1:10]]
iters = int(sys.argv[10]) if len(sys.argv) > 10 else 5
head2 = int(sys.argv[11]) if len(sys.argv) > 11 else 0        # K joints of a fused final 1x1 (deconv3 + final_layer kernel)
chain = int(sys.argv[12]) if len(sys.argv) > 12 else 0        # N2 of a chained next-layer 1x1 conv (conv3 + conv1 kernel)
dev = torch.device("cuda:0")
layer = conv.ConvLayer(kind, B, H, W, Cin, Cout, stride, act, bool(res), dev, head2=head2, chain=chain)
if chain:
    layer.load_chain(torch.randn(chain, 256, 1, 1) * 0.05, torch.rand(chain) + 0.5, torch.randn(chain) * 0.1)
if head2:
    layer.load_head2(torch.randn(head2, 256, 1, 1) * 0.05, torch.zeros(head2))
if kind == conv.CONV_DECONV4:
    w = torch.randn(Cin, Cout, 4, 4) * 0.02
elif kind == conv.CONV_3x3:
    w = torch.randn(Cout, Cin, 3, 3) * 0.02
else:
    w = torch.randn(Cout, Cin, 1, 1) * 0.05
layer.load(w, torch.rand(Cout) + 0.5, torch.randn(Cout) * 0.1)
x = torch.randn(B, H, W, Cin, device=dev).to(torch.bfloat16)
r = torch.randn(layer.out_shape(), device=dev).to(torch.bfloat16) if res else None
y = layer.alloc_out()
y2 = torch.empty((B, layer.Ho, layer.Wo, chain), dtype=torch.bfloat16, device=dev) if chain else None
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ts = []
noflush = os.environ.get("RUN_LAYER_NOFLUSH", "0") == "1"      # inputs stay L2-resident between iterations (small B)
for i in range(iters):
    if not noflush:
        flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); layer(x, r, y, None, *((y2,) if chain else ())); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
print(f"kind {kind} {Cin}->{Cout} s{stride} res{res} @{H}x{W} B={B} noflush={int(noflush)} EPI2={os.environ.get('HPE_B200_EPI2','-')}: min {min(ts)*1e3:.1f} us  median {sorted(ts)[len(ts)//2]*1e3:.1f} us")
