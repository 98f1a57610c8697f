import importlib, sys, os, time, torch
sys.path.insert(0, '/root/repo')
import hpe_b200 as hpe
models = importlib.import_module("2d-human-pose-estimation_b200.models")
dev = torch.device("cuda:0")
m = models.HeatmapPose(17, "resnet50", pretrained=False).to(dev).eval()
B,H,W=256,256,192
gen = torch.Generator().manual_seed(1)
h_u8 = [torch.randint(0,256,(B,3,H,W),generator=gen,dtype=torch.uint8).pin_memory() for _ in range(2)]
h_f32 = [t.float().div_(255.0).pin_memory() for t in h_u8]
center = torch.rand(B, 2, generator=gen) * torch.tensor([640.0, 480.0]); sw = torch.rand(B, 1, generator=gen) * 2 + 0.5
h_aff = hpe.affine_from_center_scale(center, torch.cat([sw, sw * 4 / 3], 1), (H // 4, W // 4)).pin_memory()
est = hpe.PoseEstimator(m, dev)
def region(batches, n):
    for _ in est.stream(((batches[i&1], h_aff) for i in range(3))): pass
    torch.cuda.synchronize()
    a,b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0=time.perf_counter(); a.record()
    for p,mv in est.stream(((batches[i&1], h_aff) for i in range(n))): float(mv[0,0,0])
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b)/n, (time.perf_counter()-t0)*1e3/n
# device-resident
x = [t.to(dev) for t in h_u8]; aff = h_aff.to(dev)
def dev_region(n):
    torch.cuda.synchronize(); a,b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); a.record()
    for i in range(n): est.xyv(x[i&1], aff)
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b)/n
dev_region(5)
for n in (20, 50, 200):
    print(n, "device u8 %.3f" % dev_region(n), "e2e u8 %.3f (wall %.3f)" % region(h_u8, n), "e2e f32 %.3f (wall %.3f)" % region(h_f32, n), "e2e u8 again %.3f (wall %.3f)" % region(h_u8, n))
# H2D alone
s = torch.cuda.Stream()
d = torch.empty_like(h_u8[0], device=dev); df = torch.empty_like(h_f32[0], device=dev)
for t, dd, name in ((h_u8[0], d, "u8"), (h_f32[0], df, "f32")):
    torch.cuda.synchronize(); a,b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); a.record()
    for _ in range(10): dd.copy_(t, non_blocking=True)
    b.record(); torch.cuda.synchronize(); ms=a.elapsed_time(b)/10
    print("H2D", name, "%.3f ms  %.1f GB/s" % (ms, t.numel()*t.element_size()/ms/1e6))
