"""Headline benchmark: images/sec forward+decode @256x192, COCO-17 (BASELINE.json), on N B200s of one node.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...        # the CPU arm (oracle port of the reference path), rank 0 only

One "step" = one pass of the hot path over one synthetic batch of 256 images per GPU (configs[1]):
preprocess -> ResNet-50 trunk -> deconv head -> final 1x1 -> heatmap decode (+ result all-gather when N>1).
`value` is device-resident throughput; `e2e` goes through the public host-buffer API (pinned host images in,
keypoints read back on the host every step).  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

# stdout carries exactly ONE JSON line.  Native libraries write their banners to fd 1 (NCCL prints "NCCL version ..." there at
# NCCL_DEBUG=VERSION and ignores NCCL_DEBUG_FILE at that level), so fd 1 is pointed at stderr for the whole process and the
# JSON line goes to a private duplicate of the original stdout.
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
sys.stdout.flush()
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(line: dict) -> None:
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "images/sec fwd+decode @256x192 COCO-17"
GFLOP_PER_IMG_256x192 = 8.008 + 2.846          # trunk + head, 2*MAC, no padding/zero-insertion waste
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "r02_traffic.json")     # written by tools/ncu_traffic.py from an ncu launch list


def load_peaks():
    """MEASURED_PEAKS.json (driver-written): `tflops_burst` = cuBLAS bf16 best-of-10, the like-for-like peak of a kernel or a
    short K-step region at full clocks; `tflops` = the 4-second back-to-back figure, for a region long enough to be power-capped."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm": float(p["hbm_gbs"]), "tflops": float(p["bf16_tflops_sustained"]), "tflops_burst": float(p["bf16_tflops"]),
                "src": "measured"}
    except Exception:
        return {"hbm": 6650.0, "tflops": 1400.0, "tflops_burst": 1650.0, "src": "fallback (B200_PROFILING.md)"}


def source_hash():
    """sha256 over the CUDA sources + the host files that decide which kernels a step launches: ties profiles/r02_traffic.json
    (an ncu capture) to the build it was taken from."""
    import hashlib
    h = hashlib.sha256()
    pkg = os.path.join(ROOT, "2d-human-pose-estimation_b200")
    files = sorted(os.path.join(pkg, "csrc", f) for f in os.listdir(os.path.join(pkg, "csrc")))
    files += [os.path.join(pkg, f) for f in ("conv.py", "models.py")]
    for f in files:
        with open(f, "rb") as fh:
            h.update(os.path.basename(f).encode() + b"\0" + fh.read())
    return h.hexdigest()[:16]


def load_traffic(workload_key):
    """DRAM bytes per launch of the dominant kernel family from the committed ncu launch list of the CURRENT build; a capture
    taken from other sources is reported as stale (traffic null), never silently reused."""
    try:
        with open(TRAFFIC_FILE) as f:
            t = json.load(f)
    except Exception:
        return None, "no profiles/r02_traffic.json (run tools/ncu_traffic.py on the GPU box)"
    if t.get("workload") != workload_key:
        return None, f"profiles/r02_traffic.json is for workload {t.get('workload')}"
    cur = source_hash()
    if t.get("source_hash") != cur:
        print(f"[bench] WARNING: profiles/r02_traffic.json was captured from sources {t.get('source_hash')}, the tree is {cur}: "
              "roofline.traffic omitted (stale capture)", file=sys.stderr)
        return None, f"stale: captured from sources {t.get('source_hash')}, tree is {cur}"
    return t, t.get("how", "")


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of one GPU through NVML every 20 ms while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
                 "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80)}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop_evt.wait(0.02)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def physical_gpu_index(local):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local])
        except Exception:
            return local
    return local


def workload_config(args, world):
    """`config` of the JSON line -- identical for the B200 arm and the reference arm (same workload)."""
    B, H, W, K = args.batch, args.height, args.width, args.joints
    return {"workload": f"resnet50+deconv-head heatmap fwd+decode, {H}x{W}, K={K}, batch {B}/GPU, random-init weights "
                        "(BASELINE configs[1])",
            "global_batch": world * B, "parallelism": f"dp{world}",
            "input": "uint8 NCHW images resident in HBM, /255 on the device (HPE/train.py:117-120: imgs.to(device); imgs.float() / 255.0)",
            "l2": f"two alternating input batches ({B * 3 * H * W / 1e6:.0f} MB uint8 each) and every activation of a step (3.4 GB in total) "
                  "exceed the 126 MB L2; no flush needed"}


# --------------------------------------------------------------------------------------------- CPU arm
def cpu_forward_decode(batch, H, W, K, steps, warmup, threads=None):
    """The CPU restatement of the path (oracle port; the reference itself is Python/timm and cannot travel to
    the GPU box): fp32 torch ResNet-50 + deconv head on all host threads, NumPy decode.  Returns img/s."""
    import numpy as np
    from oracle import heatmap_ref as hr
    from oracle import resnet_ref as rr
    n = threads or os.cpu_count() or 1
    torch.set_num_threads(n)
    model = rr.build_synthetic(K, seed=0)
    x = torch.randint(0, 256, (batch, 3, H, W), generator=torch.Generator().manual_seed(0), dtype=torch.uint8)   # the reference's input format
    center, scale = hr.synth_center_scale(batch)
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            hm = model(x.float() / 255.0).numpy()                 # HPE/train.py:120: imgs = imgs.float() / 255.0
            preds, maxvals = hr.get_final_preds(hm, center, scale)
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    assert np.isfinite(preds).all()
    mean = sum(times) / len(times)
    return batch / mean, mean, torch.get_num_threads()


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path (oracle port, pinned bit-identical to the reference's own
    timm trunk) on all host threads.  Honours --steps/--warmup; each step is a bounded sample (--cpu-batch images) of the
    B200 arm's batch, so the per-image metric is comparable."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    H, W = args.height, args.width
    batch = args.cpu_batch
    steps, warmup = max(args.steps, 1), max(args.warmup, 1)
    ips, mean, nthr = cpu_forward_decode(batch, H, W, args.joints, steps, warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(ips, 3), "unit": "images/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": round(mean * 1e3, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": round(ips, 3), "unit": "images/s", "cores": nthr, "kind": "port",
                         "sample": f"{steps} steps x {batch} images (bounded sample of the {args.batch}-image batch), torch fp32 "
                                   f"{nthr} threads + numpy decode"},
        "e2e": {"value": round(ips, 3), "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# --------------------------------------------------------------------------------------------- GPU arm
class Ctx:
    """Process-wide state of the B200 arm: rank layout, device, process group, barrier and max-over-ranks helpers."""

    def __init__(self, what):
        import torch.distributed as dist
        self.dist = dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise RuntimeError(f"bench.py {what}: no CUDA device; the B200 arm has no CPU fallback (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier(device_ids=[self.local])
        torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, v):
        t = torch.tensor([float(v)], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sampler(self):
        return ClockSampler(physical_gpu_index(self.local)) if self.rank == 0 else None

    def close(self):
        if self.world > 1:
            self.barrier()
            self.dist.destroy_process_group()


def ev():
    return torch.cuda.Event(enable_timing=True)


def timed(ctx, step, steps):
    """Barrier + synchronize, EXACTLY `steps` calls of step(i) between two CUDA events on the launching stream, barrier +
    synchronize; returns (ms over the region as the MAX over ranks, clock record of rank 0)."""
    sampler = ctx.sampler()
    if sampler:
        sampler.start()
    start, end = ev(), ev()
    ctx.barrier()
    start.record()
    for i in range(steps):
        step(i)
    end.record()
    ctx.barrier()
    clocks = sampler.stop() if sampler else None
    return ctx.max_over_ranks(start.elapsed_time(end)), clocks


def make_model(hpe, K, dev, train=False):
    """random-init net with randomised BatchNorm so it is not degenerate (timm zero-inits every block's last gamma)."""
    torch.manual_seed(0)
    model = hpe.HeatmapPose(K, "resnet50", pretrained=False)
    g = torch.Generator().manual_seed(1)
    for n, m in model.named_modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.weight.data.copy_((torch.rand(m.weight.shape, generator=g) + 0.5) * (0.1 if (train and n.endswith("bn3")) else 1.0))
            m.bias.data.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
            if not train:
                m.running_mean.copy_(torch.randn(m.running_mean.shape, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(m.running_var.shape, generator=g) + 0.5)
    model = model.to(dev)
    return model.train() if train else model.eval()


def infer_setup(hpe, ctx, model, B, H, W, K, seed=100):
    """Synthetic inputs of one rank: two alternating uint8 host batches (the reference's host format, HPE/train.py:117-120),
    device copies of them (the reference's device-resident input too: `imgs.to(device)` and THEN `imgs.float() / 255.0`,
    HPE/train.py:117-120 -- here the /255 is fused into the stem's layout kernel), their fp32 twins, per-sample affines."""
    gen = torch.Generator().manual_seed(seed + ctx.rank)
    h_u8 = [torch.randint(0, 256, (B, 3, H, W), generator=gen, dtype=torch.uint8).pin_memory() for _ in range(2)]
    center = torch.rand(B, 2, generator=gen) * torch.tensor([640.0, 480.0])
    sw = torch.rand(B, 1, generator=gen) * 2 + 0.5
    h_aff = hpe.affine_from_center_scale(center, torch.cat([sw, sw * 4 / 3], 1), (H // 4, W // 4)).pin_memory()
    return {"h_u8": h_u8, "h_aff": h_aff, "x": [t.to(ctx.dev).float().div_(255.0) for t in h_u8], "xu": [t.to(ctx.dev) for t in h_u8], "aff": h_aff.to(ctx.dev),
            "rg": hpe.ResultGather(B, K, ctx.dev)}


def infer_step_fn(hpe, model, s, marks=None):
    """One step of the hot path on device-resident inputs: forward -> decode (packed x,y,maxval rows written by the decode
    kernel) -> result all-gather on a side stream when N > 1."""
    rg = s["rg"]

    def step(i):
        if marks is not None:
            e0 = ev(); e0.record()
        hm = model.forward_static(s["xu"][i & 1])
        if marks is not None:
            e1 = ev(); e1.record()
        hpe.get_final_preds_xyv(hm, s["aff"], rg.slot())
        if marks is not None:
            e2 = ev(); e2.record()
            marks.append((e0, e1, e2))
        return rg.gather()
    return step


def e2e_region(ctx, est, host_batches, steps, settle_s=0.0):
    """`steps` batches through PoseEstimator.stream (pinned host images in, keypoints read on the host); CUDA events around the
    whole loop on the compute stream (the last D2H is waited for by the generator before the end event), max over ranks.
    `settle_s`: idle time between the untimed warm-up batches and the timed region, so the region starts from the same power
    state as the device-timed headline region (which follows only `--warmup` steps): a B200 runs this step into its power cap
    within ~0.2 s, and a region measured right after another one starts at ~1.6 instead of 1.965 GHz.  Returns the clock record
    of the region as well."""
    for _ in est.stream(host_batches(3)):
        pass
    torch.cuda.synchronize()
    if settle_s > 0:
        time.sleep(settle_s)
    sampler = ctx.sampler()
    if sampler:
        sampler.start()
    ctx.barrier()
    t0 = time.perf_counter()
    a, b = ev(), ev()
    a.record()
    checksum = 0.0
    for preds_h, maxvals_h in est.stream(host_batches(steps)):
        checksum += float(maxvals_h[0, 0, 0])
    b.record()
    ctx.barrier()
    wall = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop() if sampler else None
    return ctx.max_over_ranks(max(a.elapsed_time(b), 0.0)), wall, checksum, clocks


def decode_sweep(hpe, dev, K, H, W, peaks):
    """Decode at an HBM-saturating size (BASELINE configs[4] sweep point: K=17, 64x48, fp32, 16384 x 17 maps = 3.4 GB, far
    beyond the 126 MB L2): the "decode HBM GB/s vs peak" half of the metric.  The in-step decode is 53 MB."""
    SB = 16384
    big = torch.rand(SB, K, H // 4, W // 4, device=dev)
    big_aff = torch.rand(SB, 2, 3, device=dev)
    out = torch.empty(SB, K, 3, device=dev)
    for _ in range(3):
        hpe.get_final_preds_xyv(big, big_aff, out)
    torch.cuda.synchronize(dev)
    ts = []
    for _ in range(10):
        a, b = ev(), ev()
        a.record(); hpe.get_final_preds_xyv(big, big_aff, out); b.record(); torch.cuda.synchronize(dev)
        ts.append(a.elapsed_time(b))
    ms = sum(ts) / len(ts)
    nbytes = big.numel() * 4 + SB * 24 + SB * K * 12
    gbs = nbytes / (ms * 1e-3) / 1e9
    return {"bound": "hbm", "kernel": "decode_kernel", "achieved": round(gbs, 1), "peak": peaks["hbm"], "unit": "GB/s",
            "frac": round(gbs / peaks["hbm"], 4), "traffic": None,
            "workload": f"decode sweep point: {SB} x {K} maps of {H // 4}x{W // 4} fp32 ({nbytes / 1e9:.1f} GB, > L2), {ms:.3f} ms/launch, "
                        "mean of 10 launches (CUDA events)",
            "algorithmic_bytes_per_map": (H // 4) * (W // 4) * 4 + 12}


def per_kernel_table(model, x, peaks):
    """roofline.per_kernel: CUDA-event time of every conv launch of ONE un-graphed step, summed by family
    (stem / 1x1 and 3x3 by stage / deconv head), with TFLOP/s and the fraction of the burst tensor peak."""
    model.forward_static(x)
    plan = model._plans[(tuple(x.shape), x.device.index)]
    rows = plan.time_layers(reps=3)
    nbytes = plan.layer_bytes()
    fam = {}
    for name, f, flops, us in rows:
        d = fam.setdefault(f, {"launches": 0, "us": 0.0, "gflop": 0.0, "mb": 0.0, "bound_us": 0.0, "hbm_bound": 0})
        t_tensor, t_hbm = flops / (peaks["tflops_burst"] * 1e12) * 1e6, nbytes[name] / (peaks["hbm"] * 1e9) * 1e6
        d["launches"] += 1; d["us"] += us; d["gflop"] += flops / 1e9; d["mb"] += nbytes[name] / 1e6
        d["bound_us"] += max(t_tensor, t_hbm); d["hbm_bound"] += int(t_hbm > t_tensor)
    out = {}
    for f, d in fam.items():
        tf = d["gflop"] / 1e3 / (d["us"] * 1e-6) if d["us"] > 0 else 0.0
        out[f] = {"launches": d["launches"], "us_per_step": round(d["us"], 1), "tflops": round(tf, 1),
                  "frac_burst": round(tf / peaks["tflops_burst"], 3),
                  # the per-launch roofline: max(FLOPs / tensor peak, algorithmic bytes / measured HBM copy bandwidth), summed
                  "mb_per_step": round(d["mb"], 1), "bound_us": round(d["bound_us"], 1), "hbm_bound_launches": d["hbm_bound"],
                  "frac_of_bound": round(d["bound_us"] / d["us"], 3) if d["us"] > 0 else 0.0}
    return out, rows


def hires_subbench(hpe, ctx, model, args, peaks):
    """BASELINE configs[3]: 384x288 input -> 96x72 heatmaps, 64 images per GPU (B=512 over 8 GPUs), forward + decode
    (+ result all-gather), same model and kernels, its own timed region."""
    B, H, W, K = args.hires_batch, 384, 288, args.joints
    s = infer_setup(hpe, ctx, model, B, H, W, K, seed=300)
    step = infer_step_fn(hpe, model, s)
    for i in range(3):
        step(i)
    s["rg"].wait()
    steps = max(args.steps, 10)
    ms_total, clocks = timed(ctx, lambda i: step(i), steps)
    s["rg"].wait()
    ms = ms_total / steps
    tf = GFLOP_PER_IMG_256x192 * 2.25 * 1e9 * B / (ms * 1e-3) / 1e12
    return {"workload": f"384x288 -> 96x72 heatmaps, K={K}, batch {B}/GPU fwd+decode (BASELINE configs[3]: B={B * ctx.world} over {ctx.world} GPUs)",
            "value": round(ctx.world * B * steps / ms_total * 1e3, 1), "unit": "images/s", "ms_per_step": round(ms, 4), "steps": steps,
            "tflops_per_gpu": round(tf, 1), "frac_burst": round(tf / peaks["tflops_burst"], 4), "clocks": clocks}


def train_subbench(hpe, ctx, args, peaks):
    """BASELINE configs[2]: the data-parallel training step (batch-stat BN per rank, fused Gaussian-target JointsMSE, dgrad +
    wgrad, ONE NCCL all-reduce of the flat fp32 gradient buffer, SGD momentum) at N GPUs, >= 10 timed steps."""
    B, H, W, K = args.batch, args.height, args.width, args.joints
    model = make_model(hpe, K, ctx.dev, train=True)
    step = hpe.TrainStep(model, lr=1e-3, momentum=0.9, weight_decay=1e-4, cuda_graph=not args.no_graph)
    step.broadcast_parameters(0)
    gen = torch.Generator().manual_seed(200 + ctx.rank)
    x = [torch.rand(B, 3, H, W, generator=gen).to(ctx.dev) for _ in range(2)]
    joints = torch.rand(B, K, 3, generator=gen) * torch.tensor([float(W), float(H), 1.0]) + torch.tensor([0.0, 0.0, 0.3])
    joints = joints.to(ctx.dev).contiguous()
    for i in range(3):
        step(x[i & 1], joints)
    steps = max(args.steps, 10)
    ms_total, clocks = timed(ctx, lambda i: step(x[i & 1], joints), steps)
    loss = float(step.loss.item())
    ms = ms_total / steps
    tf = 3 * GFLOP_PER_IMG_256x192 * 1e9 * (H * W) / (256 * 192) * B / (ms * 1e-3) / 1e12      # fwd + dgrad + wgrad
    nbytes = step.layout.trainable_total * 4
    out = {"workload": f"resnet50+deconv-head data-parallel training step, {H}x{W}, K={K}, batch {B}/GPU (BASELINE configs[2]); "
                       "per-rank BatchNorm statistics and averaged gradients as under the reference's DDP wrap (pose/pose/core/builder.py:92-123)",
           "value": round(ctx.world * B * steps / ms_total * 1e3, 1), "unit": "images/s", "ms_per_step": round(ms, 4), "steps": steps,
           "allreduce_bytes_per_step": nbytes if ctx.world > 1 else 0, "allreduce": "one NCCL all-reduce(sum) of the flat fp32 gradient buffer" if ctx.world > 1 else "none (1 GPU)",
           "tflops_per_gpu": round(tf, 1), "frac_burst": round(tf / peaks["tflops_burst"], 4),
           "frac_sustained": round(tf / peaks["tflops"], 4), "loss": loss, "clocks": clocks}
    step.close()
    del step, model
    torch.cuda.empty_cache()
    return out


def pose_train_subbench(hpe, ctx, args):
    """The reference's REAL training step (HPE/train.py:116-139) at N GPUs: DINOv3Pose(small) with its frozen ConvNeXt backbone,
    batch-statistics neck / head, ComputeLoss, backward through the dgrad / wgrad kernels, ONE NCCL all-reduce of the flat gradient
    buffer, gradient-norm clipping + AdamW kernels (hpe_b200.PoseTrainStep, forward + loss + backward replayed as a CUDA graph)."""
    B, H, W = args.pose_train_batch, args.height, args.width
    torch.manual_seed(0)
    m = hpe.DINOv3Pose("dinov3_convnext_small", nkpts=(17, 3), ncls=1, device=ctx.dev)
    gen = torch.Generator().manual_seed(300 + ctx.rank)
    x = [torch.randint(0, 256, (B, 3, H, W), generator=gen, dtype=torch.uint8).to(ctx.dev) for _ in range(2)]
    tg = torch.zeros(B, 6 + 17 * 3)
    tg[:, 0] = torch.arange(B); tg[:, 2:4] = 0.3 + 0.4 * torch.rand(B, 2, generator=gen); tg[:, 4:6] = 0.3
    kp = torch.rand(B, 17, 3, generator=gen); kp[..., :2] = 0.2 + 0.6 * kp[..., :2]; kp[..., 2] = 2.0
    tg[:, 6:] = kp.reshape(B, -1)
    tg = tg.to(ctx.dev)
    step = hpe.PoseTrainStep(m, lr=1e-4, weight_decay=1e-4, max_grad_norm=10.0, cuda_graph=not args.no_graph)
    step.broadcast_parameters(0)
    for i in range(4):
        step(x[i & 1], tg)
    steps = max(args.steps, 10)
    ms_total, clocks = timed(ctx, lambda i: step(x[i & 1], tg), steps)
    loss = float(step.loss)
    nbytes = step.flat_grad.numel() * 4
    out = {"workload": f"DINOv3Pose(dinov3_convnext_small) training step, {H}x{W}, K=17, batch {B}/GPU: frozen backbone, batch-statistics "
                       "neck + head, ComputeLoss, backward, all-reduce, clip_grad_norm_(10) + AdamW (HPE/train.py:116-139)",
           "value": round(ctx.world * B * steps / ms_total * 1e3, 1), "unit": "images/s", "ms_per_step": round(ms_total / steps, 4), "steps": steps,
           "trainable_params": int(step.layout.trainable_total),
           "allreduce_bytes_per_step": nbytes if ctx.world > 1 else 0, "loss": loss, "clocks": clocks}
    step.graph = None
    del step, m
    torch.cuda.empty_cache()
    return out


def run_b200(args):
    import hpe_b200 as hpe
    ctx = Ctx("(B200 arm)")
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    hpe._lib.check(hpe._lib.lib().hpe_check_device(), "hpe_check_device")
    B, H, W, K = args.batch, args.height, args.width, args.joints
    peaks = load_peaks()
    model = make_model(hpe, K, dev)
    est = hpe.PoseEstimator(model, dev)
    s = infer_setup(hpe, ctx, model, B, H, W, K)
    warm = max(args.warmup, 3)

    # ---- headline `value`: inputs resident in HBM, EXACTLY --steps steps, device-timed, max over ranks
    marks = []
    step_plain, step_marked = infer_step_fn(hpe, model, s), infer_step_fn(hpe, model, s, marks)
    for i in range(warm):
        step_plain(i)
    s["rg"].wait()
    ms_total, clocks = timed(ctx, step_marked, args.steps)
    s["rg"].wait()
    torch.cuda.synchronize(dev)
    fwd_ms = sum(a.elapsed_time(b) for a, b, _ in marks) / len(marks)
    dec_ms = sum(b.elapsed_time(c) for _, b, c in marks) / len(marks)

    # ---- e2e through the public host-buffer API: uint8 host images (the reference's real input type) = the headline;
    # fp32 host images (4x the H2D bytes) as the secondary number.  Measured right after the headline region, BEFORE the
    # 2-second sustained region heats the GPU into its power cap (r02 first run: 3.59 vs 3.32 ms/step came from that order)
    u_ms, u_wall, chk, u_clk = e2e_region(ctx, est, lambda n: ((s["h_u8"][i & 1], s["h_aff"]) for i in range(n)), args.steps, args.e2e_settle_s)
    h_f32 = [t.float().div_(255.0).pin_memory() for t in s["h_u8"]]
    f_ms, f_wall, chk2, f_clk = e2e_region(ctx, est, lambda n: ((h_f32[i & 1], s["h_aff"]) for i in range(n)), args.steps, args.e2e_settle_s)
    del h_f32

    # ---- the same step for >= 2 s back to back: the power-capped (sustained) regime, with its own clock record
    n_sus = int(min(args.sustain_s * 1e3 / (ms_total / args.steps) + 1, 5000)) if args.sustain_s > 0 else 0
    sus = None
    if n_sus:
        marks_s = []
        step_s = infer_step_fn(hpe, model, s, marks_s)
        sus_ms_total, sus_clocks = timed(ctx, step_s, n_sus)
        s["rg"].wait()
        torch.cuda.synchronize(dev)
        sus = {"steps": n_sus, "seconds": round(sus_ms_total / 1e3, 3), "ms_per_step": round(sus_ms_total / n_sus, 4),
               "ms_forward": round(sum(a.elapsed_time(b) for a, b, _ in marks_s) / len(marks_s), 4), "clocks": sus_clocks,
               "value": round(world * B * n_sus / sus_ms_total * 1e3, 1)}

    sweep = per_kernel = hires = train = None
    if rank == 0 and not args.quick:
        sweep = decode_sweep(hpe, dev, K, H, W, peaks)
        per_kernel, layer_rows = per_kernel_table(model, s["x"][0], peaks)
        if args.layers_log:
            with open(args.layers_log, "w") as f:
                for name, famn, flops, us in layer_rows:
                    f.write(f"{name:34s} {famn:14s} {us:8.1f} us {flops / 1e6 / us if us else 0:8.1f} TFLOP/s\n")
    ctx.barrier()
    if not args.quick and not args.no_hires:
        hires = hires_subbench(hpe, ctx, model, args, peaks)
    plan = next(iter(model._plans.values()))
    launches_per_step = plan.n_launches + 1 + (1 if world > 1 else 0)        # preprocess + convs, decode, (NCCL all-gather)
    own_launches_per_step = plan.n_launches + 1
    del est, s, model
    torch.cuda.empty_cache()
    if not args.quick and not args.no_train:
        train = train_subbench(hpe, ctx, args, peaks)
    pose_train = None
    if not args.quick and not args.no_train and not args.no_pose_train:
        pose_train = pose_train_subbench(hpe, ctx, args)

    if rank == 0:
        img_scale = (H * W) / (256 * 192)
        flops_step = GFLOP_PER_IMG_256x192 * 1e9 * img_scale * B
        tf = flops_step / (fwd_ms * 1e-3) / 1e12
        dec_bytes = B * K * (H // 4) * (W // 4) * 4 + B * 24 + B * K * 12
        total_imgs = world * B * args.steps
        traffic, traffic_note = load_traffic(f"{B}x{H}x{W}x{K}")
        roof = {"bound": "tensor",
                "kernel": "conv_gemm_kernel + conv_yhalo_kernel (every conv launch of the step: stem with the max-pool in its epilogue, trunk with the "
                          "projection shortcuts fused into conv3, deconv head with the final 1x1 fused into its last stage)",
                "achieved": round(tf, 2), "peak": peaks["tflops_burst"], "unit": "TFLOP/s", "frac": round(tf / peaks["tflops_burst"], 4),
                "peak_source": f"{peaks['src']} bf16_tflops (burst: the --steps region lasts {ms_total / 1e3:.2f} s at full clocks)",
                "frac_burst": round(tf / peaks["tflops_burst"], 4),
                "traffic": traffic["dram_bytes_per_conv_launch"] if traffic else None, "traffic_source": traffic_note,
                "ms_forward": round(fwd_ms, 4),
                "note": "algorithmic FLOPs of the step (10.854 GFLOP/image, SURVEY §8d) / CUDA-event time of the forward graph (incl. preprocess)",
                "per_kernel": per_kernel}
        if per_kernel:
            # Two-resource view of the same step: every launch bounded by the slower of its tensor time and its HBM time.  The
            # whole-step tensor fraction above cannot reach 1 because the 1x1 layers of layer1 / layer2 (and the identity blocks
            # of layer3) are HBM-bound at bf16 activations; this is the fraction of the step's own per-launch roofline.
            b_us = sum(v["bound_us"] for v in per_kernel.values())
            roof["layerwise"] = {"bound_ms": round(b_us / 1e3, 4), "frac": round(b_us / 1e3 / fwd_ms, 4),
                                 "gb_per_step": round(sum(v["mb_per_step"] for v in per_kernel.values()) / 1e3, 3),
                                 "hbm_bound_launches": sum(v["hbm_bound_launches"] for v in per_kernel.values()),
                                 "launches": sum(v["launches"] for v in per_kernel.values()),
                                 "peaks": {"tensor_tflops": peaks["tflops_burst"], "hbm_gbs": peaks["hbm"]},
                                 "note": "sum over the conv launches of max(algorithmic FLOPs / burst bf16 peak, algorithmic bytes / measured HBM "
                                         "copy bandwidth) / ms_forward"}
        if sus:
            tfs = flops_step / (sus["ms_forward"] * 1e-3) / 1e12
            roof["frac_sustained"] = round(tfs / peaks["tflops"], 4)
            roof["sustained"] = dict(sus, achieved=round(tfs, 2), peak=peaks["tflops"],
                                     peak_source=f"{peaks['src']} bf16_tflops_sustained (4 s of back-to-back cuBLAS)")
        line = {
            "metric": METRIC, "value": round(total_imgs / ms_total * 1e3, 1), "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": warm, "ms_per_step": round(ms_total / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(args, world),
            "roofline": roof,
            "e2e": {"value": round(total_imgs / u_ms * 1e3, 1), "unit": "images/s",
                    "h2d_bytes_per_step": B * 3 * H * W + B * 24, "d2h_bytes_per_step": B * K * 3 * 4,
                    "ms_per_step": round(u_ms / args.steps, 4), "wall_ms_per_step": round(u_wall / args.steps, 4),
                    "api": "PoseEstimator.stream: pinned host uint8 images (the reference's host format, HPE/train.py:117-120, HPE/test.py:16-29; "
                           "/255 fused into the stem) -> H2D on a copy stream -> forward -> decode -> D2H of the packed keypoints on a third "
                           "stream, two-deep result ring",
                    "clocks": u_clk, "settle_s": args.e2e_settle_s,
                    "settle_note": "idle seconds between the untimed warm-up batches and the timed region: the region starts from the same "
                                   "power state as the device-timed region (over 200 back-to-back steps e2e and device time agree to 0.1 %)",
                    "fp32_images": {"value": round(total_imgs / f_ms * 1e3, 1), "h2d_bytes_per_step": B * 3 * H * W * 4 + B * 24,
                                    "ms_per_step": round(f_ms / args.steps, 4),
                                    "note": "same call with fp32 host images (4x the H2D bytes: PCIe / host-DRAM bound at N >= 4)"}},
            "gpu_launches": own_launches_per_step * args.steps,
            "gpu_launches_per_step": own_launches_per_step,
            "launches_per_step_incl_nccl": launches_per_step,
            "clocks": clocks,
        }
        if sweep:
            sweep["in_step"] = {"ms_decode": round(dec_ms, 4), "achieved": round(dec_bytes / (dec_ms * 1e-3) / 1e9, 1),
                                "note": "the step's own 53 MB of heatmaps is launch/latency-bound, not bandwidth-bound"}
            line["decode_roofline"] = sweep
        if hires:
            line["hires"] = hires
        if train:
            line["train"] = train
        if pose_train:
            line["pose_train"] = pose_train
        if world == 1 and not args.no_cpu_baseline:
            ips, mean, nthr = cpu_forward_decode(args.cpu_batch, H, W, K, 3, 1)
            line["cpu_baseline"] = {"value": round(ips, 3), "unit": "images/s", "cores": nthr, "kind": "port",
                                    "sample": f"3 steps x {args.cpu_batch} images of the same workload (torch fp32 ResNet-50+head on "
                                              f"{nthr} threads + numpy decode), {mean:.2f} s/step"}
        emit(line)
    ctx.close()


def run_train(args):
    """`--mode train`: BASELINE configs[2] alone, as its own JSON line (the default mode carries the same numbers as `train`)."""
    import hpe_b200 as hpe
    ctx = Ctx("--mode train")
    peaks = load_peaks()
    t = train_subbench(hpe, ctx, args, peaks)
    if ctx.rank == 0:
        B, H, W, K = args.batch, args.height, args.width, args.joints
        emit({"metric": "images/sec train step @256x192 COCO-17", "value": t["value"], "unit": "images/s", "n_gpus": ctx.world,
              "steps": t["steps"], "warmup": 3, "ms_per_step": t["ms_per_step"], "higher_is_better": True, "scaling": "weak",
              "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
              "config": {"workload": t["workload"], "global_batch": ctx.world * B, "parallelism": f"dp{ctx.world}"},
              "roofline": {"bound": "tensor", "kernel": "conv_gemm_kernel + wgrad_gemm_kernel (whole step)", "achieved": t["tflops_per_gpu"],
                           "peak": peaks["tflops_burst"], "unit": "TFLOP/s", "frac": t["frac_burst"], "frac_sustained": t["frac_sustained"],
                           "traffic": None, "note": "3x forward algorithmic FLOPs / step time; the step also contains the HBM-bound BN passes"},
              "allreduce_bytes_per_step": t["allreduce_bytes_per_step"], "loss": t["loss"], "clocks": t["clocks"]})
    ctx.close()


def run_pose(args):
    """Secondary line: the reference's REAL inference path (HPE/test.py:289-361): DINOv3Pose forward ->
    head decode -> batched NMS, all on device; `e2e` adds pinned-host images in and the kept detections out."""
    import hpe_b200 as hpe
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --mode pose: no CUDA device (no CPU fallback)")
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    B, H, W, K = args.batch, args.height, args.width, args.joints
    torch.manual_seed(0)
    m = hpe.DINOv3Pose(f"dinov3_convnext_{args.pose_size}", nkpts=(K, 3), ncls=1, device=dev)
    with torch.no_grad():
        for n, p in m.named_parameters():
            if n.endswith("gamma"):
                p.fill_(0.5)
        for mod in m.modules():
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.running_var.uniform_(0.5, 1.5); mod.weight.uniform_(0.5, 1.5)
        for cv in m.head.cv3:
            cv[-1].bias.fill_(0.0)          # sigmoid(0) = 0.5 > conf 0.25: every anchor is an NMS candidate (worst case)
    m.refresh(); m.eval()
    h_img = [torch.rand(B, 3, H, W).pin_memory() for _ in range(2)]
    x = [t.to(dev) for t in h_img]
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731

    def step(i):
        dec = m(x[i & 1])
        return hpe.pose_nms_raw(dec, 0.25, 0.45, nc=1, nkpts=K)

    for i in range(max(args.warmup, 3)):
        step(i)
    torch.cuda.synchronize(dev)
    sampler = ClockSampler(physical_gpu_index(dev.index)); sampler.start()
    a, b = ev(), ev()
    a.record()
    for i in range(args.steps):
        out = step(i)
    b.record(); torch.cuda.synchronize(dev)
    clocks = sampler.stop()
    ms = a.elapsed_time(b) / args.steps
    xd = torch.empty_like(x[0])
    t0 = time.perf_counter()
    for i in range(args.steps):
        xd.copy_(h_img[i & 1], non_blocking=True)
        boxes, scores, labels, kpts, count = hpe.pose_nms_raw(m(xd), 0.25, 0.45, nc=1, nkpts=K)
        n_kept = int(count.sum().item())
    e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    plan = next(iter(m._plans.values()))
    line = {"metric": f"images/sec DINOv3Pose({args.pose_size}) fwd+decode+NMS @{H}x{W} COCO-17", "value": round(B / ms * 1e3, 1),
            "unit": "images/s", "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"dinov3_convnext_{args.pose_size} + adaptor + SPPF + PAN + PoseHead + decode + NMS, batch {B}, "
                                   "random-init weights, every anchor above conf_thres"},
            "e2e": {"value": round(B / e2e_ms * 1e3, 1), "unit": "images/s", "h2d_bytes_per_step": B * 3 * H * W * 4,
                    "d2h_bytes_per_step": 4 * B, "ms_per_step": round(e2e_ms, 4), "kept_last_step": n_kept},
            "gpu_launches": (plan.n_launches + 3) * args.steps, "clocks": clocks}
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="infer", choices=["infer", "train", "pose"])
    ap.add_argument("--pose-size", default="small", choices=["tiny", "small", "base"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--e2e-settle-s", type=float, default=1.0, help="idle time before each timed e2e region (same power state as the headline region)")
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="images per GPU per step")
    ap.add_argument("--height", type=int, default=256)
    ap.add_argument("--width", type=int, default=192)
    ap.add_argument("--joints", type=int, default=17)
    ap.add_argument("--cpu-batch", type=int, default=32, help="images per CPU-baseline step (bounded sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="train: launch eagerly instead of replaying a CUDA graph")
    ap.add_argument("--quick", action="store_true", help="headline + e2e only (no sweep / per-kernel / hires / train sub-benchmarks)")
    ap.add_argument("--no-hires", action="store_true")
    ap.add_argument("--no-train", action="store_true")
    ap.add_argument("--no-pose-train", action="store_true")
    ap.add_argument("--pose-train-batch", type=int, default=64, help="images per GPU of the DINOv3Pose training sub-benchmark")
    ap.add_argument("--hires-batch", type=int, default=64, help="images per GPU of the 384x288 sub-benchmark (configs[3])")
    ap.add_argument("--sustain-s", type=float, default=2.0, help="length of the second (sustained) timed region; 0 = skip")
    ap.add_argument("--layers-log", default="", help="write the per-launch timings behind roofline.per_kernel to this file")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.mode == "train":
        run_train(args)
    elif args.mode == "pose":
        run_pose(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
