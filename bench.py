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
# ncu (profiles/r01_launches_step_v4.csv): DRAM bytes of one B=256 256x192 step's conv launches / 52 launches
NCU_DRAM_BYTES_PER_CONV_LAUNCH = round((6703 + 2672 + 407 + 204) * 1e6 / 52)


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm": float(p["hbm_gbs"]), "tflops": float(p["bf16_tflops_sustained"]), "src": "measured"}
    except Exception:
        return {"hbm": 6650.0, "tflops": 1400.0, "src": "fallback"}


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
    x = torch.rand(batch, 3, H, W, generator=torch.Generator().manual_seed(0))
    center, scale = hr.synth_center_scale(batch)
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            hm = model(x).numpy()
            preds, maxvals = hr.get_final_preds(hm, center, scale)
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    assert np.isfinite(preds).all()
    mean = sum(times) / len(times)
    return batch / mean, mean, torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    H, W = args.height, args.width
    batch = args.cpu_batch
    steps, warmup = min(args.steps, 10), min(args.warmup, 2)
    ips, mean, nthr = cpu_forward_decode(batch, H, W, args.joints, steps, max(warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": round(ips, 3), "unit": "images/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": max(warmup, 1), "ms_per_step": round(mean * 1e3, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"resnet50+deconv-head heatmap fwd+decode, {H}x{W}, K={args.joints}, CPU sample of {batch} images/step "
                               f"(GPU arm: {args.batch}/GPU), random-init weights"},
        "cpu_baseline": {"value": round(ips, 3), "unit": "images/s", "cores": nthr, "kind": "port",
                         "sample": f"{steps} steps x {batch} images, torch fp32 {nthr} threads + numpy decode"},
        "e2e": {"value": round(ips, 3), "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# --------------------------------------------------------------------------------------------- GPU arm
def run_b200(args):
    import torch.distributed as dist
    import hpe_b200 as hpe

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    hpe._lib.check(hpe._lib.lib().hpe_check_device(), "hpe_check_device")

    B, H, W, K = args.batch, args.height, args.width, args.joints
    torch.manual_seed(0)
    model = hpe.HeatmapPose(K, "resnet50", pretrained=False)
    g = torch.Generator().manual_seed(1)
    for m in model.modules():               # random-init net with randomised BN so it is not degenerate
        if isinstance(m, torch.nn.BatchNorm2d):
            m.weight.data.copy_(torch.rand(m.weight.shape, generator=g) + 0.5)
            m.bias.data.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
            m.running_mean.copy_(torch.randn(m.running_mean.shape, generator=g) * 0.1)
            m.running_var.copy_(torch.rand(m.running_var.shape, generator=g) + 0.5)
    model = model.to(dev).eval()
    est = hpe.PoseEstimator(model, dev)

    gen = torch.Generator().manual_seed(100 + rank)
    h_img = [torch.rand(B, 3, H, W, generator=gen).pin_memory() for _ in range(2)]
    x = [t.to(dev) for t in h_img]
    from_center = torch.rand(B, 2, generator=gen) * torch.tensor([640.0, 480.0])
    sw = torch.rand(B, 1, generator=gen) * 2 + 0.5
    h_aff = hpe.affine_from_center_scale(from_center, torch.cat([sw, sw * 4 / 3], 1), (H // 4, W // 4)).pin_memory()
    aff = h_aff.to(dev)
    gathered = torch.empty((world * B, K, 3), dtype=torch.float32, device=dev) if world > 1 else None

    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    marks = []

    def step(i, record=False):
        if record:
            e0 = ev(); e0.record()
        hm = model(x[i & 1])
        if record:
            e1 = ev(); e1.record()
        preds, maxvals = hpe.get_final_preds_affine(hm, aff)
        if record:
            e2 = ev(); e2.record()
            marks.append((e0, e1, e2))
        if world > 1:
            return hpe.gather_keypoints_equal(preds, maxvals, gathered)
        return preds

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize(dev)

    for i in range(max(args.warmup, 3)):
        step(i)
    barrier()
    sampler = ClockSampler(physical_gpu_index(local)) if rank == 0 else None
    if sampler:
        sampler.start()
    start, end = ev(), ev()
    barrier()
    start.record()
    for i in range(args.steps):
        step(i, record=True)
    end.record()
    barrier()
    clocks = sampler.stop() if sampler else None
    ms_total = start.elapsed_time(end)
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    fwd_ms = sum(a.elapsed_time(b) for a, b, _ in marks) / len(marks)
    dec_ms = sum(b.elapsed_time(c) for _, b, c in marks) / len(marks)

    # ---- e2e: pinned host images in, keypoints read on the host every step (public API, H2D overlapped)
    def host_batches(n):
        for i in range(n):
            yield h_img[i & 1], h_aff
    for _ in est.stream(host_batches(3)):
        pass
    barrier()
    t0 = time.perf_counter()
    e_start, e_end = ev(), ev()
    e_start.record()
    checksum = 0.0
    for preds_h, maxvals_h in est.stream(host_batches(args.steps)):
        checksum += float(maxvals_h[0, 0, 0])
    e_end.record()
    barrier()
    e2e_ms = e_start.elapsed_time(e_end)
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3
    te = torch.tensor([max(e2e_ms, 0.0)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_ms = float(te.item())

    # ---- decode at an HBM-saturating size (BASELINE configs[4] sweep point: K=17, 64x48, fp32, 16384 maps x 17 joints = 3.4 GB,
    # far beyond the 126 MB L2): the "decode HBM GB/s vs peak" half of the metric.  The in-step decode above is 53 MB.
    sweep = None
    if rank == 0:
        SB = 16384
        big = torch.rand(SB, K, H // 4, W // 4, device=dev)
        big_aff = torch.rand(SB, 2, 3, device=dev)
        for _ in range(3):
            hpe.get_final_preds_affine(big, big_aff)
        torch.cuda.synchronize(dev)
        ts = []
        for _ in range(10):
            a, b = ev(), ev()
            a.record(); hpe.get_final_preds_affine(big, big_aff); b.record(); torch.cuda.synchronize(dev)
            ts.append(a.elapsed_time(b))
        sweep_ms = sum(ts) / len(ts)
        sweep_bytes = big.numel() * 4 + SB * 24 + SB * K * 16
        sweep = {"B": SB, "ms": sweep_ms, "gbs": sweep_bytes / (sweep_ms * 1e-3) / 1e9}
        del big, big_aff

    # ---- same e2e with uint8 host images (the reference's real input, HPE/train.py:117-120: `imgs.float() / 255.0` fused into the stem)
    h_u8 = [(t * 255).to(torch.uint8).pin_memory() for t in h_img]

    def host_batches_u8(n):
        for i in range(n):
            yield h_u8[i & 1], h_aff
    for _ in est.stream(host_batches_u8(3)):
        pass
    barrier()
    u_start, u_end = ev(), ev()
    u_start.record()
    for preds_h, maxvals_h in est.stream(host_batches_u8(args.steps)):
        checksum += float(maxvals_h[0, 0, 0])
    u_end.record()
    barrier()
    tu = torch.tensor([u_start.elapsed_time(u_end)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tu, op=dist.ReduceOp.MAX)
    e2e_u8_ms = float(tu.item())

    if rank == 0:
        peaks = load_peaks()
        plan = next(iter(model._plans.values()))
        launches_per_step = plan.n_launches + 1
        img_scale = (H * W) / (256 * 192)
        flops_step = GFLOP_PER_IMG_256x192 * 1e9 * img_scale * B
        dec_bytes = B * K * (H // 4) * (W // 4) * 4 + B * 24 + B * K * 16
        total_imgs = world * B * args.steps
        line = {
            "metric": METRIC, "value": round(total_imgs / ms_total * 1e3, 1), "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms_total / args.steps, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"resnet50+deconv-head heatmap fwd+decode, {H}x{W}, K={K}, batch {B}/GPU, bf16 (fp32 accum), "
                                   "random-init weights (BASELINE configs[1])",
                       "global_batch": world * B, "parallelism": f"dp{world}",
                       "l2": "inputs (151 MB/step, 2 alternating batches) and every activation exceed the 126 MB L2; no flush needed"},
            "roofline": {"bound": "tensor", "kernel": "conv_gemm_kernel + conv_yhalo_kernel (52 launches/step: stem with the max-pool in its epilogue, trunk with the projection "
                                                      "shortcuts fused into conv3, deconv head with the final 1x1 fused into its last stage; 256-wide tiles as cta_group::2 pairs)",
                         "achieved": round(flops_step / (fwd_ms * 1e-3) / 1e12, 2), "peak": peaks["tflops"], "unit": "TFLOP/s",
                         "frac": round(flops_step / (fwd_ms * 1e-3) / 1e12 / peaks["tflops"], 4),
                         "traffic": NCU_DRAM_BYTES_PER_CONV_LAUNCH if (B, H, W, K) == (256, 256, 192, 17) else None,
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum averaged over the step's 52 conv launches "
                                           "(9.99 GB / step vs 11.8 GB of per-layer algorithmic input+output+residual+weight bytes: part of each layer's input is still in L2, no re-reads), profiles/r01_launches_step_v4.csv",
                         "peak_source": f"{peaks['src']} bf16_tflops_sustained (kernel timed inside a long step)",
                         "ms_forward": round(fwd_ms, 4),
                         "note": "algorithmic FLOPs of the step / CUDA-event time of the forward graph (incl. preprocess; the max-pool runs in the stem's epilogue)"},
            "decode_roofline": {"bound": "hbm", "kernel": "decode_kernel", "achieved": round(sweep["gbs"], 1), "peak": peaks["hbm"],
                                "unit": "GB/s", "frac": round(sweep["gbs"] / peaks["hbm"], 4), "traffic": None,
                                "workload": f"decode sweep point: {sweep['B']} x {K} maps of {H // 4}x{W // 4} fp32 (3.4 GB, > L2), "
                                            f"{sweep['ms']:.3f} ms/launch, mean of 10 launches (CUDA events)",
                                "algorithmic_bytes_per_map": (H // 4) * (W // 4) * 4 + 16,
                                "in_step": {"ms_decode": round(dec_ms, 4), "achieved": round(dec_bytes / (dec_ms * 1e-3) / 1e9, 1),
                                            "note": "the step's own 53 MB of heatmaps is launch/latency-bound, not bandwidth-bound"}},
            "e2e": {"value": round(total_imgs / e2e_ms * 1e3, 1), "unit": "images/s",
                    "h2d_bytes_per_step": B * 3 * H * W * 4 + B * 24, "d2h_bytes_per_step": B * K * 3 * 4,
                    "ms_per_step": round(e2e_ms / args.steps, 4), "wall_ms_per_step": round(e2e_wall_ms / args.steps, 4),
                    "api": "PoseEstimator.stream (pinned host fp32 images -> H2D overlapped on a side stream -> forward -> decode -> D2H)",
                    "uint8_images": {"value": round(total_imgs / e2e_u8_ms * 1e3, 1), "h2d_bytes_per_step": B * 3 * H * W + B * 24,
                                     "note": "same call with uint8 host images (/255 fused into the stem, HPE/train.py:117-120)"}},
            "gpu_launches": launches_per_step * args.steps,
            "gpu_launches_per_step": launches_per_step,
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            ips, mean, nthr = cpu_forward_decode(args.cpu_batch, H, W, K, 3, 1)
            line["cpu_baseline"] = {"value": round(ips, 3), "unit": "images/s", "cores": nthr, "kind": "port",
                                    "sample": f"3 steps x {args.cpu_batch} images of the same workload (torch fp32 ResNet-50+head on "
                                              f"{nthr} threads + numpy decode), {mean:.2f} s/step"}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_train(args):
    """BASELINE configs[2]: data-parallel training step (forward with batch-stat BN, fused Gaussian-target MSE loss,
    backward through dgrad/wgrad, ONE NCCL all-reduce of the flat fp32 gradient buffer, SGD) at N GPUs."""
    import torch.distributed as dist
    import hpe_b200 as hpe

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --mode train: no CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    B, H, W, K = args.batch, args.height, args.width, args.joints
    torch.manual_seed(0)
    model = hpe.HeatmapPose(K, "resnet50", pretrained=False)
    g = torch.Generator().manual_seed(1)
    for n, m in model.named_modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.weight.data.copy_((torch.rand(m.weight.shape, generator=g) + 0.5) * (0.1 if n.endswith("bn3") else 1.0))
            m.bias.data.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
    model = model.to(dev).train()
    step = hpe.TrainStep(model, lr=1e-3, momentum=0.9, weight_decay=1e-4, cuda_graph=not args.no_graph)
    step.broadcast_parameters(0)
    gen = torch.Generator().manual_seed(100 + rank)
    x = [torch.rand(B, 3, H, W, generator=gen).to(dev) for _ in range(2)]
    joints = torch.rand(B, K, 3, generator=gen) * torch.tensor([float(W), float(H), 1.0]) + torch.tensor([0.0, 0.0, 0.3])
    joints = joints.to(dev).contiguous()

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize(dev)

    for i in range(max(args.warmup, 3)):
        step(x[i & 1], joints)
    barrier()
    sampler = ClockSampler(physical_gpu_index(local)) if rank == 0 else None
    if sampler:
        sampler.start()
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    start, end = ev(), ev()
    barrier()
    start.record()
    for i in range(args.steps):
        loss = step(x[i & 1], joints)
    end.record()
    barrier()
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([start.elapsed_time(end)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    if rank == 0:
        peaks = load_peaks()
        flops_step = 3 * GFLOP_PER_IMG_256x192 * 1e9 * (H * W) / (256 * 192) * B      # fwd + dgrad + wgrad
        ms = ms_total / args.steps
        line = {
            "metric": "images/sec train step @256x192 COCO-17", "value": round(world * B * args.steps / ms_total * 1e3, 1),
            "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"resnet50+deconv-head data-parallel training step (batch-stat BN, fused Gaussian-target JointsMSE, "
                                   f"dgrad+wgrad, NCCL all-reduce of {step.layout.total * 4 / 1e6:.0f} MB fp32 grads, SGD momentum), "
                                   f"{H}x{W}, K={K}, batch {B}/GPU (BASELINE configs[2])",
                       "global_batch": world * B, "parallelism": f"dp{world}"},
            "roofline": {"bound": "tensor", "kernel": "conv_gemm_kernel + wgrad_gemm_kernel (whole step)",
                         "achieved": round(flops_step / (ms * 1e-3) / 1e12, 2), "peak": peaks["tflops"], "unit": "TFLOP/s",
                         "frac": round(flops_step / (ms * 1e-3) / 1e12 / peaks["tflops"], 4), "traffic": None,
                         "note": "3x forward algorithmic FLOPs / step time; the step also contains the HBM-bound BN passes"},
            "loss": float(loss.item()), "clocks": clocks,
        }
        emit(line)
    if world > 1:
        # a live captured graph that contains the NCCL all-reduce has been seen to block communicator teardown:
        # drop the graph first, and leave without the teardown once every rank is done
        step.graph = None
        barrier()
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(0)


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
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="images per GPU per step")
    ap.add_argument("--height", type=int, default=256)
    ap.add_argument("--width", type=int, default=192)
    ap.add_argument("--joints", type=int, default=17)
    ap.add_argument("--cpu-batch", type=int, default=32, help="images per CPU-baseline step (bounded sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="train mode: launch eagerly instead of replaying a CUDA graph")
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
