"""Data-parallel DINOv3Pose training-step check, run under torchrun on >= 2 GPUs (launched by tests/test_dist_gpu.py): every rank
takes its own shard; after ONE PoseTrainStep (ComputeLoss, backward, NCCL all-reduce of the flat gradient buffer, clip + AdamW)
  * AdamW's first moment equals (1 - beta1) * clip * mean over ranks of the LOCAL gradients,
  * all ranks hold bit-identical parameters, the frozen backbone is untouched,
  * BatchNorm running statistics stay per-rank (plain BatchNorm2d under DDP, pose/pose/core/builder.py:92-123).
Prints 'DIST_POSETRAIN_OK' on rank 0."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import hpe_b200 as hpe  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    B = 4

    def make():
        torch.manual_seed(0)
        return hpe.DINOv3Pose("dinov3_convnext_tiny", nkpts=(17, 3), ncls=1, device=dev).train()

    gen = torch.Generator().manual_seed(100 + rank)
    x = torch.rand(B, 3, 128, 96, generator=gen).to(dev)
    tg = torch.zeros(B, 6 + 17 * 3)
    tg[:, 0] = torch.arange(B); tg[:, 2:4] = 0.3 + 0.4 * torch.rand(B, 2, generator=gen); tg[:, 4:6] = 0.3
    kp = torch.rand(B, 17, 3, generator=gen); kp[..., :2] = 0.2 + 0.6 * kp[..., :2]; kp[..., 2] = 2.0
    tg[:, 6:] = kp.reshape(B, -1)
    tg = tg.to(dev)

    # local gradients through plain autograd (no collective)
    m0 = make()
    hpe.ComputeLoss(m0)(m0(x), tg)[0].backward()
    local_flat = torch.cat([p.grad.flatten() for p in m0.parameters() if p.requires_grad])
    gathered = [torch.empty_like(local_flat) for _ in range(world)]
    dist.all_gather(gathered, local_flat)
    mean_grad = torch.stack(gathered).mean(0)

    m1 = make()
    step = hpe.PoseTrainStep(m1, lr=1e-3, weight_decay=1e-2, max_grad_norm=10.0)
    step.broadcast_parameters(0)
    bb0 = [p.detach().clone() for p in m1.backbone.parameters()]
    step(x, tg)
    torch.cuda.synchronize()
    clip = min(1.0, 10.0 / (mean_grad.norm().item() + 1e-6))
    got = torch.cat([step.exp_avg.narrow(0, step.layout.offsets[n], p.numel()) for n, p in m1.named_parameters() if p.requires_grad])
    want = 0.1 * clip * mean_grad
    err = ((got - want).norm() / want.norm()).item()
    assert err <= 1e-3, f"rank {rank}: averaged gradient mismatch {err:.3g}"
    ps = [torch.empty_like(step.flat_param) for _ in range(world)]
    dist.all_gather(ps, step.flat_param)
    for q in ps[1:]:
        assert torch.equal(q, ps[0]), "ranks hold different parameters after the step"
    assert all(torch.equal(a, b) for a, b in zip(bb0, m1.backbone.parameters()))
    rm = m1.sppf.conv1.bn.running_mean.clone()
    rms = [torch.empty_like(rm) for _ in range(world)]
    dist.all_gather(rms, rm)
    assert not torch.equal(rms[0], rms[1]), "running statistics should be per-rank (different shards)"
    dist.barrier()
    if rank == 0:
        print("DIST_POSETRAIN_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
