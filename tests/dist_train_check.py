"""Data-parallel training-step check, run under torchrun on >= 2 GPUs (launched by tests/test_dist_gpu.py):
every rank takes its own shard; after ONE fused step (NCCL all-reduce of the flat gradient buffer + SGD)
  * the SGD momentum buffer equals mean over ranks of the LOCAL gradients + wd * p   (averaged gradients),
  * all ranks hold bit-identical parameters,
  * BatchNorm running statistics stay per-rank (the reference uses plain BatchNorm2d under DDP, never SyncBN).
Prints 'DIST_TRAIN_OK' on rank 0."""
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
    B, H, W, K = 4, 128, 96, 17
    lr, mom, wd = 0.05, 0.9, 1e-4

    def make():
        torch.manual_seed(0)
        m = hpe.HeatmapPose(K, "resnet50", pretrained=False)
        g = torch.Generator().manual_seed(1)
        for n, mod in m.named_modules():
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.weight.data.copy_((torch.rand(mod.weight.shape, generator=g) + 0.5) * (0.1 if n.endswith("bn3") else 1.0))
                mod.bias.data.copy_(torch.randn(mod.bias.shape, generator=g) * 0.1)
        return m.to(dev).train()

    gen = torch.Generator().manual_seed(100 + rank)
    x = torch.rand(B, 3, H, W, generator=gen).to(dev)
    joints = (torch.rand(B, K, 3, generator=gen) * torch.tensor([float(W), float(H), 1.0]) + torch.tensor([0., 0., 0.3])).to(dev).contiguous()

    def note(msg):
        print(f"[rank {rank}] {msg}", file=sys.stderr, flush=True)

    # local gradients through the drop-in autograd path (no collective)
    m0 = make()
    loss0 = hpe.JointsMSELoss()(m0(x), joints, None)
    loss0.backward()
    layout = hpe.train.ParamLayout(m0)
    local_flat = torch.zeros(layout.total, device=dev)
    lv = layout.views(local_flat)
    for n, p in m0.named_parameters():
        lv[n].copy_(p.grad)
    gathered = [torch.empty_like(local_flat) for _ in range(world)]
    dist.all_gather(gathered, local_flat)
    mean_grad = torch.stack(gathered).mean(0)
    note("local gradients gathered")

    m1 = make()
    p_before = torch.cat([p.detach().flatten() for p in m1.parameters()]).clone()
    step = hpe.TrainStep(m1, lr=lr, momentum=mom, weight_decay=wd, cuda_graph=False)
    step.broadcast_parameters(0)
    before = step.flat_param.clone()
    loss = step(x, joints)
    torch.cuda.synchronize()
    note("eager DP step done")
    want_m = mean_grad + wd * before
    err = ((step.flat_mom - want_m).norm() / want_m.norm()).item()
    assert err <= 1e-4, f"rank {rank}: averaged gradient mismatch {err:.3g}"
    assert torch.allclose(step.flat_param, before - lr * step.flat_mom, rtol=1e-6, atol=1e-9)
    ps = [torch.empty_like(step.flat_param) for _ in range(world)]
    dist.all_gather(ps, step.flat_param)
    for q in ps[1:]:
        assert torch.equal(q, ps[0]), "ranks hold different parameters after the step"
    rm = m1.backbone.bn1.running_mean.clone()
    rms = [torch.empty_like(rm) for _ in range(world)]
    dist.all_gather(rms, rm)
    assert not torch.equal(rms[0], rms[1]), "running statistics should be per-rank (different shards)"
    note("parameters identical, running stats per rank")
    # graph replay: the compute launches are captured, the all-reduce + optimizer kernel follow the replay eagerly
    step_g = hpe.TrainStep(make(), lr=lr, momentum=mom, weight_decay=wd, cuda_graph=True)
    for _ in range(3):
        lg = step_g(x, joints)
    torch.cuda.synchronize()
    assert step_g.graph is not None and torch.isfinite(lg)
    ps = [torch.empty_like(step_g.flat_param) for _ in range(world)]
    dist.all_gather(ps, step_g.flat_param)
    for q in ps[1:]:
        assert torch.equal(q, ps[0]), "graph replay: ranks diverged"
    dist.barrier(device_ids=[local])
    if rank == 0:
        print(f"DIST_TRAIN_OK world={world} loss={loss.item():.6f} avg-grad rel err {err:.2e}", flush=True)
    torch.cuda.synchronize()
    # clean teardown: the captured graph holds no collective (TrainStep keeps the all-reduce outside it), so the
    # communicator can be destroyed normally -- a long-lived trainer that rebuilds its graph must be able to do this
    step_g.close(); step.close()
    dist.barrier(device_ids=[local])
    dist.destroy_process_group()
    note("process group destroyed")
    if rank == 0:
        print("DIST_TEARDOWN_OK", flush=True)


if __name__ == "__main__":
    try:
        main()
    except BaseException:                      # a failed rank must not leave its peer waiting in a collective
        import traceback
        traceback.print_exc()
        sys.stderr.flush()
        os._exit(1)
