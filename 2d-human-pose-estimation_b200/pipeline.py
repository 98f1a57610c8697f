"""User-facing inference path: images -> model forward -> heatmap decode -> keypoints.

Mirrors what a caller of the reference does around ``output = model(data)`` in ``Trainer.test`` /
``_epoch`` (/root/reference/HPE/engine/trainer.py:63-65, 86-90, 110-113) and in the inference script
(HPE/test.py:333-361: preprocess -> ``model(img)`` -> decode -> scale to image coordinates): host batch
in, keypoints in image pixels out.  Data parallelism follows the reference (one process per GPU under
torchrun, batch sharded by rank: pose/pose/core/builder.py:92-123); inference needs only a result gather.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .decode import affine_from_center_scale, get_final_preds_affine


def shard_range(n: int, rank: int, world: int):
    """Contiguous, near-even shard of ``n`` images for ``rank`` (rank r takes [r*n/W, (r+1)*n/W))."""
    lo = (n * rank) // world
    hi = (n * (rank + 1)) // world
    return lo, hi


def gather_keypoints(preds: torch.Tensor, maxvals: torch.Tensor, group=None):
    """All-gather the per-rank ``preds[b,K,2]`` / ``maxvals[b,K,1]`` into ``[sum b, K, 3]`` on every rank.
    One collective of (b*K*3*4) bytes per rank (52 KB for 256x17); equal shard sizes use the single-tensor
    NCCL all-gather, ragged shards fall back to a padded gather."""
    local = torch.cat([preds, maxvals], dim=2).contiguous()
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    sizes = [torch.zeros(1, dtype=torch.int64, device=local.device) for _ in range(world)]
    n_local = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    dist.all_gather(sizes, n_local, group=group)
    sizes = [int(s.item()) for s in sizes]
    if len(set(sizes)) == 1:
        out = torch.empty((world * sizes[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local, group=group)
        return out
    m = max(sizes)
    pad = torch.zeros((m,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    return torch.cat([b[:s] for b, s in zip(bufs, sizes)], dim=0)


def gather_keypoints_equal(preds: torch.Tensor, maxvals: torch.Tensor, out: torch.Tensor = None, group=None):
    """Fast path when every rank holds the same number of images (the benchmark's weak-scaling case):
    no size exchange, one ``all_gather_into_tensor``."""
    local = torch.cat([preds, maxvals], dim=2).contiguous()
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    if out is None:
        out = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, local, group=group)
    return out


class PoseEstimator:
    """``estimator(images, center, scale) -> preds[B,K,2] (image px), maxvals[B,K,1]``.

    * device tensors in -> device tensors out (``__call__``);
    * ``infer_host``: pinned host tensors in, pinned host tensors out (H2D + forward + decode + D2H);
    * ``stream``: the same over an iterable of host batches with the next batch's H2D copy overlapped with the
      current batch's kernels (second CUDA stream, double-buffered device staging).
    """

    def __init__(self, model, device="cuda"):
        self.model = model.eval()
        self.device = torch.device(device)
        self._stage = {}
        self._copy_stream = None

    def __call__(self, images: torch.Tensor, center=None, scale=None, affine: torch.Tensor = None):
        hm = self.model(images)
        if affine is None:
            affine = affine_from_center_scale(torch.as_tensor(center), torch.as_tensor(scale), hm.shape[2:])
            affine = affine.to(hm.device, non_blocking=True).contiguous()
        return get_final_preds_affine(hm, affine)

    # ---- host-buffer paths ---------------------------------------------------------------------------
    def _buffers(self, shape, slot, dtype=torch.float32):
        key = (tuple(shape), slot, dtype)
        b = self._stage.get(key)
        if b is None:
            B = shape[0]
            K = self.model.njoints
            b = {"img": torch.empty(shape, dtype=dtype, device=self.device),      # uint8 images: 4x less H2D, /255 fused in the stem
                 "aff": torch.empty((B, 2, 3), dtype=torch.float32, device=self.device),
                 "preds": torch.empty((B, K, 2), dtype=torch.float32).pin_memory(),
                 "maxvals": torch.empty((B, K, 1), dtype=torch.float32).pin_memory(),
                 "ready": torch.cuda.Event(), "done": torch.cuda.Event()}
            self._stage[key] = b
        return b

    def infer_host(self, images: torch.Tensor, affine: torch.Tensor):
        """One batch, fully synchronous: returns pinned host ``preds, maxvals``."""
        b = self._buffers(images.shape, 0, images.dtype)
        b["img"].copy_(images, non_blocking=True)
        b["aff"].copy_(affine, non_blocking=True)
        preds, maxvals = self(b["img"], affine=b["aff"])
        b["preds"].copy_(preds, non_blocking=True)
        b["maxvals"].copy_(maxvals, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return b["preds"], b["maxvals"]

    def stream(self, batches):
        """Generator over ``(images_host, affine_host)`` pairs -> ``(preds_host, maxvals_host)`` in order.
        The H2D copy of batch i+1 runs on a side stream while batch i computes."""
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(self.device)
        cs = self._copy_stream
        main = torch.cuda.current_stream(self.device)
        it = iter(batches)

        def upload(item, slot):
            images, affine = item
            b = self._buffers(images.shape, slot, images.dtype)
            cs.wait_event(b["done"])                 # previous use of this slot's device buffers has finished
            with torch.cuda.stream(cs):
                b["img"].copy_(images, non_blocking=True)
                b["aff"].copy_(affine, non_blocking=True)
                b["ready"].record(cs)
            return b

        try:
            cur = upload(next(it), 0)
        except StopIteration:
            return
        slot = 0
        while cur is not None:
            try:
                nxt = upload(next(it), slot ^ 1)
            except StopIteration:
                nxt = None
            main.wait_event(cur["ready"])
            preds, maxvals = self(cur["img"], affine=cur["aff"])
            cur["preds"].copy_(preds, non_blocking=True)
            cur["maxvals"].copy_(maxvals, non_blocking=True)
            cur["done"].record(main)
            cur["done"].synchronize()                # the step's result is read on the host
            yield cur["preds"], cur["maxvals"]
            cur, slot = nxt, slot ^ 1
