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

from .decode import affine_from_center_scale, get_final_preds_affine, get_final_preds_xyv


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


def gather_keypoints_equal(preds: torch.Tensor, maxvals: torch.Tensor = None, out: torch.Tensor = None, group=None):
    """Fast path when every rank holds the same number of images (the benchmark's weak-scaling case): no size
    exchange, one ``all_gather_into_tensor``.  ``preds`` may already be the packed ``xyv[b,K,3]`` the decode kernel
    writes (``maxvals=None``): then nothing is concatenated either."""
    local = preds if maxvals is None else torch.cat([preds, maxvals], dim=2).contiguous()
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    if out is None:
        out = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, local, group=group)
    return out


class ResultGather:
    """The inference path's only collective: every rank's packed ``xyv[b,K,3]`` rows -> ``[world*b,K,3]``
    on every rank, issued on a SIDE stream so the compute stream goes straight on to the next batch's forward (the latency
    of one small all-gather per step otherwise sits on the critical path of every rank, and the step time is the max over
    ranks).  Two result slots: slot ``i & 1`` of step ``i`` is rewritten by step ``i + 2``, whose decode first waits for the
    gather of step ``i``."""

    def __init__(self, B, K, device, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.device = torch.device(device)
        self.xyv = [torch.empty((B, K, 3), dtype=torch.float32, device=self.device) for _ in range(2)]
        self.out = [torch.empty((self.world * B, K, 3), dtype=torch.float32, device=self.device) for _ in range(2)] \
            if self.world > 1 else self.xyv
        self.side = torch.cuda.Stream(self.device) if self.world > 1 else None
        self.decoded = [torch.cuda.Event() for _ in range(2)]
        self.gathered = [torch.cuda.Event() for _ in range(2)]
        self.i = 0

    def slot(self):
        """Packed decode output of the coming step (call before the decode kernel: waits for the slot's previous gather)."""
        s = self.i & 1
        if self.world > 1:
            torch.cuda.current_stream(self.device).wait_event(self.gathered[s])
        return self.xyv[s]

    def gather(self):
        """Issue the all-gather of the slot just decoded (side stream) and return the buffer it fills; ``wait()`` before
        reading it on the compute stream."""
        s = self.i & 1
        self.i += 1
        if self.world == 1:
            return self.out[s]
        main = torch.cuda.current_stream(self.device)
        self.decoded[s].record(main)
        with torch.cuda.stream(self.side):
            self.side.wait_event(self.decoded[s])
            dist.all_gather_into_tensor(self.out[s], self.xyv[s], group=self.group)
            self.gathered[s].record(self.side)
        return self.out[s]

    def wait(self):
        if self.world > 1:
            torch.cuda.current_stream(self.device).wait_stream(self.side)


class PoseEstimator:
    """``estimator(images, center, scale) -> preds[B,K,2] (image px), maxvals[B,K,1]``.

    * device tensors in -> device tensors out (``__call__``);
    * ``infer_host``: pinned host tensors in, pinned host tensors out (H2D + forward + decode + D2H);
    * ``stream``: the same over an iterable of host batches: the H2D copy of batch i+1 (side stream, double-buffered
      device staging) and the D2H read-back of batch i-1's keypoints both overlap batch i's kernels.

    Host images are ``uint8`` [B,3,H,W] -- the reference's real input format (``imgs.to(device).float() / 255.0``,
    /root/reference/HPE/train.py:117-120; ``preprocess_simple``, HPE/test.py:16-29), with the /255 fused into the stem's
    layout kernel -- or ``float32``.
    """

    def __init__(self, model, device="cuda"):
        self.model = model.eval()
        self.device = torch.device(device)
        self._stage = {}
        self._copy_stream = None
        self._d2h_stream = None

    def _heatmaps(self, images):
        fs = getattr(self.model, "forward_static", None)      # zero-copy: the plan's persistent heatmap buffer
        return fs(images) if fs is not None else self.model(images)

    def __call__(self, images: torch.Tensor, center=None, scale=None, affine: torch.Tensor = None):
        hm = self._heatmaps(images)
        if affine is None:
            affine = affine_from_center_scale(torch.as_tensor(center), torch.as_tensor(scale), hm.shape[2:])
            affine = affine.to(hm.device, non_blocking=True).contiguous()
        return get_final_preds_affine(hm, affine)

    def xyv(self, images: torch.Tensor, affine: torch.Tensor, out: torch.Tensor = None):
        """Device path into the packed ``[B,K,3] = (x, y, maxval)`` layout (what ``ResultGather`` ships)."""
        return get_final_preds_xyv(self._heatmaps(images), affine, out)

    # ---- host-buffer paths ---------------------------------------------------------------------------
    def _buffers(self, shape, slot, dtype=torch.float32):
        key = (tuple(shape), slot, dtype)
        b = self._stage.get(key)
        if b is None:
            B = shape[0]
            K = self.model.njoints
            h_xyv = torch.empty((B, K, 3), dtype=torch.float32).pin_memory()
            b = {"img": torch.empty(shape, dtype=dtype, device=self.device),      # uint8 images: 4x less H2D, /255 fused in the stem
                 "aff": torch.empty((B, 2, 3), dtype=torch.float32, device=self.device),
                 "xyv": torch.empty((B, K, 3), dtype=torch.float32, device=self.device),
                 "h_xyv": h_xyv, "preds": h_xyv[:, :, :2], "maxvals": h_xyv[:, :, 2:],
                 "ready": torch.cuda.Event(), "computed": torch.cuda.Event(), "done": torch.cuda.Event()}
            self._stage[key] = b
        return b

    def infer_host(self, images: torch.Tensor, affine: torch.Tensor):
        """One batch, fully synchronous: returns pinned host ``preds, maxvals`` (views of one packed buffer)."""
        b = self._buffers(images.shape, 0, images.dtype)
        b["img"].copy_(images, non_blocking=True)
        b["aff"].copy_(affine, non_blocking=True)
        self.xyv(b["img"], b["aff"], b["xyv"])
        b["h_xyv"].copy_(b["xyv"], non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return b["preds"], b["maxvals"]

    def stream(self, batches):
        """Generator over ``(images_host, affine_host)`` pairs -> ``(preds_host, maxvals_host)`` in order.

        Three streams: the H2D copy of batch i+1 and the D2H read-back of batch i run beside batch i+1's kernels, and the
        host only ever waits for a result that was launched one batch earlier (two-deep ring: batch i's keypoints are
        yielded after batch i+1 has been enqueued), so the device never idles on the host.  The yielded tensors are
        views of a pinned slot that is rewritten two batches later: consume (or copy) them before asking for the batch
        after the next."""
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(self.device)
            self._d2h_stream = torch.cuda.Stream(self.device)
        cs, ds = self._copy_stream, self._d2h_stream
        main = torch.cuda.current_stream(self.device)
        it = iter(batches)

        def upload(item, slot):
            images, affine = item
            b = self._buffers(images.shape, slot, images.dtype)
            cs.wait_event(b["computed"])             # the forward that last read this slot's device image has finished
            with torch.cuda.stream(cs):
                b["img"].copy_(images, non_blocking=True)
                b["aff"].copy_(affine, non_blocking=True)
                b["ready"].record(cs)
            return b

        def launch(b):
            main.wait_event(b["ready"])
            main.wait_event(b["done"])               # the slot's previous keypoints have left the device buffer
            self.xyv(b["img"], b["aff"], b["xyv"])
            b["computed"].record(main)
            with torch.cuda.stream(ds):
                ds.wait_event(b["computed"])
                b["h_xyv"].copy_(b["xyv"], non_blocking=True)
                b["done"].record(ds)

        try:
            cur = upload(next(it), 0)
        except StopIteration:
            return
        slot, prev = 0, None
        while cur is not None:
            try:
                nxt = upload(next(it), slot ^ 1)
            except StopIteration:
                nxt = None
            launch(cur)
            if prev is not None:
                prev["done"].synchronize()           # launched one batch ago: normally already complete
                yield prev["preds"], prev["maxvals"]
            prev, cur, slot = cur, nxt, slot ^ 1
        prev["done"].synchronize()
        yield prev["preds"], prev["maxvals"]
