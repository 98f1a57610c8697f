"""Host-side mirror of the reference's pose-head post-processing, loss and metrics.

Same names, argument meaning and return shapes as the reference:
  * ``pose_decode``               <- ``PoseHead._inference_pose`` (/root/reference/HPE/models/nn/modules/head.py:185-211)
  * ``non_max_suppression_pose``  <- /root/reference/HPE/test.py:69-185
  * ``ComputeLoss``               <- /root/reference/HPE/loss.py:228-448
  * ``OKS`` / ``PCK`` / ``KeypointLoss`` <- HPE/core/metric/oksap.py:5-74, pck.py:5-104, HPE/core/loss/keypoint.py:5-21
Every call runs the sm_100a kernels of libhpe_b200.so (csrc/pose_head.cu, compute_loss.cu, metrics.cu); tensors must live
on the device and there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

OKS_SIGMAS = torch.tensor([0.26, 0.25, 0.25, 0.35, 0.35, 0.79, 0.79, 0.72, 0.72, 0.62, 0.62, 1.07, 1.07, 0.87, 0.87,
                           0.89, 0.89], dtype=torch.float32)          # HPE/core/metric/__init__.py:13-18


def _ptr_array(tensors):
    return (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


def _int_array(vals):
    return (C.c_int * len(vals))(*[int(v) for v in vals])


def pose_decode(levels, nc, kpt_shape=(17, 3), strides=(8.0, 16.0, 32.0)):
    """Raw head maps ``[B, nc+nk, H_l, W_l]`` (fp32 or bf16, NCHW) -> ``[B, nc+nk, A]`` fp32 (head.py:185-211)."""
    levels = [l.contiguous() for l in levels]
    _lib.require_cuda(*levels)
    B, Cc = levels[0].shape[:2]
    nk = kpt_shape[0] * kpt_shape[1]
    if Cc != nc + nk or any(l.shape[:2] != (B, Cc) or l.dtype != levels[0].dtype for l in levels):
        raise ValueError(f"levels must all be [B, {nc + nk}, H, W] of one dtype")
    A = sum(l.shape[2] * l.shape[3] for l in levels)
    out = torch.empty((B, Cc, A), dtype=torch.float32, device=levels[0].device)
    st = (C.c_float * len(levels))(*[float(s) for s in strides[:len(levels)]])
    with torch.cuda.device(out.device):
        rc = _lib.lib().hpe_pose_decode(_ptr_array(levels), _int_array([l.shape[2] for l in levels]),
                                        _int_array([l.shape[3] for l in levels]), st, len(levels), _lib.dtype_code(levels[0]),
                                        B, nc, kpt_shape[0], kpt_shape[1], _lib.ptr(out), _lib.stream_ptr())
    _lib.check(rc, "hpe_pose_decode")
    return out


def pose_nms_raw(prediction, conf_thres=0.25, iou_thres=0.45, max_det=300, nc=7, nkpts=4):
    """Device-resident result of the NMS: (boxes[B,max_det,4], scores[B,max_det], labels[B,max_det] i32,
    keypoints[B,max_det,nkpts,3], count[B] i32) -- no host sync."""
    _lib.require_cuda(prediction)
    if prediction.dim() != 3:
        raise ValueError("prediction must be [B, nc+nk, A] or [B, A, nc+nk]")
    if prediction.shape[1] >= prediction.shape[2]:           # test.py:87-88: (B, A, C) input is transposed first
        prediction = prediction.transpose(1, 2)
    pred = prediction.float().contiguous()
    B, Cc, A = pred.shape
    dev = pred.device
    scratch = torch.empty(_lib.lib().hpe_pose_nms_scratch_elems(B, A), dtype=torch.float32, device=dev)
    boxes = torch.zeros((B, max_det, 4), dtype=torch.float32, device=dev)
    scores = torch.zeros((B, max_det), dtype=torch.float32, device=dev)
    labels = torch.zeros((B, max_det), dtype=torch.int32, device=dev)
    kpts = torch.zeros((B, max_det, nkpts, 3), dtype=torch.float32, device=dev)
    count = torch.zeros((B,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().hpe_pose_nms(_lib.ptr(pred), B, Cc, A, nc, nkpts, float(conf_thres), float(iou_thres), int(max_det),
                                     _lib.ptr(scratch), _lib.ptr(boxes), _lib.ptr(scores), _lib.ptr(labels), _lib.ptr(kpts),
                                     _lib.ptr(count), _lib.stream_ptr())
    _lib.check(rc, "hpe_pose_nms")
    return boxes, scores, labels, kpts, count


def non_max_suppression_pose(prediction, conf_thres=0.25, iou_thres=0.45, max_det=300, nc=7, nkpts=4):
    """Reference signature and return type (test.py:69): a list of per-image dicts with ``boxes``, ``scores``,
    ``labels`` (int64) and ``keypoints``; the only host sync is the read of the per-image counts."""
    boxes, scores, labels, kpts, count = pose_nms_raw(prediction, conf_thres, iou_thres, max_det, nc, nkpts)
    out = []
    for b, n in enumerate(count.tolist()):
        out.append({"boxes": boxes[b, :n], "scores": scores[b, :n], "labels": labels[b, :n].long(), "keypoints": kpts[b, :n]})
    return out


class _ComputeLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, crit, targets, *preds):
        out, grads = crit._run(preds, targets, any(p.requires_grad for p in preds))
        ctx.grads = grads
        ctx.mark_non_differentiable(out)
        total = out[0:1].clone()
        ctx.parts = out
        return total, out

    @staticmethod
    def backward(ctx, g_total, _g_parts):
        if ctx.grads is None:
            return (None, None) + tuple(None for _ in range(len(ctx.needs_input_grad) - 2))
        return (None, None) + tuple(g * g_total for g in ctx.grads)


class ComputeLoss:
    """``ComputeLoss(model)(preds, targets) -> (total_loss[1], (cls, kpt, vis))`` (loss.py:228-448).

    ``model`` needs ``model.head.nc`` / ``model.head.nk`` like the reference (loss.py:231-234); alternatively pass
    ``nc=``/``nkpts=``.  ``total_loss`` is autograd-connected to ``preds`` (the fused kernel also produces the gradient).
    ``items=False`` returns the three parts as a device tensor instead of Python floats (no host sync)."""

    def __init__(self, model=None, use_focal=True, use_vari_focal=True, nc=None, nkpts=None, items=True):
        if not (use_focal and use_vari_focal):
            raise NotImplementedError("only the reference's default class term (VarifocalLoss) is built")
        if model is not None:
            nc, nkpts = model.head.nc, model.head.nk // 3
        self.nc, self.num_kpts, self.items = int(nc), int(nkpts), items
        self.sigmas = (OKS_SIGMAS / 10.0) if self.num_kpts == 17 else torch.ones(self.num_kpts) * 0.05   # loss.py:250-257
        self._ws = None

    def _run(self, preds, targets, want_grad):
        preds = [p.float().contiguous() for p in preds]
        _lib.require_cuda(*preds)
        dev = preds[0].device
        targets = targets.to(dev, torch.float32).contiguous()
        B = preds[0].shape[0]
        Hs, Ws = _int_array([p.shape[2] for p in preds]), _int_array([p.shape[3] for p in preds])
        n = targets.shape[0]
        if self.sigmas.device != dev:                 # moved once (a host->device copy per call would also break CUDA-graph capture)
            self.sigmas = self.sigmas.to(dev)
        sig = self.sigmas
        lib = _lib.lib()
        need = lib.hpe_compute_loss_workspace_bytes(len(preds), B, Hs, Ws, self.nc, n)
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        out = torch.empty(4, dtype=torch.float32, device=dev)
        grads = [torch.empty_like(p) for p in preds] if want_grad else None
        with torch.cuda.device(dev):
            rc = lib.hpe_compute_loss(_ptr_array(preds), Hs, Ws, len(preds), B, self.nc, self.num_kpts,
                                      _lib.ptr(targets) if n else 0, n, _lib.ptr(sig), _lib.ptr(self._ws), need, _lib.ptr(out),
                                      _ptr_array(grads) if grads else None, 1.0, _lib.stream_ptr())
        _lib.check(rc, "hpe_compute_loss")
        return out, grads

    def __call__(self, preds, targets):
        total, out = _ComputeLossFn.apply(self, targets, *preds)
        parts = out[1:4]
        if self.items:
            c, k, v = parts.tolist()
            return total, (c, k, v)
        return total, parts


class OKS:
    """``OKS(sigmas).compute(gt[N,K,3], pred[N,K,2], areas[N]) -> oks[N]`` (oksap.py:5-74)."""

    def __init__(self, sigmas, in_vis_thres: float = 0.0, eps: float = 1e-10):
        self.sigmas = torch.as_tensor(sigmas, dtype=torch.float32)
        self.in_vis_thres, self.eps = in_vis_thres, eps

    @torch.no_grad()
    def compute(self, gt_keypoints, pred_keypoints, areas):
        gt = gt_keypoints.float().contiguous(); pred = pred_keypoints.float().contiguous(); areas = areas.float().contiguous()
        _lib.require_cuda(gt, pred, areas)
        N, K, D = gt.shape
        assert D == 3 and pred.shape == (N, K, 2) and self.sigmas.shape[0] == K and areas.shape[0] == N
        out = torch.empty(N, dtype=torch.float32, device=gt.device)
        sig = self.sigmas.to(gt.device)
        with torch.cuda.device(gt.device):
            rc = _lib.lib().hpe_oks(_lib.ptr(gt), _lib.ptr(pred), _lib.ptr(areas), _lib.ptr(sig), N, K, self.in_vis_thres,
                                    self.eps, _lib.ptr(out), _lib.stream_ptr())
        _lib.check(rc, "hpe_oks")
        return out


class PCK:
    """``PCK(alpha).compute(gt[N,K,2], pred[N,K,2], visible, normalizer) -> float percentage`` (pck.py:5-104)."""

    def __init__(self, alpha: float = 0.5, eps: float = 1e-6):
        self.alpha, self.eps = alpha, eps

    @torch.no_grad()
    def counts(self, gt_keypoints, pred_keypoints, visible=None, normalizer=None):
        gt = gt_keypoints.float().contiguous(); pred = pred_keypoints.float().contiguous()
        _lib.require_cuda(gt, pred)
        N, K, _ = gt.shape
        dev = gt.device
        vis = None if visible is None else visible.to(dev).bool().to(torch.uint8).contiguous()
        norm = None
        if normalizer is not None:
            norm = torch.as_tensor(normalizer, dtype=torch.float32, device=dev).reshape(-1).contiguous()
            assert norm.numel() in (1, N)
        scratch = torch.empty(2 * N, dtype=torch.int32, device=dev)
        counts = torch.empty(2, dtype=torch.int64, device=dev)
        with torch.cuda.device(dev):
            rc = _lib.lib().hpe_pck(_lib.ptr(gt), _lib.ptr(pred), _lib.ptr(vis), _lib.ptr(norm), 0 if norm is None else norm.numel(),
                                    N, K, self.alpha, self.eps, _lib.ptr(scratch), _lib.ptr(counts), _lib.stream_ptr())
        _lib.check(rc, "hpe_pck")
        return counts

    def compute(self, gt_keypoints, pred_keypoints, visible=None, normalizer=None) -> float:
        correct, total = self.counts(gt_keypoints, pred_keypoints, visible, normalizer).tolist()
        return 0.0 if total == 0 else 100.0 * correct / total


class KeypointLoss(torch.nn.Module):
    """``KeypointLoss(sigmas)(pred_kpts[N,K,2+], gt_kpts[N,K,2+], kpt_mask[N,K], area[N,1]) -> scalar`` (keypoint.py:5-21)."""

    def __init__(self, sigmas):
        super().__init__()
        self.sigmas = torch.as_tensor(sigmas, dtype=torch.float32)

    @torch.no_grad()
    def forward(self, pred_kpts, gt_kpts, kpt_mask, area):
        p = pred_kpts.float().contiguous(); g = gt_kpts.float().contiguous()
        m = kpt_mask.float().contiguous(); a = area.float().reshape(-1).contiguous()
        _lib.require_cuda(p, g, m, a)
        N, K = m.shape
        scratch = torch.empty(N, dtype=torch.float32, device=p.device)
        out = torch.empty(1, dtype=torch.float32, device=p.device)
        sig = self.sigmas.to(p.device)
        with torch.cuda.device(p.device):
            rc = _lib.lib().hpe_keypoint_loss(_lib.ptr(p), p.shape[-1], _lib.ptr(g), g.shape[-1], _lib.ptr(m), _lib.ptr(a),
                                              _lib.ptr(sig), N, K, _lib.ptr(scratch), _lib.ptr(out), _lib.stream_ptr())
        _lib.check(rc, "hpe_keypoint_loss")
        return out[0]
