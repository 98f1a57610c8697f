"""CPU oracle (NumPy, float32) for the stages the reference implements around its pose head.

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.  Unlike ``heatmap_ref.py`` these functions follow
code that EXISTS in the reference, so parity is **pinned**: ``tests/golden/make_golden_detpose.py`` runs the
reference's own functions (imported from /root/reference) on seeded inputs and stores their outputs in
``tests/golden/detpose_golden.npz``; ``tests/test_detpose_oracle_cpu.py`` checks every function below
against those vectors.

Rows of SURVEY.md §8(a) covered here:
  a7  ``PoseHead._inference_pose`` / ``kpts_decode`` / ``make_anchors``
      (HPE/models/nn/modules/head.py:185-211, HPE/models/utils.py:15-27)
  a8  ``non_max_suppression_pose`` (HPE/test.py:69-185) incl. torchvision's greedy NMS
  a9  ``ComputeLoss.__call__`` and helpers (HPE/loss.py:9-27, 41-84, 171-222, 277-448)
  a10 ``KeypointLoss`` (HPE/core/loss/keypoint.py:5-21), ``OKS.compute`` (HPE/core/metric/oksap.py:32-74),
      ``PCK.compute`` (HPE/core/metric/pck.py:51-104)
"""
from __future__ import annotations

import numpy as np

F32 = np.float32

OKS_SIGMAS = np.array([0.26, 0.25, 0.25, 0.35, 0.35, 0.79, 0.79, 0.72, 0.72, 0.62, 0.62, 1.07, 1.07, 0.87, 0.87,
                       0.89, 0.89], dtype=F32)          # HPE/core/metric/__init__.py:13-18


def _f32(a):
    if hasattr(a, "detach"):
        a = a.detach().to("cpu").float().numpy()
    return np.ascontiguousarray(a, dtype=F32)


def sigmoid(x):
    x = _f32(x)
    return (F32(1.0) / (F32(1.0) + np.exp(-x))).astype(F32)


# --------------------------------------------------------------------------- a7
def make_anchors(shapes, strides, offset=0.5):
    """HPE/models/utils.py:15-27.  shapes = [(H_l, W_l)], strides = [s_l] -> anchors[A,2] (x,y), stride[A]."""
    pts, st = [], []
    for (h, w), s in zip(shapes, strides):
        sx = np.arange(w, dtype=F32) + F32(offset)
        sy = np.arange(h, dtype=F32) + F32(offset)
        yy, xx = np.meshgrid(sy, sx, indexing="ij")
        pts.append(np.stack((xx, yy), -1).reshape(-1, 2))
        st.append(np.full((h * w,), s, dtype=F32))
    return np.concatenate(pts).astype(F32), np.concatenate(st).astype(F32)


def pose_decode(levels, nc, kpt_shape=(17, 3), strides=(8.0, 16.0, 32.0)):
    """head.py:185-211.  levels = list of raw head maps [B, nc+nk, H_l, W_l] -> [B, nc+nk, A]:
    cls -> sigmoid; x = (raw - 0.5 + ax) * s; y = (raw - 0.5 + ay) * s; v -> sigmoid (ndim == 3)."""
    levels = [_f32(l) for l in levels]
    B, C = levels[0].shape[:2]
    nk = kpt_shape[0] * kpt_shape[1]
    assert C == nc + nk
    x_cat = np.concatenate([l.reshape(B, C, -1) for l in levels], axis=2)
    anchors, st = make_anchors([l.shape[2:] for l in levels], strides)
    out = np.empty_like(x_cat)
    out[:, :nc] = sigmoid(x_cat[:, :nc])
    kpt = x_cat[:, nc:].copy()
    nd = kpt_shape[1]
    kpt[:, 0::nd] = ((kpt[:, 0::nd] - F32(0.5)) + anchors[:, 0]) * st
    kpt[:, 1::nd] = ((kpt[:, 1::nd] - F32(0.5)) + anchors[:, 1]) * st
    if nd == 3:
        kpt[:, 2::nd] = sigmoid(kpt[:, 2::nd])
    out[:, nc:] = kpt
    return out


# --------------------------------------------------------------------------- a8
def greedy_nms(boxes, scores, iou_thres):
    """torchvision.ops.nms on CPU (the call at HPE/test.py:170): stable descending sort by score, greedy
    suppression with ``inter / (area_i + area_j - inter) > iou_thres``; float32 arithmetic, the threshold is a
    C double.  Returns kept indices in descending-score order."""
    boxes = _f32(boxes)
    scores = _f32(scores)
    n = boxes.shape[0]
    x1, y1, x2, y2 = boxes[:, 0], boxes[:, 1], boxes[:, 2], boxes[:, 3]
    areas = ((x2 - x1) * (y2 - y1)).astype(F32)
    order = np.argsort(-scores, kind="stable")
    suppressed = np.zeros(n, dtype=bool)
    keep = []
    thr = float(iou_thres)
    for _i in range(n):
        i = order[_i]
        if suppressed[i]:
            continue
        keep.append(i)
        rest = order[_i + 1:]
        if rest.size == 0:
            continue
        xx1 = np.maximum(x1[i], x1[rest]); yy1 = np.maximum(y1[i], y1[rest])
        xx2 = np.minimum(x2[i], x2[rest]); yy2 = np.minimum(y2[i], y2[rest])
        w = np.maximum(F32(0), (xx2 - xx1).astype(F32)); h = np.maximum(F32(0), (yy2 - yy1).astype(F32))
        inter = (w * h).astype(F32)
        with np.errstate(divide="ignore", invalid="ignore"):
            ovr = (inter / ((areas[i] + areas[rest]).astype(F32) - inter).astype(F32)).astype(F32)
        suppressed[rest[ovr.astype(np.float64) > thr]] = True
    return np.asarray(keep, dtype=np.int64)


def non_max_suppression_pose(prediction, conf_thres=0.25, iou_thres=0.45, max_det=300, nc=7, nkpts=4):
    """HPE/test.py:69-185.  prediction [B, nc+nk, A] (or [B, A, nc+nk]) -> list of dicts with
    boxes[n,4] (min/max of the keypoint xy), scores[n], labels[n] (int64), keypoints[n,nkpts,3]."""
    p = _f32(prediction)
    if p.shape[1] < p.shape[2]:
        p = p.transpose(0, 2, 1)
    nk = p.shape[2] - nc
    kd = nk // nkpts
    out = []
    for x in p:
        res = {"boxes": np.zeros((0, 4), F32), "scores": np.zeros((0,), F32), "labels": np.zeros((0,), np.int64),
               "keypoints": np.zeros((0, nkpts, 3), F32)}
        conf = x[:, :nc].max(axis=1)
        x = x[conf > F32(conf_thres)]
        if x.shape[0] == 0:
            out.append(res); continue
        scores = x[:, :nc].max(axis=1)
        labels = x[:, :nc].argmax(axis=1)
        kp = x[:, nc:].reshape(-1, nkpts, kd)
        if kd == 2:
            kp = np.concatenate([kp, np.ones((kp.shape[0], nkpts, 1), F32)], axis=2)
        xy = kp[..., :2]
        boxes = np.concatenate([xy.min(axis=1), xy.max(axis=1)], axis=1).astype(F32)
        off = (labels.astype(F32) * F32(7680))[:, None]
        keep = greedy_nms((boxes + off).astype(F32), scores, iou_thres)[:max_det]
        res = {"boxes": boxes[keep], "scores": scores[keep], "labels": labels[keep].astype(np.int64), "keypoints": kp[keep]}
        out.append(res)
    return out


# --------------------------------------------------------------------------- a10
def oks(gt_keypoints, pred_keypoints, areas, sigmas=OKS_SIGMAS, in_vis_thres=0.0, eps=1e-10):
    """HPE/core/metric/oksap.py:32-74 -> oks[N]."""
    gt = _f32(gt_keypoints); pred = _f32(pred_keypoints); areas = _f32(areas); sigmas = _f32(sigmas)
    vis = (gt[..., 2] > F32(in_vis_thres)).astype(F32)
    d2 = ((pred - gt[..., :2]) ** 2).sum(-1).astype(F32)
    denom = (F32(2.0) * areas[:, None] * ((sigmas * F32(2.0)) ** 2)[None, :] + F32(eps)).astype(F32)
    return ((np.exp(-d2 / denom) * vis).sum(1) / (vis.sum(1) + F32(eps))).astype(F32)


def pck(gt_keypoints, pred_keypoints, visible=None, normalizer=None, alpha=0.5, eps=1e-6):
    """HPE/core/metric/pck.py:14-104 -> (percentage, n_correct, n_visible)."""
    gt = _f32(gt_keypoints); pred = _f32(pred_keypoints)
    N, K, _ = gt.shape
    vis = np.ones((N, K), bool) if visible is None else np.asarray(visible).astype(bool)
    if normalizer is None:
        xs = np.where(vis, gt[..., 0], np.inf); ys = np.where(vis, gt[..., 1], np.inf)
        xm = np.where(vis, gt[..., 0], -np.inf); ym = np.where(vis, gt[..., 1], -np.inf)
        with np.errstate(invalid="ignore"):
            L = np.maximum(xm.max(1) - xs.min(1), ym.max(1) - ys.min(1)).astype(F32)
    else:
        L = np.broadcast_to(_f32(normalizer), (N,)).astype(F32)
    with np.errstate(invalid="ignore"):
        L = np.maximum(L, F32(eps))          # torch.clamp(min): NaN stays NaN
    dist = np.sqrt(((gt - pred) ** 2).sum(-1)).astype(F32)
    with np.errstate(invalid="ignore"):
        correct = (dist <= F32(alpha) * L[:, None]) & vis
    nc_, nv = int(correct.sum()), int(vis.sum())
    return (0.0 if nv == 0 else 100.0 * nc_ / nv), nc_, nv


def keypoint_loss(pred_kpts, gt_kpts, kpt_mask, area, sigmas):
    """HPE/core/loss/keypoint.py:13-21 (identical to HPE/loss.py:29-39 with area pre-unsqueezed) -> scalar."""
    p = _f32(pred_kpts); g = _f32(gt_kpts); m = _f32(kpt_mask); a = _f32(area); s = _f32(sigmas)
    if a.ndim == 1:
        a = a[:, None]
    d = (p[..., 0] - g[..., 0]) ** 2 + (p[..., 1] - g[..., 1]) ** 2
    fac = F32(m.shape[1]) / ((m != 0).sum(1).astype(F32) + F32(1e-9))
    e = d / ((F32(2) * s) ** 2 * (a + F32(1e-9)) * F32(2))
    return F32((fac[:, None] * ((F32(1) - np.exp(-e)) * m)).mean())


# --------------------------------------------------------------------------- a9
COCO_LOSS_SIGMAS = (OKS_SIGMAS / F32(10.0)).astype(F32)            # HPE/loss.py:250-254


def _robust_box_xywh(xy, vis):
    """HPE/loss.py:277-315 for ONE instance: xy[K,2] torch, vis[K] -> (cx, cy, w, h) torch scalars.
    Visible points only (all points when < 2 are visible); > 4 points: 10 % trimmed min/max of each axis."""
    import torch
    pts = xy[vis > 0]
    if pts.shape[0] < 2:
        pts = xy
    n = pts.shape[0]
    if n > 4:
        xs, _ = torch.sort(pts[:, 0]); ys, _ = torch.sort(pts[:, 1])
        t = max(1, int(n * 0.1))
        x0, x1, y0, y1 = xs[t], xs[n - t - 1], ys[t], ys[n - t - 1]
    else:
        x0, x1, y0, y1 = pts[:, 0].min(), pts[:, 0].max(), pts[:, 1].min(), pts[:, 1].max()
    w = (x1 - x0).clamp(min=1.0); h = (y1 - y0).clamp(min=1.0)
    return torch.stack([(x0 + x1) * 0.5, (y0 + y1) * 0.5, w, h])


def ciou_xywh(b1, b2, eps=1e-7):
    """HPE/loss.py:317-347 (CIoU=True): b1, b2 [N,4] (cx,cy,w,h) torch -> ciou[N]."""
    import torch
    b1x1, b1x2 = b1[:, 0] - b1[:, 2] / 2, b1[:, 0] + b1[:, 2] / 2
    b1y1, b1y2 = b1[:, 1] - b1[:, 3] / 2, b1[:, 1] + b1[:, 3] / 2
    b2x1, b2x2 = b2[:, 0] - b2[:, 2] / 2, b2[:, 0] + b2[:, 2] / 2
    b2y1, b2y2 = b2[:, 1] - b2[:, 3] / 2, b2[:, 1] + b2[:, 3] / 2
    inter = (torch.min(b1x2, b2x2) - torch.max(b1x1, b2x1)).clamp(0) * (torch.min(b1y2, b2y2) - torch.max(b1y1, b2y1)).clamp(0)
    w1, h1, w2, h2 = b1[:, 2], b1[:, 3], b2[:, 2], b2[:, 3]
    union = w1 * h1 + w2 * h2 - inter + eps
    iou = inter / union
    cw = torch.max(b1x2, b2x2) - torch.min(b1x1, b2x1)
    ch = torch.max(b1y2, b2y2) - torch.min(b1y1, b2y1)
    c2 = cw ** 2 + ch ** 2 + eps
    rho2 = ((b1x1 + b1x2 - b2x1 - b2x2) ** 2 + (b1y1 + b1y2 - b2y1 - b2y2) ** 2) / 4
    v = (4 / (3.141592 ** 2)) * torch.pow(torch.atan(w2 / (h2 + eps)) - torch.atan(w1 / (h1 + eps)), 2)
    alpha = (v / (v - iou + (1 + eps))).detach()
    return torch.nan_to_num(iou - (rho2 / c2 + v * alpha), nan=0.0)


def compute_loss(preds, targets, nc, num_kpts, sigmas=None):
    """HPE/loss.py:349-448 with the default VarifocalLoss(gamma 2, alpha 0.75) class term and HybridKeypointLoss.

    preds: list of torch [B, nc+3*num_kpts, H_l, W_l] (may require grad); targets: torch [n, 6+3*num_kpts] rows
    (batch, class, cx, cy, w, h, (x, y, v)*K), all normalised to [0,1].
    Returns (total[1], (cls, kpt, vis) floats)."""
    import torch
    import torch.nn.functional as Fn
    if sigmas is None:
        sigmas = torch.from_numpy(COCO_LOSS_SIGMAS) if num_kpts == 17 else torch.ones(num_kpts) * 0.05
    l_cls = torch.zeros(1); l_kpt = torch.zeros(1); l_vis = torch.zeros(1)
    npos_total = 0
    for pred in preds:
        B, C, H, W = pred.shape
        p = pred.permute(0, 2, 3, 1)
        tcls = torch.zeros(B, H, W, nc)
        tmask = torch.zeros(B, H, W, dtype=torch.bool)
        if targets.shape[0] > 0:
            box = targets[:, 2:6].clone(); kp = targets[:, 6:].clone()
            box[:, [0, 2]] *= W; box[:, [1, 3]] *= H
            kp[:, 0::3] *= W; kp[:, 1::3] *= H
            bi = targets[:, 0].long(); ci = targets[:, 1].long()
            gx = box[:, 0].long().clamp(0, W - 1); gy = box[:, 1].long().clamp(0, H - 1)
            tmask[bi, gy, gx] = True
            npos = int(tmask.sum())
            npos_total += npos
            if npos > 0:
                boff = box.clone(); boff[:, 0] -= gx.float(); boff[:, 1] -= gy.float()
                kp[:, 0::3] -= gx.unsqueeze(1).float(); kp[:, 1::3] -= gy.unsqueeze(1).float()
                pk = p[bi, gy, gx][:, nc:].reshape(-1, num_kpts, 3)
                pxy, pconf = pk[..., :2], pk[..., 2]
                g = kp.reshape(-1, num_kpts, 3)
                gxy = g[..., :2]; m = (g[..., 2] > 0).float()
                area = box[:, 2] * box[:, 3]
                pbox = torch.stack([_robust_box_xywh(pxy[i], m[i]) for i in range(pxy.shape[0])])
                iou = ciou_xywh(pbox, boff)
                # HybridKeypointLoss (loss.py:171-222) = wing-OKS (41-84) + 0.5 L1 + 0.1 smoothness
                a = area.unsqueeze(-1).clamp(min=1.0)
                d = (pxy[..., 0] - gxy[..., 0]).pow(2) + (pxy[..., 1] - gxy[..., 1]).pow(2)
                e = d / ((2 * sigmas).pow(2) * a * 2 + 1e-6)
                wing = torch.where(e < 2.0, 10.0 * torch.log(1 + e / 10.0), e - 2.0 + 10.0 * torch.log(1 + 2.0 / 10.0 + 0.0 * e))
                nvis = m.sum().clamp(min=1.0)
                l_oks = (wing * m).sum() / nvis
                l_l1 = ((pxy - gxy).abs().sum(-1) * m).sum() / nvis
                l_sm = (pxy[:, 1:] - pxy[:, :-1]).pow(2).sum() / pxy.shape[0] if num_kpts > 1 else torch.tensor(0.0)
                l_kpt = l_kpt + (l_oks + 0.5 * l_l1 + 0.1 * l_sm)
                l_vis = l_vis + Fn.binary_cross_entropy_with_logits(pconf, m, reduction="sum")
                tcls[bi, gy, gx, ci] = iou.detach().clamp(0, 1)          # duplicates: last row wins (CPU index_put)
        pc = p[..., :nc].reshape(-1, nc); tc = tcls.reshape(-1, nc); lab = (tc > 0).float()
        wgt = 0.75 * pc.sigmoid().pow(2.0) * (1 - lab) + tc * lab            # VarifocalLoss, loss.py:9-27
        l_cls = l_cls + (Fn.binary_cross_entropy_with_logits(pc.float(), tc.float(), reduction="none") * wgt).sum()
    npos_total = max(npos_total, 1)
    l_cls = l_cls / npos_total * 1.0
    l_kpt = l_kpt / npos_total * 10.0
    l_vis = l_vis / npos_total * 5.0
    return l_cls + l_kpt + l_vis, (l_cls.item(), l_kpt.item(), l_vis.item())


# --------------------------------------------------------------------------- seeded synthetic inputs
def synth_head_levels(B, nc, nkpts, shapes, seed=11, pos_rate=0.03):
    """Raw head maps: class logits mostly negative (sigmoid ~ 0.02) with a few confident anchors,
    keypoint offsets N(0, 2), visibility logits N(0, 1.5)."""
    rng = np.random.default_rng(seed)
    out = []
    for (h, w) in shapes:
        cls = rng.normal(-4.0, 1.0, size=(B, nc, h, w))
        hot = rng.random((B, nc, h, w)) < pos_rate
        cls = np.where(hot, rng.normal(1.5, 1.0, size=cls.shape), cls)
        kp = rng.normal(0.0, 2.0, size=(B, nkpts, 3, h, w))
        kp[:, :, 2] = rng.normal(0.0, 1.5, size=(B, nkpts, h, w))
        out.append(np.concatenate([cls, kp.reshape(B, nkpts * 3, h, w)], axis=1).astype(F32))
    return out


def synth_targets(B, nkpts, n_per_image=2, seed=12, nc=1):
    """Target rows [b, cls, cx, cy, w, h, (x, y, v)*K], normalised; includes duplicates of one cell, a box centre
    at the image border and instances with 0 / 1 visible keypoints."""
    rng = np.random.default_rng(seed)
    rows = []
    for b in range(B):
        for j in range(n_per_image):
            cx, cy = rng.uniform(0.1, 0.9, 2); w, h = rng.uniform(0.1, 0.5, 2)
            kx = rng.uniform(cx - w / 2, cx + w / 2, nkpts); ky = rng.uniform(cy - h / 2, cy + h / 2, nkpts)
            v = rng.choice([0.0, 1.0, 2.0], size=nkpts, p=[0.2, 0.2, 0.6])
            rows.append([b, rng.integers(0, nc), cx, cy, w, h] + np.stack([kx, ky, v], 1).reshape(-1).tolist())
    rows.append(list(rows[0]))                      # exact duplicate cell
    r = list(rows[1]); r[2] = 1.0; r[3] = 0.999; rows.append(r)     # centre on the border -> clamped cell
    r = list(rows[2]); r[8::3] = [0.0] * nkpts; rows.append(r)      # nothing visible
    r = list(rows[3]); r[8::3] = [0.0] * nkpts; r[8] = 2.0; rows.append(r)   # one visible
    return np.asarray(rows, dtype=F32)
