"""CPU oracle (NumPy) for the heatmap stages the reference does not implement.

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.  **Parity unpinned**: the reference
(`/root/reference`) has no ``get_max_preds`` / ``generate_target`` / ``JointsMSELoss`` and no
golden vectors for them.  Every function below restates one paragraph of
SURVEY.md Appendix A, which fixes the public Simple-Baselines / HRNet conventions that
``BASELINE.json:north_star`` names.  The call *shape* of the criterion follows the reference's
trainer: ``criterion(output, target, target_weight[B,K,1])`` (HPE/engine/trainer.py:92).

All arithmetic is float32 NumPy so it is value-comparable with the fp32 CUDA kernels.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def _as_f32(hm) -> np.ndarray:
    """Accept numpy arrays or torch tensors (bf16 is upcast value-preservingly, App. A.2)."""
    if hasattr(hm, "detach"):
        hm = hm.detach().to("cpu").float().numpy()
    return np.ascontiguousarray(hm, dtype=F32)


# --------------------------------------------------------------------------- A.2
def get_max_preds(hm):
    """SURVEY.md App. A.2.  hm[B,K,H,W] -> coords[B,K,2] f32, maxvals[B,K,1] f32, idx[B,K] i32.

    First-max tie rule = ``np.argmax`` (smallest flat index among equal maxima);
    coordinates are zeroed where ``maxval <= 0``.
    """
    hm = _as_f32(hm)
    B, K, H, W = hm.shape
    flat = hm.reshape(B, K, H * W)
    idx = np.argmax(flat, axis=2)
    maxvals = np.take_along_axis(flat, idx[..., None], axis=2).astype(F32)
    coords = np.empty((B, K, 2), dtype=F32)
    coords[..., 0] = (idx % W).astype(F32)
    coords[..., 1] = (idx // W).astype(F32)
    coords *= (maxvals > 0.0).astype(F32)
    return coords, maxvals, idx.astype(np.int32)


# --------------------------------------------------------------------------- A.3
def quarter_pixel_refine(hm, coords):
    """SURVEY.md App. A.3.  +-0.25 px toward the larger neighbour, strictly interior peaks only
    (1 < px < W-1 and 1 < py < H-1), sign(0) = 0, differences in fp32.  Uses the *masked*
    integer coordinates, so masked joints sit at (0,0) and are never refined."""
    hm = _as_f32(hm)
    B, K, H, W = hm.shape
    out = coords.astype(F32).copy()
    px = coords[..., 0].astype(np.int64)
    py = coords[..., 1].astype(np.int64)
    interior = (px > 1) & (px < W - 1) & (py > 1) & (py < H - 1)
    b, k = np.nonzero(interior)
    x, y = px[b, k], py[b, k]
    dx = hm[b, k, y, x + 1] - hm[b, k, y, x - 1]
    dy = hm[b, k, y + 1, x] - hm[b, k, y - 1, x]
    out[b, k, 0] += F32(0.25) * np.sign(dx).astype(F32)
    out[b, k, 1] += F32(0.25) * np.sign(dy).astype(F32)
    return out


# --------------------------------------------------------------------------- A.4
def affine_from_center_scale(center, scale, heatmap_hw):
    """SURVEY.md App. A.4.  center[B,2] (cx,cy), scale[B,2] (sw,sh in units of 200 px), rot=0 ->
    M[B,2,3] f32 with r = 200*sw/W,  M = [[r,0,cx-r*W/2],[0,r,cy-r*H/2]]  (width-only scale)."""
    H, W = heatmap_hw
    center = np.asarray(center, dtype=F32).reshape(-1, 2)
    scale = np.asarray(scale, dtype=F32).reshape(-1, 2)
    r = (F32(200.0) * scale[:, 0] / F32(W)).astype(F32)
    M = np.zeros((center.shape[0], 2, 3), dtype=F32)
    M[:, 0, 0] = r
    M[:, 1, 1] = r
    M[:, 0, 2] = center[:, 0] - r * F32(W) * F32(0.5)
    M[:, 1, 2] = center[:, 1] - r * F32(H) * F32(0.5)
    return M


def apply_affine(coords, M):
    """[x_img, y_img]^T = M_b . [x, y, 1]^T for every joint (fp32)."""
    coords = np.asarray(coords, dtype=F32)
    M = np.asarray(M, dtype=F32)
    x, y = coords[..., 0], coords[..., 1]
    out = np.empty_like(coords)
    out[..., 0] = M[:, None, 0, 0] * x + M[:, None, 0, 1] * y + M[:, None, 0, 2]
    out[..., 1] = M[:, None, 1, 0] * x + M[:, None, 1, 1] * y + M[:, None, 1, 2]
    return out


def get_final_preds_affine(hm, M):
    """A.2 -> A.3 -> A.4 with an explicit per-sample 2x3 matrix.  Returns preds, maxvals, idx."""
    coords, maxvals, idx = get_max_preds(hm)
    coords = quarter_pixel_refine(hm, coords)
    return apply_affine(coords, M), maxvals, idx


def get_final_preds(hm, center, scale):
    """SURVEY.md App. A.4 top-level: ``preds[B,K,2], maxvals[B,K,1]``."""
    hm = _as_f32(hm)
    M = affine_from_center_scale(center, scale, hm.shape[2:])
    preds, maxvals, _ = get_final_preds_affine(hm, M)
    return preds, maxvals


# --------------------------------------------------------------------------- A.5
def generate_target(joints, heatmap_hw, sigma, stride=(4.0, 4.0)):
    """SURVEY.md App. A.5.  joints[B,K,3] = (x, y in input-image px, visibility v) ->
    target[B,K,H,W] f32, target_weight[B,K,1] f32.

    mu = int(x/stride + 0.5) with C truncation toward zero; support +-3*sigma; weight 0 when the
    support is fully outside the map or v <= 0; peak exactly 1.0 at the rounded centre.
    """
    joints = np.asarray(joints, dtype=F32)
    B, K, _ = joints.shape
    H, W = heatmap_hw
    sx, sy = F32(stride[0]), F32(stride[1])
    target = np.zeros((B, K, H, W), dtype=F32)
    weight = (joints[..., 2] > 0).astype(F32)
    mu_x = np.trunc(joints[..., 0] / sx + F32(0.5)).astype(np.int64)
    mu_y = np.trunc(joints[..., 1] / sy + F32(0.5)).astype(np.int64)
    t = int(3 * sigma)
    inv2s2 = F32(1.0) / F32(2.0 * sigma * sigma)
    for b in range(B):
        for k in range(K):
            mx, my = int(mu_x[b, k]), int(mu_y[b, k])
            ul = (mx - t, my - t)
            br = (mx + t + 1, my + t + 1)
            if ul[0] >= W or ul[1] >= H or br[0] < 0 or br[1] < 0:
                weight[b, k] = 0.0
                continue
            if weight[b, k] == 0.0:
                continue
            x0, x1 = max(0, ul[0]), min(W, br[0])
            y0, y1 = max(0, ul[1]), min(H, br[1])
            if x1 <= x0 or y1 <= y0:
                continue
            xs = np.arange(x0, x1, dtype=F32) - F32(mx)
            ys = np.arange(y0, y1, dtype=F32) - F32(my)
            d2 = (ys[:, None] * ys[:, None] + xs[None, :] * xs[None, :]).astype(F32)
            target[b, k, y0:y1, x0:x1] = np.exp(-(d2 * inv2s2)).astype(F32)
    return target, weight[..., None]


# --------------------------------------------------------------------------- A.6
def joints_mse_loss(pred, target, target_weight):
    """SURVEY.md App. A.6.  L = 0.5/(K*B*HW) * sum w^2 (p-g)^2 ; returns (loss f32, grad[B,K,H,W] f32)
    with grad = w^2 (p-g) / (K*B*HW).  Accumulation in float64 so the oracle is the *exact*
    value the fp32 kernels are held to 1e-4 relative against."""
    p = _as_f32(pred)
    g = _as_f32(target)
    w = np.asarray(target_weight, dtype=F32).reshape(p.shape[0], p.shape[1], 1, 1)
    B, K, H, W = p.shape
    d = (p - g).astype(np.float64) * w.astype(np.float64)
    n = float(K * B * H * W)
    loss = 0.5 * float(np.sum(d * d)) / n
    grad = (d * w.astype(np.float64) / n).astype(F32)
    return F32(loss), grad


def joints_mse_loss_from_joints(pred, joints, sigma, stride=(4.0, 4.0)):
    """Fused-mode criterion: targets generated from joints (A.5) then A.6."""
    p = _as_f32(pred)
    target, w = generate_target(joints, p.shape[2:], sigma, stride)
    return joints_mse_loss(p, target, w)


# --------------------------------------------------------------------------- §8(d) synthetic inputs
def synth_heatmaps(B, K, H, W, seed=3, adversarial=True):
    """SURVEY.md §8(d) decode-sweep heatmaps: Gaussian blob (sigma 2, amplitude U(0.3,1)) at a random
    sub-pixel centre + N(0,0.01) noise, with the mandatory adversarial cases mixed in (~1 % each):
    exact two-way ties, plateau ties, all-zero map, all-negative map, border peaks, equal neighbours."""
    rng = np.random.default_rng(seed)
    ys = np.arange(H, dtype=F32)[:, None]
    xs = np.arange(W, dtype=F32)[None, :]
    hm = np.empty((B, K, H, W), dtype=F32)
    cx = rng.uniform(0, W - 1, size=(B, K)).astype(F32)
    cy = rng.uniform(0, H - 1, size=(B, K)).astype(F32)
    amp = rng.uniform(0.3, 1.0, size=(B, K)).astype(F32)
    for b in range(B):
        d2 = (xs[None] - cx[b][:, None, None]) ** 2 + (ys[None] - cy[b][:, None, None]) ** 2
        hm[b] = amp[b][:, None, None] * np.exp(-d2 / F32(8.0))
    hm += rng.normal(0, 0.01, size=hm.shape).astype(F32)
    if adversarial:
        n = B * K
        kinds = rng.integers(0, 100, size=n)
        for j in np.nonzero(kinds < 7)[0]:
            b, k = divmod(int(j), K)
            m = hm[b, k]
            kind = int(kinds[j])
            if kind == 0:      # exact two-way tie: later duplicate of the max must lose
                i = int(np.argmax(m))
                j2 = int(rng.integers(0, H * W))
                m.reshape(-1)[j2] = m.reshape(-1)[i]
            elif kind == 1:    # plateau
                y0 = int(rng.integers(0, H - 3)); x0 = int(rng.integers(0, W - 3))
                m[y0:y0 + 3, x0:x0 + 3] = m.max() + F32(0.5)
            elif kind == 2:    # all zero
                m[...] = 0.0
            elif kind == 3:    # all negative
                m[...] = -np.abs(m) - F32(0.1)
            elif kind == 4:    # peak on/near the border in x
                x = int(rng.choice([0, 1, W - 2, W - 1])); y = int(rng.integers(0, H))
                m[y, x] = m.max() + F32(1.0)
            elif kind == 5:    # peak on/near the border in y
                y = int(rng.choice([0, 1, H - 2, H - 1])); x = int(rng.integers(0, W))
                m[y, x] = m.max() + F32(1.0)
            else:              # equal neighbours -> sign(0) = 0
                y = int(rng.integers(2, H - 2)); x = int(rng.integers(2, W - 2))
                m[y, x] = m.max() + F32(1.0)
                m[y, x - 1] = m[y, x + 1] = F32(0.25)
                m[y - 1, x] = m[y + 1, x] = F32(0.125)
    return hm


def synth_center_scale(B, seed=4):
    """§8(d): centre ~U over a 640x480 image, scale_w ~U(0.5,2.5) (units of 200 px), rot = 0."""
    rng = np.random.default_rng(seed)
    center = np.stack([rng.uniform(0, 640, B), rng.uniform(0, 480, B)], axis=1).astype(F32)
    sw = rng.uniform(0.5, 2.5, B).astype(F32)
    scale = np.stack([sw, sw * F32(4.0 / 3.0)], axis=1).astype(F32)
    return center, scale


def synth_joints(B, K, in_hw, seed=5):
    """§8(d): x~U(-8,W_in+8), y~U(-8,H_in+8), v in {0,1,2} with P=(0.15,0.15,0.7)."""
    rng = np.random.default_rng(seed)
    Hin, Win = in_hw
    j = np.empty((B, K, 3), dtype=F32)
    j[..., 0] = rng.uniform(-8, Win + 8, (B, K))
    j[..., 1] = rng.uniform(-8, Hin + 8, (B, K))
    j[..., 2] = rng.choice([0, 1, 2], size=(B, K), p=[0.15, 0.15, 0.7])
    return j
