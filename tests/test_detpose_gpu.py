"""GPU parity for the reference's real pose-head decode / NMS / ComputeLoss / metrics:
the sm_100a kernels (through the C ABI) against oracle/detpose_ref.py AND against the golden vectors produced by
the reference's own code (tests/golden/detpose_golden.npz)."""
import os

import numpy as np
import pytest
import torch

from oracle import detpose_ref as dr

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "detpose_golden.npz"))
SHAPES = [(32, 24), (16, 12), (8, 6)]
NC, K = 1, 17


def _levels(B, nc, k, seed, dev, **kw):
    lv = dr.synth_head_levels(B, nc, k, SHAPES, seed=seed, **kw)
    return lv, [torch.from_numpy(l).to(dev) for l in lv]


def test_pose_decode_vs_reference_golden(hpe, cuda_dev):
    lv, dl = _levels(4, NC, K, 11, cuda_dev)
    out = hpe.pose_decode(dl, NC, (K, 3)).cpu().numpy()
    assert out.shape == (4, 52, 1008)
    kp = out[:, NC:].reshape(4, K, 3, -1); gk = GOLD["decoded"][:, NC:].reshape(4, K, 3, -1)
    np.testing.assert_array_equal(kp[:, :, :2], gk[:, :, :2])                 # coordinates: bit-exact
    np.testing.assert_allclose(out, GOLD["decoded"], rtol=0, atol=2e-7)       # sigmoids: 2 ulp of fp32 at [0,1]


def test_pose_decode_bf16_and_two_dims(hpe, cuda_dev):
    lv = dr.synth_head_levels(3, 2, 5, [(16, 20), (8, 10)], seed=5)
    lv = [l[:, :2 + 10] for l in lv]                                           # nc=2, 5 keypoints x (x,y)
    dl = [torch.from_numpy(np.ascontiguousarray(l)).to(cuda_dev).bfloat16() for l in lv]
    want = dr.pose_decode([d.float().cpu().numpy() for d in dl], 2, (5, 2), strides=(8.0, 16.0))
    got = hpe.pose_decode(dl, 2, (5, 2), strides=(8.0, 16.0)).cpu().numpy()
    np.testing.assert_allclose(got, want, rtol=0, atol=2e-7)


def _check_nms(res, want):
    assert len(res) == len(want)
    for r, w in zip(res, want):
        assert r["labels"].dtype == torch.int64
        np.testing.assert_array_equal(r["scores"].cpu().numpy(), w["scores"])
        np.testing.assert_array_equal(r["labels"].cpu().numpy(), w["labels"])
        np.testing.assert_array_equal(r["boxes"].cpu().numpy(), w["boxes"])
        np.testing.assert_array_equal(r["keypoints"].cpu().numpy(), w["keypoints"])


def test_nms_vs_reference_golden(hpe, cuda_dev):
    dec = torch.from_numpy(GOLD["decoded"]).to(cuda_dev)
    res = hpe.non_max_suppression_pose(dec, 0.25, 0.45, nc=NC, nkpts=K)
    np.testing.assert_array_equal([r["scores"].shape[0] for r in res], GOLD["nms_count"])
    cat = lambda k: np.concatenate([r[k].cpu().numpy() for r in res])        # noqa: E731
    np.testing.assert_array_equal(cat("scores"), GOLD["nms_scores"])
    np.testing.assert_array_equal(cat("boxes"), GOLD["nms_boxes"])
    np.testing.assert_array_equal(cat("labels"), GOLD["nms_labels"])
    np.testing.assert_array_equal(cat("keypoints"), GOLD["nms_keypoints"])


@pytest.mark.parametrize("B,nc,k,shapes,pos,seed", [
    (2, 3, 4, SHAPES, 0.15, 21),                         # multi-class offset 7680, > max_det candidates
    (8, 1, 17, SHAPES, 0.03, 31),
    (3, 2, 17, [(80, 80), (40, 40), (20, 20)], 0.05, 41),   # 640x640: 8400 anchors
    (2, 1, 17, [(16, 12)], 0.5, 51),                  # single level (A must exceed C: the reference infers the layout from it)
])
def test_nms_vs_oracle(hpe, cuda_dev, B, nc, k, shapes, pos, seed):
    lv = dr.synth_head_levels(B, nc, k, shapes, seed=seed, pos_rate=pos)
    dec = hpe.pose_decode([torch.from_numpy(l).to(cuda_dev) for l in lv], nc, (k, 3))
    want = dr.non_max_suppression_pose(dec.cpu().numpy(), 0.25, 0.45, nc=nc, nkpts=k)
    _check_nms(hpe.non_max_suppression_pose(dec, 0.25, 0.45, nc=nc, nkpts=k), want)
    # (B, A, C) layout is accepted like the reference does (test.py:87-88)
    _check_nms(hpe.non_max_suppression_pose(dec.transpose(1, 2).contiguous(), 0.25, 0.45, nc=nc, nkpts=k), want)


def test_nms_ties_empty_and_max_det(hpe, cuda_dev):
    # exact score ties (stable order = anchor order), identical boxes, an empty image, max_det truncation
    B, A = 3, 1008
    pred = np.zeros((B, NC + 3 * K, A), np.float32)
    rng = np.random.default_rng(7)
    pred[0, 0, :] = 0.9                                                       # all tied
    pred[0, 1:] = rng.uniform(0, 200, (3 * K, A)).astype(np.float32)
    pred[0, 1:, 1::2] = pred[0, 1:, 0::2]                                     # pairs of identical boxes
    pred[2, 0, :] = rng.uniform(0.3, 1.0, A).astype(np.float32)
    pred[2, 1:] = (rng.uniform(0, 3, (3 * K, A)) + np.arange(A) * 10.0).astype(np.float32)   # disjoint boxes -> all kept
    want = dr.non_max_suppression_pose(pred, 0.25, 0.45, max_det=50, nc=NC, nkpts=K)
    got = hpe.non_max_suppression_pose(torch.from_numpy(pred).to(cuda_dev), 0.25, 0.45, max_det=50, nc=NC, nkpts=K)
    assert got[1]["scores"].shape[0] == 0 and got[1]["keypoints"].shape == (0, K, 3)
    assert got[2]["scores"].shape[0] == 50
    _check_nms(got, want)


def test_compute_loss_vs_reference_golden(hpe, cuda_dev):
    lv, dl = _levels(4, NC, K, 11, cuda_dev)
    preds = [d.clone().requires_grad_(True) for d in dl]
    targets = torch.from_numpy(dr.synth_targets(4, K, seed=12, nc=NC)).to(cuda_dev)
    crit = hpe.ComputeLoss(nc=NC, nkpts=K)
    total, parts = crit(preds, targets)
    assert total.shape == (1,) and total.requires_grad
    total.backward()
    np.testing.assert_allclose(total.item(), GOLD["loss_total"], rtol=1e-4)
    np.testing.assert_allclose(parts, GOLD["loss_parts"], rtol=1e-4)
    for i, p in enumerate(preds):
        g = GOLD[f"loss_grad{i}"]
        np.testing.assert_allclose(p.grad.cpu().numpy(), g, rtol=2e-4, atol=1e-6 * np.abs(g).max())
    total0, parts0 = crit(dl, torch.zeros(0, 6 + 3 * K, device=cuda_dev))
    np.testing.assert_allclose(total0.item(), GOLD["loss0_total"], rtol=1e-4)
    np.testing.assert_allclose(parts0, GOLD["loss0_parts"], rtol=1e-4, atol=1e-7)


@pytest.mark.parametrize("B,nc,k,n_per,seed", [(16, 1, 17, 3, 61), (5, 3, 4, 2, 62), (2, 1, 17, 1, 63)])
def test_compute_loss_vs_oracle(hpe, cuda_dev, B, nc, k, n_per, seed):
    lv = dr.synth_head_levels(B, nc, k, SHAPES, seed=seed)
    t = dr.synth_targets(B, k, n_per_image=n_per, seed=seed + 100, nc=nc)
    ref_p = [torch.from_numpy(l.copy()).requires_grad_(True) for l in lv]
    want, wparts = dr.compute_loss(ref_p, torch.from_numpy(t), nc, k)
    want.backward()
    preds = [torch.from_numpy(l).to(cuda_dev).requires_grad_(True) for l in lv]
    total, parts = hpe.ComputeLoss(nc=nc, nkpts=k)(preds, torch.from_numpy(t).to(cuda_dev))
    (total * 2.0).backward()                                                  # upstream gradient is honoured
    np.testing.assert_allclose(total.item(), want.item(), rtol=1e-4)
    np.testing.assert_allclose(parts, wparts, rtol=1e-4)
    for p, r in zip(preds, ref_p):
        g = 2.0 * r.grad.numpy()
        np.testing.assert_allclose(p.grad.cpu().numpy(), g, rtol=2e-4, atol=1e-6 * np.abs(g).max())


def _metric_inputs(N, seed):
    rng = np.random.default_rng(seed)
    gt = np.zeros((N, K, 3), np.float32)
    gt[..., 0] = rng.uniform(0, 192, (N, K)); gt[..., 1] = rng.uniform(0, 256, (N, K))
    gt[..., 2] = rng.choice([0.0, 1.0, 2.0], size=(N, K), p=[0.2, 0.2, 0.6])
    gt[0, :, 2] = 0.0
    pred = (gt[..., :2] + rng.normal(0, 6.0, (N, K, 2))).astype(np.float32)
    areas = rng.uniform(400, 20000, N).astype(np.float32)
    return gt, pred, areas


def test_metrics_vs_reference_golden(hpe, cuda_dev):
    gt, pred, areas = _metric_inputs(64, 13)
    d = lambda a: torch.from_numpy(a).to(cuda_dev)                            # noqa: E731
    oks = hpe.OKS(hpe.OKS_SIGMAS).compute(d(gt), d(pred), d(areas)).cpu().numpy()
    np.testing.assert_allclose(oks, GOLD["oks"], rtol=5e-6, atol=1e-7)
    vis = gt[..., 2] > 0
    assert abs(hpe.PCK(alpha=0.05).compute(d(gt[..., :2].copy()), d(pred), d(vis)) - float(GOLD["pck"])) < 1e-9
    assert abs(hpe.PCK(alpha=0.5).compute(d(gt[..., :2].copy()), d(pred), None, torch.tensor(20.0)) - float(GOLD["pck_norm"])) < 1e-9
    kl = hpe.KeypointLoss(hpe.OKS_SIGMAS)(d(pred), d(gt[..., :2].copy()), d(vis.astype(np.float32)), d(areas).unsqueeze(-1))
    np.testing.assert_allclose(kl.item(), GOLD["keypoint_loss"], rtol=1e-5)
    assert hpe.PCK().compute(d(gt[:1, :, :2].copy()), d(pred[:1]), d(vis[:1])) == 0.0


def test_metrics_large_vs_oracle(hpe, cuda_dev):
    gt, pred, areas = _metric_inputs(5000, 17)
    d = lambda a: torch.from_numpy(a).to(cuda_dev)                            # noqa: E731
    np.testing.assert_allclose(hpe.OKS(hpe.OKS_SIGMAS).compute(d(gt), d(pred), d(areas)).cpu().numpy(),
                               dr.oks(gt, pred, areas), rtol=5e-6, atol=1e-7)
    vis = gt[..., 2] > 0
    pct, c, v = dr.pck(gt[..., :2], pred, vis, alpha=0.05)
    cnt = hpe.PCK(alpha=0.05).counts(d(gt[..., :2].copy()), d(pred), d(vis)).tolist()
    assert cnt == [c, v]


def test_decoded_keypoints_give_identical_oks(hpe, cuda_dev):
    # end-to-end acceptance: kernel-decoded vs oracle-decoded keypoints -> same OKS / PCK
    from oracle import heatmap_ref as hr
    hm = hr.synth_heatmaps(32, 17, 64, 48, seed=3)
    center, scale = hr.synth_center_scale(32, seed=4)
    po, _ = hr.get_final_preds(hm, center, scale)
    pk, _ = hpe.decode(torch.from_numpy(hm).to(cuda_dev), torch.from_numpy(center), torch.from_numpy(scale))
    rng = np.random.default_rng(0)
    gt = np.concatenate([po + rng.normal(0, 3, po.shape), np.ones((32, 17, 1))], -1).astype(np.float32)
    areas = np.full(32, 5000, np.float32)
    a = hpe.OKS(hpe.OKS_SIGMAS).compute(torch.from_numpy(gt).to(cuda_dev), pk.contiguous(), torch.from_numpy(areas).to(cuda_dev))
    np.testing.assert_allclose(a.cpu().numpy(), dr.oks(gt, po, areas), rtol=1e-4, atol=1e-6)
