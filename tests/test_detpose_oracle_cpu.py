"""CPU suite: oracle/detpose_ref.py against golden vectors produced by the REFERENCE's own code
(tests/golden/make_golden_detpose.py) -- rows a7-a10 of SURVEY.md §8 are parity-pinned by these."""
import os

import numpy as np
import torch

from oracle import detpose_ref as dr

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "detpose_golden.npz"))
SHAPES = [(32, 24), (16, 12), (8, 6)]
NC, K = 1, 17


def _cat(res, key):
    return np.concatenate([r[key] for r in res])


def test_pose_decode_matches_reference():
    levels = dr.synth_head_levels(4, NC, K, SHAPES, seed=11)
    dec = dr.pose_decode(levels, NC, (K, 3))
    assert dec.shape == GOLD["decoded"].shape == (4, 52, 1008)
    np.testing.assert_allclose(dec, GOLD["decoded"], rtol=0, atol=2e-6)
    # coordinates are plain fp32 add/mul: bit-exact
    kp = dec[:, NC:].reshape(4, K, 3, -1); gk = GOLD["decoded"][:, NC:].reshape(4, K, 3, -1)
    np.testing.assert_array_equal(kp[:, :, :2], gk[:, :, :2])


def test_make_anchors_layout():
    a, s = dr.make_anchors(SHAPES, (8.0, 16.0, 32.0))
    assert a.shape == (1008, 2) and s.shape == (1008,)
    np.testing.assert_array_equal(a[:3], [[0.5, 0.5], [1.5, 0.5], [2.5, 0.5]])        # row-major, x fastest
    np.testing.assert_array_equal(a[24], [0.5, 1.5])
    assert s[767] == 8 and s[768] == 16 and s[-1] == 32


def test_nms_matches_reference():
    res = dr.non_max_suppression_pose(GOLD["decoded"], 0.25, 0.45, nc=NC, nkpts=K)
    np.testing.assert_array_equal([r["scores"].shape[0] for r in res], GOLD["nms_count"])
    np.testing.assert_array_equal(_cat(res, "scores"), GOLD["nms_scores"])
    np.testing.assert_array_equal(_cat(res, "boxes"), GOLD["nms_boxes"])
    np.testing.assert_array_equal(_cat(res, "labels"), GOLD["nms_labels"])
    np.testing.assert_array_equal(_cat(res, "keypoints"), GOLD["nms_keypoints"])


def test_nms_multiclass_max_det_matches_reference():
    lv3 = dr.synth_head_levels(2, 3, 4, SHAPES, seed=21, pos_rate=0.15)
    dec3 = dr.pose_decode(lv3, 3, (4, 3))
    res = dr.non_max_suppression_pose(dec3, 0.25, 0.45, nc=3, nkpts=4)
    np.testing.assert_array_equal([r["scores"].shape[0] for r in res], GOLD["nms3_count"])     # capped at 300
    np.testing.assert_allclose(_cat(res, "scores"), GOLD["nms3_scores"], rtol=0, atol=1e-6)
    np.testing.assert_array_equal(_cat(res, "labels"), GOLD["nms3_labels"])
    np.testing.assert_allclose(_cat(res, "boxes"), GOLD["nms3_boxes"], rtol=0, atol=1e-4)


def test_nms_empty_and_transposed_input():
    z = np.zeros((2, 52, 1008), np.float32)
    res = dr.non_max_suppression_pose(z, nc=NC, nkpts=K)
    assert all(r["scores"].shape == (0,) and r["keypoints"].shape == (0, K, 3) for r in res)
    a = dr.non_max_suppression_pose(GOLD["decoded"][:1], nc=NC, nkpts=K)
    b = dr.non_max_suppression_pose(GOLD["decoded"][:1].transpose(0, 2, 1), nc=NC, nkpts=K)
    np.testing.assert_array_equal(a[0]["scores"], b[0]["scores"])


def test_compute_loss_matches_reference():
    levels = dr.synth_head_levels(4, NC, K, SHAPES, seed=11)
    targets = torch.from_numpy(dr.synth_targets(4, K, seed=12, nc=NC))
    preds = [torch.from_numpy(l.copy()).requires_grad_(True) for l in levels]
    total, parts = dr.compute_loss(preds, targets, NC, K)
    total.backward()
    np.testing.assert_allclose(total.item(), GOLD["loss_total"], rtol=1e-6)
    np.testing.assert_allclose(parts, GOLD["loss_parts"], rtol=1e-6)
    for i, p in enumerate(preds):
        g = GOLD[f"loss_grad{i}"]
        np.testing.assert_allclose(p.grad.numpy(), g, rtol=1e-5, atol=1e-7 * np.abs(g).max())
    total0, parts0 = dr.compute_loss([torch.from_numpy(l) for l in levels], torch.zeros(0, 6 + 3 * K), NC, K)
    np.testing.assert_allclose(total0.item(), GOLD["loss0_total"], rtol=1e-6)
    np.testing.assert_allclose(parts0, GOLD["loss0_parts"], rtol=1e-6)


def _metric_inputs(N=64, seed=13):
    rng = np.random.default_rng(seed)
    gt = np.zeros((N, K, 3), np.float32)
    gt[..., 0] = rng.uniform(0, 192, (N, K)); gt[..., 1] = rng.uniform(0, 256, (N, K))
    gt[..., 2] = rng.choice([0.0, 1.0, 2.0], size=(N, K), p=[0.2, 0.2, 0.6])
    gt[0, :, 2] = 0.0
    pred = gt[..., :2] + rng.normal(0, 6.0, (N, K, 2)).astype(np.float32)
    areas = rng.uniform(400, 20000, N).astype(np.float32)
    return gt, pred.astype(np.float32), areas


def test_metrics_match_reference():
    gt, pred, areas = _metric_inputs()
    np.testing.assert_allclose(dr.oks(gt, pred, areas), GOLD["oks"], rtol=2e-6, atol=1e-7)
    vis = gt[..., 2] > 0
    assert abs(dr.pck(gt[..., :2], pred, vis, alpha=0.05)[0] - float(GOLD["pck"])) < 1e-9
    assert abs(dr.pck(gt[..., :2], pred, None, 20.0, alpha=0.5)[0] - float(GOLD["pck_norm"])) < 1e-9
    kl = dr.keypoint_loss(pred, gt[..., :2], vis.astype(np.float32), areas, dr.OKS_SIGMAS)
    np.testing.assert_allclose(kl, GOLD["keypoint_loss"], rtol=1e-5)
    assert dr.pck(gt[:1, :, :2], pred[:1], vis[:1])[0] == 0.0          # nothing visible -> 0.0 (pck.py:101-102)
