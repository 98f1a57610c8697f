"""CPU suite: the oracle against the committed golden vectors and its own edge-case contracts."""
import os

import numpy as np
import pytest
import torch

from oracle import heatmap_ref as hr
from oracle import resnet_ref as rr

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_resnet50_oracle_matches_reference_golden():
    # golden = output of the REFERENCE's own DeepPose/timm code (tests/golden/make_golden.py)
    g = np.load(os.path.join(GOLD, "resnet50_ref_features.npz"))
    model = rr.build_synthetic(17, seed=0)
    x = torch.rand(2, 3, 128, 96, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        feat = model.backbone.forward_features(x).numpy()
    assert feat.shape == g["features"].shape == (2, 2048, 4, 3)
    np.testing.assert_allclose(feat, g["features"], rtol=0, atol=1e-5)


@pytest.mark.skipif(not os.path.isdir("/root/reference/HPE"), reason="reference tree only exists in the build container")
def test_resnet50_oracle_matches_reference_live():
    import sys
    ref = "/root/reference/HPE"
    sys.path[:0] = [ref, ref + "/thirdparty/pytorch-image-models"]
    sys.dont_write_bytecode = True
    try:
        from models.reference.bottom_up.regression.models.DeepPose import DeepPose
    finally:
        del sys.path[:2]
    model = rr.build_synthetic(17, seed=0)
    ref_model = DeepPose(17, "resnet50", pretrained=False).eval()
    ref_model.backbone.load_state_dict(model.backbone.state_dict(), strict=False)
    # state_dict key compatibility (checkpoints of the reference load into the oracle and vice versa)
    ref_keys = {k for k in ref_model.state_dict() if not k.startswith("backbone.fc.")}
    our_keys = {k for k in model.state_dict() if k.startswith("backbone.")}
    assert ref_keys == our_keys
    x = torch.rand(1, 3, 64, 64, generator=torch.Generator().manual_seed(7))
    with torch.no_grad():
        a = ref_model.backbone.forward_features(x)
        b = model.backbone.forward_features(x)
    assert torch.equal(a, b)


def test_heatmap_stage_golden():
    g = np.load(os.path.join(GOLD, "heatmap_stage_golden.npz"))
    hm = hr.synth_heatmaps(8, 17, 64, 48, seed=3)
    center, scale = hr.synth_center_scale(8, seed=4)
    preds, maxvals = hr.get_final_preds(hm, center, scale)
    _, _, idx = hr.get_max_preds(hm)
    np.testing.assert_array_equal(idx, g["idx"])
    np.testing.assert_allclose(preds, g["preds"], atol=1e-4)
    np.testing.assert_array_equal(maxvals, g["maxvals"])
    joints = hr.synth_joints(8, 17, (256, 192), seed=5)
    target, tw = hr.generate_target(joints, (64, 48), 2.0)
    np.testing.assert_array_equal(tw, g["target_weight"])
    np.testing.assert_allclose(target.sum(axis=(2, 3)), g["target_sum"], rtol=1e-6)
    loss, grad = hr.joints_mse_loss(hm, target, tw)
    assert abs(float(loss) - float(g["loss"])) <= 1e-7 * abs(float(g["loss"])) + 1e-12
    assert abs(float(np.abs(grad).sum()) - float(g["grad_abs_sum"])) < 1e-6 * float(g["grad_abs_sum"])
    # full arrays (first two samples), not only sums
    np.testing.assert_array_equal(joints, g["joints"])
    np.testing.assert_array_equal(hm[:2], g["heatmaps_first2"])
    np.testing.assert_allclose(target[:2], g["target_first2"], rtol=0, atol=1e-7)
    np.testing.assert_allclose(grad[:2], g["grad_first2"], rtol=1e-6, atol=1e-12)


def test_get_max_preds_tie_and_mask_rules():
    hm = np.zeros((1, 4, 6, 5), dtype=np.float32)
    hm[0, 0, 2, 3] = 1.0; hm[0, 0, 4, 1] = 1.0          # two-way tie -> first (row-major) wins
    hm[0, 1] = -1.0                                      # all negative -> idx 0, coords zeroed
    hm[0, 2, 5, 4] = 0.5                                 # last element
    # map 3 stays all zero -> maxval 0 -> masked
    coords, maxvals, idx = hr.get_max_preds(hm)
    assert idx.tolist() == [[2 * 5 + 3, 0, 5 * 5 + 4, 0]]
    assert coords[0, 0].tolist() == [3.0, 2.0]
    assert coords[0, 1].tolist() == [0.0, 0.0] and maxvals[0, 1, 0] == -1.0
    assert coords[0, 2].tolist() == [4.0, 5.0]
    assert coords[0, 3].tolist() == [0.0, 0.0] and maxvals[0, 3, 0] == 0.0


def test_quarter_pixel_strict_interior():
    H, W = 8, 8
    hm = np.zeros((1, 3, H, W), dtype=np.float32)
    hm[0, 0, 3, 3] = 1.0; hm[0, 0, 3, 4] = 0.5; hm[0, 0, 2, 3] = 0.2     # -> +0.25 x, -0.25 y
    hm[0, 1, 3, 1] = 1.0; hm[0, 1, 3, 2] = 0.5                           # column 1: NOT refined
    hm[0, 2, 3, 3] = 1.0                                                 # equal neighbours: sign(0)=0
    c, _, _ = hr.get_max_preds(hm)
    r = hr.quarter_pixel_refine(hm, c)
    assert r[0, 0].tolist() == [3.25, 2.75]
    assert r[0, 1].tolist() == [1.0, 3.0]
    assert r[0, 2].tolist() == [3.0, 3.0]


def test_generate_target_rules():
    # centre rounding: int(x/4 + 0.5) with truncation toward zero (negative coordinates!)
    j = np.array([[[10.0, 10.0, 1.0],      # mu = (3,3)  (2.5+0.5 = 3)
                   [-3.0, 20.0, 2.0],      # x: int(-0.75+0.5) = int(-0.25) = 0 (truncation, not floor)
                   [1000.0, 10.0, 1.0],    # far right -> ul_x >= W -> weight 0
                   [10.0, 10.0, 0.0],      # invisible
                   [-40.0, 10.0, 1.0]]],   # mu_x = int(-9.5) = -9, br_x = -9+6+1 = -2 < 0 -> weight 0
                 dtype=np.float32)
    t, w = hr.generate_target(j, (64, 48), 2.0)
    assert w[0, :, 0].tolist() == [1.0, 1.0, 0.0, 0.0, 0.0]
    assert t[0, 0, 3, 3] == 1.0 and abs(t[0, 0, 3, 4] - np.exp(-1 / 8)) < 1e-7
    assert t[0, 0, 3, 3 + 7] == 0.0 and t[0, 0, 3, 3 + 6] > 0            # support is +-3 sigma
    assert t[0, 1, 5, 0] == 1.0
    assert not t[0, 2].any() and not t[0, 3].any() and not t[0, 4].any()


def test_joints_mse_loss_formula():
    rng = np.random.default_rng(0)
    p = rng.normal(size=(3, 5, 8, 6)).astype(np.float32)
    g = rng.normal(size=(3, 5, 8, 6)).astype(np.float32)
    w = rng.integers(0, 2, size=(3, 5, 1)).astype(np.float32)
    loss, grad = hr.joints_mse_loss(p, g, w)
    # torch restatement of JointsMSELoss(use_target_weight=True): mean over joints of 0.5*MSE(w*p, w*g)
    tp, tg, tw = torch.tensor(p, requires_grad=True), torch.tensor(g), torch.tensor(w)
    B, K = 3, 5
    hp = tp.reshape(B, K, -1); hg = tg.reshape(B, K, -1)
    tl = sum(0.5 * torch.nn.functional.mse_loss(hp[:, k] * tw[:, k], hg[:, k] * tw[:, k]) for k in range(K)) / K
    tl.backward()
    assert abs(float(loss) - float(tl)) < 1e-6 * abs(float(tl))
    np.testing.assert_allclose(grad, tp.grad.numpy(), atol=1e-8)


def test_bf16_emulated_forward_close_to_fp32():
    m = rr.build_synthetic(17, seed=0)
    x = torch.rand(1, 3, 64, 64, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        a = m(x)
    b = rr.forward_bf16_emulated(m, x)
    assert a.shape == b.shape == (1, 17, 16, 16)
    assert (a - b).abs().max().item() < 0.05 * a.abs().max().item() + 1e-3


def test_uint8_scale_by_reciprocal_equals_division_in_bf16():
    """`imgs.float() / 255.0` (HPE/train.py:120) followed by the bf16 rounding of the stem's input: the uint8 layout kernel
    multiplies by fp32(1/255) instead of dividing.  The fp32 values differ in the last bit for about half of the 256 inputs, the
    bf16 values for none -- checked here for every possible byte."""
    import torch
    b = torch.arange(256, dtype=torch.float32)
    want = (b / 255.0).to(torch.bfloat16)
    got = (b * torch.tensor(1.0 / 255.0, dtype=torch.float32)).to(torch.bfloat16)
    assert torch.equal(want, got)
