"""CPU suite: oracle/dinov3pose_ref.py against the golden outputs of the REFERENCE's own DINOv3Pose
(tests/golden/make_golden_dinov3pose.py) -- pins rows a4-a6 of SURVEY.md §8."""
import os

import numpy as np
import torch

from oracle import detpose_ref as dr
from oracle import dinov3pose_ref as dp

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "dinov3pose_golden.npz"))


def test_dinov3pose_oracle_matches_reference_golden():
    torch.set_num_threads(4)
    m = dp.build_synthetic("tiny", (17, 3), 1, seed=0)
    x = torch.rand(2, 3, 128, 96, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        lv = m(x)
    assert [tuple(l.shape) for l in lv] == [(2, 52, 16, 12), (2, 52, 8, 6), (2, 52, 4, 3)]
    for i, l in enumerate(lv):
        np.testing.assert_allclose(l.numpy(), GOLD[f"level{i}"], rtol=0, atol=1e-5)
    # eval-mode output of the reference = decode of the same level maps (head.py:185-199)
    dec = dr.pose_decode([l.numpy() for l in lv], 1, (17, 3))
    np.testing.assert_allclose(dec, GOLD["decoded"], rtol=0, atol=2e-5)


def test_state_dict_layout():
    m = dp.build_synthetic("tiny", (17, 3), 1, seed=0)
    keys = set(m.state_dict())
    for k in ("backbone.downsample_layers.0.0.weight", "backbone.stages.2.8.gamma", "adaptor.adapters.0.1.conv.weight",
              "sppf.conv2.bn.running_var", "pan.fpn_conv2.m.1.cv2.conv.weight", "pan.pan_conv2.bn.weight",
              "head.cv3.0.0.0.conv.weight", "head.cv4.2.2.bias", "head.stride"):
        assert k in keys, k
    assert m.head.cv3[0][0][0].conv.groups == 192 and m.head.cv4[0][0].conv.out_channels == 51       # head.py:139-140
