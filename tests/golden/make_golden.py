"""Generate the golden fixtures under tests/golden/ .  Run ONCE in the build container, where the
reference is mounted read-only at /root/reference:

    python tests/golden/make_golden.py

* ``resnet50_ref_features.npz`` -- output of the REFERENCE's own model code
  (``DeepPose(17,'resnet50',pretrained=False).backbone.forward_features`` =
  vendored timm ResNet-50, HPE/models/reference/bottom_up/regression/models/DeepPose.py:4-13,
  HPE/thirdparty/pytorch-image-models/timm/models/resnet.py:742-756) on a seeded input, with the
  seeded synthetic weights of ``oracle.resnet_ref.build_synthetic`` loaded through
  ``load_state_dict`` (key-compatible).  This is what pins ``oracle/resnet_ref.py``.
* ``heatmap_stage_golden.npz`` -- outputs of the App.-A oracle (``oracle/heatmap_ref.py``) for the
  fixed seeds of SURVEY.md §8(d).  The reference has no implementation of these stages, so these
  vectors are self-generated ("parity unpinned"); they freeze the oracle so that any later edit
  of it is visible.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True

from oracle import heatmap_ref as hr  # noqa: E402
from oracle import resnet_ref as rr   # noqa: E402


def load_reference_deeppose():
    ref = "/root/reference/HPE"
    sys.path[:0] = [ref, ref + "/thirdparty/pytorch-image-models"]
    # HPE/models/__init__.py star-imports the regression package only -> DeepPose
    from models.reference.bottom_up.regression.models.DeepPose import DeepPose
    return DeepPose


def main():
    torch.set_num_threads(8)
    # ---- backbone: reference code, oracle weights
    oracle_model = rr.build_synthetic(17, seed=0)
    DeepPose = load_reference_deeppose()
    ref_model = DeepPose(17, "resnet50", pretrained=False).eval()
    sd = {k: v for k, v in oracle_model.backbone.state_dict().items()}
    missing, unexpected = ref_model.backbone.load_state_dict(sd, strict=False)
    assert set(missing) == {"fc.weight", "fc.bias"} and not unexpected, (missing, unexpected)
    g = torch.Generator().manual_seed(0)
    x = torch.rand(2, 3, 128, 96, generator=g)
    with torch.no_grad():
        feat_ref = ref_model.backbone.forward_features(x)
        feat_orc = oracle_model.backbone.forward_features(x)
    err = (feat_ref - feat_orc).abs().max().item()
    print("reference vs oracle ResNet-50 features: max-abs diff", err, "shape", tuple(feat_ref.shape))
    assert err < 1e-5
    np.savez_compressed(os.path.join(HERE, "resnet50_ref_features.npz"),
                        features=feat_ref.numpy(), input_seed=0, weight_seed=0)

    # ---- heatmap stages (self-golden)
    hm = hr.synth_heatmaps(8, 17, 64, 48, seed=3)
    center, scale = hr.synth_center_scale(8, seed=4)
    preds, maxvals = hr.get_final_preds(hm, center, scale)
    _, _, idx = hr.get_max_preds(hm)
    joints = hr.synth_joints(8, 17, (256, 192), seed=5)
    target, tw = hr.generate_target(joints, (64, 48), 2.0)
    loss, grad = hr.joints_mse_loss(hm, target, tw)
    np.savez_compressed(os.path.join(HERE, "heatmap_stage_golden.npz"),
                        preds=preds, maxvals=maxvals, idx=idx, target_weight=tw,
                        target_sum=target.sum(axis=(2, 3)), loss=np.float32(loss),
                        grad_abs_sum=np.float64(np.abs(grad).sum()),
                        # FULL arrays for the first two samples (sums alone would let compensating errors through)
                        joints=joints, heatmaps_first2=hm[:2], target_first2=target[:2], grad_first2=grad[:2])
    print("wrote golden fixtures; loss", loss)


if __name__ == "__main__":
    main()
