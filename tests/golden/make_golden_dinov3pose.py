"""Golden vectors for the reference's real pose model.

Run ONCE in the build container (reference mounted read-only at /root/reference):

    python tests/golden/make_golden_dinov3pose.py

Builds the REFERENCE's own ``DINOv3Pose('dinov3_convnext_tiny', nkpts=(17,3), ncls=1)`` (HPE/models/pose.py:9-65)
with the shims of SURVEY.md App. C (stubbed ``torch.hub.load``, fabricated local checkpoint), loads the seeded
``state_dict`` of ``oracle.dinov3pose_ref.build_synthetic('tiny')`` into it (strict: the key sets are identical),
and stores its train-mode output (3 raw level maps) and eval-mode output ([B,52,A] decoded) on a seeded image.
"""
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True
REF = "/root/reference/HPE"

from oracle import dinov3pose_ref as dp  # noqa: E402


def load_reference_pose(size="tiny"):
    sys.path[:0] = [REF, REF + "/thirdparty/dinov3", REF + "/thirdparty/pytorch-image-models"]
    from dinov3.models.convnext import ConvNeXt, convnext_sizes
    torch.hub.load = lambda *a, **k: ConvNeXt(**convnext_sizes["tiny"])           # only its __class__ is used
    scratch = tempfile.mkdtemp(prefix="hpe_ref_")
    os.chdir(scratch); os.environ["TORCH_HOME"] = scratch + "/torchhome"
    from models.backbones.dinov3convnext import convnext_ckps
    m = ConvNeXt(**convnext_sizes[size]); m.init_weights()
    os.makedirs(os.path.dirname(convnext_ckps[size]), exist_ok=True)
    torch.save(m.state_dict(), convnext_ckps[size])
    from models.pose import DINOv3Pose
    return DINOv3Pose(f"dinov3_convnext_{size}", nkpts=(17, 3), ncls=1, device="cpu")


def main():
    torch.set_num_threads(8)
    oracle = dp.build_synthetic("tiny", (17, 3), 1, seed=0)
    ref = load_reference_pose("tiny")
    rk, ok = set(ref.state_dict().keys()), set(oracle.state_dict().keys())
    assert rk == ok, (sorted(rk - ok)[:5], sorted(ok - rk)[:5])
    ref.load_state_dict(oracle.state_dict(), strict=True)
    x = torch.rand(2, 3, 128, 96, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        ref.train()
        for mod in ref.modules():                     # train-mode OUTPUT FORMAT with eval-mode BatchNorm statistics
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.eval()
        lv = ref(x)
        ref.eval()
        dec = ref(x)
        mine = oracle(x)
    for a, b in zip(lv, mine):
        print("level", tuple(a.shape), "reference vs oracle max-abs", (a - b).abs().max().item(), "scale", a.abs().max().item())
    np.savez_compressed(os.path.join(HERE, "dinov3pose_golden.npz"), level0=lv[0].numpy(), level1=lv[1].numpy(),
                        level2=lv[2].numpy(), decoded=dec.numpy())
    print("wrote dinov3pose_golden.npz; decoded", tuple(dec.shape))


if __name__ == "__main__":
    main()
