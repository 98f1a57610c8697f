"""Golden vectors for the reference's REAL decode / NMS / loss / metric code.

Run ONCE in the build container (reference mounted read-only at /root/reference):

    python tests/golden/make_golden_detpose.py

Every array in ``detpose_golden.npz`` is an output of the reference's own functions, imported from
/root/reference/HPE, on the seeded inputs of ``oracle.detpose_ref.synth_*``:
  * ``PoseHead._inference_pose`` (HPE/models/nn/modules/head.py:185-211) called on raw level maps;
  * ``non_max_suppression_pose`` (HPE/test.py:69-185; extracted by AST because importing test.py opens a window);
  * ``ComputeLoss.__call__`` (HPE/loss.py:349-448) forward and autograd gradients;
  * ``OKS.compute`` / ``PCK.compute`` / ``KeypointLoss`` (HPE/core/metric/oksap.py, pck.py, HPE/core/loss/keypoint.py).
These pin ``oracle/detpose_ref.py`` (tests/test_detpose_oracle_cpu.py).
"""
import ast
import contextlib
import io
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True
REF = "/root/reference/HPE"

from oracle import detpose_ref as dr  # noqa: E402

SHAPES = [(32, 24), (16, 12), (8, 6)]          # 256x192 input, strides 8/16/32
NC, K = 1, 17


def load_reference():
    sys.path[:0] = [REF, REF + "/thirdparty/dinov3", REF + "/thirdparty/pytorch-image-models"]
    from models.nn.modules.head import PoseHead
    import loss as ref_loss
    from core.metric.oksap import OKS
    from core.metric.pck import PCK
    from core.loss.keypoint import KeypointLoss
    src = open(REF + "/test.py").read()
    tree = ast.parse(src)
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "non_max_suppression_pose"][0]
    ns = {}
    import torchvision
    exec(compile(ast.Module([fn], []), "ref_test_nms", "exec"), {"torch": torch, "torchvision": torchvision, "np": np}, ns)
    return PoseHead, ref_loss, OKS, PCK, KeypointLoss, ns["non_max_suppression_pose"]


def synth_metric_inputs(N=64, seed=13):
    rng = np.random.default_rng(seed)
    gt = np.zeros((N, K, 3), np.float32)
    gt[..., 0] = rng.uniform(0, 192, (N, K)); gt[..., 1] = rng.uniform(0, 256, (N, K))
    gt[..., 2] = rng.choice([0.0, 1.0, 2.0], size=(N, K), p=[0.2, 0.2, 0.6])
    gt[0, :, 2] = 0.0                                        # instance with nothing visible
    pred = gt[..., :2] + rng.normal(0, 6.0, (N, K, 2)).astype(np.float32)
    areas = rng.uniform(400, 20000, N).astype(np.float32)
    return gt, pred.astype(np.float32), areas


def main():
    torch.set_num_threads(4)
    PoseHead, ref_loss, OKS, PCK, KeypointLoss, ref_nms = load_reference()
    out = {}
    # ---- a7 decode
    B = 4
    levels = dr.synth_head_levels(B, NC, K, SHAPES, seed=11)
    head = PoseHead(NC, (K, 3), (192, 384, 768)).eval()
    with torch.no_grad():
        dec = head._inference_pose([torch.from_numpy(l.copy()) for l in levels]).numpy()
    out["decoded"] = dec
    mine = dr.pose_decode(levels, NC, (K, 3))
    print("decode: oracle vs reference max-abs", np.abs(mine - dec).max())
    # ---- a8 NMS
    with contextlib.redirect_stdout(io.StringIO()):
        res = ref_nms(torch.from_numpy(dec), conf_thres=0.25, iou_thres=0.45, nc=NC, nkpts=K)
    out["nms_count"] = np.array([r["scores"].shape[0] for r in res], np.int64)
    out["nms_boxes"] = np.concatenate([r["boxes"].numpy() for r in res]).astype(np.float32)
    out["nms_scores"] = np.concatenate([r["scores"].numpy() for r in res]).astype(np.float32)
    out["nms_labels"] = np.concatenate([r["labels"].numpy() for r in res]).astype(np.int64)
    out["nms_keypoints"] = np.concatenate([r["keypoints"].numpy() for r in res]).astype(np.float32)
    print("nms: kept per image", out["nms_count"])
    # a multi-class case (class offset 7680) on a denser map
    lv3 = dr.synth_head_levels(2, 3, 4, SHAPES, seed=21, pos_rate=0.15)
    head3 = PoseHead(3, (4, 3), (192, 384, 768)).eval()
    with torch.no_grad():
        dec3 = head3._inference_pose([torch.from_numpy(l.copy()) for l in lv3]).numpy()
    with contextlib.redirect_stdout(io.StringIO()):
        res3 = ref_nms(torch.from_numpy(dec3), conf_thres=0.25, iou_thres=0.45, nc=3, nkpts=4)
    out["nms3_count"] = np.array([r["scores"].shape[0] for r in res3], np.int64)
    out["nms3_scores"] = np.concatenate([r["scores"].numpy() for r in res3]).astype(np.float32)
    out["nms3_labels"] = np.concatenate([r["labels"].numpy() for r in res3]).astype(np.int64)
    out["nms3_boxes"] = np.concatenate([r["boxes"].numpy() for r in res3]).astype(np.float32)
    print("nms3: kept per image", out["nms3_count"])
    # ---- a9 ComputeLoss
    class _M(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.head = head
    crit = ref_loss.ComputeLoss(_M())
    targets = torch.from_numpy(dr.synth_targets(B, K, seed=12, nc=NC))
    preds = [torch.from_numpy(l.copy()).requires_grad_(True) for l in levels]
    total, parts = crit(preds, targets)
    total.backward()
    out["loss_total"] = np.float32(total.item()); out["loss_parts"] = np.asarray(parts, np.float32)
    for i, p in enumerate(preds):
        out[f"loss_grad{i}"] = p.grad.numpy()
    print("loss:", total.item(), parts)
    # empty-target case
    total0, parts0 = crit([torch.from_numpy(l.copy()) for l in levels], torch.zeros(0, 6 + 3 * K))
    out["loss0_total"] = np.float32(total0.item()); out["loss0_parts"] = np.asarray(parts0, np.float32)
    # ---- a10 metrics
    gt, pred, areas = synth_metric_inputs()
    out["oks"] = OKS(torch.from_numpy(dr.OKS_SIGMAS)).compute(torch.from_numpy(gt), torch.from_numpy(pred), torch.from_numpy(areas)).numpy()
    vis = gt[..., 2] > 0
    out["pck"] = np.float64(PCK(alpha=0.05).compute(torch.from_numpy(gt[..., :2]), torch.from_numpy(pred), torch.from_numpy(vis)))
    out["pck_norm"] = np.float64(PCK(alpha=0.5).compute(torch.from_numpy(gt[..., :2]), torch.from_numpy(pred), None, torch.tensor(20.0)))
    kl = KeypointLoss(torch.from_numpy(dr.OKS_SIGMAS))
    out["keypoint_loss"] = np.float32(kl(torch.from_numpy(pred), torch.from_numpy(gt[..., :2]), torch.from_numpy(vis.astype(np.float32)),
                                         torch.from_numpy(areas).unsqueeze(-1)).item())
    print("oks mean", out["oks"].mean(), "pck", out["pck"], out["pck_norm"], "kptloss", out["keypoint_loss"])
    np.savez_compressed(os.path.join(HERE, "detpose_golden.npz"), **out)
    print("wrote detpose_golden.npz")


if __name__ == "__main__":
    main()
