"""B200-native top-down heatmap pose hot path (sm_100a CUDA behind a C ABI; no CPU fallback).

The directory name is not a Python identifier; import it with
``importlib.import_module("2d-human-pose-estimation_b200")`` or through the root-level alias
module ``hpe_b200``.

Public surface mirrors the reference's duck-typed plugin API:
``models.__all__``-style registry + ``Model(nkpts, backbone, pretrained)`` (HPE/core/builder.py:15-19),
``criterion(output, target, target_weight)`` (HPE/engine/trainer.py:92), ``decode(output, center, scale)``.
"""
from . import _lib
from .config import C as default_cfg, CfgNode, load_config, parse_args, update_config
from .decode import (HostDecoder, affine_from_center_scale, decode, get_final_preds, get_final_preds_affine,
                     get_final_preds_xyv, get_max_preds)
from .loss import JointsMSELoss, generate_target
from . import models
from .models import DeepPose, HeatmapPose, build_model
from . import train
from .train import TrainPlan, TrainStep, adamw_step, grad_sumsq
from .detpose import (OKS, OKS_SIGMAS, PCK, ComputeLoss, KeypointLoss, non_max_suppression_pose, pose_decode,
                      pose_nms_raw)
from .posemodel import DINOv3Pose
from .posetrain import PoseTrainStep
from .pipeline import PoseEstimator, ResultGather, gather_keypoints, gather_keypoints_equal, shard_range

__version__ = "0.1.0"
