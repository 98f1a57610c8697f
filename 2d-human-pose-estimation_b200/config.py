"""Config loader accepting the reference's keys.

The reference uses a yacs ``CfgNode`` with defaults in /root/reference/HPE/configs/default.py:4-110,
a YAML overlay per method (HPE/configs/method/deep_pose.yaml) and the CLI flags
``--cfg --test --gpus --ckpt`` (HPE/configs/parser.py:3-28; merge in default.py:114-126).  yacs is
not a dependency here; this is a small attribute-dict with the same keys, the same defaults for
everything the heatmap path reads, and the same override rules.  Keys the heatmap path needs and the
reference lacks (image size, heatmap size, sigma) are ADDED with defaults, never required.
"""
from __future__ import annotations

import argparse
import copy

import yaml


class CfgNode(dict):
    """Attribute-style nested dict (subset of yacs.CfgNode: attribute access, merge, freeze)."""

    def __init__(self, d=None):
        super().__init__()
        object.__setattr__(self, "_frozen", False)
        for k, v in (d or {}).items():
            self[k] = CfgNode(v) if isinstance(v, dict) else v

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        if self._frozen:
            raise AttributeError(f"cannot set {k}: config is frozen")
        self[k] = v

    def defrost(self):
        object.__setattr__(self, "_frozen", False)
        for v in self.values():
            if isinstance(v, CfgNode):
                v.defrost()

    def freeze(self):
        object.__setattr__(self, "_frozen", True)
        for v in self.values():
            if isinstance(v, CfgNode):
                v.freeze()

    def merge(self, other: dict):
        for k, v in other.items():
            if isinstance(v, dict) and isinstance(self.get(k), CfgNode):
                self[k].merge(v)
            else:
                self[k] = CfgNode(v) if isinstance(v, dict) else v

    def clone(self):
        return copy.deepcopy(self)


def _defaults() -> CfgNode:
    return CfgNode({
        # reference keys (HPE/configs/default.py:4-110), same defaults
        "title": "HPE_Experiment", "description": "", "seed": 42, "use_deterministic": True,
        "gpus": (0,), "saveDir": "",
        "model": {"name": "DeepPose", "backbone": "resnet50", "iou_threshold": 0.5, "nkpts": 17,
                  "checkpoint": "", "pretrained": True,
                  # added for the heatmap path (not in the reference; defaults per SURVEY.md §8b)
                  "image_size": (256, 192), "heatmap_size": (64, 48), "sigma": 2.0},
        "train": {"batch_size": 32, "num_epochs": 100, "mini_batch_count": 1, "optimizer": "SGD",
                  "shuffle": True, "lr": 2.5e-4, "lr_step": 30, "lr_gamma": 0.1, "momentum": 0.9,
                  "weight_decay": 5e-4, "save_freq": 10, "dropLR": 5, "dropMag": 0.7},
        "valid": {"val_interval": 5, "save_interval": 5},      # default.py:62-63 (they override the 3 / 5 set at :57-58)
        "test": {"batch_size": 32, "shuffle": False},
        "data": {"dataset": "MPII", "data_dir": "./data/MPII/images", "train_split": "train",
                 "val_split": "valid", "test_split": "test", "num_workers": 4, "pin_memory": True,
                 "drop_last": False},
        "metric": {"classify": "CMet", "detect": "DMet", "pose": "PMet"},
        "hardware": {"device": "cuda", "mixed_precision": False},
    })


C = _defaults()


def parse_args(argv=None):
    """Same flags as HPE/configs/parser.py:3-28."""
    p = argparse.ArgumentParser(description="b200 heatmap pose")
    p.add_argument("--cfg", default=None, type=str, help="Path to the configuration file")
    p.add_argument("--test", default=False, help="model test")
    p.add_argument("--gpus", "-g", default=None, type=str, help="GPUs to use, e.g. 0,1,2,3")
    p.add_argument("--ckpt", default=None, type=str, help="Path to the checkpoint")
    return p.parse_known_args(argv)[0]


def update_config(cfg: CfgNode, args) -> CfgNode:
    """Same rules as HPE/configs/default.py:114-126: merge YAML, then --gpus / --ckpt override, freeze.
    (``gpus`` accepts multi-digit ids, unlike the reference's char-by-char parse.)"""
    cfg.defrost()
    if getattr(args, "cfg", None):
        with open(args.cfg) as f:
            cfg.merge(yaml.safe_load(f) or {})
    if isinstance(cfg.gpus, str):
        cfg.gpus = tuple(int(g) for g in cfg.gpus.replace(" ", "").split(",") if g != "")
    if getattr(args, "gpus", None) is not None:
        cfg.gpus = tuple(int(g) for g in str(args.gpus).split(",") if g != "")
    if getattr(args, "ckpt", None) is not None:
        cfg.model.checkpoint = args.ckpt
    cfg.freeze()
    return cfg


def load_config(path=None, **overrides) -> CfgNode:
    cfg = _defaults()
    update_config(cfg, argparse.Namespace(cfg=path, gpus=overrides.pop("gpus", None), ckpt=overrides.pop("ckpt", None)))
    return cfg
