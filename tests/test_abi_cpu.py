"""CPU suite: the C-ABI library loads and exports every symbol include/hpe_b200.h declares; host logic."""
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "hpe_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hpe_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(hpe):
    lib = hpe._lib.lib()
    names = declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/hpe_b200.h but not exported"
        assert n in hpe._lib.SIGNATURES, f"{n} has no ctypes signature"
    assert lib.hpe_abi_version() == 1


def test_no_cpu_fallback(hpe):
    hm = torch.zeros(1, 2, 8, 8)
    with pytest.raises(RuntimeError, match="no CPU path"):
        hpe.get_max_preds(hm)
    with pytest.raises(RuntimeError, match="no CPU path"):
        hpe.JointsMSELoss()(hm, torch.zeros(1, 2, 3), torch.ones(1, 2, 1))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "2d-human-pose-estimation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, f


def test_affine_matches_oracle(hpe):
    from oracle import heatmap_ref as hr
    c, s = hr.synth_center_scale(5)
    M = hpe.affine_from_center_scale(torch.tensor(c), torch.tensor(s), (64, 48)).numpy()
    assert abs(M - hr.affine_from_center_scale(c, s, (64, 48))).max() < 1e-4


def test_config_reference_keys(hpe, tmp_path):
    cfg = hpe.load_config()
    assert cfg.model.name == "DeepPose" and cfg.model.backbone == "resnet50" and cfg.model.nkpts == 17
    assert cfg.train.mini_batch_count == 1 and cfg.model.heatmap_size == (64, 48)
    y = tmp_path / "m.yaml"
    y.write_text("title: 'DeepPose'\ngpus: 0, 1, 2, 3\nmodel:\n  nkpts: 133\n  checkpoint:\ntrain:\n  optimizer: 'RMSprop'\n")
    cfg = hpe.load_config(str(y), ckpt="x.pt")
    assert cfg.gpus == (0, 1, 2, 3) and cfg.model.nkpts == 133 and cfg.model.checkpoint == "x.pt"
    assert cfg.train.optimizer == "RMSprop" and cfg.train.lr == 2.5e-4
    with pytest.raises(AttributeError):
        cfg.seed = 1


def test_direct_gradient_target_is_granted_once_per_step():
    """posetrain._direct_target: a module's parameter may be written straight into the flat gradient buffer by the ONE kernel that
    produces its gradient; a second use of the same module in the same step (shared weights) must fall back to autograd's
    accumulation, and nothing is granted outside a PoseTrainStep."""
    from importlib import import_module
    import torch
    pt = import_module("2d-human-pose-estimation_b200.posetrain")
    m = torch.nn.Conv2d(8, 8, 1)
    view = torch.zeros_like(m.weight)
    assert pt._direct_target(m, "weight") is None                 # no step in flight, nothing registered
    m.__dict__["_hpe_grad"] = {"weight": view}
    assert pt._direct_target(m, "weight") is None                 # registered, but no step in flight
    pt._DIRECT[0] = ("step", 1)
    try:
        assert pt._direct_target(m, "weight") is view
        assert pt._direct_target(m, "weight") is None             # second use in the same step
        assert pt._direct_target(m, "bias") is None               # never registered
        pt._DIRECT[0] = ("step", 2)
        assert pt._direct_target(m, "weight") is view             # next step
    finally:
        pt._DIRECT[0] = 0
