"""Target-weighted joints MSE with in-kernel Gaussian targets.

Call shape = the reference trainer's: ``loss = criterion(output, target, target_weight)`` with
``target_weight = meta1.float().unsqueeze(-1)`` -> ``[B,K,1]`` (/root/reference/HPE/engine/trainer.py:92),
returning a 0-dim tensor that supports ``.item()`` and ``.backward()`` (trainer.py:93,97).  The
reference's ``Builder.loss()`` is a stub (HPE/core/builder.py:26-27); semantics follow SURVEY.md
App. A.5 / A.6.

``target`` may be
  * a materialised heatmap ``[B,K,H,W]`` (drop-in mode, ``hpe_joints_mse_dense``), or
  * joints ``[B,K,2|3]`` in input-image pixels (fused mode, ``hpe_joints_mse_fused``: the Gaussian is
    evaluated inside the kernel and never stored).  With ``[B,K,2]`` the visibility comes from
    ``target_weight``.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib


class _FusedMSE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, joints, sigma, stride_x, stride_y):
        _lib.require_cuda(pred, joints)
        B, K, H, W = pred.shape
        partials = torch.empty(B * K, dtype=torch.float32, device=pred.device)
        loss = torch.empty((), dtype=torch.float32, device=pred.device)
        grad = torch.empty_like(pred) if pred.requires_grad else None
        with torch.cuda.device(pred.device):
            rc = _lib.lib().hpe_joints_mse_fused(_lib.ptr(pred), _lib.dtype_code(pred), _lib.ptr(joints), B, K, H, W,
                                                 float(sigma), float(stride_x), float(stride_y), _lib.ptr(partials),
                                                 _lib.ptr(loss), _lib.ptr(grad), 1.0, 0, _lib.stream_ptr())
        _lib.check(rc, "hpe_joints_mse_fused")
        ctx.grad = grad
        return loss

    @staticmethod
    def backward(ctx, gout):
        g = ctx.grad
        if g is None:
            return None, None, None, None, None
        return g * gout.to(g.dtype), None, None, None, None


class _DenseMSE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target, weight):
        _lib.require_cuda(pred, target, weight)
        B, K, H, W = pred.shape
        partials = torch.empty(B * K, dtype=torch.float32, device=pred.device)
        loss = torch.empty((), dtype=torch.float32, device=pred.device)
        grad = torch.empty_like(pred) if pred.requires_grad else None
        with torch.cuda.device(pred.device):
            rc = _lib.lib().hpe_joints_mse_dense(_lib.ptr(pred), _lib.dtype_code(pred), _lib.ptr(target), _lib.ptr(weight),
                                                 B, K, H, W, _lib.ptr(partials), _lib.ptr(loss), _lib.ptr(grad), 1.0,
                                                 _lib.stream_ptr())
        _lib.check(rc, "hpe_joints_mse_dense")
        ctx.grad = grad
        return loss

    @staticmethod
    def backward(ctx, gout):
        g = ctx.grad
        if g is None:
            return None, None, None
        return g * gout.to(g.dtype), None, None


def generate_target(joints: torch.Tensor, heatmap_hw, sigma: float = 2.0, stride=(4.0, 4.0)):
    """App. A.5 on the device: joints ``[B,K,3]`` -> ``target[B,K,H,W] f32, target_weight[B,K,1] f32``."""
    _lib.require_cuda(joints)
    joints = joints.to(torch.float32).contiguous()
    B, K, _ = joints.shape
    H, W = heatmap_hw
    target = torch.empty((B, K, H, W), dtype=torch.float32, device=joints.device)
    weight = torch.empty((B, K, 1), dtype=torch.float32, device=joints.device)
    with torch.cuda.device(joints.device):
        rc = _lib.lib().hpe_generate_target(_lib.ptr(joints), B, K, H, W, float(sigma), float(stride[0]),
                                            float(stride[1]), _lib.ptr(target), _lib.ptr(weight), _lib.stream_ptr())
    _lib.check(rc, "hpe_generate_target")
    return target, weight


class JointsMSELoss(nn.Module):
    """``criterion(output, target, target_weight)`` -- App. A.6 with ``use_target_weight=True`` semantics."""

    def __init__(self, use_target_weight: bool = True, sigma: float = 2.0, stride=(4.0, 4.0)):
        super().__init__()
        self.use_target_weight = use_target_weight
        self.sigma = float(sigma)
        self.stride = (float(stride[0]), float(stride[1]))

    def forward(self, output: torch.Tensor, target: torch.Tensor, target_weight: torch.Tensor = None):
        if output.dim() != 4:
            raise ValueError(f"output must be heatmaps [B,K,H,W], got {tuple(output.shape)}")
        B, K, H, W = output.shape
        output = output.contiguous()
        if target.dim() == 4:                                   # drop-in mode: materialised heatmaps
            if tuple(target.shape) != (B, K, H, W):
                raise ValueError(f"target {tuple(target.shape)} does not match output {tuple(output.shape)}")
            if self.use_target_weight and target_weight is not None:
                w = target_weight.to(torch.float32).reshape(B, K).contiguous()
            else:
                w = torch.ones((B, K), dtype=torch.float32, device=output.device)
            return _DenseMSE.apply(output, target.to(torch.float32).contiguous(), w)
        if target.dim() == 3 and target.shape[:2] == (B, K) and target.shape[2] in (2, 3):   # fused mode
            joints = target.to(torch.float32)
            if joints.shape[2] == 2:
                if target_weight is None:
                    vis = torch.ones((B, K, 1), dtype=torch.float32, device=output.device)
                else:
                    vis = target_weight.to(torch.float32).reshape(B, K, 1)
                joints = torch.cat([joints, vis], dim=2)
            elif target_weight is not None and self.use_target_weight:
                joints = joints.clone()
                joints[..., 2] = joints[..., 2] * (target_weight.to(torch.float32).reshape(B, K) > 0)
            return _FusedMSE.apply(output, joints.contiguous(), self.sigma, self.stride[0], self.stride[1])
        raise ValueError(f"target must be [B,K,H,W] heatmaps or [B,K,2|3] joints, got {tuple(target.shape)}")
