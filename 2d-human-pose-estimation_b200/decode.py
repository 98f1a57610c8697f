"""Host-side heatmap decode: ``preds, maxvals = decode(output, center, scale)``.

This is the call a metric / test loop makes on ``output = model(data)``
(/root/reference/HPE/engine/trainer.py:90, :110-113).  The reference itself ships no heatmap decode
(SURVEY.md §0); names and semantics follow SURVEY.md App. A.2-A.4 (``get_max_preds``,
``get_final_preds``).  All device work happens in ``hpe_decode_heatmaps`` (csrc/decode.cu).
"""
from __future__ import annotations

import torch

from . import _lib


def affine_from_center_scale(center: torch.Tensor, scale: torch.Tensor, heatmap_hw) -> torch.Tensor:
    """App. A.4 host-side matrix build: r = 200*sw/W, M = [[r,0,cx-r*W/2],[0,r,cy-r*H/2]] (fp32, [B,2,3]).
    Tiny per-sample metadata; computed on whatever device ``center`` lives on."""
    H, W = heatmap_hw
    center = torch.as_tensor(center, dtype=torch.float32).reshape(-1, 2)
    scale = torch.as_tensor(scale, dtype=torch.float32, device=center.device).reshape(-1, 2)
    r = 200.0 * scale[:, 0] / float(W)
    M = torch.zeros(center.shape[0], 2, 3, dtype=torch.float32, device=center.device)
    M[:, 0, 0] = r
    M[:, 1, 1] = r
    M[:, 0, 2] = center[:, 0] - r * float(W) * 0.5
    M[:, 1, 2] = center[:, 1] - r * float(H) * 0.5
    return M


def _decode(hm: torch.Tensor, affine, refine: bool, want_idx: bool):
    if hm.dim() != 4:
        raise ValueError(f"heatmaps must be [B,K,H,W], got {tuple(hm.shape)}")
    _lib.require_cuda(hm, affine)
    B, K, H, W = hm.shape
    if B * K == 0:
        z = hm.new_zeros((B, K, 2), dtype=torch.float32)
        return z, hm.new_zeros((B, K, 1), dtype=torch.float32), hm.new_zeros((B, K), dtype=torch.int32)
    if affine is not None:
        if affine.dtype != torch.float32 or tuple(affine.shape) != (B, 2, 3):
            raise ValueError(f"affine must be float32 [B,2,3], got {affine.dtype} {tuple(affine.shape)}")
    coords = torch.empty((B, K, 2), dtype=torch.float32, device=hm.device)
    maxvals = torch.empty((B, K, 1), dtype=torch.float32, device=hm.device)
    idx = torch.empty((B, K), dtype=torch.int32, device=hm.device) if want_idx else None
    with torch.cuda.device(hm.device):
        rc = _lib.lib().hpe_decode_heatmaps(_lib.ptr(hm), _lib.dtype_code(hm), B, K, H, W, _lib.ptr(affine),
                                            int(refine), _lib.ptr(coords), _lib.ptr(maxvals), _lib.ptr(idx),
                                            _lib.stream_ptr())
    _lib.check(rc, "hpe_decode_heatmaps")
    return coords, maxvals, idx


def get_max_preds(heatmaps: torch.Tensor):
    """App. A.2: ``coords[B,K,2] f32, maxvals[B,K,1] f32, idx[B,K] i32`` (first-max ties, masked at <= 0)."""
    return _decode(heatmaps, None, False, True)


def get_final_preds_affine(heatmaps: torch.Tensor, affine: torch.Tensor, return_idx: bool = False):
    """A.2 -> A.3 -> A.4 with an explicit ``[B,2,3]`` fp32 inverse-affine (any rotation)."""
    coords, maxvals, idx = _decode(heatmaps, affine, True, return_idx)
    return (coords, maxvals, idx) if return_idx else (coords, maxvals)


def get_final_preds_xyv(heatmaps: torch.Tensor, affine: torch.Tensor, out: torch.Tensor = None, refine: bool = True):
    """A.2 -> A.3 -> A.4 into ONE packed tensor ``xyv[B,K,3] = (x, y, maxval)``: the row layout the data-parallel
    result gather ships, written by the decode kernel itself -- no ``torch.cat`` afterwards."""
    if heatmaps.dim() != 4:
        raise ValueError(f"heatmaps must be [B,K,H,W], got {tuple(heatmaps.shape)}")
    _lib.require_cuda(heatmaps, affine, out)
    B, K, H, W = heatmaps.shape
    if affine is not None and (affine.dtype != torch.float32 or tuple(affine.shape) != (B, 2, 3)):
        raise ValueError(f"affine must be float32 [B,2,3], got {affine.dtype} {tuple(affine.shape)}")
    if out is None:
        out = torch.empty((B, K, 3), dtype=torch.float32, device=heatmaps.device)
    elif out.dtype != torch.float32 or tuple(out.shape) != (B, K, 3):
        raise ValueError(f"out must be float32 [B,K,3], got {out.dtype} {tuple(out.shape)}")
    if B * K == 0:
        return out
    with torch.cuda.device(heatmaps.device):
        rc = _lib.lib().hpe_decode_heatmaps_xyv(_lib.ptr(heatmaps), _lib.dtype_code(heatmaps), B, K, H, W, _lib.ptr(affine),
                                                int(refine), _lib.ptr(out), 0, _lib.stream_ptr())
    _lib.check(rc, "hpe_decode_heatmaps_xyv")
    return out


def get_final_preds(heatmaps: torch.Tensor, center, scale):
    """App. A.4 top level: ``preds[B,K,2]`` in image pixels and ``maxvals[B,K,1]``."""
    M = affine_from_center_scale(torch.as_tensor(center), torch.as_tensor(scale), heatmaps.shape[2:])
    M = M.to(heatmaps.device, non_blocking=True).contiguous()
    return get_final_preds_affine(heatmaps, M)


decode = get_final_preds


class HostDecoder:
    """Reference-facing call with HOST buffers (the e2e-benchmark path): pinned staging in, pinned results
    out, copies + kernel issued by ``hpe_decode_heatmaps_host`` on one stream."""

    def __init__(self, B, K, H, W, dtype=torch.float32, device="cuda"):
        self.shape = (B, K, H, W)
        self.dtype = dtype
        self.device = torch.device(device)
        code = _lib.HPE_BF16 if dtype == torch.bfloat16 else _lib.HPE_F32
        self.code = code
        nbytes = _lib.lib().hpe_decode_host_scratch_bytes(code, B, K, H, W)
        self.scratch = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        self.coords = torch.empty((B, K, 2), dtype=torch.float32).pin_memory()
        self.maxvals = torch.empty((B, K, 1), dtype=torch.float32).pin_memory()
        self.idx = torch.empty((B, K), dtype=torch.int32).pin_memory()

    def __call__(self, h_heatmaps: torch.Tensor, h_affine: torch.Tensor, refine: bool = True):
        B, K, H, W = self.shape
        assert not h_heatmaps.is_cuda and tuple(h_heatmaps.shape) == self.shape and h_heatmaps.dtype == self.dtype
        assert h_heatmaps.is_contiguous()
        with torch.cuda.device(self.device):
            rc = _lib.lib().hpe_decode_heatmaps_host(
                h_heatmaps.data_ptr(), self.code, B, K, H, W,
                0 if h_affine is None else h_affine.data_ptr(), int(refine),
                self.coords.data_ptr(), self.maxvals.data_ptr(), self.idx.data_ptr(),
                self.scratch.data_ptr(), self.scratch.numel(), _lib.stream_ptr())
        _lib.check(rc, "hpe_decode_heatmaps_host")
        return self.coords, self.maxvals, self.idx
