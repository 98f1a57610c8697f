"""ctypes binding of libhpe_b200.so (the C ABI declared in include/hpe_b200.h).

There is deliberately no fallback: if the shared library is missing or a call fails, a
``RuntimeError`` is raised.  PyTorch is only used by callers for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "lib", "libhpe_b200.so")

HPE_F32, HPE_BF16 = 0, 1
ACT_NONE, ACT_RELU, ACT_SILU, ACT_GELU, ACT_GELU_LN = 0, 1, 2, 3, 4
CONV_1x1, CONV_3x3, CONV_STEM7x7, CONV_DECONV4, CONV_FINAL = 0, 1, 2, 3, 4
CONV_DGRAD3x3S2, CONV_DGRAD1x1S2, CONV_4x4S2, CONV_2x2S2 = 5, 6, 7, 8


class ConvDesc(C.Structure):
    _fields_ = [("kind", C.c_int), ("B", C.c_int), ("H", C.c_int), ("W", C.c_int), ("Cin", C.c_int),
                ("Cout", C.c_int), ("stride", C.c_int), ("act", C.c_int), ("has_residual", C.c_int),
                ("in_cstride", C.c_int), ("out_cstride", C.c_int), ("Cin2", C.c_int), ("stride2", C.c_int), ("head2", C.c_int), ("pool", C.c_int), ("chain", C.c_int)]


_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t

# name -> (restype, argtypes); must list every symbol include/hpe_b200.h declares
SIGNATURES = {
    "hpe_abi_version": (_i, []),
    "hpe_last_error": (C.c_char_p, []),
    "hpe_check_device": (_i, []),
    "hpe_decode_heatmaps": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _vp, _vp, _vp, _vp]),
    "hpe_decode_heatmaps_xyv": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _vp, _vp, _vp]),
    "hpe_decode_host_scratch_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "hpe_decode_heatmaps_host": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "hpe_joints_mse_fused": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _f, _f, _f, _vp, _vp, _vp, _f, _vp, _vp]),
    "hpe_static_loss_max_maps": (_i, []),
    "hpe_joints_mse_fused_noalloc": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _f, _f, _f, _vp, _vp, _f, _vp, _vp]),
    "hpe_joints_mse_dense": (_i, [_vp, _i, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _f, _vp]),
    "hpe_generate_target": (_i, [_vp, _i, _i, _i, _i, _f, _f, _f, _vp, _vp, _vp]),
    "hpe_conv_plan_create": (_i, [C.POINTER(ConvDesc), C.POINTER(_vp)]),
    "hpe_conv_plan_destroy": (None, [_vp]),
    "hpe_conv_packed_weight_elems": (_sz, [C.POINTER(ConvDesc)]),
    "hpe_conv_pack_weight": (_i, [C.POINTER(ConvDesc), _vp, _vp, _vp]),
    "hpe_conv_run": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "hpe_conv_pack_weight2": (_i, [C.POINTER(ConvDesc), _vp, _vp, _vp, _vp]),
    "hpe_conv_run2": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "hpe_conv_run_head2": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "hpe_conv_run_chain": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "hpe_conv_set_schedule": (_i, [_vp, _i, _i, _i, _i]),
    "hpe_conv_run_ln": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _f, _vp, _vp]),
    "hpe_pack_job_bytes": (_sz, []),
    "hpe_conv_pack_job_fill": (_i, [C.POINTER(ConvDesc), _i, _vp, _vp, _i, _vp, C.POINTER(_i)]),
    "hpe_conv_pack_jobs_run": (_i, [_vp, _i, _i, _vp]),
    "hpe_conv_out_hw": (_i, [C.POINTER(ConvDesc), C.POINTER(_i), C.POINTER(_i)]),
    "hpe_preprocess_out_elems": (_sz, [_i, _i, _i]),
    "hpe_preprocess": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "hpe_preprocess_u8": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "hpe_maxpool3x3s2": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
    "hpe_conv_dgrad_desc": (_i, [C.POINTER(ConvDesc), C.POINTER(ConvDesc)]),
    "hpe_conv_pack_weight_dgrad": (_i, [C.POINTER(ConvDesc), _vp, _vp, _vp]),
    "hpe_bn_stats": (_i, [_vp, C.c_longlong, _i, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "hpe_bn_apply": (_i, [_vp, C.c_longlong, _i, _vp, _vp, _vp, _i, _vp, _vp]),
    "hpe_bn_scratch_elems": (_sz, [_i]),
    "hpe_bn_backward": (_i, [_vp, _vp, _vp, C.c_longlong, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "hpe_bn_apply_bits": (_i, [_vp, C.c_longlong, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp]),
    "hpe_bn_backward_bits": (_i, [_vp, _vp, _vp, C.c_longlong, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "hpe_bn_backward_zmask": (_i, [_vp, _vp, _vp, C.c_longlong, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "hpe_maxpool3x3s2_train": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _vp]),
    "hpe_maxpool3x3s2_backward": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "hpe_heatmap_grad_nhwc": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _vp, _vp]),
    "hpe_sgd_step": (_i, [_vp, _vp, _vp, C.c_longlong, _f, _f, _f, _f, _vp]),
    "hpe_grad_sumsq_scratch_elems": (_sz, []),
    "hpe_grad_sumsq": (_i, [_vp, C.c_longlong, _vp, _vp, _i, _vp]),
    "hpe_adamw_step": (_i, [_vp, _vp, _vp, _vp, C.c_longlong, _f, _f, _f, _f, _f, _i, _f, _vp, _f, _vp]),
    "hpe_wgrad_plan_create": (_i, [C.POINTER(ConvDesc), C.POINTER(_vp)]),
    "hpe_wgrad_plan_destroy": (None, [_vp]),
    "hpe_wgrad_workspace_elems": (_sz, [_vp]),
    "hpe_wgrad_run": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "hpe_dwconv": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _f, _vp, _i, _vp]),
    "hpe_dwconv7_bias": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _i, _vp]),
    "hpe_dwconv7_stats_elems": (_sz, [_i, _i, _i, _i]),
    "hpe_dwconv7_bias_stats": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _i, _vp, _vp]),
    "hpe_layernorm_nhwc": (_i, [_vp, _i, C.c_longlong, _i, _vp, _vp, _f, _vp, _i, _vp]),
    "hpe_bnact_scratch_elems": (_sz, [_i]),
    "hpe_bnact_stats": (_i, [_vp, C.c_longlong, _i, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "hpe_bnact_apply": (_i, [_vp, C.c_longlong, _i, _vp, _vp, _i, _vp, _vp]),
    "hpe_bnact_backward": (_i, [_vp, _vp, C.c_longlong, _i, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "hpe_dwconv3_wgrad_scratch_elems": (_sz, [_i]),
    "hpe_dwconv3_wgrad": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp]),
    "hpe_maxpool5_train": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _vp]),
    "hpe_maxpool5_backward": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "hpe_upsample_bilinear_backward": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "hpe_patchify4": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "hpe_patchify4_u8": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "hpe_maxpool5": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _vp]),
    "hpe_upsample_bilinear": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _vp, _i, _vp]),
    "hpe_pose_decode": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "hpe_pose_nms_scratch_elems": (_sz, [_i, _i]),
    "hpe_pose_nms": (_i, [_vp, _i, _i, _i, _i, _i, _f, C.c_double, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "hpe_compute_loss_workspace_bytes": (_sz, [_i, _i, _vp, _vp, _i, _i]),
    "hpe_compute_loss": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _i, _vp, _vp, _sz, _vp, _vp, _f, _vp]),
    "hpe_oks": (_i, [_vp, _vp, _vp, _vp, _i, _i, _f, _f, _vp, _vp]),
    "hpe_pck": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _f, _f, _vp, _vp, _vp]),
    "hpe_keypoint_loss": (_i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp]),
}

_lib = None


def lib() -> C.CDLL:
    """Load (once) and return the shared library; raises loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python 2d-human-pose-estimation_b200/build.py` "
                "(or __graft_entry__.build()).  There is no CPU / PyTorch fallback for this path.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)          # AttributeError if a declared symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int, what: str = "libhpe_b200") -> None:
    if rc != 0:
        msg = lib().hpe_last_error()
        raise RuntimeError(f"{what} failed (code {rc}): {msg.decode() if msg else '?'}")


def dtype_code(t) -> int:
    import torch
    if t.dtype == torch.float32:
        return HPE_F32
    if t.dtype == torch.bfloat16:
        return HPE_BF16
    raise TypeError(f"unsupported dtype {t.dtype}: the sm_100a kernels take float32 or bfloat16")


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


def stream_ptr() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


def require_cuda(*tensors) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("libhpe_b200 has no CPU path: tensors must live on a CUDA (sm_100a) device")
        if t is not None and not t.is_contiguous():
            raise RuntimeError("libhpe_b200 needs contiguous tensors")
