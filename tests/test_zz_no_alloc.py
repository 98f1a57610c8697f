"""GPU test of the allocation-free loss entry point `hpe_joints_mse_fused_noalloc` (include/hpe_b200.h): its per-map partial sums
live in a device array built into the library instead of caller-provided scratch.  Called through the C ABI, compared with the
existing `hpe_joints_mse_fused` on the same inputs, two calls in a row (the second smaller, so stale contents of the shared array
would show).  Bar: loss within 1e-4 relative (the bar of tests/test_loss_gpu.py); gradients and weights elementwise equal -- the
two entry points launch the same kernels, only the scratch pointer differs.

Named test_zz_* so that pytest collects it last: a failure here cannot hide another test under `pytest -x`."""
import pytest
import torch

pytestmark = pytest.mark.gpu
LOSS_RTOL = 1e-4


def _both(hpe, dev, B, K, H, W, dtype, seed, with_grad=True):
    lib, _lib = hpe._lib.lib(), hpe._lib
    g = torch.Generator().manual_seed(seed)
    pred = torch.rand(B, K, H, W, generator=g).to(dev).to(dtype).contiguous()
    joints = torch.rand(B, K, 3, generator=g)
    joints[..., 0] = joints[..., 0] * (W * 4 + 40) - 20          # some joints outside the map
    joints[..., 1] = joints[..., 1] * (H * 4 + 40) - 20
    joints[..., 2] = (joints[..., 2] > 0.2).float()              # some invisible joints (weight 0 maps)
    joints = joints.to(dev).contiguous()
    code = _lib.dtype_code(pred)
    st = torch.cuda.current_stream().cuda_stream
    # every buffer either call needs is allocated BEFORE the calls; the new entry point gets inputs and outputs only
    partials = torch.full((B * K,), float("nan"), device=dev)
    loss_a, loss_b = torch.zeros(1, device=dev), torch.zeros(1, device=dev)
    grad_a, grad_b = torch.empty_like(pred), torch.empty_like(pred)
    w_a, w_b = torch.empty(B, K, device=dev), torch.empty(B, K, device=dev)
    ga, gb = (_lib.ptr(grad_a), _lib.ptr(grad_b)) if with_grad else (None, None)
    torch.cuda.synchronize()
    mem0 = torch.cuda.memory_allocated(dev)
    rc = lib.hpe_joints_mse_fused(_lib.ptr(pred), code, _lib.ptr(joints), B, K, H, W, 2.0, 4.0, 4.0,
                                  _lib.ptr(partials), _lib.ptr(loss_a), ga, 1.0, _lib.ptr(w_a), st)
    _lib.check(rc, "hpe_joints_mse_fused")
    rc = lib.hpe_joints_mse_fused_noalloc(_lib.ptr(pred), code, _lib.ptr(joints), B, K, H, W, 2.0, 4.0, 4.0,
                                          _lib.ptr(loss_b), gb, 1.0, _lib.ptr(w_b), st)
    _lib.check(rc, "hpe_joints_mse_fused_noalloc")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(dev) == mem0
    la, lb = loss_a.item(), loss_b.item()
    assert la > 0 and abs(la - lb) <= LOSS_RTOL * abs(la), (la, lb)
    assert torch.equal(w_a, w_b)
    if with_grad:
        assert torch.equal(grad_a, grad_b)
    return la


def test_noalloc_loss_matches_existing_entry_point_twice(hpe, cuda_dev):
    # first call: many maps; second call right behind it: fewer maps and another dtype -- entries the first call left in the
    # shared array beyond the second call's B*K must not leak into its loss
    l1 = _both(hpe, cuda_dev, 64, 17, 64, 48, torch.float32, seed=1)
    l2 = _both(hpe, cuda_dev, 3, 17, 64, 48, torch.bfloat16, seed=2)
    l3 = _both(hpe, cuda_dev, 64, 17, 64, 48, torch.float32, seed=1, with_grad=False)
    assert l1 == l3 and l1 != l2


def test_noalloc_loss_rejects_sizes_beyond_the_builtin_scratch(hpe, cuda_dev):
    lib, _lib = hpe._lib.lib(), hpe._lib
    cap = lib.hpe_static_loss_max_maps()
    assert cap == 1 << 17
    t = torch.zeros(8, device=cuda_dev)
    rc = lib.hpe_joints_mse_fused_noalloc(_lib.ptr(t), _lib.HPE_F32, _lib.ptr(t), cap + 1, 1, 1, 1, 2.0, 4.0, 4.0,
                                          _lib.ptr(t), None, 1.0, None, torch.cuda.current_stream().cuda_stream)
    assert rc == -3                                              # HPE_EUNSUPPORTED, nothing launched
    assert b"exceed the built-in scratch" in lib.hpe_last_error()
