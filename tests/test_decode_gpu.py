"""GPU parity: hpe_decode_heatmaps (through the C ABI) vs the App.-A oracle on the same tensors.
Bar: argmax indices bit-exact, maxvals bit-exact, keypoints within 0.05 px."""
import numpy as np
import pytest
import torch

from oracle import heatmap_ref as hr

pytestmark = pytest.mark.gpu
KP_TOL = 0.05   # px, north_star


def run_both(hpe, dev, hm_np, dtype, B):
    center, scale = hr.synth_center_scale(B, seed=4)
    hm_t = torch.from_numpy(hm_np).to(dev).to(dtype).contiguous()
    ref_in = hm_t.float().cpu().numpy()            # oracle sees exactly the values the kernel sees
    M = hr.affine_from_center_scale(center, scale, hm_np.shape[2:])
    preds_o, maxv_o, idx_o = hr.get_final_preds_affine(ref_in, M)
    preds, maxv, idx = hpe.get_final_preds_affine(hm_t, torch.from_numpy(M).to(dev), return_idx=True)
    torch.cuda.synchronize()
    return (preds.cpu().numpy(), maxv.cpu().numpy(), idx.cpu().numpy()), (preds_o, maxv_o, idx_o)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,K,H,W", [(64, 17, 64, 48), (16, 17, 96, 72), (8, 133, 64, 48), (4, 133, 128, 96),
                                     (3, 5, 7, 9), (1, 1, 1, 1), (2, 3, 5, 6)])
def test_decode_matches_oracle(hpe, cuda_dev, B, K, H, W, dtype):
    hm = hr.synth_heatmaps(B, K, H, W, seed=3, adversarial=min(H, W) >= 5)
    (p, m, i), (po, mo, io) = run_both(hpe, cuda_dev, hm, dtype, B)
    np.testing.assert_array_equal(i, io)
    np.testing.assert_array_equal(m, mo)
    assert np.abs(p - po).max() <= KP_TOL


def test_decode_get_max_preds_and_edge_cases(hpe, cuda_dev):
    H, W = 64, 48
    hm = np.zeros((2, 6, H, W), dtype=np.float32)
    hm[0, 0, 10, 20] = 1.0; hm[0, 0, 40, 5] = 1.0            # exact tie: first wins
    hm[0, 1] = -0.5                                           # all negative
    hm[0, 2, 20:23, 30:33] = 0.7                              # plateau
    hm[0, 3, H - 1, W - 1] = 0.3                              # last element
    hm[0, 4, 0, 0] = 0.3                                      # first element
    hm[1, 0, 5, 1] = 1.0; hm[1, 0, 5, 2] = 0.9                # column 1 -> no refinement
    hm[1, 1, 5, 2] = 1.0; hm[1, 1, 5, 3] = 0.9; hm[1, 1, 4, 2] = 0.1   # column 2 -> refined
    t = torch.from_numpy(hm).to(cuda_dev)
    c, m, i = hpe.get_max_preds(t)
    co, mo, io = hr.get_max_preds(hm)
    np.testing.assert_array_equal(i.cpu().numpy(), io)
    np.testing.assert_array_equal(c.cpu().numpy(), co)
    np.testing.assert_array_equal(m.cpu().numpy(), mo)
    ident = torch.tensor([[[1., 0, 0], [0, 1., 0]]], device=cuda_dev).repeat(2, 1, 1).contiguous()
    p, _ = hpe.get_final_preds_affine(t, ident)
    po = hr.quarter_pixel_refine(hm, co)
    np.testing.assert_array_equal(p.cpu().numpy(), po)
    assert p[1, 0].tolist() == [1.0, 5.0] and p[1, 1].tolist() == [2.25, 4.75]


def test_decode_full_size_properties(hpe, cuda_dev):
    """BASELINE config sizes (B=1024, K=17, 64x48): size-independent properties instead of the slow oracle:
    (1) planted unique maxima are recovered exactly, (2) decode is invariant to a positive per-map scale,
    (3) bf16 and fp32 agree on planted peaks."""
    B, K, H, W = 1024, 17, 64, 48
    g = torch.Generator(device="cpu").manual_seed(11)
    hm = torch.rand(B, K, H * W, generator=g) * 0.5
    where = torch.randint(0, H * W, (B, K), generator=g)
    hm.scatter_(2, where[..., None], 2.0)
    hm = hm.reshape(B, K, H, W).to(cuda_dev)
    c, m, i = hpe.get_max_preds(hm)
    assert torch.equal(i.cpu().long(), where)
    assert torch.all(m == 2.0)
    c2, _, i2 = hpe.get_max_preds(hm * 4.0)
    assert torch.equal(i, i2) and torch.equal(c, c2)
    c3, _, i3 = hpe.get_max_preds(hm.to(torch.bfloat16))
    assert torch.equal(i, i3)
    assert torch.equal(c[..., 0].cpu().long(), where % W) and torch.equal(c[..., 1].cpu().long(), where // W)


def test_decode_host_buffer_entry_point(hpe, cuda_dev):
    B, K, H, W = 32, 17, 64, 48
    hm = hr.synth_heatmaps(B, K, H, W, seed=5)
    center, scale = hr.synth_center_scale(B)
    M = hr.affine_from_center_scale(center, scale, (H, W))
    dec = hpe.HostDecoder(B, K, H, W, torch.float32, cuda_dev)
    c, m, i = dec(torch.from_numpy(hm).pin_memory(), torch.from_numpy(M).pin_memory())
    po, mo, io = hr.get_final_preds_affine(hm, M)
    np.testing.assert_array_equal(i.numpy(), io)
    np.testing.assert_array_equal(m.numpy(), mo)
    assert np.abs(c.numpy() - po).max() <= KP_TOL


def test_decode_bad_args_fail_loudly(hpe, cuda_dev):
    with pytest.raises(TypeError):
        hpe.get_max_preds(torch.zeros(1, 1, 4, 4, dtype=torch.float16, device=cuda_dev))
    with pytest.raises(ValueError):
        hpe.get_max_preds(torch.zeros(4, 4, device=cuda_dev))
    c, m, i = hpe.get_max_preds(torch.zeros(0, 17, 64, 48, device=cuda_dev))     # empty batch
    assert c.shape == (0, 17, 2) and m.shape == (0, 17, 1)


def test_packed_xyv_output_equals_separate_outputs(hpe, cuda_dev):
    """hpe_decode_heatmaps_xyv (the layout the result gather ships) = the same kernel writing (x, y, maxval) rows."""
    hm = torch.from_numpy(hr.synth_heatmaps(9, 17, 64, 48, seed=8)).to(cuda_dev)
    aff = torch.rand(9, 2, 3, device=cuda_dev)
    p, v = hpe.get_final_preds_affine(hm, aff)
    xyv = hpe.get_final_preds_xyv(hm, aff)
    assert xyv.shape == (9, 17, 3)
    assert torch.equal(xyv[..., :2], p) and torch.equal(xyv[..., 2:], v)
    out = torch.empty(9, 17, 3, device=cuda_dev)
    assert hpe.get_final_preds_xyv(hm.to(torch.bfloat16), aff, out) is out
    p16, v16 = hpe.get_final_preds_affine(hm.to(torch.bfloat16), aff)
    assert torch.equal(out[..., :2], p16) and torch.equal(out[..., 2:], v16)


def _sweep_corner_heatmaps(B, K, H, W, dev, seed):
    """Full-size synthetic heatmaps generated ON the device (the NumPy generator would take minutes at these sizes): a third of
    the maps are quantised to 1/32 steps (thousands of exact ties per map -> the first-max rule decides), a third smooth noise
    with a planted peak, the rest noise; 1 % of the maps all-negative, 1 % all-zero (maxval <= 0 -> zero coordinates)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    hm = torch.rand(B, K, H, W, device=dev, generator=g)
    kind = torch.randint(0, 100, (B, K), device=dev, generator=g)
    q = (kind < 33)[..., None, None]
    hm = torch.where(q, torch.round(hm * 32) / 32, hm)
    peak = torch.randint(0, H * W, (B, K), device=dev, generator=g)
    planted = (kind >= 33) & (kind < 66)
    flat = hm.view(B, K, H * W)
    cur = flat.gather(2, peak[..., None])
    flat.scatter_(2, peak[..., None], torch.where(planted[..., None], cur + 1.5, cur))
    hm = torch.where((kind == 98)[..., None, None], -hm - 0.1, hm)
    hm = torch.where((kind == 99)[..., None, None], torch.zeros_like(hm), hm)
    return hm.contiguous()


@pytest.mark.parametrize("B,K,H,W,dtype,chunk", [(1024, 133, 128, 96, torch.float32, 128), (1024, 133, 128, 96, torch.bfloat16, 128),
                                                  (65536, 17, 64, 48, torch.float32, 8192)])
def test_decode_sweep_corners_against_the_oracle(hpe, cuda_dev, B, K, H, W, dtype, chunk):
    """BASELINE configs[4] corners compared with oracle/heatmap_ref.py itself (not with properties): ONE kernel launch over the
    whole tensor (6.7 GB fp32 / 3.3 GB bf16 at K=133 128x96 B=1024; 13.7 GB at K=17 64x48 B=65536), the oracle over host chunks of
    the same values.  Bar: argmax indices and maxvals bit-exact, keypoints within 0.05 px."""
    hm = _sweep_corner_heatmaps(B, K, H, W, cuda_dev, seed=21).to(dtype)
    g = torch.Generator().manual_seed(22)
    M = torch.rand(B, 2, 3, generator=g) * 4 - 2
    preds, maxv, idx = hpe.get_final_preds_affine(hm, M.to(cuda_dev), return_idx=True)
    torch.cuda.synchronize()
    preds, maxv, idx = preds.cpu().numpy(), maxv.cpu().numpy(), idx.cpu().numpy()
    n_ties = 0
    for b0 in range(0, B, chunk):
        sl = slice(b0, b0 + chunk)
        ref_in = hm[sl].float().cpu().numpy()              # the exact values the kernel saw (bf16 upcast is value-preserving)
        po, mo, io = hr.get_final_preds_affine(ref_in, M[sl].numpy())
        np.testing.assert_array_equal(idx[sl], io)
        np.testing.assert_array_equal(maxv[sl], mo)
        assert np.abs(preds[sl] - po).max() <= KP_TOL
        if b0 == 0:
            flat = ref_in.reshape(chunk * K, -1)
            n_ties = int(((flat == flat.max(1, keepdims=True)).sum(1) > 1).sum())
    assert n_ties > chunk * K // 10, "the corner case must contain tied maxima"
    assert (maxv <= 0).any() and (idx > 0).any()
