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
