"""GPU parity of the whole forward (preprocess -> stem -> maxpool -> 16 bottlenecks -> 3 deconvs -> final 1x1)
against the CPU oracle with identical weights, then decode on the produced heatmaps.

Tolerances.  The synthetic model is normalised so that heatmaps live in [-1,1] like a trained net's
(oracle.resnet_ref.build_synthetic(normalize=True)).  north_star's example figure (max-abs 1e-2 in bf16) is not
reachable by ANY bf16 evaluation of this random-init 53-layer net: the CPU oracle evaluated with bf16 weights and
bf16 stored activations (fp32 accumulate) itself differs from the fp32 oracle by ~1.5e-2 max-abs (printed below,
measured on every run), and two bf16 evaluations with different accumulation order differ from each other by the
same amount.  Stated tolerances on the normalised scale:
  * max-abs <= 3e-2 and RMS <= 5e-3 vs the fp32 oracle, and max-abs no worse than 2x the CPU bf16 emulation's own
    deviation from fp32 (+5e-3) -- i.e. the kernels add no error beyond the bf16 noise floor;
  * the per-layer tests (tests/test_conv_gpu.py) hold every kernel flavour to ONE bf16 rounding of the output;
  * argmax bit-exact on the SAME tensor; keypoints <= 0.05 px on the same tensor; end-to-end argmax only where
    the top-2 margin exceeds the heatmap tolerance."""
import numpy as np
import pytest
import torch

from oracle import heatmap_ref as hr
from oracle import resnet_ref as rr

pytestmark = pytest.mark.gpu
HM_TOL_MAXABS = 3e-2      # normalised heatmaps, vs fp32 and vs bf16-emulating oracle
HM_TOL_RMS = 5e-3


@pytest.fixture(scope="module")
def pair(hpe, cuda_dev):
    from importlib import import_module
    models = import_module("2d-human-pose-estimation_b200.models")
    ref = rr.build_synthetic(17, seed=0, normalize=True)
    m = models.HeatmapPose(17, "resnet50", pretrained=False)
    missing, unexpected = m.load_state_dict(ref.state_dict())
    assert not missing and not unexpected            # state_dict keys are timm/oracle compatible
    return m.to(cuda_dev).eval(), ref


@pytest.mark.parametrize("B,H,W", [(4, 256, 192), (2, 384, 288), (3, 64, 96)])
def test_forward_matches_oracle(pair, cuda_dev, B, H, W):
    m, ref = pair
    x = torch.rand(B, 3, H, W, generator=torch.Generator().manual_seed(0))
    hm = m(x.to(cuda_dev)).float().cpu()
    torch.cuda.synchronize()
    assert tuple(hm.shape) == (B, 17, H // 4, W // 4)
    with torch.no_grad():
        want32 = ref(x)
    want16, feat16 = rr.forward_bf16_emulated(ref, x, return_features=True)
    scale = want32.abs().max().item()
    err32 = (hm - want32).abs().max().item()
    err16 = (hm - want16).abs().max().item()
    feat = m.forward_features(x.to(cuda_dev)).cpu()
    ferr = (feat - feat16).abs().max().item() / feat16.abs().max().item()
    print(f"[{B}x{H}x{W}] heatmap scale {scale:.3f}  err vs fp32 {err32:.2e}  vs bf16-emulated {err16:.2e}  trunk rel {ferr:.2e}")
    emu32 = (want16 - want32).abs().max().item()
    print(f"      (CPU bf16-emulation itself differs from fp32 by {emu32:.2e})")
    rms32 = (hm - want32).pow(2).mean().sqrt().item()
    print(f"      rms vs fp32 {rms32:.2e}")
    assert err16 <= HM_TOL_MAXABS, f"max-abs {err16} vs bf16-emulating oracle exceeds {HM_TOL_MAXABS}"
    assert err32 <= HM_TOL_MAXABS, f"max-abs {err32} vs fp32 oracle exceeds {HM_TOL_MAXABS}"
    assert rms32 <= HM_TOL_RMS
    assert err32 <= 2 * emu32 + 5e-3, "kernel error exceeds the bf16 noise floor of this network"
    assert ferr <= 3e-2


def test_forward_is_deterministic_and_graph_equals_eager(pair, cuda_dev, monkeypatch):
    m, _ = pair
    x = torch.rand(2, 3, 256, 192, generator=torch.Generator().manual_seed(3)).to(cuda_dev)
    a = m(x).clone()
    b = m(x).clone()
    assert torch.equal(a, b)
    monkeypatch.setenv("HPE_B200_CUDA_GRAPH", "0")
    c = m(x).clone()
    assert torch.equal(a, c)


@pytest.mark.parametrize("B", [1, 3, 5, 8, 33])
def test_fused_paths_equal_unfused_layers(pair, cuda_dev, monkeypatch, B):
    """The shipped forward (projection shortcuts accumulated into conv3, cta_group::2 tile pairs incl. odd tile counts, deconv3 +
    final 1x1 in one kernel) against the same library with every fusion and the pair kernels switched off: same bf16
    activations up to accumulation order and the fp32-before-rounding fold of the shortcut's BN scale.  Also run twice: the
    persistent schedules are static, so the result is bit-reproducible."""
    from importlib import import_module
    models = import_module("2d-human-pose-estimation_b200.models")
    m, ref = pair
    x = torch.rand(B, 3, 256, 192, generator=torch.Generator().manual_seed(40 + B)).to(cuda_dev)
    hm1 = m(x).clone()
    hm2 = m(x).clone()
    assert torch.equal(hm1, hm2)
    for k in ("HPE_B200_FUSE_SHORTCUT", "HPE_B200_FUSE_HEAD", "HPE_B200_PAIR", "HPE_B200_FUSE_POOL", "HPE_B200_FUSE_CHAIN"):
        monkeypatch.setenv(k, "0")
    plain = models.HeatmapPose(17, "resnet50", pretrained=False)
    plain.load_state_dict(ref.state_dict())
    plain = plain.to(cuda_dev).eval()
    want = plain(x)
    torch.cuda.synchronize()
    plan = next(iter(plain._plans.values()))
    assert plan.fused_head is None and not plan.fused_shortcut and not plan.fused_pool
    scale = want.abs().max().item()
    assert (hm1 - want).abs().max().item() <= 2e-2 * scale, (hm1 - want).abs().max().item() / scale
    assert ((hm1 - want) ** 2).mean().sqrt().item() <= 3e-3 * scale


@pytest.mark.parametrize("B", [3, 32])
def test_tile_schedule_is_bit_identical(pair, cuda_dev, monkeypatch, B):
    """hpe_conv_set_schedule (reversed tile walk, L2 eviction hints on stores / shortcut loads) changes WHEN a tile is computed and
    how long its lines stay in L2, never its bits: heatmaps with the default schedule == heatmaps with HPE_B200_L2_SCHED=0, and the
    default plan really reverses some launches."""
    from importlib import import_module
    models = import_module("2d-human-pose-estimation_b200.models")
    m, ref = pair
    x = torch.rand(B, 3, 256, 192, generator=torch.Generator().manual_seed(70 + B)).to(cuda_dev)
    monkeypatch.setenv("HPE_B200_L2_SCHED", "7")          # direction + store / shortcut / input hints
    hinted = models.HeatmapPose(17, "resnet50", pretrained=False)
    hinted.load_state_dict(ref.state_dict())
    hinted = hinted.to(cuda_dev).eval()
    hm = hinted(x).clone()
    plan = next(iter(hinted._plans.values()))
    assert all(any(s[i] for s in plan.schedule.values()) for i in range(4))
    assert torch.equal(hm, m(x))                          # the default schedule (direction only)
    assert any(s[0] for s in m._plans[(tuple(x.shape), cuda_dev.index)].schedule.values())
    monkeypatch.setenv("HPE_B200_L2_SCHED", "0")
    plain = models.HeatmapPose(17, "resnet50", pretrained=False)
    plain.load_state_dict(ref.state_dict())
    plain = plain.to(cuda_dev).eval()
    want = plain(x)
    torch.cuda.synchronize()
    assert not next(iter(plain._plans.values())).schedule
    assert torch.equal(hm, want)


@pytest.mark.parametrize("B", [2, 5, 32])
def test_chained_conv1_is_bit_identical(pair, cuda_dev, monkeypatch, B):
    """desc.chain: conv3 (+ shortcut + ReLU) of a layer1 Bottleneck and conv1 + BN + ReLU of the next block in one kernel (the
    second GEMM reads the staged output chunks).  Same bf16 A operand, same k order, same epilogue arithmetic as the separate
    conv1 launch: every activation of the chained convs and the heatmaps are bit-identical with the fusion switched off."""
    from importlib import import_module
    models = import_module("2d-human-pose-estimation_b200.models")
    m, ref = pair
    x = torch.rand(B, 3, 256, 192, generator=torch.Generator().manual_seed(90 + B)).to(cuda_dev)
    hm = m(x).clone()
    plan = m._plans[(tuple(x.shape), cuda_dev.index)]
    assert len(plan.chained) == 5, plan.chained          # layer1.{1,2}.conv1, layer2.{0,2,3}.conv1
    t1 = {k: plan.bufs[k.replace("conv1", "t1")].clone() for k in plan.chained}
    monkeypatch.setenv("HPE_B200_FUSE_CHAIN", "0")
    plain = models.HeatmapPose(17, "resnet50", pretrained=False)
    plain.load_state_dict(ref.state_dict())
    plain = plain.to(cuda_dev).eval()
    want = plain(x)
    torch.cuda.synchronize()
    pplan = next(iter(plain._plans.values()))
    assert not pplan.chained
    for k, got in t1.items():
        assert torch.equal(got, pplan.bufs[k.replace("conv1", "t1")]), k
    assert torch.equal(hm, want)


def test_full_size_batch_equals_small_batches(pair, cuda_dev):
    """BASELINE configs[1] size (B=256 @256x192): a size-independent property instead of the (slow) oracle -- an image's heatmaps do not
    depend on the batch it sits in or on its position in it.  The 256-batch runs different bricks, tile pairs, waves and persistent
    schedules than the 32-batch the oracle tests cover, but every output element is the same sequence of MMAs: bit-identical."""
    m, _ = pair
    x32 = torch.rand(32, 3, 256, 192, generator=torch.Generator().manual_seed(77)).to(cuda_dev)
    hm32 = m(x32).clone()
    hm256 = m(x32.repeat(8, 1, 1, 1))
    torch.cuda.synchronize()
    assert hm256.shape == (256, 17, 64, 48)
    assert torch.isfinite(hm256).all()
    assert torch.equal(hm256.view(8, 32, 17, 64, 48), hm32.unsqueeze(0).expand(8, -1, -1, -1, -1))


def test_forward_then_decode_end_to_end(hpe, pair, cuda_dev):
    m, ref = pair
    B = 4
    x = torch.rand(B, 3, 256, 192, generator=torch.Generator().manual_seed(1))
    center, scale = hr.synth_center_scale(B)
    hm = m(x.to(cuda_dev))
    preds, maxvals = hpe.decode(hm, torch.from_numpy(center), torch.from_numpy(scale))
    _, _, idx = hpe.get_max_preds(hm)
    # (1) same tensor: bit-exact indices, keypoints within 0.05 px
    hm_cpu = hm.float().cpu().numpy()
    po, mo = hr.get_final_preds(hm_cpu, center, scale)
    _, _, io = hr.get_max_preds(hm_cpu)
    np.testing.assert_array_equal(idx.cpu().numpy(), io)
    assert np.abs(preds.cpu().numpy() - po).max() <= 0.05
    np.testing.assert_array_equal(maxvals.cpu().numpy(), mo)
    # (2) end to end vs the fp32 oracle network, only where the top-2 margin exceeds the heatmap tolerance
    with torch.no_grad():
        want = ref(x).numpy()
    flat = np.sort(want.reshape(B, 17, -1), axis=2)
    clear = (flat[..., -1] - flat[..., -2]) > 2 * HM_TOL_MAXABS
    _, _, iw = hr.get_max_preds(want)
    assert np.array_equal(idx.cpu().numpy()[clear], iw[clear])


def test_model_boundary(hpe, cuda_dev):
    from importlib import import_module
    models = import_module("2d-human-pose-estimation_b200.models")
    assert "DeepPose" in models.__all__ and models.DeepPose is models.HeatmapPose
    cfg = hpe.load_config()
    cfg.defrost(); cfg.model.pretrained = False; cfg.freeze()
    m = models.build_model(cfg)                       # Builder.model(): positional (nkpts, backbone, pretrained)
    assert m.njoints == 17
    with pytest.raises(RuntimeError, match="no CPU path"):
        m.eval()(torch.zeros(1, 3, 64, 64))
    # model.train() runs the batch-statistics forward and returns an autograd-connected heatmap (trainer.py:77-97)
    out = m.to(cuda_dev).train()(torch.rand(2, 3, 64, 64, device=cuda_dev))
    assert out.shape == (2, 17, 16, 16) and out.requires_grad
    m = models.build_model(cfg)
    with pytest.raises(RuntimeError, match="pretrained"):
        models.HeatmapPose(17, "resnet50", pretrained=True)


def test_uint8_input_is_float_div_255(hpe, cuda_dev):
    """uint8 images take the fused `imgs.float() / 255.0` path (HPE/train.py:117-120, HPE/test.py:16-29): identical heatmaps."""
    torch.manual_seed(3)
    m = hpe.HeatmapPose(17, "resnet50", pretrained=False)
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm2d):
            mod.weight.data.uniform_(0.5, 1.5); mod.running_var.uniform_(0.5, 1.5)
    m = m.to(cuda_dev).eval()
    u8 = torch.randint(0, 256, (3, 3, 128, 96), dtype=torch.uint8, device=cuda_dev)
    a = m(u8).clone()
    b = m(u8.float() / 255.0)
    assert torch.equal(a, b)
    p = hpe.DINOv3Pose("dinov3_convnext_tiny", nkpts=(17, 3), ncls=1, device=cuda_dev).eval()
    u8 = torch.randint(0, 256, (2, 3, 64, 64), dtype=torch.uint8, device=cuda_dev)
    a = p(u8).clone()
    assert torch.equal(a, p(u8.float() / 255.0))


# ------------------------------------------------------------------------------------------------ teacher-forced parity
def _teacher_forced_errors(m, ref, x, cuda_dev):
    """Per launch of the eval forward: the kernel's output against the fp32 oracle layer evaluated on the kernel's OWN stored
    bf16 inputs (oracle.resnet_ref.teacher_forced_report)."""
    m.forward_static(x.to(cuda_dev))
    torch.cuda.synchronize()
    plan = m._plans[(tuple(x.shape), cuda_dev.index)]
    return rr.teacher_forced_report(ref, x, plan.export_acts(), plan.fused_head, plan.fused_pool)


@pytest.mark.parametrize("weights", ["synthetic", "calibrated"])
def test_every_layer_teacher_forced_within_one_bf16_rounding(hpe, cuda_dev, weights):
    """End-to-end inference parity WITHOUT the chaotic amplification of a random-init 53-layer net: every launch of the shipped
    forward (pooled stem, fused projection shortcuts, CTA-pair kernels, fused deconv3+final head) is compared with the fp32
    oracle layer fed the kernel's own stored inputs.  Bars: each bf16 output within ONE bf16 rounding of the fp32 value
    (2^-8 |ref| + 2e-3 max|ref|), i.e. far inside north_star's max-abs 1e-2 on the layer's own scale; the fp32 heatmaps within
    1e-4 of max (1e-3 through the fused head, whose bf16 mid-tensor is rounded inside the kernel)."""
    from importlib import import_module
    models = import_module("2d-human-pose-estimation_b200.models")
    ref = rr.build_synthetic(17, seed=0, normalize=True) if weights == "synthetic" else rr.build_calibrated(17, seed=0)
    m = models.HeatmapPose(17, "resnet50", pretrained=False)
    m.load_state_dict(ref.state_dict())
    m = m.to(cuda_dev).eval()
    x = torch.rand(3, 3, 256, 192, generator=torch.Generator().manual_seed(11))
    rows = _teacher_forced_errors(m, ref, x, cuda_dev)
    assert len(rows) >= 52
    worst = max(rows, key=lambda r: r[1])
    print(f"[{weights}] {len(rows)} launches teacher-forced; worst max-abs/scale {worst[1]:.2e} at {worst[0]}")
    for name, rel, frac_out in rows:
        assert frac_out == 0.0, f"{name}: {frac_out:.2e} of the outputs are outside one bf16 rounding (max-abs/scale {rel:.2e})"
        assert rel <= 1e-2, f"{name}: max-abs {rel:.2e} of the layer's scale exceeds north_star's 1e-2"


def test_calibrated_weights_end_to_end(hpe, cuda_dev):
    """The well-conditioned weight set (BN statistics calibrated on synthetic input, activations O(1), heatmaps max-normalised):
    the kernels stay inside the bf16 noise floor of the network measured by the CPU bf16 emulation, and the TRUNK features --
    before the three un-normalised deconv stages -- meet 1e-2 of their scale in RMS."""
    from importlib import import_module
    models = import_module("2d-human-pose-estimation_b200.models")
    ref = rr.build_calibrated(17, seed=0)
    m = models.HeatmapPose(17, "resnet50", pretrained=False)
    m.load_state_dict(ref.state_dict())
    m = m.to(cuda_dev).eval()
    x = torch.rand(4, 3, 256, 192, generator=torch.Generator().manual_seed(0))
    hm = m(x.to(cuda_dev)).cpu()
    with torch.no_grad():
        want32 = ref(x)
    want16 = rr.forward_bf16_emulated(ref, x)
    err32, err16, emu = (hm - want32).abs().max().item(), (hm - want16).abs().max().item(), (want16 - want32).abs().max().item()
    print(f"calibrated: heatmap scale {want32.abs().max().item():.3f}; GPU vs fp32 {err32:.2e}, vs bf16 emulation {err16:.2e}; "
          f"emulation vs fp32 {emu:.2e}")
    assert err32 <= 1.5 * emu + 2e-3, "kernel error exceeds the bf16 noise floor of this network"
    assert err16 <= 1.5 * emu + 2e-3


# ------------------------------------------------------------------------------------------------ drop-in contract
def test_eval_after_dropin_training_step_uses_new_weights(hpe, cuda_dev):
    """train(); forward/backward/opt.step(); eval(); model(x) must see the UPDATED parameters and running statistics (the eval
    plan's packed weights / folded BN are refreshed), = a fresh model loaded from the updated state_dict."""
    torch.manual_seed(5)
    m = hpe.HeatmapPose(17, "resnet50", pretrained=False)
    for n, mod in m.named_modules():
        if isinstance(mod, torch.nn.BatchNorm2d):
            mod.weight.data.uniform_(0.5, 1.5)
            if n.endswith("bn3"):
                mod.weight.data.mul_(0.2)
    m = m.to(cuda_dev).eval()
    x = torch.rand(4, 3, 128, 96, device=cuda_dev)
    before = m(x)
    m.train()
    opt = torch.optim.SGD(m.parameters(), lr=0.05)
    out = m(x)
    out.square().mean().backward()
    opt.step()
    nbt = m.backbone.bn1.num_batches_tracked.item()
    assert nbt == 1, "training forward must bump num_batches_tracked like torch BatchNorm"
    m.eval()
    after = m(x)
    assert not torch.equal(after, before), "eval plan kept the previous weights"
    fresh = hpe.HeatmapPose(17, "resnet50", pretrained=False)
    fresh.load_state_dict(m.state_dict())
    fresh = fresh.to(cuda_dev).eval()
    assert torch.equal(fresh(x), after)


def test_forward_returns_a_fresh_tensor_and_to_keeps_plans(hpe, pair, cuda_dev):
    m, _ = pair
    x = torch.rand(2, 3, 128, 96, device=cuda_dev)
    a = m(x)
    b = m(x.flip(3))                                  # flip-test pattern: `a` must survive the second call
    assert a.data_ptr() != b.data_ptr() and not torch.equal(a, b)
    assert torch.equal(a, m(x))
    plans = dict(m._plans)
    m2 = m.to(cuda_dev)                               # the reference trainer does this every batch (trainer.py:87): a no-op move
    assert m2 is m and m._plans == plans and all(m._plans[k] is plans[k] for k in plans)
    assert torch.equal(a, m(x))
    s = m.forward_static(x)
    assert s.data_ptr() == m.forward_static(x.flip(3)).data_ptr()      # the zero-copy path does alias (documented)


@pytest.mark.parametrize("n_batches,dtype", [(1, torch.uint8), (2, torch.float32), (5, torch.uint8)])
def test_stream_ring_matches_synchronous_calls(hpe, pair, cuda_dev, n_batches, dtype):
    """PoseEstimator.stream (three streams, results yielded one batch late) returns, in order, exactly what one synchronous
    infer_host call per batch returns."""
    m, _ = pair
    est = hpe.PoseEstimator(m, cuda_dev)
    g = torch.Generator().manual_seed(9)
    B = 3
    imgs = [torch.randint(0, 256, (B, 3, 128, 96), generator=g, dtype=torch.uint8) for _ in range(n_batches)]
    if dtype == torch.float32:
        imgs = [t.float() / 255 for t in imgs]
    imgs = [t.pin_memory() for t in imgs]
    affs = [torch.rand(B, 2, 3, generator=g).pin_memory() for _ in range(n_batches)]
    want = []
    for im, af in zip(imgs, affs):
        p, v = est.infer_host(im, af)
        want.append((p.clone(), v.clone()))
    got = [(p.clone(), v.clone()) for p, v in est.stream(zip(imgs, affs))]
    assert len(got) == n_batches
    for (p, v), (pw, vw) in zip(got, want):
        assert p.shape == (B, 17, 2) and v.shape == (B, 17, 1)
        assert torch.equal(p, pw) and torch.equal(v, vw)
