"""Module-level parity of the drop-in ResUNet against (a) the golden vectors the reference's own ResUNet produced
(tests/golden/, made by oracle/make_golden.py) and (b) the fp32 oracle restatement run on the same seeded inputs.

Tolerances. Logits: max-abs <= 2e-2 against the reference's fp32 logits (BASELINE.json north_star). Masks: Dice
agreement >= 0.999 outside a 5e-3 band around the threshold (see _mask_agreement). Loss within 2e-3.
Gradients: storing activations in bf16 perturbs the forward by ~1 bf16 ulp per layer, which flips ReLU masks of
near-zero pre-activations; on this seeded random-init network (logits ~ 0, heavy cancellation in the Dice gradient)
that alone moves deep-layer gradients by 10-35 % relative L2 from the fp32 ones (measured on the oracle itself:
fp32 arithmetic with bf16-rounded storage vs pure fp32). The test is therefore self-calibrating: per parameter, the
product may be no farther from the fp32 oracle than 1.5 x the bf16-storage-emulating oracle is, plus 2e-2. Kernel-level
correctness of every backward kernel is pinned separately and tightly (1e-3..2^-7) in test_conv_gpu.py /
test_elementwise_gpu.py."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _gold(name):
    return torch.load(os.path.join(GOLD, name), weights_only=False)


def _build(nc, size, train):
    import warnings
    from oracle import resunet_oracle as R
    from pdac_pathological_image_segmentation_b200.resunet import ResUNet
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m = ResUNet(num_classes=nc, output_size=size)
    m.load_state_dict(R.make_state(nc, seed=41), strict=True)
    m.to(DEV)
    m.train(train)
    return m


def _mask_agreement(a, b, band=0.0):
    """Dice between the thresholded masks of logits a (ours) and b (reference), over the pixels whose reference logit
    is farther than `band` from the decision threshold. With the seeded random-init weights the logits sit around 0
    (|logit| < 0.7, thousands of pixels within 1e-3 of the threshold), so a 1e-3 logit error legitimately flips ~0.3 %
    of the pixels; the >= 0.999 agreement north_star asks for is asserted outside a band of 5e-3 (a quarter of the
    logit tolerance), and the raw number is bounded too."""
    keep = b.abs() > band
    pa, pb = ((a.sigmoid() >= 0.5) & keep).float(), ((b.sigmoid() >= 0.5) & keep).float()
    return ((2 * (pa * pb).sum() + 1e-6) / (pa.sum() + pb.sum() + 1e-6)).item()


def _oracle_grads(nc, x, target, training, emulate):
    from oracle import resunet_oracle as R
    from oracle import util_oracle as U
    sd = R.grad_state(R.make_state(nc, seed=41))
    o = R.forward(sd, x, x.shape[-1], training=training, emulate_bf16=emulate)
    U.dice_loss(torch.squeeze(o), target).backward()
    return {k: v.grad.double() for k, v in sd.items() if v.requires_grad and v.grad is not None}


def _check_grads(named, g_fp32, g_emu):
    bad = []
    for k, p in named.items():
        if k.startswith("encoder.fc"):
            continue
        assert p.grad is not None, k
        a = g_fp32[k]
        rel = ((p.grad.cpu().double() - a).norm() / (a.norm() + 1e-30)).item()
        cal = ((g_emu[k] - a).norm() / (a.norm() + 1e-30)).item()
        if not rel <= 1.5 * cal + 2e-2:
            bad.append((k, round(rel, 4), round(cal, 4)))
    assert not bad, f"gradients farther from fp32 than bf16 storage explains (param, ours, calibrated): {bad}"


def _check_masks(out, ref):
    assert _mask_agreement(out, ref) >= 0.99
    assert _mask_agreement(out, ref, band=5e-3) >= 0.999


@pytest.mark.parametrize("nc", [1, 2])
def test_resunet_train_step_matches_reference_golden(nc):
    from oracle import resunet_oracle as R
    from oracle import util_oracle as U
    from pdac_pathological_image_segmentation_b200.util import DiceLoss, dice_score, sigmoid_mask
    fix = _gold(f"resunet_nc{nc}.pt")
    img, mask = R.synthetic_batch(2, 64, seed=41)
    x = R.normalize(img).to(DEV)
    m = _build(nc, 64, train=True)
    out = m(x)
    assert out.shape == fix["logits_train"].shape and out.dtype == torch.float32
    ref = fix["logits_train"].to(DEV)
    err = (out - ref).abs().max().item()
    assert err <= 2e-2, f"logits max-abs err {err}"
    _check_masks(out, ref)
    sq = torch.squeeze(out)
    mask_d = mask.to(DEV)
    target = mask_d.float() if nc == 1 else torch.stack([1 - mask_d, mask_d], 1).float()
    pred = sigmoid_mask(sq)
    assert torch.equal(pred, torch.ge(sq.sigmoid(), 0.5).float())          # bit-exact mask on identical logits
    score = dice_score(pred, mask_d if nc == 1 else target)
    loss = DiceLoss()(sq, target)
    assert abs(loss.item() - fix["loss"].item()) <= 2e-3
    assert abs(score.item() - U.dice_score(pred.cpu(), (mask if nc == 1 else target.cpu())).item()) <= 1e-6
    loss.backward()
    # gradient oracles on CPU (the restatement test_oracle_cpu pins to the reference): pure fp32, and fp32 arithmetic
    # with bf16-rounded storage
    named = dict(m.named_parameters())
    assert named["encoder.fc.weight"].grad is None and named["encoder.fc.bias"].grad is None
    _check_grads(named, _oracle_grads(nc, x.cpu(), target.cpu(), True, False), _oracle_grads(nc, x.cpu(), target.cpu(), True, True))
    r = fix["grad_conv4_weight"].double()            # the reference's own gradient where bf16 noise is still small
    assert ((named["conv4.weight"].grad.cpu().double() - r).norm() / r.norm()).item() <= 1e-2
    sdm = m.state_dict()
    assert torch.allclose(sdm["encoder.bn1.running_mean"].cpu(), fix["bn1_running_mean"], rtol=2e-2, atol=2e-3)
    assert torch.allclose(sdm["encoder.bn1.running_var"].cpu(), fix["bn1_running_var"], rtol=2e-2, atol=2e-3)
    assert torch.allclose(sdm["encoder.layer4.1.bn2.running_var"].cpu(), fix["l4_bn2_running_var"], rtol=5e-2, atol=5e-3)
    assert int(sdm["encoder.bn1.num_batches_tracked"]) == fix["num_batches_tracked"]


@pytest.mark.parametrize("nc", [1, 2])
def test_resunet_eval_matches_reference_golden(nc):
    from oracle import resunet_oracle as R
    fix = _gold(f"resunet_nc{nc}.pt")
    img, _ = R.synthetic_batch(2, 64, seed=41)
    m = _build(nc, 64, train=False)
    with torch.no_grad():
        out = m(R.normalize(img).to(DEV))
    ref = fix["logits_eval"].to(DEV)
    assert (out - ref).abs().max().item() <= 2e-2
    _check_masks(out, ref)
    # eval-mode forward WITH grad (reference quirk: epochs >= 2 train in eval mode)
    out2 = m(R.normalize(img).to(DEV))
    out2.sum().backward()
    # the no-grad forward folds BatchNorm into the conv epilogue (x is never rounded to bf16), the grad-enabled one
    # keeps x for backward: same arithmetic up to that one rounding per layer
    assert (out2 - out).abs().max().item() <= 1e-2
    assert (out2 - ref).abs().max().item() <= 2e-2
    assert m.conv4.weight.grad is not None and torch.isfinite(m.encoder.conv1.weight.grad).all()


def test_resunet_eval_backward_matches_oracle():
    from oracle import resunet_oracle as R
    from oracle import util_oracle as U
    img, mask = R.synthetic_batch(2, 64, seed=41)
    x = R.normalize(img)
    m = _build(1, 64, train=False)
    out = m(x.to(DEV))
    from pdac_pathological_image_segmentation_b200.util import DiceLoss
    DiceLoss()(torch.squeeze(out), mask.to(DEV).float()).backward()
    named = dict(m.named_parameters())
    t = mask.float()
    _check_grads(named, _oracle_grads(1, x, t, False, False), _oracle_grads(1, x, t, False, True))


def test_checkpoint_roundtrip_and_module_prefix(tmp_path):
    """train_worker.py:113-125,186-201: save {'model': state_dict with 'module.' prefix}, strip and reload strictly."""
    m = _build(1, 64, train=False)
    sd = {"module." + k: v for k, v in m.state_dict().items()}
    path = tmp_path / "latest.pth"
    torch.save({"model": sd, "epoch": 3, "previous_best": 0.5}, path)
    ck = torch.load(path, map_location=DEV)
    stripped = {k.replace("module.", ""): v for k, v in ck["model"].items()}
    m2 = _build(1, 64, train=False)
    with torch.no_grad():
        for p in m2.parameters():
            p.zero_()
    m2.load_state_dict(stripped, strict=True)
    x = torch.randn(1, 3, 64, 64, device=DEV)
    with torch.no_grad():
        assert torch.equal(m(x), m2(x))
    opt = torch.optim.Adam(m.parameters(), lr=1e-4)
    assert len(opt.state_dict()["param_groups"][0]["params"]) == 76      # 76 tensors incl. the two unused fc ones


def test_no_cpu_fallback():
    m = _build(1, 64, train=False)
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.randn(1, 3, 64, 64))
    with pytest.raises(RuntimeError, match="divisible"):
        m(torch.randn(1, 3, 48, 48, device=DEV))
