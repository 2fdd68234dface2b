"""csrc/augment.cu (uint8 patches -> Normalize -> ColorJitter -> OneOf(flip / rot90) -> stem input, drawn on the device)
against the executable restatement oracle/augment_oracle.py: the random draws bit-for-bit (same counter-based hash), the
colour arithmetic to bf16 rounding, the geometry (image and label mask) exactly; plus draw statistics on the device and
the graph-replay property (a fresh draw per step from the device-side step counter)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _unpack(z, H, W):
    """space-to-depth stem input (N, H/2, W/2, 16) -> float32 (N, 3, H, W)."""
    n = z.shape[0]
    v = z.float()[..., :12].reshape(n, H // 2, W // 2, 2, 2, 3)           # (n, i, j, p, q, c)
    return v.permute(0, 5, 1, 3, 2, 4).reshape(n, 3, H, W)


def test_draws_match_the_oracle_bit_for_bit():
    from oracle import augment_oracle as A
    from pdac_pathological_image_segmentation_b200 import ops
    img = torch.randint(0, 256, (64, 32, 32, 3), dtype=torch.uint8, device=DEV)
    step = torch.tensor(12345, dtype=torch.int64, device=DEV)
    cfg = ops.Augment(seed=987654321)
    _, _, params = ops.stem_prep_u8_augment(img, None, cfg, step=step)
    want = A.draw_params(64, seed=987654321, step=12345)
    got = params.cpu().numpy()
    for k in (A.AP_APPLY, A.AP_ORDER, A.AP_GEOM):
        assert np.array_equal(got[:, k], want[:, k]), k
    for k in (A.AP_BRIGHT, A.AP_CONTRAST, A.AP_SAT, A.AP_HUE):
        assert np.abs(got[:, k] - want[:, k]).max() <= 2e-7, k             # one fused multiply-add of difference at most
    # no step pointer == step 0
    _, _, p0 = ops.stem_prep_u8_augment(img, None, cfg)
    assert np.array_equal(p0.cpu().numpy()[:, A.AP_ORDER], A.draw_params(64, seed=987654321, step=0)[:, A.AP_ORDER])


@pytest.mark.parametrize("size", [64, 96])
def test_augmented_patches_match_the_oracle(size):
    from oracle import augment_oracle as A
    from pdac_pathological_image_segmentation_b200 import ops
    g = torch.Generator().manual_seed(3)
    n = 96
    img = torch.randint(0, 256, (n, size, size, 3), dtype=torch.uint8, generator=g)
    mask = torch.randint(0, 2, (n, size, size), dtype=torch.uint8, generator=g)
    cfg = ops.Augment(seed=42)
    z, mo, params = ops.stem_prep_u8_augment(img.to(DEV), mask.to(DEV), cfg)
    p = params.cpu().numpy()
    assert {int(c) for c in p[:, A.AP_GEOM]} >= {0, 1, 2} and 0 < p[:, A.AP_APPLY].sum() < n   # the batch exercises both
    want_x, want_m = A.apply(img, mask, p)
    assert torch.equal(mo.cpu(), want_m)                                  # label geometry: exact
    got = _unpack(z, size, size).cpu()
    err = (got - want_x).abs()
    tol = 2.0 ** -8 * want_x.abs().clamp_min(1.0) + 2e-3                   # bf16 storage + fp32 hue / blend rounding
    bad = (err > tol).float().mean().item()
    assert bad <= 1e-4, f"{bad:.2e} of the values off by more than bf16 rounding (max err {err.max().item():.4f})"
    # the contrast step's per-patch gray mean
    for i in range(n):
        if p[i, A.AP_APPLY]:
            sc = torch.tensor([1 / (255 * s) for s in ops.IMAGENET_STD])
            sh = torch.tensor([-m / s for m, s in zip(ops.IMAGENET_MEAN, ops.IMAGENET_STD)])
            m = A.contrast_mean(img[i].float() * sc + sh, p[i]).item()
            assert abs(p[i, A.AP_MEAN] - m) <= 1e-5 * max(1.0, abs(m)), (i, p[i, A.AP_MEAN], m)
    # un-augmented configuration == the plain preparation kernel
    z0, m0, _ = ops.stem_prep_u8_augment(img.to(DEV), mask.to(DEV), ops.Augment(p_color=0.0, p_geom=0.0))
    z1, m1 = ops.stem_prep_u8(img.to(DEV), mask.to(DEV), None)
    assert torch.equal(z0, z1) and torch.equal(m0, m1)


def test_device_draw_statistics():
    from oracle import augment_oracle as A
    from pdac_pathological_image_segmentation_b200 import ops
    img = torch.zeros((8192, 2, 2, 3), dtype=torch.uint8, device=DEV)
    _, _, params = ops.stem_prep_u8_augment(img, None, ops.Augment(seed=5))
    p = params.cpu().numpy()
    assert abs(p[:, A.AP_APPLY].mean() - 0.5) < 0.02
    g = p[:, A.AP_GEOM].astype(int)
    for code, prob in {0: 0.725, 1: 0.1, 2: 0.1, 3: 0.025, 4: 0.025, 5: 0.025}.items():
        assert abs((g == code).mean() - prob) < 0.012, (code, (g == code).mean())
    for k in (A.AP_BRIGHT, A.AP_CONTRAST, A.AP_SAT):
        assert 0.7 - 1e-6 <= p[:, k].min() and p[:, k].max() <= 1.3 + 1e-6 and abs(p[:, k].mean() - 1) < 0.01


def test_trainstep_augments_per_step_under_graph_replay():
    """TrainStep(augment=True, graph=True): the captured graph draws new parameters every replay (device step counter),
    equal to the oracle's draws for that step; eval-mode validation is never augmented."""
    from oracle import augment_oracle as A
    from pdac_pathological_image_segmentation_b200.smp import Unet
    from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
    from pdac_pathological_image_segmentation_b200.trainer import TrainStep
    torch.manual_seed(1)
    ts = TrainStep(Unet(encoder_weights=None).to(DEV).train(), lr=1e-4, graph=True, augment=True)
    img, mask = (t.to(DEV) for t in synthetic_batch(4, 64, seed=2))
    seen = []
    for it in range(5):
        loss, _ = ts.step(img, mask)
        p = ts.last_aug_params.cpu().numpy()
        want = A.draw_params(4, seed=ts.augment.seed, step=it)
        assert np.array_equal(p[:, [A.AP_APPLY, A.AP_ORDER, A.AP_GEOM]], want[:, [A.AP_APPLY, A.AP_ORDER, A.AP_GEOM]]), it
        seen.append(p[:, :7].copy())
        assert torch.isfinite(loss)
    assert ts._g is not None
    assert any(not np.array_equal(seen[0], s) for s in seen[1:])
    assert 0.0 <= ts.validate(img, mask).item() <= 1.0
