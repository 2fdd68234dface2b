"""The augmentation oracle (oracle/augment_oracle.py) against closed-form cases and numpy/torch geometry, on the CPU:
it is the checker of csrc/augment.cu in tests/test_augment_gpu.py."""
import numpy as np
import torch

from oracle import augment_oracle as A


def test_order_decoding_enumerates_all_permutations():
    perms = {tuple(A.decode_order(i)) for i in range(24)}
    assert len(perms) == 24 and all(sorted(p) == [0, 1, 2, 3] for p in perms)
    assert A.decode_order(0) == [0, 1, 2, 3] and A.decode_order(23) == [3, 2, 1, 0]


def test_draw_statistics_and_ranges():
    p = A.draw_params(4000, seed=1234, step=7)
    assert abs(p[:, A.AP_APPLY].mean() - 0.5) < 0.03                      # ColorJitter p = 0.5
    for k in (A.AP_BRIGHT, A.AP_CONTRAST, A.AP_SAT):
        assert p[:, k].min() >= 0.7 - 1e-6 and p[:, k].max() <= 1.3 + 1e-6 and abs(p[:, k].mean() - 1.0) < 0.01
        assert abs(p[:, k].std() - 0.6 / 12 ** 0.5) < 0.01                # uniform on [0.7, 1.3]
    assert p[:, A.AP_HUE].min() >= -0.3 and p[:, A.AP_HUE].max() <= 0.3 and abs(p[:, A.AP_HUE].mean()) < 0.01
    g = p[:, A.AP_GEOM].astype(int)
    # OneOf p = 0.3, three members with equal weight; RandomRotate90's k = 0 is the identity
    want = {1: 0.1, 2: 0.1, 3: 0.025, 4: 0.025, 5: 0.025}
    for code, prob in want.items():
        assert abs((g == code).mean() - prob) < 0.015, (code, (g == code).mean())
    assert abs((g == 0).mean() - 0.725) < 0.025
    assert set(np.unique(p[:, A.AP_ORDER]).astype(int)) == set(range(24))
    # another step counter, other draws; same (seed, step), same draws
    assert not np.array_equal(p, A.draw_params(4000, seed=1234, step=8))
    assert np.array_equal(p[:50], A.draw_params(50, seed=1234, step=7))


def test_colour_operations_closed_form():
    x = torch.rand(8, 8, 3)
    ident = np.array([1, 0, 1.0, 1.0, 1.0, 0.0, 0, 0], np.float32)
    assert torch.allclose(A.color_jitter(x, ident), x, atol=2e-6)          # unit factors, zero shift: identity on [0,1] data
    p = ident.copy()
    p[A.AP_BRIGHT] = 1.3
    assert torch.allclose(A.color_jitter(x, p), (x * 1.3).clamp(0, 1), atol=1e-6)
    p = ident.copy()
    p[A.AP_SAT] = 0.0                                                      # saturation 0 -> gray
    y = A.color_jitter(x, p)
    assert torch.allclose(y[..., 0], y[..., 1], atol=1e-6) and torch.allclose(y[..., 0], A._gray(x), atol=1e-6)
    p = ident.copy()
    p[A.AP_CONTRAST] = 0.0                                                 # contrast 0 -> the gray mean everywhere
    y = A.color_jitter(x, p)
    assert torch.allclose(y, torch.full_like(y, float(A._gray(x).mean())), atol=1e-6)
    # hue: a shift by one third of the circle rotates the channels of a pure colour
    red = torch.tensor([[[1.0, 0.0, 0.0]]])
    assert torch.allclose(A._hue(red, 1.0 / 3.0), torch.tensor([[[0.0, 1.0, 0.0]]]), atol=1e-6)
    assert torch.allclose(A._hue(x, 1.0), x, atol=1e-5)                    # full turn
    # the quirk: normalised (zero-mean) data is cut to [0, 1] by the first clipping operation
    z = torch.randn(8, 8, 3)
    p = ident.copy()
    assert A.color_jitter(z, p).min() >= 0


def test_geometry_matches_numpy():
    t = torch.arange(5 * 5 * 3).reshape(5, 5, 3)
    assert torch.equal(A.geometry(t, 1), torch.from_numpy(np.ascontiguousarray(t.numpy()[:, ::-1])))
    assert torch.equal(A.geometry(t, 2), torch.from_numpy(np.ascontiguousarray(t.numpy()[::-1])))
    for k in (1, 2, 3):
        assert torch.equal(A.geometry(t, 2 + k), torch.from_numpy(np.ascontiguousarray(np.rot90(t.numpy(), k))))
