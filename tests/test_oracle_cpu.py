"""Pins the oracle (oracle/*.py, the CPU restatement) against golden vectors produced by the REFERENCE's own code
(oracle/make_golden.py imported /root/reference in the build container; fixtures under tests/golden/). CPU only."""
import os
import struct

import numpy as np
import pytest
import torch

from oracle import resunet_oracle as R
from oracle import util_oracle as U

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return torch.load(os.path.join(GOLD, name), weights_only=False)


def test_state_dict_layout_matches_reference():
    ref = _load("state_keys.pt")
    ours = [(k, tuple(s), str(d)) for k, s, d in R.state_keys(1)]
    assert ours == [tuple(r) for r in ref]
    assert len(ours) == 136
    sd = R.make_state(1)
    assert [k for k in sd] == [r[0] for r in ref]
    n_params = sum(v.numel() for k, v in sd.items()
                   if v.dtype == torch.float32 and not k.endswith(("running_mean", "running_var")))
    assert n_params == 13926889


@pytest.mark.parametrize("nc", [1, 2])
def test_resunet_restatement_matches_reference_train_step(nc):
    fix = _load(f"resunet_nc{nc}.pt")
    torch.set_num_threads(1)
    img, mask = R.synthetic_batch(2, 64, seed=41)
    x = R.normalize(img)
    sd = R.grad_state(R.make_state(nc, seed=41))
    out = R.forward(sd, x, 64, training=True)
    assert torch.allclose(out, fix["logits_train"], rtol=1e-4, atol=1e-5)
    sq = U.squeeze_all(out)
    target = mask.float() if nc == 1 else torch.stack([1 - mask, mask], 1).float()
    pred = U.sigmoid_mask(sq)
    score = U.dice_score(pred, mask if nc == 1 else target)
    loss = U.dice_loss(sq, target)
    assert abs(loss.item() - fix["loss"].item()) < 1e-6
    assert abs(score.item() - fix["score"].item()) < 1e-6
    loss.backward()
    for k in fix["params_without_grad"]:
        assert sd[k].grad is None
    for k, (s, a) in fix["grad_checksums"].items():
        g = sd[k].grad.double()
        assert abs(g.abs().sum().item() - a) <= 2e-3 * a + 1e-9, k
        assert abs(g.sum().item() - s) <= 2e-3 * a + 1e-9, k
    for key, gk in (("conv4.weight", "grad_conv4_weight"), ("conv3.weight", "grad_conv3_weight"),
                    ("encoder.conv1.weight", "grad_stem_weight")):
        ref = fix[gk]
        assert (sd[key].grad - ref).abs().max().item() <= 2e-3 * ref.abs().max().item() + 1e-9, key
    assert torch.allclose(sd["encoder.bn1.running_mean"], fix["bn1_running_mean"], rtol=1e-4, atol=1e-6)
    assert torch.allclose(sd["encoder.bn1.running_var"], fix["bn1_running_var"], rtol=1e-4, atol=1e-6)
    assert torch.allclose(sd["encoder.layer4.1.bn2.running_var"], fix["l4_bn2_running_var"], rtol=1e-4, atol=1e-6)
    assert int(sd["encoder.bn1.num_batches_tracked"]) == fix["num_batches_tracked"] == 1


@pytest.mark.parametrize("nc", [1, 2])
def test_resunet_restatement_matches_reference_eval(nc):
    fix = _load(f"resunet_nc{nc}.pt")
    torch.set_num_threads(1)
    img, _ = R.synthetic_batch(2, 64, seed=41)
    with torch.no_grad():
        out = R.forward(R.make_state(nc, seed=41), R.normalize(img), 64, training=False)
    assert torch.allclose(out, fix["logits_eval"], rtol=1e-4, atol=1e-5)


def test_util_restatement_matches_reference():
    fix = _load("util.pt")
    for c in fix["dice_cases"]:
        x = c["logits"].clone().requires_grad_(True)
        loss = U.dice_loss(x, c["target"])
        loss.backward()
        assert torch.equal(loss.detach(), c["loss"])
        assert torch.allclose(x.grad, c["grad"], rtol=1e-6, atol=1e-12)
        score = U.dice_score(U.sigmoid_mask(c["logits"]), c["target"].to(torch.uint8))
        assert torch.equal(score, c["score"])
    edge = torch.tensor([struct.unpack("<f", struct.pack("<I", u))[0] for u in fix["edge_bits"]], dtype=torch.float32)
    assert torch.equal(U.sigmoid_mask(edge), fix["edge_mask_cpu"])
    # SURVEY 8a-9 [CPU fp32]: last true at 0xB43FFFFE, first false at 0xB43FFFFF
    bits = fix["edge_bits"]
    assert fix["edge_mask_cpu"][bits.index(0xB43FFFFE)] == 1 and fix["edge_mask_cpu"][bits.index(0xB43FFFFF)] == 0
    for shp, sq in fix["squeeze_shapes"].items():
        assert tuple(U.squeeze_all(torch.zeros(shp)).shape) == sq
    assert U.dice_score(torch.zeros(10), torch.zeros(10, dtype=torch.uint8)).item() == fix["score_empty"].item() == 1.0


def test_closed_form_kats():
    n = 1000
    loss = U.dice_loss(torch.zeros(n), torch.ones(n))
    assert abs(loss.item() - (1 - (2 * 0.5 * n + 1e-6) / (0.5 * n + n + 1e-6))) < 1e-6
    x = torch.tensor([[[[0.0, 0.0], [1.0, 2.0]], [[0.0, -1.0], [1.0, 3.0]]]])     # (1,2,2,2)
    assert U.argmax_mask(x).tolist() == [[[0, 0], [0, 1]]]                         # ties -> lowest index


def test_stitch_oracle_properties():
    rng = np.random.default_rng(0)
    T, th, tw = 6, 4, 4
    tiles = rng.integers(0, 2, (T, th, tw), dtype=np.uint8)
    cells = rng.permutation(12)[:T]
    rows, cols = (cells // 4).astype(np.int32), (cells % 4).astype(np.int32)
    full = U.stitch_tiles(tiles, rows, cols, 3 * th, 4 * tw)
    for t in range(T):
        assert np.array_equal(full[rows[t] * th:(rows[t] + 1) * th, cols[t] * tw:(cols[t] + 1) * tw], tiles[t])
    assert full.sum() == tiles.sum()
    bands = [U.stitch_tiles(tiles, rows, cols, th, 4 * tw, row0=b * th) for b in range(3)]
    assert np.array_equal(np.concatenate(bands, 0), full)          # row-band sharding == whole slide
    assert U.stitch_tiles(tiles[:0], rows[:0], cols[:0], 8, 8).sum() == 0     # empty input
    rgb = rng.integers(0, 256, (8, 8, 3), dtype=np.uint8)
    ov = U.overlay_blend(rgb, np.zeros((8, 8), np.uint8))
    assert np.array_equal(ov, rgb)
