"""Dice loss / dice score / sigmoid mask / argmax / stitching / Adam through the C ABI against the oracle restatement
(oracle/util_oracle.py, which follows util.py:6-26 and train_worker.py:72 of the reference)."""
import struct

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _bits(u):
    return struct.unpack("<f", struct.pack("<I", u))[0]


@pytest.mark.parametrize("n", [1, 7, 4096, 512 * 512 * 3 + 5])
@pytest.mark.parametrize("tdtype", [torch.float32, torch.uint8])
def test_dice_loss_fwd_bwd(n, tdtype):
    from oracle import util_oracle as O
    from pdac_pathological_image_segmentation_b200 import ops
    g = torch.Generator().manual_seed(n)
    x = (torch.randn(n, generator=g) * 3).to(DEV)
    t = (torch.rand(n, generator=g) < 0.3).to(tdtype).to(DEV)
    xr = x.clone().requires_grad_(True)
    ref = O.dice_loss(xr, t.float())
    ref.backward()
    loss, sums = ops.dice_fwd(x, t)
    assert abs(loss.item() - ref.item()) <= 2e-6, (loss.item(), ref.item())
    gx = ops.dice_bwd(x, t, sums, torch.ones((), device=DEV))
    scale = xr.grad.abs().max().item()
    assert (gx - xr.grad).abs().max().item() <= 1e-4 * scale + 1e-12
    gx2 = ops.dice_bwd(x, t, sums, torch.full((), 0.5, device=DEV))
    assert torch.allclose(gx2, gx * 0.5, rtol=1e-6, atol=0)


def test_dice_bwd_nhwc_equals_dice_bwd_then_layout():
    """The fused 1-class gradient (bf16 NHWC, class in channel 0 of 16) is bit-identical to dice_bwd followed by the
    fp32 NCHW -> bf16 NHWC layout pass it replaces in the train step."""
    from pdac_pathological_image_segmentation_b200 import ops
    g = torch.Generator().manual_seed(3)
    x = (torch.randn(3, 1, 40, 56, generator=g) * 2).to(DEV)
    t = (torch.rand(3, 40, 56, generator=g) < 0.4).to(torch.uint8).to(DEV)
    _, sums = ops.dice_fwd(x, t)
    want = ops.nchw_to_nhwc(ops.dice_bwd(x, t, sums, None), c_pad=16)
    got = ops.dice_bwd_nhwc(x, t, sums, 16)
    assert got.shape == (3, 40, 56, 16) and torch.equal(got, want)
    assert got[..., 0].float().abs().max().item() > 0 and got[..., 1:].abs().max().item() == 0


def test_dice_closed_forms():
    """SURVEY 8c KATs: zero logits vs all-ones target; empty prediction vs empty target."""
    from pdac_pathological_image_segmentation_b200 import ops
    n = 1000
    x = torch.zeros(n, device=DEV)
    t = torch.ones(n, device=DEV)
    loss, _ = ops.dice_fwd(x, t)
    expect = 1 - (2 * 0.5 * n + 1e-6) / (0.5 * n + n + 1e-6)
    assert abs(loss.item() - expect) < 1e-6
    z = torch.zeros(n, device=DEV)
    score, counts = ops.dice_score(z, z.to(torch.uint8))
    assert score.item() == 1.0 and counts.tolist() == [0, 0, 0]


def test_sigmoid_mask_threshold_edge_is_bit_exact():
    """train_worker.py:72 `torch.ge(out.sigmoid(), 0.5)`: the boundary sits just below zero in fp32.
    The product must agree with torch's own CUDA sigmoid on every float around the edge and on random logits."""
    from pdac_pathological_image_segmentation_b200 import ops
    edge = [_bits(u) for u in range(0xB43FFFF0, 0xB4400010)] + [0.0, -0.0, 1e-30, -1e-30, -1e-7, -2e-7, -3e-7,
                                                                  float("inf"), float("-inf"), 88.0, -88.0, -104.0]
    x = torch.tensor(edge, dtype=torch.float32, device=DEV)
    x = torch.cat([x, torch.randn(100000, device=DEV) * 1e-6, torch.randn(100000, device=DEV) * 4])
    ref = torch.ge(x.sigmoid(), 0.5)
    got_f = ops.sigmoid_mask(x)
    got_u = ops.sigmoid_mask(x, as_uint8=True)
    assert torch.equal(got_f, ref.float())
    assert torch.equal(got_u, ref.to(torch.uint8))


@pytest.mark.parametrize("n", [5, 512 * 512 * 4])
def test_dice_score_matches_oracle(n):
    from oracle import util_oracle as O
    from pdac_pathological_image_segmentation_b200 import ops
    g = torch.Generator().manual_seed(3)
    logits = torch.randn(n, generator=g).to(DEV)
    tgt = (torch.rand(n, generator=g) < 0.4).to(torch.uint8).to(DEV)
    pred = torch.ge(logits.sigmoid(), 0.5).float()
    ref = O.dice_score(pred, tgt)
    s0, c0 = ops.dice_score(pred, tgt)
    s1, c1 = ops.dice_score(logits, tgt, pred_is_logits=True)
    s2, c2 = ops.dice_score(pred.to(torch.uint8), tgt)
    inter = int((pred.bool() & tgt.bool()).sum())
    assert c0.tolist() == [inter, int(pred.sum()), int(tgt.sum())] == c1.tolist() == c2.tolist()
    for s in (s0, s1, s2):
        assert abs(s.item() - ref.item()) <= 1e-6


def test_argmax_ties_pick_lowest_index():
    from pdac_pathological_image_segmentation_b200 import ops
    x = torch.randn(3, 4, 16, 16, device=DEV)
    x[:, 2] = x[:, 1]                       # exact ties
    x[0, :, 0, 0] = 0.0
    got = ops.argmax_nchw(x)
    assert torch.equal(got.long(), x.argmax(1))
    assert torch.equal(got.long(), x.softmax(1).argmax(1)) or True   # softmax may merge near-ties; logits rule is the contract


def test_stitch_tiles_bit_exact():
    from oracle import util_oracle as O
    from pdac_pathological_image_segmentation_b200 import ops
    T, th, tw = 12, 16, 16
    g = torch.Generator().manual_seed(1)
    tiles = torch.randint(0, 2, (T, th, tw), dtype=torch.uint8, generator=g)
    rows = torch.randint(0, 6, (T,), dtype=torch.int32, generator=g)
    cols = torch.randint(0, 5, (T,), dtype=torch.int32, generator=g)
    # unique cells so the scatter is order-independent
    cells = torch.randperm(30, generator=g)[:T]
    rows, cols = (cells // 5).int(), (cells % 5).int()
    ref = O.stitch_tiles(tiles.numpy(), rows.numpy(), cols.numpy(), 6 * th, 5 * tw, row0=0)
    slide = torch.zeros(6 * th, 5 * tw, dtype=torch.uint8, device=DEV)
    ops.stitch_tiles(tiles.to(DEV), rows.to(DEV), cols.to(DEV), slide)
    assert torch.equal(slide.cpu(), torch.from_numpy(ref))
    # a row band [2*th, 4*th) as a rank would own it
    band = torch.zeros(2 * th, 5 * tw, dtype=torch.uint8, device=DEV)
    ops.stitch_tiles(tiles.to(DEV), rows.to(DEV), cols.to(DEV), band, row0=2 * th)
    assert torch.equal(band.cpu(), torch.from_numpy(ref[2 * th:4 * th]))


def test_adam_step_matches_torch():
    from pdac_pathological_image_segmentation_b200 import ops
    n = 10007
    p = torch.randn(n, device=DEV)
    ref = p.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref], lr=1e-4)
    m = torch.zeros(n, device=DEV)
    v = torch.zeros(n, device=DEV)
    for step in range(1, 4):
        gr = torch.randn(n, device=DEV)
        ref.grad = gr.clone()
        opt.step()
        ops.adam_step(p, gr, m, v, 1e-4, step)
    assert torch.allclose(p, ref.detach(), rtol=1e-6, atol=1e-7)
