"""Whole-slide inference (BASELINE config 5): tile masks and the stitched slide mask are bit-exact against the oracle on
identical logits (sigmoid >= 0.5 rule / argmax ties), for one rank and for a 2-way band split."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _tiles(n, size, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, (n, size, size, 3), dtype=torch.uint8, generator=g)


@pytest.mark.parametrize("world", [1, 2])
def test_slide_mask_bit_exact_on_identical_logits(world):
    from oracle import util_oracle as U
    from pdac_pathological_image_segmentation_b200.smp import Unet
    from pdac_pathological_image_segmentation_b200.wsi import SlideInference, band_of
    torch.manual_seed(41)
    size, gr, gc = 64, 3, 4
    model = Unet(encoder_weights=None).to(DEV)
    rows = torch.tensor([r for r in range(gr) for _ in range(gc)], dtype=torch.int32)
    cols = torch.tensor([c for _ in range(gr) for c in range(gc)], dtype=torch.int32)
    perm = torch.randperm(gr * gc, generator=torch.Generator().manual_seed(1))     # arbitrary TSV order
    rows, cols = rows[perm], cols[perm]
    img = _tiles(gr * gc, size, 7)
    full = np.zeros((gr * size, gc * size), dtype=np.uint8)
    bands = []
    for rank in range(world):
        inf = SlideInference(model, gr, gc, tile=size, rank=rank, world=world)
        own = inf.owns(rows)
        idx = torch.nonzero(own).flatten()
        for lo in range(0, len(idx), 5):                                            # ragged last batch
            sel = idx[lo:lo + 5]
            x, r, c = img[sel].to(DEV), rows[sel].to(DEV), cols[sel].to(DEV)
            logits = inf.logits(x)
            mask = inf.run_batch(x, r, c)
            want = U.sigmoid_mask(logits.cpu()).to(torch.uint8).view(len(sel), size, size)
            assert torch.equal(mask.cpu(), want), "tile mask must be bit-exact on identical logits"
            r0, _ = band_of(rank, world, gr)
            for t in range(len(sel)):
                full[int(r[t]) * size:(int(r[t]) + 1) * size, int(c[t]) * size:(int(c[t]) + 1) * size] = want[t].numpy()
        bands.append(inf.band.cpu().numpy())
    got = np.concatenate(bands, 0)
    assert got.shape == full.shape and np.array_equal(got, full)
    # and against the oracle's stitcher on the same tile masks
    tiles = np.stack([full[int(r) * size:(int(r) + 1) * size, int(c) * size:(int(c) + 1) * size] for r, c in zip(rows, cols)])
    assert np.array_equal(U.stitch_tiles(tiles, rows.numpy(), cols.numpy(), gr * size, gc * size), full)
