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


@pytest.mark.parametrize("npix,offset", [(0, 0), (1, 0), (15, 0), (16, 0), (4099, 0), (64 * 96, 0), (1000, 1), (515, 16)])
@pytest.mark.parametrize("color,alpha", [((255, 0, 0), 102), ((17, 250, 99), 0), ((0, 255, 255), 256), ((255, 255, 255), 255)])
def test_overlay_blend_bit_exact(npix, offset, color, alpha):
    """Tumour overlay = integer blend (oracle/util_oracle.py overlay_blend), bit-exact for aligned, ragged and
    misaligned buffers, in place and out of place."""
    from oracle import util_oracle as U
    from pdac_pathological_image_segmentation_b200 import ops
    g = torch.Generator().manual_seed(npix * 7 + alpha)
    rgb_buf = torch.randint(0, 256, (offset * 3 + npix * 3,), dtype=torch.uint8, generator=g)
    mask_buf = (torch.randint(0, 3, (offset + npix,), generator=g) * 127).to(torch.uint8)    # 0, 127, 254: any non-zero counts
    if npix >= 64:
        mask_buf[offset + 16:offset + 48] = 0                                                 # an all-background group
    rgb = rgb_buf[offset * 3:].view(1, npix, 3)
    mask = mask_buf[offset:].view(1, npix)
    want = U.overlay_blend(rgb.numpy(), mask.numpy(), color, alpha)
    d_rgb = rgb_buf.to(DEV)[offset * 3:].view(1, npix, 3)
    d_mask = mask_buf.to(DEV)[offset:].view(1, npix)
    got = ops.overlay_blend(d_rgb, d_mask, color, alpha)
    assert np.array_equal(got.cpu().numpy(), want)
    assert torch.equal(d_rgb.cpu(), rgb), "out-of-place call must leave the input untouched"
    ops.overlay_blend(d_rgb, d_mask, color, alpha, out=d_rgb)
    assert np.array_equal(d_rgb.cpu().numpy(), want)


def test_overlay_blend_rejects_bad_arguments():
    from pdac_pathological_image_segmentation_b200 import ops
    rgb = torch.zeros((4, 4, 3), dtype=torch.uint8, device=DEV)
    mask = torch.zeros((4, 4), dtype=torch.uint8, device=DEV)
    with pytest.raises(RuntimeError):
        ops.overlay_blend(rgb, mask, (255, 0, 0), 257)
    with pytest.raises(RuntimeError):
        ops.overlay_blend(rgb, mask, (256, 0, 0), 100)


def test_slide_overlay_matches_oracle():
    """SlideInference.overlay: the stitched band mask blended over the band's RGB thumbnail at tile resolution."""
    from oracle import util_oracle as U
    from pdac_pathological_image_segmentation_b200.smp import Unet
    from pdac_pathological_image_segmentation_b200.wsi import SlideInference
    torch.manual_seed(5)
    size, gr, gc = 64, 2, 3
    model = Unet(encoder_weights=None).to(DEV)
    inf = SlideInference(model, gr, gc, tile=size)
    rows = torch.tensor([r for r in range(gr) for _ in range(gc)], dtype=torch.int32, device=DEV)
    cols = torch.tensor([c for _ in range(gr) for c in range(gc)], dtype=torch.int32, device=DEV)
    img = _tiles(gr * gc, size, 11).to(DEV)
    inf.run_batch(img, rows, cols)
    slide_rgb = img.view(gr, gc, size, size, 3).permute(0, 2, 1, 3, 4).reshape(gr * size, gc * size, 3).contiguous()
    got = inf.overlay(slide_rgb)
    want = U.overlay_blend(slide_rgb.cpu().numpy(), inf.band.cpu().numpy())
    assert np.array_equal(got.cpu().numpy(), want)
