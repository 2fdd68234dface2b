"""Whole-slide inference (BASELINE.json config 5; test.py:139-144 per tile + the missing visualize/predict_overlay.ipynb
per configs/visualize_config.yaml:3-15): uint8 HWC tiles -> normalise on the GPU -> eval forward on the hand-written
kernels -> sigmoid >= 0.5 mask (1 class) or argmax (>= 2 classes) -> scatter into the slide mask by the tiles' grid
coordinates (PyHIST tile_selection.tsv Row / Column). Shards naturally: rank r owns a band of grid rows and its slab of
the slide mask — no collective; the host concatenates the bands."""
from __future__ import annotations

import torch

from . import _lib, ops
from .engine import Tape


def band_of(rank, world, grid_rows):
    """Grid rows [r0, r1) of the slide owned by `rank` (contiguous bands, remainder to the first ranks)."""
    base, rem = divmod(grid_rows, world)
    r0 = rank * base + min(rank, rem)
    return r0, r0 + base + (1 if rank < rem else 0)


class SlideInference:
    def __init__(self, model, grid_rows, grid_cols, tile=512, rank=0, world=1, device=None):
        p = next(model.parameters())
        if p.device.type != "cuda":
            raise RuntimeError("SlideInference needs the model on a CUDA device; there is no CPU path")
        self.model = model.eval()
        self.device = device or p.device
        self.tile = tile
        self.grid_rows, self.grid_cols = grid_rows, grid_cols
        self.r0, self.r1 = band_of(rank, world, grid_rows)
        self.band = torch.zeros(((self.r1 - self.r0) * tile, grid_cols * tile), dtype=torch.uint8, device=self.device)

    def owns(self, rows):
        """Boolean mask of the tiles (by grid row) this rank's band holds: arbitrary TSV order is bucketed by owner."""
        return (rows >= self.r0) & (rows < self.r1)

    @torch.no_grad()
    def logits(self, img_u8):
        z, _ = ops.stem_prep_u8(img_u8, None, None)
        out, _ = self.model._run(Tape(self.model._packs, False, None), z)
        return out

    @torch.no_grad()
    def run_batch(self, img_u8, rows, cols):
        """img_u8 (B, tile, tile, 3) uint8 on the device; rows / cols (B,) int32 grid coordinates. Returns the tile masks."""
        _lib.peek_device_flag()          # non-blocking: raises if an earlier batch's kernels reported a pipeline time-out
        out = self.logits(img_u8)
        if out.shape[1] == 1:
            mask = ops.sigmoid_mask(out, as_uint8=True).view(out.shape[0], out.shape[2], out.shape[3])
        else:
            mask = ops.argmax_nchw(out)
        ops.stitch_tiles(mask.contiguous(), rows, cols, self.band, row0=self.r0 * self.tile)
        return mask

    def overlay(self, band_rgb, color=(255, 0, 0), alpha_256=102, out=None):
        """Tumour overlay of this rank's band (README.md:28-35; the notebook that drew it is a missing blob, the
        integer blend is specified in oracle/util_oracle.py overlay_blend). band_rgb: uint8 (band rows, slide
        columns, 3) on the device, same pixel grid as self.band; out=band_rgb blends in place."""
        assert tuple(band_rgb.shape) == (*self.band.shape, 3), "band_rgb must cover this rank's band"
        return ops.overlay_blend(band_rgb, self.band, color, alpha_256, out=out)
