"""Seeded synthetic inputs of the reference's data path (QuPath 512x512 RGB uint8 tiles + binary tumour masks,
QuPath_WSI_to_Patch.groovy:30-33; dataset.py:24-30) for benchmarks and demos: there is no dataset on the box."""
from __future__ import annotations

import torch
import torch.nn.functional as F


def synthetic_batch(batch, size, seed=41, blobby=True):
    """-> (img uint8 (B,S,S,3) HWC, mask uint8 (B,S,S) in {0,1}); CPU generator, identical on every box."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    img = torch.randint(0, 256, (batch, size, size, 3), dtype=torch.uint8, generator=g)
    if blobby:
        low = torch.rand(batch, 1, max(size // 32, 1), max(size // 32, 1), generator=g)
        mask = (F.interpolate(low, size=(size, size), mode="bilinear", align_corners=False)[:, 0] > 0.6)
    else:
        mask = torch.rand(batch, size, size, generator=g) < 0.5
    return img, mask.to(torch.uint8)
