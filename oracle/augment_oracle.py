"""ORACLE — test infrastructure only (tests/, __graft_entry__.smoke() and bench.py's CPU leg may import this; the product
never does).

Executable restatement of the reference's training augmentation (train_main.py:70-80):
    A.Normalize(mean, std) -> A.ColorJitter(0.3, 0.3, 0.3, 0.3) -> A.OneOf([HorizontalFlip, RandomRotate90, VerticalFlip], p=0.3)
albumentations is a third-party dependency that is absent from /root/reference, not installed here and not version-pinned
by the reference: PARITY UNPINNED. The arithmetic below restates the library's published float-image code path
(brightness / contrast / saturation as clipped linear blends, hue through HSV, operations in a random order, the whole
transform with probability 0.5; OneOf picks exactly one member with equal weights; RandomRotate90 rotates by k x 90
degrees counter-clockwise, k uniform in 0..3) and is what csrc/augment.cu is checked against, bit-for-bit for the random
draws (a counter-based splitmix64 hash) and to bf16 rounding for the colour arithmetic.

Reference quirk kept: ColorJitter runs AFTER Normalize, on zero-mean data, and the float path clips to [0, 1]."""
from __future__ import annotations

import numpy as np
import torch

M64 = (1 << 64) - 1
AP_APPLY, AP_ORDER, AP_BRIGHT, AP_CONTRAST, AP_SAT, AP_HUE, AP_GEOM, AP_MEAN = range(8)


def splitmix64(z: int) -> int:
    z = (z + 0x9E3779B97F4A7C15) & M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
    return z ^ (z >> 31)


def draw_uniform(key: int, idx: int) -> np.float32:
    return np.float32(splitmix64((key + idx * 0xD1342543DE82EF95) & M64) >> 40) * np.float32(1.0 / 16777216.0)


def decode_order(idx: int):
    pool, ops, div = [0, 1, 2, 3], [], 6
    for k in range(4):
        sel = idx // div
        idx -= sel * div
        ops.append(pool.pop(sel))
        div = div // (3 - k) if k < 2 else 1
    return ops


def draw_params(n_patches, seed, step=0, brightness=0.3, contrast=0.3, saturation=0.3, hue=0.3, p_color=0.5, p_geom=0.3):
    """-> float32 [N, 8] per-patch parameters exactly as aug_draw_kernel draws them (entry 7, the gray mean, is 0)."""
    key = splitmix64((seed & M64) ^ splitmix64(step & M64))
    out = np.zeros((n_patches, 8), np.float32)
    f = np.float32
    for n in range(n_patches):
        b = n * 16
        out[n, AP_APPLY] = 1.0 if draw_uniform(key, b + 0) < f(p_color) else 0.0
        out[n, AP_ORDER] = min(int(draw_uniform(key, b + 1) * f(24.0)), 23)
        out[n, AP_BRIGHT] = f(1) + (f(2) * draw_uniform(key, b + 2) - f(1)) * f(brightness)
        out[n, AP_CONTRAST] = f(1) + (f(2) * draw_uniform(key, b + 3) - f(1)) * f(contrast)
        out[n, AP_SAT] = f(1) + (f(2) * draw_uniform(key, b + 4) - f(1)) * f(saturation)
        out[n, AP_HUE] = (f(2) * draw_uniform(key, b + 5) - f(1)) * f(hue)
        geom = 0
        if draw_uniform(key, b + 6) < f(p_geom):
            which = min(int(draw_uniform(key, b + 7) * f(3.0)), 2)
            if which == 0:
                geom = 1
            elif which == 2:
                geom = 2
            else:
                k = min(int(draw_uniform(key, b + 8) * f(4.0)), 3)
                geom = 0 if k == 0 else 2 + k
        out[n, AP_GEOM] = geom
    return out


def _gray(x):
    return 0.299 * x[..., 0] + 0.587 * x[..., 1] + 0.114 * x[..., 2]


def _hue(x, f):
    r, g, b = x[..., 0], x[..., 1], x[..., 2]
    v = x.max(-1).values
    mn = x.min(-1).values
    d = v - mn
    s = torch.where(v > 0, d / torch.where(v > 0, v, torch.ones_like(v)), torch.zeros_like(v))
    dd = torch.where(d > 0, d, torch.ones_like(d))
    h = torch.where(v == r, 60.0 * (g - b) / dd, torch.where(v == g, 120.0 + 60.0 * (b - r) / dd, 240.0 + 60.0 * (r - g) / dd))
    h = torch.where(d > 0, h, torch.zeros_like(h))
    h = torch.where(h < 0, h + 360.0, h)
    h = h + 360.0 * f
    h = h - 360.0 * torch.floor(h * (1.0 / 360.0))
    h = torch.where(h >= 360.0, h - 360.0, h)
    sect = h * (1.0 / 60.0)
    i = torch.floor(sect)
    fr = sect - i
    p, q, t = v * (1 - s), v * (1 - s * fr), v * (1 - s * (1 - fr))
    i = i.long().clamp(0, 5)
    table = [(v, t, p), (q, v, p), (p, v, t), (p, q, v), (t, p, v), (v, p, q)]
    out = torch.zeros_like(x)
    for k, (a, bb, c) in enumerate(table):
        m = (i == k).unsqueeze(-1)
        out = torch.where(m, torch.stack([a, bb, c], -1), out)
    return out


def color_jitter(x, p, upto=4, mean=None):
    """x: float32 (H, W, 3) normalised patch; p: its parameter row. Applies the first `upto` operations of the order."""
    ops = decode_order(int(p[AP_ORDER]))
    for op in ops[:upto]:
        if op == 0:
            x = (x * float(p[AP_BRIGHT])).clamp(0, 1)
        elif op == 1:
            m = _gray(x).double().mean().float() if mean is None else mean
            f = float(p[AP_CONTRAST])
            x = (x * f + m * (1 - f)).clamp(0, 1)
        elif op == 2:
            f = float(p[AP_SAT])
            x = (x * f + _gray(x).unsqueeze(-1) * (1 - f)).clamp(0, 1)
        else:
            x = _hue(x, float(p[AP_HUE]))
    return x


def contrast_mean(x, p):
    """The gray mean the contrast step uses: of the patch after the operations that precede contrast in its order."""
    ops = decode_order(int(p[AP_ORDER]))
    return _gray(color_jitter(x, p, upto=ops.index(1))).double().mean().float()


def geometry(t, code):
    """t: (H, W, ...) tensor; code 0 none, 1 hflip, 2 vflip, 3/4/5 = np.rot90 with k = 1/2/3 (counter-clockwise)."""
    if code == 1:
        return t.flip(1)
    if code == 2:
        return t.flip(0)
    if code >= 3:
        return torch.rot90(t, code - 2, (0, 1))
    return t


def apply(img_u8, mask_u8, params, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225)):
    """img_u8 (N,H,W,3) uint8, mask_u8 (N,H,W) uint8 or None, params (N,8) as drawn -> (x float32 (N,3,H,W), mask)."""
    sc = torch.tensor([1.0 / (255.0 * s) for s in std], dtype=torch.float32)
    sh = torch.tensor([-m / s for m, s in zip(mean, std)], dtype=torch.float32)
    xs, ms = [], []
    for n in range(img_u8.shape[0]):
        x = img_u8[n].float() * sc + sh
        p = params[n]
        if p[AP_APPLY] != 0:
            x = color_jitter(x, p)
        code = int(p[AP_GEOM])
        xs.append(geometry(x, code).permute(2, 0, 1))
        if mask_u8 is not None:
            ms.append(geometry(mask_u8[n], code))
    return torch.stack(xs), (torch.stack(ms) if mask_u8 is not None else None)
