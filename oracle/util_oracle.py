"""ORACLE — test infrastructure only (tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this; the product package never does).

CPU restatement (torch fp32 / numpy) of the reference's loss, metric and mask arithmetic:

  dice_loss      <- util.py:10-19   (DiceLoss.forward)
  dice_score     <- util.py:22-26
  sigmoid_mask   <- train_worker.py:24,72 and test.py:144   (torch.ge(out.sigmoid(), 0.5).float())
  squeeze_all    <- train_worker.py:23,25,71,73              (torch.squeeze drops every size-1 dim)
  train_step     <- train_worker.py:68-80                    (zero_grad / forward / mask / score / loss / backward / step)
  argmax_mask    <- BASELINE config 1/5 (softmax-argmax; not in the reference code: argmax of logits, ties -> lowest)
  stitch_tiles   <- visualize/predict_overlay.ipynb is a missing blob (.MISSING_LARGE_BLOBS:1); spec from
                    configs/visualize_config.yaml:3-15 + BASELINE.json config 5 -> PARITY UNPINNED for stitching.

Pinning: tests/test_oracle_cpu.py checks dice_loss / dice_score / sigmoid_mask / train_step against the golden vectors
in tests/golden/ that oracle/make_golden.py produced by importing the reference's own util.py in the build container.
"""
from __future__ import annotations

import numpy as np
import torch


def dice_loss(inputs: torch.Tensor, targets: torch.Tensor, smooth: float = 1e-6) -> torch.Tensor:
    """1 - (2*sum(p*t) + s) / (sum(p) + sum(t) + s) with p = sigmoid(inputs), ONE scalar over the whole batch."""
    p = torch.sigmoid(inputs).reshape(-1)
    t = targets.reshape(-1)
    inter = (p * t).sum()
    return 1 - (2.0 * inter + smooth) / (p.sum() + t.sum() + smooth)


def dice_score(pred: torch.Tensor, target: torch.Tensor, smooth: float = 1e-6) -> torch.Tensor:
    """Hard Dice over the whole batch; `target` may be uint8 (promoted by the multiply, util.py:23)."""
    inter = (pred * target).sum()
    union = pred.sum() + target.sum()
    return (2.0 * inter + smooth) / (union + smooth)


def sigmoid_mask(logits: torch.Tensor) -> torch.Tensor:
    return torch.ge(logits.sigmoid(), 0.5).float()


def squeeze_all(t: torch.Tensor) -> torch.Tensor:
    return torch.squeeze(t)


def argmax_mask(logits_nchw: torch.Tensor) -> torch.Tensor:
    """argmax_c softmax(x)_c == argmax_c x_c; torch.argmax returns the first maximal index."""
    return logits_nchw.argmax(dim=1).to(torch.uint8)


def train_step(model, optimizer, img, label):
    """One iteration of the reference hot loop (train_worker.py:68-80) on whatever device the tensors live on.
    Returns (score, loss, logits)."""
    optimizer.zero_grad()
    out = model(img)
    sq = squeeze_all(out)
    pred = sigmoid_mask(sq)
    lab = squeeze_all(label)
    score = dice_score(pred, lab)
    loss = dice_loss(sq, lab.float())
    loss.backward()
    optimizer.step()
    return score.detach(), loss.detach(), out.detach()


def stitch_tiles(tiles: np.ndarray, rows: np.ndarray, cols: np.ndarray, slide_h: int, slide_w: int,
                 row0: int = 0) -> np.ndarray:
    """tiles uint8 (T, th, tw); tile t covers slide rows rows[t]*th .. and columns cols[t]*tw ..; returns the band of
    `slide_h` rows starting at absolute row `row0` (pixels outside the band / slide are dropped)."""
    T, th, tw = tiles.shape
    out = np.zeros((slide_h, slide_w), dtype=np.uint8)
    for t in range(T):
        r0 = int(rows[t]) * th - row0
        c0 = int(cols[t]) * tw
        for i in range(th):
            r = r0 + i
            if r < 0 or r >= slide_h:
                continue
            lo, hi = max(c0, 0), min(c0 + tw, slide_w)
            if hi > lo:
                out[r, lo:hi] = tiles[t, i, lo - c0:hi - c0]
    return out


def overlay_blend(rgb: np.ndarray, mask: np.ndarray, color=(255, 0, 0), alpha_256: int = 102) -> np.ndarray:
    """Integer overlay: where mask != 0, out = (rgb*(256-a) + color*a) >> 8 (a = alpha_256); elsewhere rgb. uint8 HWC.
    (README.md:28-35 shows a tumour overlay; the notebook is missing -> PARITY UNPINNED, integer spec is ours.)"""
    out = rgb.astype(np.uint32)
    col = np.asarray(color, dtype=np.uint32)
    blended = (out * (256 - alpha_256) + col * alpha_256) >> 8
    m = (mask != 0)[..., None]
    return np.where(m, blended, out).astype(np.uint8)
