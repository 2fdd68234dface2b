"""Drop-in replacements of the reference's loss / metric (util.py:6-26) on fused CUDA reductions."""
from __future__ import annotations

import torch
from torch import nn
from torch.autograd.function import once_differentiable

from . import ops


class _DiceFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets, smooth):
        loss, sums = ops.dice_fwd(logits, targets, smooth)
        ctx.save_for_backward(logits, targets, sums)
        ctx.smooth = smooth
        return loss

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        logits, targets, sums = ctx.saved_tensors
        return ops.dice_bwd(logits, targets, sums, g, ctx.smooth).view_as(logits), None, None


class DiceLoss(nn.Module):
    """util.py:6-19 — 1 - (2*sum(sigmoid(x)*t) + s) / (sum(sigmoid(x)) + sum(t) + s), one scalar over the batch."""

    def forward(self, inputs, targets, smooth=1e-6):
        if not inputs.is_cuda:
            raise RuntimeError("DiceLoss (pdac_b200) needs CUDA tensors; there is no CPU path")
        if targets.device != inputs.device:
            targets = targets.to(inputs.device)      # train_worker.py:77 round-trips the label through the host
        if inputs.dtype != torch.float32:
            inputs = inputs.float()
        if targets.dtype not in (torch.float32, torch.uint8, torch.bool):
            targets = targets.float()
        if inputs.numel() != targets.numel():
            raise RuntimeError(f"The size of tensor a ({inputs.numel()}) must match the size of tensor b "
                               f"({targets.numel()}) at non-singleton dimension 0")
        return _DiceFn.apply(inputs, targets, float(smooth))


def dice_score(pred, target, smooth=1e-6):
    """util.py:22-26 on hard masks; counts are accumulated as integers."""
    if target.device != pred.device:
        target = target.to(pred.device)
    if target.dtype not in (torch.float32, torch.uint8, torch.bool):
        target = target.to(torch.uint8)
    score, _ = ops.dice_score(pred, target, float(smooth))
    return score


def sigmoid_mask(logits):
    """train_worker.py:24,72 / test.py:144 — torch.ge(out.sigmoid(), 0.5).float(), bit-exact."""
    return ops.sigmoid_mask(logits)
