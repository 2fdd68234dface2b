"""ORACLE — test infrastructure only. bf16-storage emulation for the torch.nn oracle modules: every module output that
the CUDA product stores as a bf16 activation is rounded to bf16 (straight-through, gradient rounded too) while all
arithmetic stays fp32. Used to CALIBRATE gradient tolerances: how far does bf16 storage alone move the gradients from
the pure fp32 ones on a given seeded network (see tests/test_resunet_gpu.py header)."""
from __future__ import annotations

import contextlib

import torch
from torch import nn

from .resunet_oracle import round_bf16

_ROUNDED = (nn.Conv2d, nn.ConvTranspose2d, nn.ReLU, nn.BatchNorm2d, nn.GroupNorm, nn.UpsamplingBilinear2d,
            nn.AdaptiveAvgPool2d, nn.Dropout, nn.Dropout2d)


@contextlib.contextmanager
def bf16_storage(model: nn.Module):
    handles = []
    for m in model.modules():
        if isinstance(m, _ROUNDED):
            handles.append(m.register_forward_hook(lambda mod, inp, out: round_bf16(out)))
    try:
        yield model
    finally:
        for h in handles:
            h.remove()


def round_parameters_to_bf16_(model: nn.Module):
    """Makes every conv / linear weight exactly bf16-representable so the product's bf16 weight packing is lossless."""
    with torch.no_grad():
        for m in model.modules():
            if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
                m.weight.copy_(m.weight.to(torch.bfloat16).float())
    return model
