"""Shared nn.Module plumbing of the drop-in segmentation networks: device checks (no CPU fallback), the bridge between
torch.autograd (one Function per network call, so callers' `loss.backward()`, optimizers and DistributedDataParallel
work unchanged — train_worker.py:70,79,184,204) and the engine's kernel tape, and the pretrained-encoder lookup."""
from __future__ import annotations

import os
import warnings

import torch
from torch import nn
from torch.autograd.function import once_differentiable

from . import _lib, ops
from .engine import GradArena, PackCache, Tape


class _NetFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, x, *params):
        tape = Tape(mod._packs, mod.training, mod._grad_arena(params))
        logits, feed = mod._run(tape, x)
        ctx.tape, ctx.feed, ctx.params = tape, feed, params
        return logits

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        tape = ctx.tape
        tape.arena.zero()
        ctx.feed(g.contiguous())
        tape.backward()
        ctx.tape = ctx.feed = None
        return (None, None) + tuple(tape.arena.slot(p) for p in ctx.params)


class SegModule(nn.Module):
    """Base of ResUNet / Unet / FPN / PSPNet / DeepLabV3Plus. Subclasses implement
    `_run(tape, x_nchw_f32) -> (logits fp32 NCHW, feed_grad)` and `_trainable()` (parameters that receive gradients,
    i.e. everything except never-used ones such as ResUNet's encoder.fc)."""

    stride_divisor = 32

    def __init__(self):
        super().__init__()
        object.__setattr__(self, "_packs", PackCache())
        object.__setattr__(self, "_arena", None)

    # -- subclasses -------------------------------------------------------------------------------------------
    def _run(self, tape, x):
        raise NotImplementedError

    def _trainable(self):
        return [p for p in self.parameters() if p.requires_grad]

    # -- plumbing ---------------------------------------------------------------------------------------------
    def _grad_arena(self, params):
        sig = tuple(id(p) for p in params) + (params[0].device,)
        a = self._arena
        if a is None or a[0] != sig:
            a = (sig, GradArena(params, params[0].device))
            object.__setattr__(self, "_arena", a)
        return a[1]

    def _check_input(self, x):
        if not isinstance(x, torch.Tensor) or not x.is_cuda:
            raise RuntimeError(f"{type(self).__name__} (pdac_b200) runs on hand-written sm_100a CUDA kernels only: the "
                               "input must be a CUDA tensor and the module moved with .to(device); there is no CPU path")
        if x.dim() != 4 or x.shape[1] != 3:
            raise RuntimeError(f"expected input of shape (B, 3, H, W), got {tuple(x.shape)}")
        h, w = x.shape[2], x.shape[3]
        d = self.stride_divisor
        if h % d or w % d:
            raise RuntimeError(f"Wrong input shape height={h}, width={w}. Expected image height and width divisible by "
                               f"{d}.")
        p = next(self.parameters())
        if p.device != x.device:
            raise RuntimeError(f"module parameters are on {p.device} but the input is on {x.device}")
        _lib.load()
        # non-blocking poll of the device pipeline error flag: anything a finished kernel reported since the last call
        # (a timed-out TMA / tcgen05 barrier) raises here instead of flowing on into losses and optimizers
        _lib.peek_device_flag()

    def forward(self, x):
        self._check_input(x)
        if x.dtype != torch.float32:
            x = x.float()
        params = self._trainable()
        if torch.is_grad_enabled() and params:
            return _NetFn.apply(self, x, *params)
        with torch.no_grad():
            logits, _ = self._run(Tape(self._packs, self.training, None), x)
        return logits


def load_imagenet_resnet18(encoder, required):
    """The reference builds its encoders with ImageNet weights (models/resunet.py:12; encoder_weights='imagenet' at
    train_worker.py:155). There is no network here: use torchvision's cached checkpoint if one exists."""
    home = os.environ.get("TORCH_HOME", os.path.join(os.path.expanduser("~"), ".cache", "torch"))
    path = os.path.join(home, "hub", "checkpoints", "resnet18-f37072fd.pth")
    if not os.path.exists(path):
        msg = (f"ImageNet ResNet18 weights not found at {path} (no network access to download them); the encoder "
               "keeps its random (kaiming-normal) initialisation — load a state_dict to get trained weights")
        if required:
            raise RuntimeError(msg)
        warnings.warn(msg, stacklevel=3)
        return False
    sd = torch.load(path, map_location="cpu")
    own = encoder.state_dict()
    encoder.load_state_dict({k: v for k, v in sd.items() if k in own}, strict=False)
    return True
