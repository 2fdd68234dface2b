"""Drop-in replacements of the segmentation_models_pytorch networks the reference trains (train_worker.py:152-182,
test.py:81-111: smp.FPN / smp.DeepLabV3Plus / smp.PSPNet) plus smp.Unet (BASELINE.json configs 2 and 5), with the
resnet18 encoder. Same keyword arguments as the reference's calls, `forward(x) -> logits (B, classes, H, W)` fp32 and
smp's state_dict keys / shapes / order (smp itself is absent here, so that layout is
"parity unpinned" upstream and checked against oracle/smp_oracle.py). Executed on libpdac_b200.so through engine.py."""
from __future__ import annotations

import math

import torch
from torch import nn

from . import engine as E
from .containers import BatchNorm2d, Conv2d, Holder, ResNet18, _Container, seq
from .module_base import SegModule, load_imagenet_resnet18


def _check_common(encoder_name, in_channels, activation, encoder_depth=5):
    if encoder_name != "resnet18":
        raise ValueError(f"pdac_b200 implements the resnet18 encoder the reference configures "
                         f"(configs/train_config.yaml:21), got encoder_name={encoder_name!r}")
    if in_channels != 3:
        raise ValueError("in_channels must be 3 (RGB pathology patches)")
    if activation is not None:
        raise ValueError("activation must be None (the reference applies sigmoid in its loss / mask code)")


def _init_encoder(enc, encoder_weights):
    if encoder_weights is None:
        return
    if encoder_weights != "imagenet":
        raise ValueError(f"encoder_weights must be None or 'imagenet', got {encoder_weights!r}")
    load_imagenet_resnet18(enc, required=True)


def _init_decoder(mod):
    """smp initialize_decoder: kaiming-uniform(fan_in, relu) convs with zero bias, norm layers 1/0."""
    for m in mod.modules():
        if isinstance(m, Conv2d):
            nn.init.kaiming_uniform_(m.weight, mode="fan_in", nonlinearity="relu")
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)


def _init_head(mod):
    for m in mod.modules():
        if isinstance(m, Conv2d):
            nn.init.xavier_uniform_(m.weight)
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)


def conv2d_relu(cin, cout, k, padding=0, use_batchnorm=True):
    """smp Conv2dReLU container: '0' = conv (bias iff no BN), '1' = BatchNorm2d."""
    return seq(Conv2d(cin, cout, k, padding=padding, bias=not use_batchnorm),
               BatchNorm2d(cout) if use_batchnorm else None)


def run_conv2d_relu(tape, block, srcs):
    conv = getattr(block, "0")
    bn = getattr(block, "1", None)
    if bn is not None:
        return E.conv_bn_act(tape, conv, bn, srcs, relu=True)
    return E.conv_bias_act(tape, conv, srcs, relu=True)


class _SmpBase(SegModule):
    """encoder / decoder / segmentation_head registration order of smp.base.SegmentationModel."""

    def _head(self, tape, x, upsampling):
        conv = getattr(self.segmentation_head, "0")
        y = E.conv_bias_act(tape, conv, [x], relu=False)
        out_hw = (x.shape[1] * upsampling, x.shape[2] * upsampling)
        return E.logits_head(tape, y, self.classes, out_hw, True)      # UpsamplingBilinear2d: align_corners=True


# ------------------------------------------------------------------------------------------------ Unet (App. B.2)
class _UnetBlock(_Container):
    def __init__(self, cin, cskip, cout, use_batchnorm):
        super().__init__()
        self.conv1 = conv2d_relu(cin + cskip, cout, 3, 1, use_batchnorm)
        self.conv2 = conv2d_relu(cout, cout, 3, 1, use_batchnorm)


class Unet(_SmpBase):
    def __init__(self, encoder_name="resnet18", encoder_depth=5, encoder_weights="imagenet",
                 decoder_use_batchnorm=True, decoder_channels=(256, 128, 64, 32, 16), decoder_attention_type=None,
                 in_channels=3, classes=1, activation=None, aux_params=None):
        super().__init__()
        _check_common(encoder_name, in_channels, activation)
        if encoder_depth != 5 or decoder_attention_type is not None or aux_params is not None:
            raise ValueError("Unet: encoder_depth=5, no attention, no aux head (smp defaults) are implemented")
        if len(decoder_channels) != 5 or any(c % 16 for c in decoder_channels):
            raise ValueError("decoder_channels must be 5 multiples of 16")
        self.classes = classes
        self.encoder = ResNet18(with_fc=False)
        enc = [512, 256, 128, 64, 64]
        ins = [enc[0]] + list(decoder_channels[:-1])
        skips = enc[1:] + [0]
        blocks = Holder()
        for i, (ci, cs, co) in enumerate(zip(ins, skips, decoder_channels)):
            blocks.add_module(str(i), _UnetBlock(ci, cs, co, decoder_use_batchnorm))
        self.decoder = Holder(blocks=blocks)
        self.segmentation_head = seq(Conv2d(decoder_channels[-1], classes, 3, padding=1))
        _init_decoder(self.decoder)
        _init_head(self.segmentation_head)
        _init_encoder(self.encoder, encoder_weights)

    def _run(self, tape, x):
        stem, l1, l2, l3, l4 = E.resnet18_features(tape, self.encoder, x)
        skips = [l3, l2, l1, stem, None]
        y = l4
        for i, skip in enumerate(skips):
            blk = getattr(self.decoder.blocks, str(i))
            y = E.nearest2x(tape, y)                                   # F.interpolate(scale_factor=2, nearest)
            srcs = [y, skip] if skip is not None else [y]              # torch.cat([x, skip], 1) = two K ranges
            y = run_conv2d_relu(tape, blk.conv1, srcs)
            y = run_conv2d_relu(tape, blk.conv2, [y])
        return self._head(tape, y, 1)
