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
from .containers import BatchNorm2d, Conv2d, GroupNorm, Holder, ResNet18, _Container, seq
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


def run_conv2d_relu(tape, block, srcs, up2=False):
    """up2: also apply the next decoder block's nearest x2 upsampling (fused into the BatchNorm pass when there is one)."""
    conv = getattr(block, "0")
    bn = getattr(block, "1", None)
    if bn is not None:
        return E.conv_bn_act(tape, conv, bn, srcs, relu=True, up2=up2)
    y = E.conv_bias_act(tape, conv, srcs, relu=True)
    return E.nearest2x(tape, y) if up2 else y


class _SmpBase(SegModule):
    """encoder / decoder / segmentation_head registration order of smp.base.SegmentationModel."""

    def _head(self, tape, x, upsampling):
        conv = getattr(self.segmentation_head, "0")
        if (upsampling == 1 and conv.k == 3 and conv.padding == 1 and conv.stride == 1 and conv.dilation == 1
                and E.ops.head3x3_supported(x.shape[3], self.classes)):
            return E.head_conv3x3(tape, conv, x)        # Unet: full-resolution head, direct kernels (head.cu)
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
        y = E.nearest2x(tape, l4)                                      # F.interpolate(scale_factor=2, nearest)
        for i, skip in enumerate(skips):
            blk = getattr(self.decoder.blocks, str(i))
            srcs = [y, skip] if skip is not None else [y]              # torch.cat([x, skip], 1) = two K ranges
            y = run_conv2d_relu(tape, blk.conv1, srcs)
            # the NEXT block's upsampling is written by this block's last normalise pass
            y = run_conv2d_relu(tape, blk.conv2, [y], up2=i + 1 < len(skips))
        return self._head(tape, y, 1)


# ------------------------------------------------------------------------------------------------ FPN (App. B.3)
class _Conv3x3GNReLU(_Container):
    def __init__(self, cin, cout, upsample):
        super().__init__()
        self.upsample = upsample
        self.block = seq(Conv2d(cin, cout, 3, padding=1, bias=False), GroupNorm(32, cout))


class _FPNBlock(_Container):
    def __init__(self, pyramid, cskip):
        super().__init__()
        self.skip_conv = Conv2d(cskip, pyramid, 1)


class _SegBlock(_Container):
    def __init__(self, cin, cout, n_upsamples):
        super().__init__()
        blocks = [_Conv3x3GNReLU(cin, cout, bool(n_upsamples))]
        for _ in range(1, n_upsamples):
            blocks.append(_Conv3x3GNReLU(cout, cout, True))
        self.block = seq(*blocks)


class FPN(_SmpBase):
    """train_worker.py:152-161 / test.py:81-90. Decoder: 1x1 lateral convs + nearest-x2 top-down adds, four
    Conv3x3-GroupNorm(32)-ReLU(-bilinear x2) stacks brought to H/4, summed, Dropout2d, 1x1 head, bilinear x4."""

    def __init__(self, encoder_name="resnet18", encoder_depth=5, encoder_weights="imagenet",
                 decoder_pyramid_channels=256, decoder_segmentation_channels=128, decoder_merge_policy="add",
                 decoder_dropout=0.2, in_channels=3, classes=1, activation=None, upsampling=4, aux_params=None):
        super().__init__()
        _check_common(encoder_name, in_channels, activation)
        if encoder_depth != 5 or decoder_merge_policy != "add" or aux_params is not None or upsampling != 4:
            raise ValueError("FPN: encoder_depth=5, merge_policy='add', upsampling=4, no aux head (smp defaults)")
        if decoder_pyramid_channels % 64 or decoder_segmentation_channels % 32:
            raise ValueError("FPN: pyramid channels must be a multiple of 64, segmentation channels of 32")
        self.classes = classes
        self.p_drop = float(decoder_dropout)
        self.encoder = ResNet18(with_fc=False)
        pyr, segc = decoder_pyramid_channels, decoder_segmentation_channels
        seg_blocks = Holder()
        for i, n in enumerate((3, 2, 1, 0)):
            seg_blocks.add_module(str(i), _SegBlock(pyr, segc, n))
        self.decoder = Holder(p5=Conv2d(512, pyr, 1), p4=_FPNBlock(pyr, 256), p3=_FPNBlock(pyr, 128),
                              p2=_FPNBlock(pyr, 64), seg_blocks=seg_blocks)
        self.segmentation_head = seq(Conv2d(segc, classes, 1))
        _init_decoder(self.decoder)
        _init_head(self.segmentation_head)
        _init_encoder(self.encoder, encoder_weights)

    def _run(self, tape, x):
        _, c2, c3, c4, c5 = E.resnet18_features(tape, self.encoder, x)
        d = self.decoder
        p = E.conv_bias_act(tape, d.p5, [c5], relu=False)
        pyramid = [p]
        for blk, skip in ((d.p4, c4), (d.p3, c3), (d.p2, c2)):
            lateral = E.conv_bias_act(tape, blk.skip_conv, [skip], relu=False)
            p = E.add(tape, E.nearest2x(tape, p), lateral)
            pyramid.append(p)
        total = None
        for i, p in enumerate(pyramid):                                   # p5, p4, p3, p2
            y = p
            sb = getattr(d.seg_blocks, str(i)).block
            for name, layer in sb.named_children():
                y = E.conv_gn_relu(tape, getattr(layer.block, "0"), getattr(layer.block, "1"), [y])
                if layer.upsample:
                    y = E.bilinear(tape, y, (2 * y.shape[1], 2 * y.shape[2]), True)
            total = y if total is None else E.add(tape, total, y)
        total = E.dropout(tape, total, self.p_drop, per_channel=True)
        return self._head(tape, total, 4)


# ------------------------------------------------------------------------------------------------ PSPNet (App. B.4)
class PSPNet(_SmpBase):
    """train_worker.py:172-182 / test.py:101-111. encoder_depth 3 (layer2 features, stride 8; layer3/4 stay registered
    and unused, as in smp); pyramid pooling (1, 2, 3, 6) -> 1x1 conv(+BN)+ReLU -> bilinear back, concat with the
    features, 1x1 conv+BN+ReLU, Dropout2d, 3x3 head, bilinear x8."""

    stride_divisor = 8

    def __init__(self, encoder_name="resnet18", encoder_weights="imagenet", encoder_depth=3, psp_out_channels=512,
                 psp_use_batchnorm=True, psp_dropout=0.2, in_channels=3, classes=1, activation=None, upsampling=8,
                 aux_params=None):
        super().__init__()
        _check_common(encoder_name, in_channels, activation)
        if encoder_depth != 3 or upsampling != 8 or aux_params is not None:
            raise ValueError("PSPNet: encoder_depth=3, upsampling=8, no aux head (smp defaults) are implemented")
        if psp_out_channels % 16:
            raise ValueError("psp_out_channels must be a multiple of 16")
        self.classes = classes
        self.p_drop = float(psp_dropout)
        self.sizes = (1, 2, 3, 6)
        self.encoder = ResNet18(with_fc=False)
        blocks = Holder()
        for i, s in enumerate(self.sizes):
            pool = Holder()
            pool.add_module("1", conv2d_relu(128, 32, 1, 0, psp_use_batchnorm and s != 1))
            blocks.add_module(str(i), Holder(pool=pool))
        self.decoder = Holder(psp=Holder(blocks=blocks), conv=conv2d_relu(256, psp_out_channels, 1, 0, psp_use_batchnorm))
        self.segmentation_head = seq(Conv2d(psp_out_channels, classes, 3, padding=1))
        _init_decoder(self.decoder)
        _init_head(self.segmentation_head)
        _init_encoder(self.encoder, encoder_weights)

    def _trainable(self):
        unused = {id(p) for li in (3, 4) for p in getattr(self.encoder, f"layer{li}").parameters()}
        return [p for p in self.parameters() if p.requires_grad and id(p) not in unused]

    def _run(self, tape, x):
        feats = E.resnet18_features(tape, self.encoder, x, depth=3)
        f = feats[-1]                                                   # 128 @ H/8
        n, h, w, _ = f.shape
        branches = torch.empty((n, h, w, 128), dtype=torch.bfloat16, device=f.device)
        parts = []
        for i, s in enumerate(self.sizes):
            blk = getattr(getattr(self.decoder.psp.blocks, str(i)).pool, "1")
            pooled = E.adaptive_avgpool(tape, f, (s, s))
            y = run_conv2d_relu(tape, blk, [pooled])
            parts.append(E.bilinear_into(tape, y, branches[..., 32 * i:32 * (i + 1)], True))
        E.concat(tape, branches, parts)
        y = run_conv2d_relu(tape, self.decoder.conv, [branches, f])     # cat([b1, b2, b3, b6, x]) = two K ranges
        y = E.dropout(tape, y, self.p_drop, per_channel=True)
        return self._head(tape, y, 8)


# ------------------------------------------------------------------------------------------------ DeepLabV3+ (App. B.5)
def _separable(cin, cout, dilation):
    """smp SeparableConv2d: '0' depthwise 3x3 (groups = cin), '1' pointwise 1x1; both bias-free here."""
    return seq(Conv2d(cin, cin, 3, padding=dilation, dilation=dilation, bias=False, groups=cin),
               Conv2d(cin, cout, 1, bias=False))


class DeepLabV3Plus(_SmpBase):
    """train_worker.py:163-171 / test.py:92-100 (decoder_atrous_rates = cfg['dilations'] = (3, 6, 9)). Encoder at output
    stride 16 (layer4 dilated x2), separable ASPP + image pooling, 1x1 projection, Dropout(0.5), separable 3x3,
    bilinear x4, concat with a 48-channel projection of layer1, separable 3x3, 1x1 head, bilinear x4."""

    stride_divisor = 16

    def __init__(self, encoder_name="resnet18", encoder_depth=5, encoder_weights="imagenet", encoder_output_stride=16,
                 decoder_channels=256, decoder_atrous_rates=(12, 24, 36), in_channels=3, classes=1, activation=None,
                 upsampling=4, aux_params=None):
        super().__init__()
        _check_common(encoder_name, in_channels, activation)
        if encoder_depth != 5 or encoder_output_stride != 16 or upsampling != 4 or aux_params is not None:
            raise ValueError("DeepLabV3Plus: encoder_depth=5, output stride 16, upsampling=4 (smp defaults)")
        if decoder_channels % 64 or len(decoder_atrous_rates) != 3:
            raise ValueError("decoder_channels must be a multiple of 64 and three atrous rates are required")
        self.classes = classes
        self.rates = tuple(int(r) for r in decoder_atrous_rates)
        self.p_drop = 0.5
        dc = decoder_channels
        self.encoder = ResNet18(with_fc=False, output_stride=16)
        convs = Holder()
        convs.add_module("0", seq(Conv2d(512, dc, 1, bias=False), BatchNorm2d(dc)))
        for i, r in enumerate(self.rates, start=1):
            convs.add_module(str(i), seq(_separable(512, dc, r), BatchNorm2d(dc)))
        convs.add_module("4", seq(None, Conv2d(512, dc, 1, bias=False), BatchNorm2d(dc)))
        aspp = Holder(convs=convs, project=seq(Conv2d(5 * dc, dc, 1, bias=False), BatchNorm2d(dc)))
        self.decoder = Holder(aspp=seq(aspp, _separable(dc, dc, 1), BatchNorm2d(dc)),
                              block1=seq(Conv2d(64, 48, 1, bias=False), BatchNorm2d(48)),
                              block2=seq(_separable(48 + dc, dc, 1), BatchNorm2d(dc)))
        self.dc = dc
        self.segmentation_head = seq(Conv2d(dc, classes, 1))
        _init_decoder(self.decoder)
        _init_head(self.segmentation_head)
        _init_encoder(self.encoder, encoder_weights)

    def _sep_bn_relu(self, tape, sep, bn, x, out=None):
        y = E.depthwise3x3(tape, getattr(sep, "0"), x)
        return E.conv_bn_act(tape, getattr(sep, "1"), bn, [y], relu=True, out=out)

    def _run(self, tape, x):
        _, c2, _, _, c5 = E.resnet18_features(tape, self.encoder, x)    # c2: 64 @ H/4, c5: 512 @ H/16
        d, dc = self.decoder, self.dc
        aspp = getattr(d.aspp, "0")
        n, h, w, _ = c5.shape
        cat = torch.empty((n, h, w, 5 * dc), dtype=torch.bfloat16, device=c5.device)
        parts = []
        b0 = getattr(aspp.convs, "0")
        parts.append(E.conv_bn_act(tape, getattr(b0, "0"), getattr(b0, "1"), [c5], relu=True, out=cat[..., 0:dc]))
        for i in (1, 2, 3):
            b = getattr(aspp.convs, str(i))
            parts.append(self._sep_bn_relu(tape, getattr(b, "0"), getattr(b, "1"), c5, out=cat[..., i * dc:(i + 1) * dc]))
        b4 = getattr(aspp.convs, "4")
        pooled = E.adaptive_avgpool(tape, c5, (1, 1))
        g = E.conv_bn_act(tape, getattr(b4, "1"), getattr(b4, "2"), [pooled], relu=True)
        parts.append(E.bilinear_into(tape, g, cat[..., 4 * dc:5 * dc], False))
        E.concat(tape, cat, parts)
        y = E.conv_bn_act(tape, getattr(aspp.project, "0"), getattr(aspp.project, "1"), [cat], relu=True)
        y = E.dropout(tape, y, self.p_drop, per_channel=False)
        y = self._sep_bn_relu(tape, getattr(d.aspp, "1"), getattr(d.aspp, "2"), y)
        hq, wq = c2.shape[1], c2.shape[2]
        cat2 = torch.empty((n, hq, wq, dc + 48), dtype=torch.bfloat16, device=c5.device)
        up = E.bilinear_into(tape, y, cat2[..., 0:dc], True)            # UpsamplingBilinear2d(x4)
        low = E.conv_bn_act(tape, getattr(d.block1, "0"), getattr(d.block1, "1"), [c2], relu=True, out=cat2[..., dc:dc + 48])
        E.concat(tape, cat2, [up, low])
        y = self._sep_bn_relu(tape, getattr(d.block2, "0"), getattr(d.block2, "1"), cat2)
        return self._head(tape, y, 4)
