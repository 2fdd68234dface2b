"""ORACLE — test infrastructure only (see oracle/util_oracle.py header for who may import this).

fp32 torch.nn restatement of the segmentation_models_pytorch networks the reference instantiates at
train_worker.py:152-182 / test.py:81-111 (smp.FPN, smp.DeepLabV3Plus, smp.PSPNet) plus smp.Unet (named by
BASELINE.json configs 2 and 5), all with the resnet18 encoder.

PARITY UNPINNED: segmentation_models_pytorch is a third-party dependency that is absent from /root/reference, not
installed here and not version-pinned by the reference (no requirements file). The topologies, default arguments,
parameter names and initialisation below restate the published smp v0.3.x source; they anchor
the product's smp-compatible modules to stock torch fp32 arithmetic on the same topology and to smp's state_dict key
layout, but cannot be checked against the library itself in this environment.

Dropout / Dropout2d draw from torch's RNG stream, which the CUDA product cannot reproduce: parity runs use eval mode
or `dropout=0` (constructor argument of these restatements; smp's default probabilities are kept otherwise)."""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn


# ------------------------------------------------------------------------------------------------ encoder (B.1)
class BasicBlock(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        self.downsample = None
        if stride != 1 or cin != cout:
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        idt = x if self.downsample is None else self.downsample(x)
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        return self.relu(out + idt)


class ResNet18Encoder(nn.Module):
    """smp ResNetEncoder(resnet18): torchvision ResNet minus fc/avgpool; stages run up to `depth`; make_dilated(16)
    turns layer4 into stride 1 / dilation 2 (DeepLabV3+)."""
    out_channels = (3, 64, 64, 128, 256, 512)

    def __init__(self, depth=5, output_stride=32):
        super().__init__()
        self.depth = depth
        self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(3, 2, 1)
        cin = 64
        for li, c in enumerate((64, 128, 256, 512), start=1):
            setattr(self, f"layer{li}", nn.Sequential(BasicBlock(cin, c, 2 if li > 1 else 1), BasicBlock(c, c, 1)))
            cin = c
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
        if output_stride == 16:
            for m in self.layer4.modules():
                if isinstance(m, nn.Conv2d):
                    m.stride = (1, 1)
                    m.dilation = (2, 2)
                    k = m.kernel_size[0]
                    m.padding = ((k // 2) * 2, (k // 2) * 2)

    def forward(self, x):
        feats = [x]
        if self.depth >= 1:
            x = self.relu(self.bn1(self.conv1(x)))
            feats.append(x)
        if self.depth >= 2:
            x = self.layer1(self.maxpool(x))
            feats.append(x)
        for li in range(2, 5):
            if self.depth >= li + 1:
                x = getattr(self, f"layer{li}")(x)
                feats.append(x)
        return feats


def conv2d_relu(cin, cout, k, padding=0, use_batchnorm=True):
    """smp Conv2dReLU: Sequential(Conv2d(bias = not use_batchnorm), BatchNorm2d | Identity, ReLU)."""
    return nn.Sequential(nn.Conv2d(cin, cout, k, padding=padding, bias=not use_batchnorm),
                         nn.BatchNorm2d(cout) if use_batchnorm else nn.Identity(), nn.ReLU(inplace=True))


def segmentation_head(cin, classes, k, upsampling):
    """smp SegmentationHead: Conv2d -> UpsamplingBilinear2d (align_corners=True) | Identity -> Identity."""
    return nn.Sequential(nn.Conv2d(cin, classes, k, padding=k // 2),
                         nn.UpsamplingBilinear2d(scale_factor=upsampling) if upsampling > 1 else nn.Identity(),
                         nn.Identity())


def _init_decoder(m):
    for mod in m.modules():
        if isinstance(mod, nn.Conv2d):
            nn.init.kaiming_uniform_(mod.weight, mode="fan_in", nonlinearity="relu")
            if mod.bias is not None:
                nn.init.constant_(mod.bias, 0)
        elif isinstance(mod, (nn.BatchNorm2d, nn.GroupNorm)):
            nn.init.constant_(mod.weight, 1)
            nn.init.constant_(mod.bias, 0)


def _init_head(m):
    for mod in m.modules():
        if isinstance(mod, nn.Conv2d):
            nn.init.xavier_uniform_(mod.weight)
            if mod.bias is not None:
                nn.init.constant_(mod.bias, 0)


class _SegModel(nn.Module):
    def forward(self, x):
        return self.segmentation_head(self.decoder(*self.encoder(x)))


# ------------------------------------------------------------------------------------------------ Unet (B.2)
class _UnetBlock(nn.Module):
    def __init__(self, cin, cskip, cout):
        super().__init__()
        self.conv1 = conv2d_relu(cin + cskip, cout, 3, 1)
        self.conv2 = conv2d_relu(cout, cout, 3, 1)

    def forward(self, x, skip=None):
        x = F.interpolate(x, scale_factor=2, mode="nearest")
        if skip is not None:
            x = torch.cat([x, skip], dim=1)
        return self.conv2(self.conv1(x))


class _UnetDecoder(nn.Module):
    def __init__(self, channels=(256, 128, 64, 32, 16)):
        super().__init__()
        enc = [512, 256, 128, 64, 64]
        ins = [enc[0]] + list(channels[:-1])
        skips = enc[1:] + [0]
        self.blocks = nn.ModuleList([_UnetBlock(i, s, o) for i, s, o in zip(ins, skips, channels)])

    def forward(self, *features):
        feats = features[1:][::-1]
        x, skips = feats[0], feats[1:]
        for i, blk in enumerate(self.blocks):
            x = blk(x, skips[i] if i < len(skips) else None)
        return x


class Unet(_SegModel):
    def __init__(self, classes=1):
        super().__init__()
        self.encoder = ResNet18Encoder(5)
        self.decoder = _UnetDecoder()
        self.segmentation_head = segmentation_head(16, classes, 3, 1)
        _init_decoder(self.decoder)
        _init_head(self.segmentation_head)


# ------------------------------------------------------------------------------------------------ FPN (B.3)
class _Conv3x3GNReLU(nn.Module):
    def __init__(self, cin, cout, upsample):
        super().__init__()
        self.upsample = upsample
        self.block = nn.Sequential(nn.Conv2d(cin, cout, 3, 1, 1, bias=False), nn.GroupNorm(32, cout),
                                   nn.ReLU(inplace=True))

    def forward(self, x):
        x = self.block(x)
        if self.upsample:
            x = F.interpolate(x, scale_factor=2, mode="bilinear", align_corners=True)
        return x


class _FPNBlock(nn.Module):
    def __init__(self, pyramid, cskip):
        super().__init__()
        self.skip_conv = nn.Conv2d(cskip, pyramid, 1)

    def forward(self, x, skip):
        return F.interpolate(x, scale_factor=2, mode="nearest") + self.skip_conv(skip)


class _SegBlock(nn.Module):
    def __init__(self, cin, cout, n_upsamples):
        super().__init__()
        blocks = [_Conv3x3GNReLU(cin, cout, bool(n_upsamples))]
        for _ in range(1, n_upsamples):
            blocks.append(_Conv3x3GNReLU(cout, cout, True))
        self.block = nn.Sequential(*blocks)

    def forward(self, x):
        return self.block(x)


class _FPNDecoder(nn.Module):
    def __init__(self, pyramid=256, seg=128, dropout=0.2):
        super().__init__()
        self.p5 = nn.Conv2d(512, pyramid, 1)
        self.p4 = _FPNBlock(pyramid, 256)
        self.p3 = _FPNBlock(pyramid, 128)
        self.p2 = _FPNBlock(pyramid, 64)
        self.seg_blocks = nn.ModuleList([_SegBlock(pyramid, seg, n) for n in (3, 2, 1, 0)])
        self.dropout = nn.Dropout2d(p=dropout, inplace=True)

    def forward(self, *features):
        c2, c3, c4, c5 = features[-4:]
        p5 = self.p5(c5)
        p4 = self.p4(p5, c4)
        p3 = self.p3(p4, c3)
        p2 = self.p2(p3, c2)
        x = sum(b(p) for b, p in zip(self.seg_blocks, (p5, p4, p3, p2)))       # merge_policy "add"
        return self.dropout(x)


class FPN(_SegModel):
    def __init__(self, classes=1, dropout=0.2):
        super().__init__()
        self.encoder = ResNet18Encoder(5)
        self.decoder = _FPNDecoder(dropout=dropout)
        self.segmentation_head = segmentation_head(128, classes, 1, 4)
        _init_decoder(self.decoder)
        _init_head(self.segmentation_head)


# ------------------------------------------------------------------------------------------------ PSPNet (B.4)
class _PSPBlock(nn.Module):
    def __init__(self, cin, cout, pool_size, use_batchnorm):
        super().__init__()
        if pool_size == 1:
            use_batchnorm = False            # BN cannot normalise a single value per channel
        self.pool = nn.Sequential(nn.AdaptiveAvgPool2d(pool_size), conv2d_relu(cin, cout, 1, 0, use_batchnorm))

    def forward(self, x):
        h, w = x.shape[2:]
        return F.interpolate(self.pool(x), size=(h, w), mode="bilinear", align_corners=True)


class _PSPModule(nn.Module):
    def __init__(self, cin, sizes=(1, 2, 3, 6), use_batchnorm=True):
        super().__init__()
        self.blocks = nn.ModuleList([_PSPBlock(cin, cin // len(sizes), s, use_batchnorm) for s in sizes])

    def forward(self, x):
        return torch.cat([b(x) for b in self.blocks] + [x], dim=1)


class _PSPDecoder(nn.Module):
    def __init__(self, cin=128, use_batchnorm=True, out_channels=512, dropout=0.2):
        super().__init__()
        self.psp = _PSPModule(cin, use_batchnorm=use_batchnorm)
        self.conv = conv2d_relu(cin * 2, out_channels, 1, 0, use_batchnorm)
        self.dropout = nn.Dropout2d(p=dropout)

    def forward(self, *features):
        return self.dropout(self.conv(self.psp(features[-1])))


class PSPNet(_SegModel):
    """encoder_depth=3: the decoder reads layer2 (128 @ H/8); layer3/layer4 stay registered but unused."""

    def __init__(self, classes=1, psp_use_batchnorm=True, dropout=0.2):
        super().__init__()
        self.encoder = ResNet18Encoder(3)
        self.decoder = _PSPDecoder(128, psp_use_batchnorm, 512, dropout)
        self.segmentation_head = segmentation_head(512, classes, 3, 8)
        _init_decoder(self.decoder)
        _init_head(self.segmentation_head)


# ------------------------------------------------------------------------------------------------ DeepLabV3+ (B.5)
def separable_conv(cin, cout, k, padding=0, dilation=1, bias=False):
    return nn.Sequential(nn.Conv2d(cin, cin, k, 1, padding, dilation, groups=cin, bias=False),
                         nn.Conv2d(cin, cout, 1, bias=bias))


class _ASPPPooling(nn.Sequential):
    def __init__(self, cin, cout):
        super().__init__(nn.AdaptiveAvgPool2d(1), nn.Conv2d(cin, cout, 1, bias=False), nn.BatchNorm2d(cout), nn.ReLU())

    def forward(self, x):
        size = x.shape[-2:]
        for mod in self:
            x = mod(x)
        return F.interpolate(x, size=size, mode="bilinear", align_corners=False)


class _ASPP(nn.Module):
    def __init__(self, cin, cout, rates, dropout=0.5):
        super().__init__()
        mods = [nn.Sequential(nn.Conv2d(cin, cout, 1, bias=False), nn.BatchNorm2d(cout), nn.ReLU())]
        for r in rates:
            mods.append(nn.Sequential(separable_conv(cin, cout, 3, padding=r, dilation=r), nn.BatchNorm2d(cout),
                                      nn.ReLU()))
        mods.append(_ASPPPooling(cin, cout))
        self.convs = nn.ModuleList(mods)
        self.project = nn.Sequential(nn.Conv2d(5 * cout, cout, 1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(),
                                     nn.Dropout(dropout))

    def forward(self, x):
        return self.project(torch.cat([c(x) for c in self.convs], dim=1))


class _DeepLabDecoder(nn.Module):
    def __init__(self, rates=(12, 24, 36), out_channels=256, dropout=0.5):
        super().__init__()
        self.aspp = nn.Sequential(_ASPP(512, out_channels, rates, dropout),
                                  separable_conv(out_channels, out_channels, 3, padding=1),
                                  nn.BatchNorm2d(out_channels), nn.ReLU())
        self.up = nn.UpsamplingBilinear2d(scale_factor=4)
        self.block1 = nn.Sequential(nn.Conv2d(64, 48, 1, bias=False), nn.BatchNorm2d(48), nn.ReLU())
        self.block2 = nn.Sequential(separable_conv(48 + out_channels, out_channels, 3, padding=1),
                                    nn.BatchNorm2d(out_channels), nn.ReLU())

    def forward(self, *features):
        a = self.up(self.aspp(features[-1]))
        h = self.block1(features[-4])
        return self.block2(torch.cat([a, h], dim=1))


class DeepLabV3Plus(_SegModel):
    def __init__(self, classes=1, decoder_atrous_rates=(12, 24, 36), dropout=0.5):
        super().__init__()
        self.encoder = ResNet18Encoder(5, output_stride=16)
        self.decoder = _DeepLabDecoder(tuple(decoder_atrous_rates), 256, dropout)
        self.segmentation_head = segmentation_head(256, classes, 1, 4)
        _init_decoder(self.decoder)
        _init_head(self.segmentation_head)


# reference constructor-call table (train_worker.py:149-182): cfg["model"] -> builder
def build(model: str, classes=1, dilations=(3, 6, 9), dropout=None):
    if model == "unet_smp":
        return Unet(classes)
    if model == "fpn":
        return FPN(classes, 0.2 if dropout is None else dropout)
    if model == "pspnet":
        return PSPNet(classes, True, 0.2 if dropout is None else dropout)
    if model == "deeplabv3+":
        return DeepLabV3Plus(classes, dilations, 0.5 if dropout is None else dropout)
    raise ValueError(model)
