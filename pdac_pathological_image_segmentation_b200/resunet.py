"""Drop-in replacement of the reference's ResUNet (models/resunet.py:7-58): same constructor arguments, the same
`forward(x) -> logits (B, num_classes, output_size, output_size)` fp32 NCHW contract and the same 136-entry
state_dict (keys, shapes, dtypes, order — incl. the never-used encoder.fc.*), executed on libpdac_b200.so."""
from __future__ import annotations

from . import engine as E
from .containers import Conv2d, ConvTranspose2d, ResNet18
from .module_base import SegModule, load_imagenet_resnet18


class ResUNet(SegModule):
    def __init__(self, num_classes, output_size):
        super().__init__()
        self.output_size = output_size
        self.num_classes = num_classes
        self.encoder = ResNet18(with_fc=True)                   # models/resunet.py:12 (fc kept, unused)
        load_imagenet_resnet18(self.encoder, required=False)
        self.upconv1 = ConvTranspose2d(512, 256)                # :15
        self.conv1 = Conv2d(512, 256, 3, padding=1)             # :16
        self.upconv2 = ConvTranspose2d(256, 128)
        self.conv2 = Conv2d(256, 128, 3, padding=1)
        self.upconv3 = ConvTranspose2d(128, 64)
        self.conv3 = Conv2d(128, 64, 3, padding=1)
        self.conv4 = Conv2d(64, num_classes, 1)                 # :21

    def _trainable(self):
        fc = {id(self.encoder.fc.weight), id(self.encoder.fc.bias)}
        return [p for p in self.parameters() if p.requires_grad and id(p) not in fc]

    def _run(self, tape, x):
        _, x2, x3, x4, x5 = E.resnet18_features(tape, self.encoder, x)          # :25-33
        y = E.convt2x2(tape, self.upconv1, x5)                                  # :36
        y = E.conv_bias_act(tape, self.conv1, [y, x4], relu=True)               # :37-38 (cat is two K ranges)
        y = E.convt2x2(tape, self.upconv2, y)
        y = E.conv_bias_act(tape, self.conv2, [y, x3], relu=True)
        y = E.convt2x2(tape, self.upconv3, y)
        y = E.conv_bias_act(tape, self.conv3, [y, x2], relu=True)
        y = E.conv_bias_act(tape, self.conv4, [y], relu=False)                  # :48
        return E.logits_head(tape, y, self.num_classes, (self.output_size, self.output_size), False)   # :51-56
