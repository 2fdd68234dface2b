"""Parameter containers. They register parameters / buffers under exactly the names, shapes, dtypes and order the
reference's torch.nn / torchvision / smp modules use (so state_dict(), load_state_dict(strict=True), optimizers and
DistributedDataParallel see the same thing), initialise them the way the reference's constructors do, and do NO
arithmetic: calling one raises. All compute runs through the engine (engine.py) on libpdac_b200.so.

Reference layouts: nn.Conv2d / nn.ConvTranspose2d / nn.BatchNorm2d as built at models/resunet.py:12-21 and
torchvision resnet.py:59-106,197-204; nn.GroupNorm / nn.Linear for smp FPN and the unused encoder.fc."""
from __future__ import annotations

import math

import torch
from torch import nn


class _Container(nn.Module):
    def forward(self, *a, **k):
        raise RuntimeError(f"{type(self).__name__} is a parameter container of pdac_b200; it is executed by its parent "
                           "segmentation module on hand-written CUDA kernels, not called directly")


def _uniform_fan_in_(w, fan_in):
    bound = 1.0 / math.sqrt(fan_in) if fan_in > 0 else 0.0
    with torch.no_grad():
        w.uniform_(-bound, bound)


class Conv2d(_Container):
    """weight (O, I/groups, R, S) [+ bias (O)], torch default init (kaiming-uniform a=sqrt(5) == U(+-1/sqrt(fan_in)))."""

    def __init__(self, cin, cout, k, stride=1, padding=0, dilation=1, bias=True, groups=1):
        super().__init__()
        self.cin, self.cout, self.k = cin, cout, k
        self.stride, self.padding, self.dilation, self.groups = stride, padding, dilation, groups
        self.weight = nn.Parameter(torch.empty(cout, cin // groups, k, k))
        fan_in = (cin // groups) * k * k
        _uniform_fan_in_(self.weight, fan_in)
        if bias:
            self.bias = nn.Parameter(torch.empty(cout))
            _uniform_fan_in_(self.bias, fan_in)
        else:
            self.register_parameter("bias", None)

    def extra_repr(self):
        return (f"{self.cin}, {self.cout}, k={self.k}, stride={self.stride}, padding={self.padding}, "
                f"dilation={self.dilation}, groups={self.groups}, bias={self.bias is not None}")


class ConvTranspose2d(_Container):
    """k=2, s=2 only (models/resunet.py:15,17,19). weight (Cin, Cout, 2, 2), bias (Cout)."""

    def __init__(self, cin, cout):
        super().__init__()
        self.cin, self.cout = cin, cout
        self.weight = nn.Parameter(torch.empty(cin, cout, 2, 2))
        self.bias = nn.Parameter(torch.empty(cout))
        fan_in = cout * 4            # torch computes fan_in from weight.size(1) for transposed convs
        _uniform_fan_in_(self.weight, fan_in)
        _uniform_fan_in_(self.bias, fan_in)


class BatchNorm2d(_Container):
    def __init__(self, c, eps=1e-5, momentum=0.1):
        super().__init__()
        self.num_features, self.eps, self.momentum = c, eps, momentum
        self.weight = nn.Parameter(torch.ones(c))
        self.bias = nn.Parameter(torch.zeros(c))
        self.register_buffer("running_mean", torch.zeros(c))
        self.register_buffer("running_var", torch.ones(c))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))


class GroupNorm(_Container):
    def __init__(self, groups, c, eps=1e-5):
        super().__init__()
        self.num_groups, self.num_channels, self.eps = groups, c, eps
        self.weight = nn.Parameter(torch.ones(c))
        self.bias = nn.Parameter(torch.zeros(c))


class Linear(_Container):
    """Only ever the unused encoder.fc of ResUNet (models/resunet.py:12 keeps torchvision's classifier)."""

    def __init__(self, cin, cout):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(cout, cin))
        self.bias = nn.Parameter(torch.empty(cout))
        _uniform_fan_in_(self.weight, cin)
        _uniform_fan_in_(self.bias, cin)


class Holder(_Container):
    """A name-space node (nn.Sequential / block stand-in) whose children are added with add_module."""

    def __init__(self, **children):
        super().__init__()
        for k, v in children.items():
            self.add_module(k, v)


def seq(*mods):
    """Children named '0', '1', ... like nn.Sequential; None entries keep their index but register nothing."""
    h = Holder()
    for i, m in enumerate(mods):
        if m is not None:
            h.add_module(str(i), m)
    return h


# ---------------------------------------------------------------------------- torchvision ResNet18 as containers
class BasicBlock(_Container):
    """torchvision resnet.py:59-106 registration order: conv1, bn1, conv2, bn2, downsample."""

    def __init__(self, cin, cout, stride, dilation=1):
        super().__init__()
        self.conv1 = Conv2d(cin, cout, 3, stride=stride, padding=dilation, dilation=dilation, bias=False)
        self.bn1 = BatchNorm2d(cout)
        self.conv2 = Conv2d(cout, cout, 3, stride=1, padding=dilation, dilation=dilation, bias=False)
        self.bn2 = BatchNorm2d(cout)
        if stride != 1 or cin != cout:
            self.downsample = seq(Conv2d(cin, cout, 1, stride=stride, bias=False), BatchNorm2d(cout))
        else:
            self.downsample = None


class ResNet18(_Container):
    """conv1, bn1, layer1..4 [, fc] in torchvision's order; kaiming-normal(fan_out, relu) conv init, BN 1/0
    (resnet.py:208-213). with_fc=True keeps the classifier the reference's ResUNet never uses."""

    def __init__(self, with_fc, in_channels=3, output_stride=32):
        super().__init__()
        if output_stride not in (16, 32):
            raise ValueError("output_stride must be 32 or 16 (smp make_dilated on layer4)")
        self.conv1 = Conv2d(in_channels, 64, 7, stride=2, padding=3, bias=False)
        self.bn1 = BatchNorm2d(64)
        cin = 64
        for li, c in enumerate((64, 128, 256, 512), start=1):
            if li == 4 and output_stride == 16:
                # smp ResNetEncoder.make_dilated: every layer4 conv gets stride 1, dilation 2, padding (k // 2) * 2
                blocks = [BasicBlock(cin, c, 1, dilation=2), BasicBlock(c, c, 1, dilation=2)]
            else:
                blocks = [BasicBlock(cin, c, 2 if li > 1 else 1), BasicBlock(c, c, 1)]
            self.add_module(f"layer{li}", seq(*blocks))
            cin = c
        if with_fc:
            self.fc = Linear(512, 1000)
        for m in self.modules():
            if isinstance(m, Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def blocks(self, li):
        layer = getattr(self, f"layer{li}")
        return [getattr(layer, "0"), getattr(layer, "1")]
