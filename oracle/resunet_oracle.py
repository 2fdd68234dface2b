"""ORACLE — test infrastructure only (see oracle/util_oracle.py header for who may import this).

Functional fp32 restatement of the reference's only in-tree model, ResUNet (models/resunet.py:7-58), over a plain
state_dict, written against torch.nn.functional so that it depends on neither the reference tree nor torchvision:

  stem / BasicBlock      <- torchvision/models/resnet.py:59-106 (BasicBlock.forward), :197-204 (stem + layers),
                            :239-243 (1x1 stride-2 downsample + BN), as called from models/resunet.py:25-33
  decoder                <- models/resunet.py:36-46 (ConvTranspose2d k2 s2 -> cat(up, skip) -> Conv3x3+bias -> ReLU) x3
  head + resize          <- models/resunet.py:48-56 (Conv1x1, bilinear to output_size, align_corners=False)
  state_dict layout      <- models/resunet.py:8-21 (136 entries incl. the unused encoder.fc.*)

Pinning: oracle/make_golden.py imports the reference's own ResUNet (ImageNet download bypassed, nothing edited),
loads make_state() into it with strict=True and stores its logits / loss / gradient checksums under tests/golden/;
tests/test_oracle_cpu.py replays this restatement against those vectors.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import torch
import torch.nn.functional as F

LAYER_CH = (64, 128, 256, 512)
BN_EPS = 1e-5
BN_MOMENTUM = 0.1


def state_keys(num_classes: int = 1):
    """[(key, shape, dtype)] in the reference's state_dict iteration order."""
    out = []

    def bn(prefix, c):
        out.extend([(f"{prefix}.weight", (c,), torch.float32), (f"{prefix}.bias", (c,), torch.float32),
                    (f"{prefix}.running_mean", (c,), torch.float32), (f"{prefix}.running_var", (c,), torch.float32),
                    (f"{prefix}.num_batches_tracked", (), torch.int64)])

    out.append(("encoder.conv1.weight", (64, 3, 7, 7), torch.float32))
    bn("encoder.bn1", 64)
    cin = 64
    for li, c in enumerate(LAYER_CH, start=1):
        for b in range(2):
            p = f"encoder.layer{li}.{b}"
            out.append((f"{p}.conv1.weight", (c, cin if b == 0 else c, 3, 3), torch.float32))
            bn(f"{p}.bn1", c)
            out.append((f"{p}.conv2.weight", (c, c, 3, 3), torch.float32))
            bn(f"{p}.bn2", c)
            if b == 0 and li > 1:
                out.append((f"{p}.downsample.0.weight", (c, cin, 1, 1), torch.float32))
                bn(f"{p}.downsample.1", c)
        cin = c
    out.append(("encoder.fc.weight", (1000, 512), torch.float32))
    out.append(("encoder.fc.bias", (1000,), torch.float32))
    for i, (ci, co) in enumerate(((512, 256), (256, 128), (128, 64)), start=1):
        out.append((f"upconv{i}.weight", (ci, co, 2, 2), torch.float32))
        out.append((f"upconv{i}.bias", (co,), torch.float32))
        out.append((f"conv{i}.weight", (co, ci, 3, 3), torch.float32))
        out.append((f"conv{i}.bias", (co,), torch.float32))
    out.append(("conv4.weight", (num_classes, 64, 1, 1), torch.float32))
    out.append(("conv4.bias", (num_classes,), torch.float32))
    return out


def make_state(num_classes: int = 1, seed: int = 41, randomize_bn: bool = True) -> "OrderedDict[str, torch.Tensor]":
    """Seeded synthetic state (CPU generator => identical on every box with the same torch). Conv weights follow the
    torchvision rule (kaiming-normal, fan_out; resnet.py:208-213), decoder layers the torch default
    (uniform +-1/sqrt(fan_in)); BN affine / running statistics are perturbed so eval-mode tests are not trivial."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    sd = OrderedDict()
    for key, shape, dtype in state_keys(num_classes):
        leaf = key.rsplit(".", 1)[1]
        if dtype == torch.int64:
            sd[key] = torch.zeros((), dtype=torch.int64)
        elif "bn" in key or "downsample.1" in key:
            c = shape[0]
            if leaf == "weight":
                sd[key] = 1.0 + (0.2 * torch.rand(c, generator=g) - 0.1 if randomize_bn else torch.zeros(c))
            elif leaf == "bias":
                sd[key] = 0.05 * torch.randn(c, generator=g) if randomize_bn else torch.zeros(c)
            elif leaf == "running_mean":
                sd[key] = 0.1 * torch.randn(c, generator=g) if randomize_bn else torch.zeros(c)
            else:
                sd[key] = 0.5 + torch.rand(c, generator=g) if randomize_bn else torch.ones(c)
        elif key.startswith("encoder.") and len(shape) == 4:
            fan_out = shape[0] * shape[2] * shape[3]
            sd[key] = torch.randn(shape, generator=g) * math.sqrt(2.0 / fan_out)
        else:
            if leaf == "weight":          # decoder convs / convT / fc: torch default, uniform +-1/sqrt(fan_in)
                fan_in = shape[1] * 4 if key.startswith("upconv") else int(torch.tensor(shape[1:]).prod())
            bound = 1.0 / math.sqrt(fan_in)   # a bias reuses its sibling weight's fan_in (the weight precedes it)
            sd[key] = (torch.rand(shape, generator=g) * 2 - 1) * bound
    return sd


class _RoundBF16(torch.autograd.Function):
    """x -> bf16 -> fp32 with a straight-through gradient that is itself rounded to bf16: emulates WHERE the product
    stores activations / gradients in bf16 while keeping every multiply-add in fp32 (quantisation-aware oracle)."""

    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).to(torch.float32)


def round_bf16(x):
    return _RoundBF16.apply(x)


def _id(x):
    return x


def _bn(x, sd, p, training):
    if training:
        sd[p + ".num_batches_tracked"] += 1
    return F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"], sd[p + ".weight"], sd[p + ".bias"],
                        training, BN_MOMENTUM, BN_EPS)


def _basic_block(x, sd, p, stride, training, q=_id):
    out = q(F.conv2d(x, q(sd[p + ".conv1.weight"]), None, stride=stride, padding=1))
    out = q(F.relu(_bn(out, sd, p + ".bn1", training)))
    out = q(F.conv2d(out, q(sd[p + ".conv2.weight"]), None, stride=1, padding=1))
    out = _bn(out, sd, p + ".bn2", training)
    if (p + ".downsample.0.weight") in sd:
        x = q(F.conv2d(x, q(sd[p + ".downsample.0.weight"]), None, stride=stride))
        x = q(_bn(x, sd, p + ".downsample.1", training))
    return q(F.relu(out + x))


def encoder_features(x, sd, training, prefix="encoder.", q=_id):
    """-> [stem (64 @ /2), layer1 (/4), layer2 (/8), layer3 (/16), layer4 (/32)]."""
    s = q(F.conv2d(q(x), q(sd[prefix + "conv1.weight"]), None, stride=2, padding=3))
    s = q(F.relu(_bn(s, sd, prefix + "bn1", training)))
    f = F.max_pool2d(s, 3, 2, 1)
    feats = [s]
    for li in range(1, 5):
        for b in range(2):
            stride = 2 if (b == 0 and li > 1) else 1
            f = _basic_block(f, sd, f"{prefix}layer{li}.{b}", stride, training, q)
        feats.append(f)
    return feats


def forward(sd, x, output_size, training=False, emulate_bf16=False):
    """ResUNet.forward (models/resunet.py:23-58). `sd` maps reference keys to tensors (leaf tensors with
    requires_grad for the gradient oracle); BN running statistics in `sd` are updated in place when training.
    emulate_bf16=True rounds weights and every stored activation (and its gradient) to bf16 at the points where the
    CUDA product stores them, with all arithmetic still fp32: the oracle for kernel-level correctness of the
    backward pass (ReLU masks then agree with the product's; the plain fp32 mode is the reference's arithmetic)."""
    q = round_bf16 if emulate_bf16 else _id
    _, x2, x3, x4, x5 = encoder_features(x, sd, training, q=q)
    y = q(F.conv_transpose2d(x5, q(sd["upconv1.weight"]), sd["upconv1.bias"], stride=2))
    y = q(F.relu(F.conv2d(torch.cat((y, x4), 1), q(sd["conv1.weight"]), sd["conv1.bias"], padding=1)))
    y = q(F.conv_transpose2d(y, q(sd["upconv2.weight"]), sd["upconv2.bias"], stride=2))
    y = q(F.relu(F.conv2d(torch.cat((y, x3), 1), q(sd["conv2.weight"]), sd["conv2.bias"], padding=1)))
    y = q(F.conv_transpose2d(y, q(sd["upconv3.weight"]), sd["upconv3.bias"], stride=2))
    y = q(F.relu(F.conv2d(torch.cat((y, x2), 1), q(sd["conv3.weight"]), sd["conv3.bias"], padding=1)))
    y = q(F.conv2d(y, q(sd["conv4.weight"]), sd["conv4.bias"]))
    return F.interpolate(y, size=(output_size, output_size), mode="bilinear", align_corners=False)


def grad_state(sd, device="cpu"):
    """Copy of `sd` on `device` whose float entries are leaves with requires_grad (BN running stats excluded)."""
    out = OrderedDict()
    for k, v in sd.items():
        t = v.detach().clone().to(device)
        if t.dtype == torch.float32 and not (k.endswith("running_mean") or k.endswith("running_var")):
            t.requires_grad_(True)
        out[k] = t
    return out


def synthetic_batch(batch, size, seed=41, blobby=True):
    """Seeded uint8 HWC patches + uint8 {0,1} masks. Returns (img_u8 (B,S,S,3), mask_u8 (B,S,S))."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    img = torch.randint(0, 256, (batch, size, size, 3), dtype=torch.uint8, generator=g)
    if blobby:
        low = torch.rand(batch, 1, max(size // 32, 1), max(size // 32, 1), generator=g)
        mask = (F.interpolate(low, size=(size, size), mode="bilinear", align_corners=False)[:, 0] > 0.6)
    else:
        mask = torch.rand(batch, size, size, generator=g) < 0.5
    return img, mask.to(torch.uint8)


IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def normalize(img_u8):
    """uint8 HWC -> fp32 NCHW, (x/255 - mean)/std (A.Normalize at train_main.py:73 + ToTensorV2)."""
    mean = torch.tensor(IMAGENET_MEAN).view(1, 1, 1, 3)
    std = torch.tensor(IMAGENET_STD).view(1, 1, 1, 3)
    return ((img_u8.float() / 255.0 - mean) / std).permute(0, 3, 1, 2).contiguous()
