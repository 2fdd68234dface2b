"""Diagnostic (GPU box): smp-Unet decoder, op by op, product vs torch on the product's own inputs (eval mode)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from oracle import smp_oracle as S, resunet_oracle as R
from oracle.emulate import round_parameters_to_bf16_
from pdac_pathological_image_segmentation_b200 import engine as E, ops
from pdac_pathological_image_segmentation_b200.smp import Unet, run_conv2d_relu

torch.manual_seed(41)
ref = round_parameters_to_bf16_(S.Unet()).eval()
ours = Unet(encoder_weights=None)
ours.load_state_dict(ref.state_dict()); ours.cuda().eval()
img, _ = R.synthetic_batch(2, 64, seed=41)
x = R.normalize(img)
nchw = lambda t: t.permute(0, 3, 1, 2).float().cpu()
def cmp(name, got, want):
    d = (nchw(got) - want).abs().max().item()
    print(f"{name:28s} shape {tuple(want.shape)} max|ref| {want.abs().max():.3f} maxerr {d:.3e}")
with torch.no_grad():
    tape = E.Tape(ours._packs, False, None)
    feats = E.resnet18_features(tape, ours.encoder, x.cuda())
    rf = ref.encoder(x)[1:]
    for n, g, r in zip(["stem", "l1", "l2", "l3", "l4"], feats, rf):
        cmp(n, g, r)
    stem, l1, l2, l3, l4 = feats
    skips = [l3, l2, l1, stem, None]
    y = l4
    for i, skip in enumerate(skips):
        blk = getattr(ours.decoder.blocks, str(i)); rb = ref.decoder.blocks[i]
        yin = y
        y = E.nearest2x(tape, y)
        cmp(f"block{i}.nearest", y, F.interpolate(nchw(yin), scale_factor=2, mode="nearest"))
        srcs = [y, skip] if skip is not None else [y]
        cat = torch.cat([nchw(s) for s in srcs], 1)
        y1 = run_conv2d_relu(tape, blk.conv1, srcs)
        cmp(f"block{i}.conv1", y1, rb.conv1(cat))
        y = run_conv2d_relu(tape, blk.conv2, [y1.clone()])
        cmp(f"block{i}.conv2", y, rb.conv2(nchw(y1)))
    logits, _ = ours._head(tape, y, 1)
    want = ref.segmentation_head(nchw(y))
    print("head maxerr", (logits.cpu() - want).abs().max().item(), "max|ref|", want.abs().max().item())
