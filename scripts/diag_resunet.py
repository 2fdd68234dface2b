"""Diagnostic (GPU box): per-parameter gradient error of the drop-in ResUNet vs the fp32 oracle, train and eval."""
import sys, os, warnings
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import resunet_oracle as R, util_oracle as U
from pdac_pathological_image_segmentation_b200.resunet import ResUNet
from pdac_pathological_image_segmentation_b200.util import DiceLoss

size = int(sys.argv[1]) if len(sys.argv) > 1 else 64
B = int(sys.argv[2]) if len(sys.argv) > 2 else 2
img, mask = R.synthetic_batch(B, size, seed=41)
x = R.normalize(img)
for train in (True, False):
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m = ResUNet(1, size)
    m.load_state_dict(R.make_state(1, 41)); m.cuda().train(train)
    out = m(x.cuda())
    DiceLoss()(torch.squeeze(out), mask.cuda().float()).backward()
    sd = R.grad_state(R.make_state(1, 41))
    o = R.forward(sd, x, size, training=train, emulate_bf16=(os.environ.get('EMU','1')=='1'))
    U.dice_loss(torch.squeeze(o), mask.float()).backward()
    print(f"train={train} logits max err {(out.cpu()-o).abs().max().item():.3e}  |logits| {o.abs().max().item():.3f}")
    rows = []
    for k, p in m.named_parameters():
        if k.startswith("encoder.fc"): continue
        g, r = p.grad.cpu().double(), sd[k].grad.double()
        rows.append((((g - r).norm() / (r.norm() + 1e-12)).item(), k, r.norm().item()))
    for rel, k, n in reversed(rows):
        if k.endswith("weight"): print(f"   {rel:.4f}  {k}  |ref|={n:.3e}")
    print("   median", sorted(r[0] for r in rows)[len(rows)//2])
