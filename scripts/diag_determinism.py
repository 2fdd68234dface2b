"""Run-to-run spread of the gradient of one smp-Unet train step on identical inputs (fp32 atomic-add order is the only
difference between launches): python scripts/diag_determinism.py [B] [HW]"""
import os, sys, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200.smp import Unet
from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
from pdac_pathological_image_segmentation_b200.trainer import TrainStep
B = int(sys.argv[1]) if len(sys.argv) > 1 else 2
hw = int(sys.argv[2]) if len(sys.argv) > 2 else 64
torch.manual_seed(3)
m0 = Unet(encoder_weights=None).cuda().train()
img, mask = synthetic_batch(B, hw, seed=7)
img, mask = img.cuda(), mask.cuda()
grads = []
for rep in range(4):
    m = copy.deepcopy(m0)
    ts = TrainStep(m, lr=1e-3)
    loss, _ = ts.step(img, mask)
    grads.append(({k: ts.arena.slot(p).clone().double() for k, p in m.named_parameters()}, loss.item()))
ref, l0 = grads[0]
for g, l in grads[1:]:
    num = sum(((g[k] - ref[k]).norm() ** 2).item() for k in ref)
    den = sum((ref[k].norm() ** 2).item() for k in ref)
    worst = max(ref, key=lambda k: ((g[k] - ref[k]).norm() / (ref[k].norm() + 1e-30)).item())
    wv = ((g[worst] - ref[worst]).norm() / (ref[worst].norm() + 1e-30)).item()
    print(f"B={B} hw={hw}: loss {l0:.6f} vs {l:.6f}; whole-gradient rel diff {(num/den)**0.5:.4f}; worst {worst} {wv:.4f}")
