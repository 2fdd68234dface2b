"""Per-layer timing table of one smp-Unet (or ResUNet) training step from the C-ABI call trace (GPU box)."""
import sys, os, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200 import _lib
from pdac_pathological_image_segmentation_b200.smp import Unet
from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
from pdac_pathological_image_segmentation_b200.trainer import TrainStep
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
torch.manual_seed(41)
m = Unet(encoder_weights=None).cuda().train()
ts = TrainStep(m)
img, mask = synthetic_batch(B, 512, seed=41)
img, mask = img.cuda(), mask.cuda()
for _ in range(3): ts.step(img, mask)
torch.cuda.synchronize()
_lib.trace = []
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); ts.step(img, mask); e1.record(); torch.cuda.synchronize()
tr, _lib.trace = _lib.trace, None
agg = collections.OrderedDict()
for name, a, b, fl, nb, tag in tr:
    g = agg.setdefault(tag, [0.0, 0, 0, 0]); g[0] += a.elapsed_time(b); g[1] += fl; g[2] += nb; g[3] += 1
print(f"step {e0.elapsed_time(e1):.2f} ms, traced {sum(v[0] for v in agg.values()):.2f} ms")
print(f"{'layer':48s} {'n':>2s} {'ms':>8s} {'TFLOP/s':>8s} {'GB/s':>8s} {'tc_ms':>7s} {'hbm_ms':>7s}")
for tag, (ms, fl, nb, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{tag:48s} {n:2d} {ms:8.3f} {fl/ms/1e9:8.1f} {nb/ms/1e6:8.1f} {fl/1400.6e9:7.3f} {nb/6531.3e6:7.3f}")
