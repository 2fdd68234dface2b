"""Time of one WSI inference batch (uint8 tiles -> eval forward -> mask -> stitch) by C-ABI entry point (GPU box)."""
import sys, os, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200 import _lib
from pdac_pathological_image_segmentation_b200.smp import Unet
from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
from pdac_pathological_image_segmentation_b200.wsi import SlideInference
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
m = Unet(encoder_weights=None).cuda().eval()
si = SlideInference(m, 16, 16, 512, 0, 1)
img, _ = synthetic_batch(B, 512, seed=1)
img = img.cuda()
rows = torch.arange(B, device="cuda", dtype=torch.int32) // 16
cols = torch.arange(B, device="cuda", dtype=torch.int32) % 16
import inspect
print(inspect.signature(si.run_batch))
for _ in range(3): si.run_batch(img, rows, cols)
torch.cuda.synchronize()
_lib.trace = []; _lib.trace_all = True
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); si.run_batch(img, rows, cols); e1.record(); torch.cuda.synchronize()
tr, _lib.trace = _lib.trace, None
agg = collections.OrderedDict()
for name, a, b, fl, nb, tag in tr:
    g = agg.setdefault(name, [0.0, 0]); g[0] += a.elapsed_time(b); g[1] += 1
print(f"batch {e0.elapsed_time(e1):.2f} ms traced")
for name, (ms, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{name:36s} {n:4d} calls {ms:8.3f} ms")
