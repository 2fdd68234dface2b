"""Eager smp-Unet train steps at the bench configuration, for ncu: python scripts/one_step.py [batch] [steps]
(2 warm-up steps + 1 profiled step by default; every step launches the same kernel sequence)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200 import _lib
from pdac_pathological_image_segmentation_b200.smp import Unet
from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
from pdac_pathological_image_segmentation_b200.trainer import TrainStep
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
torch.manual_seed(41)
ts = TrainStep(Unet(encoder_weights=None).cuda().train())
img, mask = (t.cuda() for t in synthetic_batch(B, 512, seed=41))
for i in range(steps):
    l0 = _lib.launches
    loss, _ = ts.step(img, mask)
    torch.cuda.synchronize()
    print(f"step {i}: loss {loss.item():.6f}, {_lib.launches - l0} C-ABI calls", flush=True)
_lib.peek_device_flag()
