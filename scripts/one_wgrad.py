"""One 3x3 weight gradient through the C ABI: python scripts/one_wgrad.py CIN COUT HW B [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200 import ops
cin, cout, hw, B = (int(a) for a in sys.argv[1:5])
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 3
x = torch.randn(B, hw, hw, cin, device="cuda").to(torch.bfloat16)
dy = torch.randn(B, hw, hw, cout, device="cuda").to(torch.bfloat16)
dw = torch.zeros(cout, cin, 3, 3, device="cuda")
for _ in range(2):
    ops.conv2d_wgrad([x], dy, 3, 3, 1, 1, 1, out=dw, oihw=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    ops.conv2d_wgrad([x], dy, 3, 3, 1, 1, 1, out=dw, oihw=True)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
fl = 2 * B * hw * hw * cout * cin * 9
by = B * hw * hw * (cin + cout) * 2
print(f"wgrad3x3 {cin}->{cout} {B}x{hw}x{hw}: {ms:.3f} ms  {fl/ms/1e9:.1f} TFLOP/s  {by/ms/1e6:.0f} GB/s")
ops.check_device_flag()
