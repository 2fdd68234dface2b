"""One 3x3 convolution through the C ABI (for ncu captures / quick timing): python scripts/one_conv.py CIN COUT HW B [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200 import ops
cin, cout, hw, B = (int(a) for a in sys.argv[1:5])
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 3
cins = [cin] if cin <= 512 else [cin // 2, cin // 2]
xs = [torch.randn(B, hw, hw, c, device="cuda").to(torch.bfloat16) for c in cins]
w = torch.randn(cout, cin, 3, 3, device="cuda") * 0.05
wp, _ = ops.pack_conv_weight(w, want_dgrad=False, cout_pad=cout)
stats = ops.new_stats(cout, torch.device("cuda"))
out = ops.empty_act(B, hw, hw, cout, "cuda")
for _ in range(2):
    ops.conv2d_fwd(xs, wp, cout, 3, 3, 1, 1, 1, stats=stats, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    ops.conv2d_fwd(xs, wp, cout, 3, 3, 1, 1, 1, stats=stats, out=out)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
fl = 2 * B * hw * hw * cout * cin * 9
by = B * hw * hw * (cin + cout) * 2
print(f"conv3x3 {cin}->{cout} {B}x{hw}x{hw}: {ms:.3f} ms  {fl/ms/1e9:.1f} TFLOP/s  {by/ms/1e6:.0f} GB/s")
ops.check_device_flag()
