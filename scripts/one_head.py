"""The three head kernels on the bench shape (for ncu / quick timing): python scripts/one_head.py [B] [C] [classes] [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200 import ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
Cc = int(sys.argv[2]) if len(sys.argv) > 2 else 16
K = int(sys.argv[3]) if len(sys.argv) > 3 else 1
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 5
x = torch.randn(B, 512, 512, Cc, device="cuda").to(torch.bfloat16)
w = (torch.randn(K, Cc, 3, 3, device="cuda") * 0.1)
b = torch.randn(K, device="cuda")
g = torch.randn(B, K, 512, 512, device="cuda")
dw = torch.zeros_like(w); db = torch.zeros_like(b)
big = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timed(fn):
    tot = 0.0
    fn(); torch.cuda.synchronize()
    for _ in range(iters):
        big.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e3
nx = x.numel() * 2
t1 = timed(lambda: ops.head3x3_fwd(x, w, b))
t2 = timed(lambda: ops.head3x3_dgrad(g, w, tuple(x.shape)))
t3 = timed(lambda: ops.head3x3_wgrad(x, g, dw, db))
print(f"head {B}x512x512x{Cc} -> {K}: fwd {t1:.1f} us ({(nx + g.numel()*4)/t1/1e3:.0f} GB/s)  dgrad {t2:.1f} us ({(nx + g.numel()*4)/t2/1e3:.0f} GB/s)  "
      f"wgrad {t3:.1f} us ({(nx + g.numel()*4)/t3/1e3:.0f} GB/s)")
ops.check_device_flag()
