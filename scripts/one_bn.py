"""BatchNorm passes through the C ABI on one tensor shape (for ncu / quick timing): python scripts/one_bn.py C HW B [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200 import ops
C, hw, B = (int(a) for a in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 20
x = torch.randn(B, hw, hw, C, device="cuda").to(torch.bfloat16)
dy = torch.randn(B, hw, hw, C, device="cuda").to(torch.bfloat16)
y = torch.empty_like(x); dx = torch.empty_like(x)
ss = torch.randn(2, C, device="cuda"); mi = torch.rand(2, C, device="cuda") + 0.5
gamma = torch.ones(C, device="cuda"); sums = ops.new_bwd_sums(C, torch.device("cuda"))
dg = torch.zeros(C, device="cuda"); db = torch.zeros(C, device="cuda")
big = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # L2 flush between timed calls
def timed(fn):
    tot = 0.0
    for _ in range(iters):
        big.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e3
n = B * hw * hw * C * 2
t1 = timed(lambda: ops.scale_shift_act(x, ss, None, True, out=y))
t2 = timed(lambda: ops.bn_bwd_reduce(dy, None, x, mi, sums, scale_shift=ss))
t3 = timed(lambda: ops.bn_bwd_apply(dy, None, x, mi, gamma, sums, B * hw * hw, True, dx=dx, dgamma=dg, dbeta=db, scale_shift=ss))
print(f"BN {B}x{hw}x{hw}x{C}: scale_shift_act {t1:.1f} us ({2*n/t1/1e3:.0f} GB/s)  bwd_reduce {t2:.1f} us ({2*n/t2/1e3:.0f} GB/s)  "
      f"bwd_apply {t3:.1f} us ({3*n/t3/1e3:.0f} GB/s)")
ops.check_device_flag()
