"""Overlay blend through the C ABI on a slide band (for ncu / quick timing): python scripts/one_overlay.py H W [tumour_frac] [iters]
Algorithmic bytes per pixel: 1 (mask) + 3 (rgb read) + 3 (written) = 7."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200 import ops
H, W = (int(a) for a in sys.argv[1:3])
frac = float(sys.argv[3]) if len(sys.argv) > 3 else 0.3
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 20
g = torch.Generator(device="cuda").manual_seed(0)
rgb = torch.randint(0, 256, (H, W, 3), dtype=torch.uint8, device="cuda", generator=g)
# tumour regions in 64-pixel runs (masks are piecewise constant at tile scale), frac of the runs set
runs = (torch.rand((H, W // 64), device="cuda", generator=g) < frac).to(torch.uint8)
mask = runs.repeat_interleave(64, dim=1).contiguous()
out = torch.empty_like(rgb)
big = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # L2 flush between timed calls
peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
def timed(fn):
    tot = 0.0
    for _ in range(iters):
        big.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e3
for _ in range(3):
    ops.overlay_blend(rgb, mask, out=out)
nbytes = 7 * H * W
t = timed(lambda: ops.overlay_blend(rgb, mask, out=out))
print(f"overlay {H}x{W} tumour {frac:.2f}: {t:.1f} us  {nbytes / t / 1e3:.0f} GB/s  ({nbytes / t / 1e3 / peak:.2f} of the measured {peak} GB/s copy peak)")
ops.check_device_flag()
