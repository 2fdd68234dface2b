"""Where conv_halo_kernel's roles wait (diagnostic build: `make -C pdac_pathological_image_segmentation_b200/csrc prof`,
run with PDAC_B200_LIB=.../libpdac_b200_prof.so). python scripts/halo_prof.py CIN COUT HW B [CIN COUT HW B ...]"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200 import ops, _lib

prof_lib = C.CDLL(_lib.LIB_PATH)
args = [int(a) for a in sys.argv[1:]]
dbg = int(os.environ.get("HALO_DBG", "0"))
prof_lib.pdac_halo_dbg_set(dbg)
print("dbg mask", dbg)
for i in range(0, len(args), 4):
    cin, cout, hw, B = args[i:i + 4]
    cins = [cin] if cin <= 512 else [cin // 2, cin // 2]
    xs = [torch.randn(B, hw, hw, c, device="cuda").to(torch.bfloat16) for c in cins]
    w = torch.randn(cout, cin, 3, 3, device="cuda") * 0.05
    wp, _ = ops.pack_conv_weight(w, want_dgrad=False, cout_pad=cout)
    stats = ops.new_stats(cout, torch.device("cuda"))
    out = ops.empty_act(B, hw, hw, cout, "cuda")
    if os.environ.get("HALO_OP") == "pool":      # data gradient CIN -> COUT channels, all of them 2x2-pooled
        wt = torch.randn(cin, cout, 3, 3, device="cuda") * 0.05
        _, wpt = ops.pack_conv_weight(wt, want_fwd=False, stride=1, pad=1, cout_pad=cin)
        run = lambda: ops.conv2d_dgrad_pool2(xs[0], wpt, cout, 0, 3, 3, 1, 1)
    else:
        run = lambda: ops.conv2d_fwd(xs, wp, cout, 3, 3, 1, 1, 1, stats=stats, out=out)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    buf = (C.c_longlong * (160 * 16))()
    prof_lib.pdac_halo_prof_read(buf)
    t = torch.tensor(list(buf), dtype=torch.float64).view(160, 16)[:148]
    m = t.mean(0)
    wall_ns = t[:, 14].max().item() - t[:, 13].min().item()
    print(f"  kernel wall {wall_ns/1e3:.1f} us (globaltimer), CTA cycles mean {m[15]:.0f} max {t[:,15].max():.0f} -> {m[15]/wall_ns:.3f} GHz if one CTA spans the kernel; "
          f"MMA loop max {t[:,0].max():.0f} min {t[:,0].min():.0f}")
    tiles = m[4].item()
    mma_per_tile = None
    print(f"conv3x3 {cin}->{cout} {B}x{hw}x{hw}: {ms:.3f} ms = {ms*1.965e6:.0f} cyc | tiles/CTA {tiles:.1f} | MMA warp total {m[0]:.0f}: "
          f"wait tmem_empty {m[1]:.0f}, full_a {m[2]:.0f}, full_b {m[3]:.0f}, issue+rest {m[0]-m[1]-m[2]-m[3]:.0f} (commits {m[9]:.0f}, syncwarp {m[10]:.0f}, fences {m[11]:.0f}, leader region {m[12]:.0f}) | "
          f"epi total {m[5]:.0f}: wait tmem_full {m[6]:.0f}, work {m[5]-m[6]:.0f} | producers wait empty_a {m[7]:.0f}, empty_b {m[8]:.0f}")
ops.check_device_flag()
