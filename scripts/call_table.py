"""Time of one smp-Unet training step by C-ABI entry point (all calls traced with CUDA events; GPU box)."""
import sys, os, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200 import _lib
from pdac_pathological_image_segmentation_b200.smp import Unet
from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
from pdac_pathological_image_segmentation_b200.trainer import TrainStep
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
torch.manual_seed(41)
_mk = os.environ.get("MODEL", "unet")       # MODEL=fpn|pspnet|deeplabv3+ for the other smp networks
if _mk == "unet":
    m = Unet(encoder_weights=None).cuda().train()
else:
    from pdac_pathological_image_segmentation_b200 import smp as _P
    if _mk == "resunet":
        import warnings
        from pdac_pathological_image_segmentation_b200.resunet import ResUNet
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            m = ResUNet(num_classes=1, output_size=512).cuda().train()
    else:
        _cls = {"fpn": _P.FPN, "pspnet": _P.PSPNet, "deeplabv3+": _P.DeepLabV3Plus}[_mk]
        m = _cls(encoder_weights=None).cuda().train()
ts = TrainStep(m)
img, mask = synthetic_batch(B, 512, seed=41)
img, mask = img.cuda(), mask.cuda()
for _ in range(3): ts.step(img, mask)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); ts.step(img, mask); e1.record(); torch.cuda.synchronize()
plain = e0.elapsed_time(e1)
_lib.trace = []; _lib.trace_all = True
e0.record(); ts.step(img, mask); e1.record(); torch.cuda.synchronize()
tr, _lib.trace = _lib.trace, None
agg = collections.OrderedDict()
for name, a, b, fl, nb, tag in tr:
    g = agg.setdefault(name, [0.0, 0]); g[0] += a.elapsed_time(b); g[1] += 1
tot = sum(v[0] for v in agg.values())
print(f"step {plain:.2f} ms untraced, {e0.elapsed_time(e1):.2f} ms traced; sum of C-ABI calls {tot:.2f} ms")
for name, (ms, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{name:36s} {n:4d} calls {ms:8.3f} ms  {100*ms/tot:5.1f}%")

if len(sys.argv) > 2:  # per-shape rows of the named entry points: python scripts/call_table.py 64 pdac_bn_bwd_apply ...
    for want in sys.argv[2:]:
        rows = collections.OrderedDict()
        for name, a, b, fl, nb, tag in tr:
            if name == want:
                g = rows.setdefault(tag, [0.0, 0]); g[0] += a.elapsed_time(b); g[1] += 1
        print(f"-- {want}")
        for tag, (ms, n) in sorted(rows.items(), key=lambda kv: -kv[1][0]):
            try:
                d = [int(x) for x in tag.split("x")]; el = d[0] * d[1] * d[2] * d[3]
            except Exception:
                el = 0
            print(f"   {tag:22s} {n:3d} calls {ms:7.3f} ms  {ms/n*1e3:7.1f} us/call  {el*2/1e6/(ms/n) if ms else 0:7.1f} GB/s per 2-byte pass")
