"""Diagnostic (GPU box): layer-by-layer forward error of the product vs the bf16-emulating oracle (eval mode)."""
import sys, os, warnings
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from oracle import resunet_oracle as R
from pdac_pathological_image_segmentation_b200 import engine as E, ops
from pdac_pathological_image_segmentation_b200.resunet import ResUNet

size, B = 64, 2
img, _ = R.synthetic_batch(B, size, seed=41)
x = R.normalize(img)
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    m = ResUNet(1, size)
m.load_state_dict(R.make_state(1, 41)); m.cuda().eval()
sd = R.make_state(1, 41)
q = R.round_bf16
with torch.no_grad():
    ref = R.encoder_features(x, sd, False, q=q)
    # walk the product by hand to get the intermediate tensors
    class KeepTape(E.Tape):
        pass
    tape = E.Tape(m._packs, False, None)
    tape.enabled = False
    feats = E.resnet18_features(tape, m.encoder, x.cuda())
    def cmp(name, got_nhwc, r):
        g = got_nhwc.permute(0, 3, 1, 2).float().cpu()
        d = (g - r).abs()
        ulp = (d > 0).float().mean().item()
        print(f"{name:10s} max|ref| {r.abs().max():.3f} maxerr {d.max():.3e} mean err {d.mean():.3e} frac!=0 {ulp:.4f}")
    # NOTE inference mode overwrites conv outputs in place, so compare block outputs only
    for n, g, r in zip(["stem", "l1", "l2", "l3", "l4"], feats, ref):
        cmp(n, g, r)
    # stem conv alone
    z = ops.stem_prep_f32(x.cuda())
    wp = m._packs.stem(m.encoder.conv1)
    s = ops.stem_conv_fwd(z, wp, 64)
    rs = q(F.conv2d(q(x), q(sd["encoder.conv1.weight"]), None, stride=2, padding=3))
    cmp("stemconv", s, rs)
    # maxpool of the reference stem
    rstem = ref[0]
    y, _ = ops.maxpool_fwd(rstem.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda(), want_idx=False)
    cmp("maxpool", y, F.max_pool2d(rstem, 3, 2, 1))
