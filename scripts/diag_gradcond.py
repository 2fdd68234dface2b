"""How well-conditioned is a whole-network gradient comparison against fp32? Prints, per network, the relative L2
distance of the product's parameter gradients from the fp32 oracle (GPU, TF32 off), of the bf16-storage emulation from
the fp32 oracle, and of the product from the emulation, for (a) the Dice loss, (b) a linear loss sum(logits * R) and
(c) the linear loss with BatchNorm shifts moved off the ReLU kink (beta += 3).  python scripts/diag_gradcond.py [size] [batch]"""
import copy, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
from oracle import resunet_oracle as R
from oracle import smp_oracle as S
from oracle import util_oracle as U
from oracle.emulate import bf16_storage, round_parameters_to_bf16_
from pdac_pathological_image_segmentation_b200 import smp as P
from pdac_pathological_image_segmentation_b200.util import DiceLoss

size = int(sys.argv[1]) if len(sys.argv) > 1 else 64
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 2
DEV = "cuda"
img, mask = R.synthetic_batch(batch, size, seed=41)
x = R.normalize(img).to(DEV)
t = mask.float().to(DEV)
Rnd = torch.randn(batch, 1, size, size, generator=torch.Generator().manual_seed(9)).to(DEV)


def grads(model, loss_kind, emulate=False):
    for p in model.parameters():
        p.grad = None
    if emulate:
        with bf16_storage(model):
            out = model(R.round_bf16(x))
    else:
        out = model(x)
    if loss_kind == "dice":
        loss = (DiceLoss() if hasattr(model, "_run") else U.dice_loss)(torch.squeeze(out), t)
    else:
        loss = (out * Rnd).sum() / Rnd.numel()
    loss.backward()
    return {k: p.grad.double().clone() for k, p in model.named_parameters() if p.grad is not None}


def rel(a, b):
    num = sum(((a[k] - b[k]).norm() ** 2).item() for k in a)
    den = sum((b[k].norm() ** 2).item() for k in a)
    worst = max(((a[k] - b[k]).norm() / (b[k].norm() + 1e-30)).item() for k in a)
    return (num / den) ** 0.5, worst


for shift in (0.0, 3.0):
    torch.manual_seed(41)
    ref = S.build("unet_smp", dropout=0.0)
    with torch.no_grad():
        for m in ref.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.bias.add_(shift)
    round_parameters_to_bf16_(ref)
    ours = P.Unet(encoder_weights=None)
    ours.load_state_dict(ref.state_dict(), strict=True)
    ref = ref.to(DEV).train()
    ours = ours.to(DEV).train()
    emu = copy.deepcopy(ref)
    for kind in ("dice", "linear"):
        g0 = grads(ref, kind)
        ge = grads(emu, kind, emulate=True)
        gp = grads(ours, kind)
        gp2 = grads(ours, kind)
        same = all(torch.equal(gp[k], gp2[k]) for k in gp)
        print(f"unet {size}x{size} B={batch} bn_shift={shift} loss={kind}: ours-vs-fp32 {rel(gp, g0)}  emu-vs-fp32 {rel(ge, g0)}  "
              f"ours-vs-emu {rel(gp, ge)}  run-to-run identical: {same}", flush=True)
