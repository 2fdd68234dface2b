"""Generates tests/golden/*.pt by running the REFERENCE's own code (imported from /root/reference, nothing edited, only
the ImageNet weight download at models/resunet.py:12 bypassed) on seeded inputs. Runs only in the build container
(/root/reference does not exist on the GPU box); the fixtures it writes are committed.

    python oracle/make_golden.py            # rewrites tests/golden/

Fixtures
  resunet_nc1.pt / resunet_nc2.pt : train-mode logits, loss, dice score, per-parameter gradient checksums, two full
                                    gradients, BN running stats after the step, eval-mode logits
                                    (input: synthetic_batch(2, 64, seed=41); state: make_state(nc, seed=41))
  util.pt                         : DiceLoss / dice_score on seeded vectors, the sigmoid>=0.5 mask around its fp32
                                    edge, torch.squeeze shapes
  state_keys.pt                   : the reference ResUNet's state_dict keys / shapes / dtypes in iteration order
"""
import os
import struct
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def import_reference():
    import torchvision.models as tvm
    orig = tvm.resnet18
    tvm.resnet18 = lambda weights=None, **kw: orig(weights=None, **kw)      # no network: skip the download only
    sys.path.insert(0, REF)
    from models.resunet import ResUNet          # noqa: E402  (models/resunet.py:7)
    from util import DiceLoss, dice_score       # noqa: E402  (util.py:6,22)
    return ResUNet, DiceLoss, dice_score


def main():
    from oracle import resunet_oracle as R
    torch.manual_seed(0)
    torch.set_num_threads(1)
    ResUNet, DiceLoss, dice_score = import_reference()
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)

    ref_model = ResUNet(num_classes=1, output_size=64)
    torch.save([(k, tuple(v.shape), str(v.dtype)) for k, v in ref_model.state_dict().items()],
               os.path.join(out_dir, "state_keys.pt"))

    img, mask = R.synthetic_batch(2, 64, seed=41)
    x = R.normalize(img)
    for nc in (1, 2):
        m = ResUNet(num_classes=nc, output_size=64)
        m.load_state_dict(R.make_state(nc, seed=41), strict=True)
        m.train()
        out = m(x)
        sq = torch.squeeze(out)
        if nc == 1:
            label = mask
            target = label.type(torch.FloatTensor)
            pred = torch.ge(sq.sigmoid(), 0.5).float()
            score = dice_score(pred, label)
        else:
            target = torch.stack([1 - mask, mask], 1).float()           # one-hot (B,2,H,W), SURVEY 8d
            pred = torch.ge(sq.sigmoid(), 0.5).float()
            score = dice_score(pred, target)
        loss = DiceLoss()(sq, target)
        loss.backward()
        checks = {k: (p.grad.double().sum().item(), p.grad.double().abs().sum().item())
                  for k, p in m.named_parameters() if p.grad is not None}
        no_grad = [k for k, p in m.named_parameters() if p.grad is None]
        sd_after = m.state_dict()
        fix = {
            "logits_train": out.detach().clone(), "loss": loss.detach().clone(), "score": score.detach().clone(),
            "grad_checksums": checks, "params_without_grad": no_grad,
            "grad_conv4_weight": m.conv4.weight.grad.clone(),
            "grad_conv3_weight": m.conv3.weight.grad.clone(),
            "grad_stem_weight": m.encoder.conv1.weight.grad.clone(),
            "bn1_running_mean": sd_after["encoder.bn1.running_mean"].clone(),
            "bn1_running_var": sd_after["encoder.bn1.running_var"].clone(),
            "l4_bn2_running_var": sd_after["encoder.layer4.1.bn2.running_var"].clone(),
            "num_batches_tracked": int(sd_after["encoder.bn1.num_batches_tracked"]),
        }
        m2 = ResUNet(num_classes=nc, output_size=64)
        m2.load_state_dict(R.make_state(nc, seed=41), strict=True)
        m2.eval()
        with torch.no_grad():
            fix["logits_eval"] = m2(x).clone()
        torch.save(fix, os.path.join(out_dir, f"resunet_nc{nc}.pt"))
        print(f"nc={nc}: loss {loss.item():.6f} score {score.item():.6f} |logits| {out.abs().max().item():.4f} "
              f"params without grad: {no_grad}")

    g = torch.Generator().manual_seed(7)
    cases = []
    for n in (1, 17, 4096, 3 * 64 * 64):
        lg = torch.randn(n, generator=g) * 3
        tg = (torch.rand(n, generator=g) < 0.3).float()
        lr = lg.clone().requires_grad_(True)
        ls = DiceLoss()(lr, tg)
        ls.backward()
        pr = torch.ge(lg.sigmoid(), 0.5).float()
        cases.append({"logits": lg, "target": tg, "loss": ls.detach().clone(), "grad": lr.grad.clone(),
                      "score": dice_score(pr, tg.to(torch.uint8)).clone()})
    bits = list(range(0xB43FFFF0, 0xB4400010)) + [0x00000000, 0x80000000, 0x00000001, 0x80000001]
    edge = torch.tensor([struct.unpack("<f", struct.pack("<I", u))[0] for u in bits], dtype=torch.float32)
    util_fix = {
        "dice_cases": cases,
        "edge_bits": bits, "edge_mask_cpu": torch.ge(edge.sigmoid(), 0.5).float(),
        "squeeze_shapes": {(2, 1, 8, 8): tuple(torch.squeeze(torch.zeros(2, 1, 8, 8)).shape),
                           (1, 1, 8, 8): tuple(torch.squeeze(torch.zeros(1, 1, 8, 8)).shape),
                           (2, 2, 8, 8): tuple(torch.squeeze(torch.zeros(2, 2, 8, 8)).shape)},
        "score_empty": dice_score(torch.zeros(10), torch.zeros(10, dtype=torch.uint8)).clone(),
    }
    torch.save(util_fix, os.path.join(out_dir, "util.pt"))
    print("edge mask first-true index:", int(util_fix["edge_mask_cpu"].argmax()))


if __name__ == "__main__":
    main()
