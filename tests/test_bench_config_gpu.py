"""Parity at the BENCHMARKED configuration: 512 x 512 patches at real batch sizes, where the halo
planner picks the tiles the bench runs (TW = 128 / four accumulators, streamed 512 -> 512 weights, the two-CTAs-per-SM
variants of the 16 / 32-channel layers at 512^2 and 256^2), the weight-gradient kernels use their full split-K width and
the BatchNorm reductions span 16.8 M values per channel. The oracle is the stock torch.nn restatement (oracle/) run ON THE
GPU in fp32 with TF32 off (tests/conftest.py), plus its bf16-storage emulation for calibration, as in test_smp_gpu.py.
smp-Unet runs at the bench's own batch of 64."""
import copy
import gc

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _free():
    gc.collect()
    torch.cuda.empty_cache()


def _inputs(size, batch):
    from oracle import resunet_oracle as R
    img, mask = R.synthetic_batch(batch, size, seed=41)
    return R.normalize(img).to(DEV), mask.to(DEV)


def _oracle_pass(model, x, t, emulate, loss_kind, rnd):
    """-> dict(logits, loss, dlogits, grads{name: fp64 cpu}, buffers) of one fp32 train-mode step of the oracle module."""
    from oracle import util_oracle as U
    from oracle.emulate import bf16_storage
    from oracle.resunet_oracle import round_bf16
    for p in model.parameters():
        p.grad = None
    if emulate:
        with bf16_storage(model):
            out = model(round_bf16(x))
    else:
        out = model(x)
    out.retain_grad()
    loss = U.dice_loss(torch.squeeze(out), t) if loss_kind == "dice" else (out * rnd).sum() / rnd.numel()
    loss.backward()
    res = {"logits": out.detach().clone(), "loss": loss.item(), "dlogits": out.grad.double().clone(),
           "grads": {k: p.grad.double().cpu() for k, p in model.named_parameters() if p.grad is not None},
           "buffers": {k: v.detach().cpu().clone() for k, v in model.state_dict().items() if "running" in k or "tracked" in k}}
    del out, loss
    for p in model.parameters():
        p.grad = None
    _free()
    return res


def _smp_pair(name):
    from test_smp_gpu import _pair          # tests/ is on sys.path (pytest rootdir insertion)
    return _pair(name)


CASES = [("unet_smp", 64), ("fpn", 8), ("pspnet", 16), ("deeplabv3+", 8)]


@pytest.mark.parametrize("name,batch", CASES)
def test_smp_train_step_parity_at_512(name, batch):
    from pdac_pathological_image_segmentation_b200.util import DiceLoss
    ref, ours = _smp_pair(name)
    x, mask = _inputs(512, batch)
    t = mask.float()
    ref = ref.to(DEV).train()
    emu = copy.deepcopy(ref)
    ours.train()
    a = _oracle_pass(ref, x, t, False, "dice", None)
    e = _oracle_pass(emu, x, t, True, "dice", None)
    del emu
    _free()
    out = ours(x)
    out.retain_grad()
    loss = DiceLoss()(torch.squeeze(out), t)
    loss.backward()
    # logits / loss / loss gradient
    err = (out.detach() - a["logits"]).abs().max().item()
    cal = (e["logits"] - a["logits"]).abs().max().item()
    assert err <= max(2e-2, 1.5 * cal), f"{name}: logits max-abs err {err}, bf16 storage alone {cal}"
    assert abs(loss.item() - a["loss"]) <= max(2e-3, 1.5 * abs(e["loss"] - a["loss"])), (loss.item(), a["loss"], e["loss"])
    gn = a["dlogits"].norm() + 1e-30
    rel_g = ((out.grad.double() - a["dlogits"]).norm() / gn).item()
    cal_g = ((e["dlogits"] - a["dlogits"]).norm() / gn).item()
    assert rel_g <= 1.5 * cal_g + 2e-2, f"{name}: dL/dlogits rel err {rel_g}, bf16 storage alone {cal_g}"
    # per-parameter gradients, calibrated per parameter by the bf16-storage emulation
    g_slack = 2.0 * max(rel_g, cal_g)
    gmax = max(g.norm().item() for g in a["grads"].values())
    bad = []
    for k, p in ours.named_parameters():
        if k not in a["grads"]:
            assert p.grad is None, f"{k}: the reference leaves this parameter without a gradient"
            continue
        assert p.grad is not None, k
        r = a["grads"][k]
        den = r.norm() + 1e-4 * gmax
        if r.numel() == 1 and k.startswith("segmentation_head"):
            den = a["dlogits"].abs().sum().cpu()
        rel = ((p.grad.double().cpu() - r).norm() / den).item()
        calp = ((e["grads"][k] - r).norm() / den).item()
        if not rel <= 2.0 * calp + 2e-2 + g_slack:
            bad.append((k, round(rel, 4), round(calp, 4)))
    assert not bad, f"{name}: gradients farther from fp32 than bf16 storage explains (param, ours, calibrated): {bad}"
    # BatchNorm running statistics and counters
    so = ours.state_dict()
    for k, v in a["buffers"].items():
        if k.endswith("num_batches_tracked"):
            assert int(so[k]) == int(v), k
        else:
            diff = (so[k].cpu() - v).abs()
            assert (diff <= 5e-3 + 5e-2 * v.abs()).all(), (k, diff.max().item())
    from pdac_pathological_image_segmentation_b200 import ops
    ops.check_device_flag()


def test_resunet_train_step_parity_at_512():
    """The reference's own network (models/resunet.py:23-58) at 512 x 512, batch 16: logits within the north_star's
    max-abs 2e-2 of the fp32 oracle, loss, loss gradient, per-parameter gradients, running statistics."""
    import warnings
    from oracle import resunet_oracle as R
    from oracle import util_oracle as U
    from pdac_pathological_image_segmentation_b200.resunet import ResUNet
    from pdac_pathological_image_segmentation_b200.util import DiceLoss
    x, mask = _inputs(512, 16)
    t = mask.float()
    state = R.make_state(1, seed=41)

    def oracle(emulate):
        sd = R.grad_state(state, device=DEV)
        o = R.forward(sd, x, 512, training=True, emulate_bf16=emulate)
        o.retain_grad()
        loss = U.dice_loss(torch.squeeze(o), t)
        loss.backward()
        res = (o.detach().clone(), loss.item(), o.grad.double().clone(),
               {k: v.grad.double().cpu() for k, v in sd.items() if v.requires_grad and v.grad is not None},
               {k: v.detach().cpu() for k, v in sd.items() if "running" in k})
        del sd, o, loss
        _free()
        return res
    a_out, a_loss, a_dl, a_g, a_buf = oracle(False)
    e_out, e_loss, e_dl, e_g, _ = oracle(True)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m = ResUNet(num_classes=1, output_size=512)
    m.load_state_dict(state, strict=True)
    m.to(DEV).train()
    out = m(x)
    out.retain_grad()
    loss = DiceLoss()(torch.squeeze(out), t)
    loss.backward()
    err = (out.detach() - a_out).abs().max().item()
    assert err <= 2e-2, f"logits max-abs err {err} (bf16 storage alone {(e_out - a_out).abs().max().item()})"
    assert abs(loss.item() - a_loss) <= 2e-3
    gn = a_dl.norm() + 1e-30
    rel_g = ((out.grad.double() - a_dl).norm() / gn).item()
    cal_g = ((e_dl - a_dl).norm() / gn).item()
    assert rel_g <= 1.5 * cal_g + 2e-2, (rel_g, cal_g)
    bad = []
    for k, p in m.named_parameters():
        if k.startswith("encoder.fc"):
            assert p.grad is None
            continue
        r = a_g[k]
        rel = ((p.grad.double().cpu() - r).norm() / (r.norm() + 1e-30)).item()
        cal = ((e_g[k] - r).norm() / (r.norm() + 1e-30)).item()
        if not rel <= 1.5 * cal + 2e-2 + 2.0 * max(rel_g, cal_g):
            bad.append((k, round(rel, 4), round(cal, 4)))
    assert not bad, f"gradients farther from fp32 than bf16 storage explains (param, ours, calibrated): {bad}"
    so = m.state_dict()
    for k, v in a_buf.items():
        diff = (so[k].cpu() - v).abs()
        assert (diff <= 5e-3 + 5e-2 * v.abs()).all(), (k, diff.max().item())


# ---- kernel cases at the exact bench shapes that test_conv_gpu.py's small cases do not reach -------------------------
BENCH_CONVS = [  # (N, H, W, [cin...], cout)
    (64, 512, 512, [16], 16),          # decoder block 4 conv2: DUAL variant, TW = 128
    (64, 512, 512, [32], 16),          # decoder block 4 conv1 (nearest-x2 input, no skip)
    (64, 256, 256, [64, 64], 32),      # decoder block 3 conv1: upsampled 64 + stem skip 64
    (64, 16, 16, [512], 512),          # layer4: streamed weights
    (64, 32, 32, [512, 256], 256),     # decoder block 0 conv1: 768 -> 256
]


@pytest.mark.parametrize("case", BENCH_CONVS, ids=lambda c: f"{c[0]}x{c[1]}x{c[2]}_{'+'.join(map(str, c[3]))}to{c[4]}")
def test_conv3x3_fwd_dgrad_wgrad_at_bench_shapes(case):
    """Forward (+ BatchNorm statistics), data gradient and weight gradient of the 3x3 stride-1 convolutions of the smp-Unet
    step at batch 64, against torch fp32 convolutions (cuDNN, TF32 off) on bf16-representable operands."""
    import torch.nn.functional as F
    from pdac_pathological_image_segmentation_b200 import ops
    N, H, W, cins, cout = case
    g = torch.Generator(device=DEV).manual_seed(3)
    xs = [torch.randn(N, H, W, c, device=DEV, generator=g).to(torch.bfloat16) for c in cins]
    ctot = sum(cins)
    w = (torch.randn(cout, ctot, 3, 3, device=DEV, generator=g) * (2.0 / (9 * ctot)) ** 0.5).to(torch.bfloat16).float()
    dy = torch.randn(N, H, W, cout, device=DEV, generator=g).to(torch.bfloat16)
    wp_f, wp_t = ops.pack_conv_weight(w.contiguous())
    stats = ops.new_stats(cout, xs[0].device)
    out = ops.conv2d_fwd(xs, wp_f, cout, 3, 3, 1, 1, 1, stats=stats)
    dx = ops.conv2d_dgrad(dy, wp_t, ctot, (H, W), 3, 3, 1, 1, 1)
    dw = ops.conv2d_wgrad(xs, dy, 3, 3, 1, 1, 1, oihw=True)
    ops.check_device_flag()
    # torch fp32 reference in image chunks (the fp32 NCHW copies of a 64 x 512 x 512 tensor are GBs)
    step = max(1, min(N, (1 << 28) // (H * W * max(ctot, cout))))
    dw_ref = torch.zeros_like(w, dtype=torch.float64)
    s_ref = torch.zeros(cout, dtype=torch.float64, device=DEV)
    q_ref = torch.zeros(cout, dtype=torch.float64, device=DEV)
    worst_f = worst_d = 0.0
    for n0 in range(0, N, step):
        sl = slice(n0, n0 + step)
        xc = torch.cat([x[sl] for x in xs], 3).permute(0, 3, 1, 2).float().requires_grad_(True)
        wr = w.clone().requires_grad_(True)
        ref = F.conv2d(xc, wr, None, 1, 1)
        ref.backward(dy[sl].permute(0, 3, 1, 2).float())
        o = out[sl].permute(0, 3, 1, 2).float()
        tol = 2.0 ** -7 * ref.abs().clamp_min(1.0)
        worst_f = max(worst_f, ((o - ref.detach()).abs() / tol).max().item())
        d = dx[sl].permute(0, 3, 1, 2).float()
        told = 2.0 ** -7 * xc.grad.abs().clamp_min(1.0)
        worst_d = max(worst_d, ((d - xc.grad).abs() / told).max().item())
        dw_ref += wr.grad.double()
        s_ref += out[sl].double().sum((0, 1, 2))
        q_ref += (out[sl].double() ** 2).sum((0, 1, 2))
        del xc, wr, ref, o, d
    assert worst_f <= 1.0, f"fwd: {worst_f} x the one-bf16-rounding bound"
    assert worst_d <= 1.0, f"dgrad: {worst_d} x the one-bf16-rounding bound"
    rel_w = ((dw.double() - dw_ref).norm() / dw_ref.norm()).item()
    assert rel_w <= 1e-3, f"wgrad rel L2 {rel_w}"
    st = stats.double().sum(0)
    assert torch.allclose(st[0], s_ref, rtol=1e-4, atol=1e-2 * q_ref.sqrt().max().item())
    assert torch.allclose(st[1], q_ref, rtol=1e-4)
