"""Kernel-level parity of the layers smp FPN / PSPNet / DeepLabV3+ add to the ResNet18 path (norm_dw.cu), against torch
fp32 on bf16-representable inputs: GroupNorm(32, C) forward / backward (Conv3x3GNReLU), depthwise 3x3 forward / data
gradient / weight gradient at dilations 1, 3, 6, 9 and the channel counts DeepLabV3+ uses (304, 512, 256), and the
counter-based Dropout / Dropout2d masks (keep rate, 1/(1-p) scaling, identical mask in forward and backward)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return (torch.randn(*shape, device=DEV, generator=g) * scale).to(torch.bfloat16)


def _nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


def _nchw(x):
    return x.permute(0, 3, 1, 2)


@pytest.mark.parametrize("shape,groups", [((2, 128, 32, 32), 32), ((3, 128, 20, 12), 32), ((2, 64, 16, 16), 32),
                                          ((2, 256, 8, 8), 32), ((1, 128, 128, 128), 32)])
@pytest.mark.parametrize("relu", [True, False])
def test_groupnorm_fwd_bwd(shape, groups, relu):
    from pdac_pathological_image_segmentation_b200 import ops
    N, C, H, W = shape
    x = _rnd(N, C, H, W, seed=1) * 1.5 + 0.3
    dy = _rnd(N, C, H, W, seed=2)
    gn = torch.nn.GroupNorm(groups, C).to(DEV)
    with torch.no_grad():
        gn.weight.copy_(torch.rand(C, device=DEV) + 0.5)
        gn.bias.copy_(torch.randn(C, device=DEV) * 0.2)
    xr = x.float().requires_grad_(True)
    ref = gn(xr)
    if relu:
        ref = ref.relu()
    ref.backward(dy.float())
    xh, dyh = _nhwc(x), _nhwc(dy)
    y, stats = ops.groupnorm_fwd(xh, groups, gn.weight.detach(), gn.bias.detach(), gn.eps, relu=relu)
    err = (_nchw(y).float() - ref).abs().max().item()
    assert err <= 2.0 ** -7 * max(ref.abs().max().item(), 1.0), f"gn fwd err {err}"
    dgamma = torch.zeros(C, device=DEV)
    dbeta = torch.zeros(C, device=DEV)
    # mask from OUR y (both sides then agree on the ReLU decision except at outputs that round to exactly 0)
    dx = ops.groupnorm_bwd(dyh, y if relu else None, xh, groups, stats, gn.weight.detach(), gn.eps, dgamma=dgamma,
                           dbeta=dbeta)
    scale = xr.grad.abs().max().item()
    assert (_nchw(dx).float() - xr.grad).abs().max().item() <= 0.03 * scale + 1e-3, "gn dx"
    assert torch.allclose(dgamma, gn.weight.grad, rtol=2e-2, atol=2e-2 * gn.weight.grad.abs().max().item())
    assert torch.allclose(dbeta, gn.bias.grad, rtol=2e-2, atol=2e-2 * gn.bias.grad.abs().max().item())
    ops.check_device_flag()


@pytest.mark.parametrize("shape", [(2, 304, 128, 128), (2, 512, 32, 32), (2, 256, 32, 32), (3, 48, 17, 23),
                                   (1, 16, 16, 16)])
@pytest.mark.parametrize("dil", [1, 3, 6, 9])
def test_depthwise3x3_fwd_dgrad_wgrad(shape, dil):
    from pdac_pathological_image_segmentation_b200 import ops
    N, C, H, W = shape
    x = _rnd(N, C, H, W, seed=1)
    w = _rnd(C, 1, 3, 3, seed=2, scale=0.3).float().contiguous()          # bf16-representable fp32 state_dict tensor
    dy = _rnd(N, C, H, W, seed=3)
    xr = x.float().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    ref = F.conv2d(xr, wr, None, 1, dil, dil, groups=C)
    ref.backward(dy.float())
    xh, dyh = _nhwc(x), _nhwc(dy)
    y = ops.dwconv3x3(xh, w, dil)
    err = (_nchw(y).float() - ref).abs().max().item()
    assert err <= 2.0 ** -7 * max(ref.abs().max().item(), 1.0), f"dw fwd err {err}"
    dx = ops.dwconv3x3(dyh, w, dil, flip=True)
    errd = (_nchw(dx).float() - xr.grad).abs().max().item()
    assert errd <= 2.0 ** -7 * max(xr.grad.abs().max().item(), 1.0), f"dw dgrad err {errd}"
    dw = torch.zeros_like(w)
    ops.dwconv3x3_wgrad(xh, dyh, dil, dw)
    errw = (dw - wr.grad).abs().max().item()
    assert errw <= 1e-3 * wr.grad.abs().max().item() + 1e-3, f"dw wgrad err {errw}"
    ops.check_device_flag()


@pytest.mark.parametrize("per_channel", [True, False])
@pytest.mark.parametrize("p", [0.2, 0.5])
def test_dropout_mask_statistics_and_replay(per_channel, p):
    """nn.Dropout2d (whole channels of an image) / nn.Dropout (elements): keep rate ~ 1 - p, kept values scaled by
    1/(1-p), and the backward call with the same seed applies the SAME mask to the gradient."""
    from pdac_pathological_image_segmentation_b200 import ops
    N, C, H, W = 16, 256, 24, 24
    x = (torch.rand(N, H, W, C, device=DEV) + 0.5).to(torch.bfloat16)
    g = (torch.rand(N, H, W, C, device=DEV) + 0.5).to(torch.bfloat16)
    y = ops.dropout(x, p, 1234, per_channel)
    kept = y != 0
    ratio = (y.float() / x.float())[kept]
    assert (ratio - 1.0 / (1.0 - p)).abs().max().item() <= 2.0 ** -7 / (1.0 - p)
    if per_channel:
        per = kept.float().mean((1, 2))                                # (N, C): every channel map is all-or-nothing
        assert ((per == 0) | (per == 1)).all()
        rate, n = per.mean().item(), N * C
    else:
        rate, n = kept.float().mean().item(), x.numel()
    assert abs(rate - (1 - p)) <= 5 * (p * (1 - p) / n) ** 0.5 + 1e-3, (rate, 1 - p)
    dg = ops.dropout(g, p, 1234, per_channel)
    assert torch.equal(dg != 0, kept)                                  # same mask in backward
    y2 = ops.dropout(x, p, 99, per_channel)
    assert not torch.equal(y2 != 0, kept)                              # another seed, another mask
    assert torch.equal(ops.dropout(x, 0.0, 5, per_channel), x)
