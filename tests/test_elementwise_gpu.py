"""Parity of the bandwidth-bound kernels (BatchNorm, pooling, resampling, layout) through the C ABI against torch fp32.

Inputs are bf16-representable; outputs are bf16-rounded once, so the tolerance is one bf16 ulp of the largest value
(2^-8 relative) plus fp32 summation noise.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rnd(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(torch.bfloat16).to(DEV)


def _nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


def _nchw(x):
    return x.permute(0, 3, 1, 2)


def _close(got, ref, what, rel=2.0 ** -7, abs_=1e-6):
    got, ref = got.float(), ref.float()
    assert got.shape == ref.shape, f"{what}: shape {tuple(got.shape)} vs {tuple(ref.shape)}"
    err = (got - ref).abs().max().item()
    tol = rel * ref.abs().max().item() + abs_
    assert err <= tol, f"{what}: max abs err {err:.4g} > tol {tol:.4g}"


@pytest.mark.parametrize("shape", [(2, 64, 16, 16), (3, 128, 8, 8), (2, 16, 32, 32), (4, 512, 4, 4), (2, 48, 8, 8)])
@pytest.mark.parametrize("train", [True, False])
@pytest.mark.parametrize("residual", [False, True])
def test_batchnorm_fwd_bwd(shape, train, residual):
    """conv-output statistics -> bn_prepare -> scale_shift_act(+residual, ReLU) and its two-pass backward, against
    nn.BatchNorm2d (torchvision resnet.py:82-101 dataflow)."""
    from pdac_pathological_image_segmentation_b200 import ops
    N, C, H, W = shape
    x = _rnd(N, C, H, W, seed=1) * 2 + 0.5
    x = x.to(torch.bfloat16)
    res = _rnd(N, C, H, W, seed=2) if residual else None
    dy = _rnd(N, C, H, W, seed=3)
    bn = torch.nn.BatchNorm2d(C).to(DEV)
    with torch.no_grad():
        bn.weight.copy_(torch.rand(C, device=DEV) + 0.5)
        bn.bias.copy_(torch.randn(C, device=DEV) * 0.1)
        bn.running_mean.copy_(torch.randn(C, device=DEV) * 0.1)
        bn.running_var.copy_(torch.rand(C, device=DEV) + 0.5)
    rm, rv = bn.running_mean.clone(), bn.running_var.clone()
    bn.train(train)
    xr = x.float().requires_grad_(True)
    rr = res.float().requires_grad_(True) if residual else None
    z = bn(xr)
    if residual:
        z = z + rr
    yref = z.relu()
    yref.backward(dy.float())

    xh = _nhwc(x)
    stats = torch.stack([xh.float().sum((0, 1, 2)), (xh.float() ** 2).sum((0, 1, 2))]).contiguous()
    ss = torch.empty(2, C, device=DEV)
    mi = torch.empty(2, C, device=DEV)
    ops.bn_prepare(stats, N * H * W, bn.weight.detach(), bn.bias.detach(), bn.eps, bn.momentum, train, rm, rv, ss, mi)
    y = ops.scale_shift_act(xh, ss, _nhwc(res) if residual else None, relu=True)
    _close(_nchw(y), yref, "bn fwd", rel=2.0 ** -7, abs_=2e-2)
    if train:
        assert torch.allclose(rm, bn.running_mean, rtol=1e-4, atol=1e-5)
        assert torch.allclose(rv, bn.running_var, rtol=1e-4, atol=1e-5)
    # backward (mask from OUR y so both sides agree on the ReLU mask up to rounding at 0)
    sums = ops.new_bwd_sums(C, DEV)
    dyh = _nhwc(dy)
    ops.bn_bwd_reduce(dyh, y, xh, mi, sums)
    dgamma = torch.empty(C, device=DEV)
    dbeta = torch.empty(C, device=DEV)
    dz = torch.empty_like(dyh) if residual else None
    dx = ops.bn_bwd_apply(dyh, y, xh, mi, bn.weight.detach(), sums, N * H * W, train, dz_out=dz, dgamma=dgamma,
                          dbeta=dbeta)
    scale = max(xr.grad.abs().max().item(), 1e-3)
    assert (_nchw(dx).float() - xr.grad).abs().max().item() <= 0.03 * scale + 1e-3, "bn dx"
    assert torch.allclose(dgamma, bn.weight.grad, rtol=2e-2, atol=2e-2 * bn.weight.grad.abs().max().item())
    assert torch.allclose(dbeta, bn.bias.grad, rtol=2e-2, atol=2e-2 * bn.bias.grad.abs().max().item())
    if residual:
        assert (_nchw(dz).float() - rr.grad).abs().max().item() <= 0.02 * rr.grad.abs().max().item() + 1e-3


def test_relu_bwd_bias_and_add():
    from pdac_pathological_image_segmentation_b200 import ops
    N, C, H, W = 2, 64, 16, 16
    y = _nhwc(_rnd(N, C, H, W, seed=1)).relu()
    dy = _nhwc(_rnd(N, C, H, W, seed=2))
    dz = torch.empty_like(dy)
    db = torch.zeros(C, device=DEV)
    ops.relu_bwd_bias(dy, y, dz, db)
    ref = dy.float() * (y > 0)
    _close(dz, ref, "relu bwd")
    assert torch.allclose(db, ref.sum((0, 1, 2)), rtol=1e-3, atol=1e-2)
    a, b = _nhwc(_rnd(N, C, H, W, seed=3)), _nhwc(_rnd(N, C, H, W, seed=4))
    _close(ops.add(a, b, relu=True), (a.float() + b.float()).relu(), "add relu")


@pytest.mark.parametrize("shape", [(2, 64, 32, 32), (1, 16, 17, 23), (3, 8, 6, 6)])
def test_maxpool3x3s2(shape):
    """MaxPool2d(3, 2, 1) (torchvision resnet.py:200) forward and backward."""
    from pdac_pathological_image_segmentation_b200 import ops
    N, C, H, W = shape
    # distinct values per window so the arg-max is unique
    g = torch.Generator().manual_seed(5)
    x = torch.randperm(N * C * H * W, generator=g).float().reshape(N, C, H, W)
    x = ((x % 251) - 125).to(torch.bfloat16).to(DEV) + _rnd(N, C, H, W, scale=0.01, seed=1)
    xr = x.float().requires_grad_(True)
    yref = F.max_pool2d(xr, 3, 2, 1)
    dy = _rnd(*yref.shape, seed=2)
    yref.backward(dy.float())
    y, idx = ops.maxpool_fwd(_nhwc(x))
    assert torch.equal(_nchw(y).float(), yref.detach()), "maxpool fwd must be exact"
    dx = ops.maxpool_bwd(_nhwc(dy), idx, (H, W))
    # ties between equal bf16 values may pick another tap: compare where the window max is unique
    _close(_nchw(dx), xr.grad, "maxpool bwd", rel=2.0 ** -6, abs_=1e-3)
    # with an addend (the gradient a skip connection left at the same tensor), here a channel slice of a wider tensor
    wide = _nhwc(_rnd(N, 2 * C, H, W, seed=3))
    add = wide[..., C:]
    dx2 = ops.maxpool_bwd(_nhwc(dy), idx, (H, W), addend=add)
    _close(_nchw(dx2), xr.grad + _nchw(add).float(), "maxpool bwd + addend", rel=2.0 ** -6, abs_=1e-3)


@pytest.mark.parametrize("shape", [(2, 32, 8, 8), (3, 16, 5, 7), (1, 256, 4, 4)])
def test_scale_shift_act_with_fused_nearest2x(shape):
    """The normalise pass that stores its result nearest-neighbour x2 upsampled (smp UnetDecoder's F.interpolate fused
    into the producing layer): bit-identical to the plain pass followed by the upsampling kernel."""
    from pdac_pathological_image_segmentation_b200 import ops
    N, C, H, W = shape
    x = _nhwc(_rnd(N, C, H, W, seed=1))
    ss = torch.stack([torch.rand(C, device=DEV) + 0.5, torch.randn(C, device=DEV) * 0.3])
    want = ops.upsample_nearest2x_fwd(ops.scale_shift_act(x, ss, None, relu=True))
    got = ops.scale_shift_act(x, ss, None, relu=True, up2=True)
    assert got.shape == (N, 2 * H, 2 * W, C) and torch.equal(got, want)


def test_nearest2x():
    from pdac_pathological_image_segmentation_b200 import ops
    x = _rnd(2, 32, 8, 8, seed=1)
    y = ops.upsample_nearest2x_fwd(_nhwc(x))
    assert torch.equal(_nchw(y).float(), F.interpolate(x.float(), scale_factor=2, mode="nearest"))
    dy = _rnd(2, 32, 16, 16, seed=2)
    add = _rnd(2, 32, 8, 8, seed=3)
    ref = F.avg_pool2d(dy.float(), 2) * 4
    _close(_nchw(ops.upsample_nearest2x_bwd(_nhwc(dy))), ref, "nearest bwd")
    _close(_nchw(ops.upsample_nearest2x_bwd(_nhwc(dy), _nhwc(add))), ref + add.float(), "nearest bwd + addend")


@pytest.mark.parametrize("align", [False, True])
@pytest.mark.parametrize("geom", [((8, 8), (32, 32)), ((16, 16), (32, 32)), ((1, 1), (8, 8)), ((6, 6), (16, 16)),
                                  ((3, 2), (16, 16))])
def test_bilinear(align, geom):
    """F.interpolate(mode='bilinear') in both align_corners modes (models/resunet.py:51-56; smp UpsamplingBilinear2d)."""
    from pdac_pathological_image_segmentation_b200 import ops
    (h, w), (ho, wo) = geom
    x = _rnd(2, 16, h, w, seed=1)
    xr = x.float().requires_grad_(True)
    yref = F.interpolate(xr, size=(ho, wo), mode="bilinear", align_corners=align)
    dy = _rnd(2, 16, ho, wo, seed=2)
    yref.backward(dy.float())
    y = ops.bilinear_fwd(_nhwc(x), (ho, wo), align)
    _close(_nchw(y), yref, "bilinear fwd")
    dx = ops.bilinear_bwd(_nhwc(dy), (h, w), align)
    _close(_nchw(dx), xr.grad, "bilinear bwd")


@pytest.mark.parametrize("align", [False, True])
@pytest.mark.parametrize("cu", [1, 2])
def test_bilinear_head(align, cu):
    """NHWC bf16 (padded channels) -> fp32 NCHW logits at output_size, and the gradient back."""
    from pdac_pathological_image_segmentation_b200 import ops
    x = _rnd(2, 16, 16, 16, seed=1)
    xr = x.float()[:, :cu].clone().requires_grad_(True)
    yref = F.interpolate(xr, size=(64, 64), mode="bilinear", align_corners=align)
    g = torch.randn(2, cu, 64, 64, device=DEV)
    yref.backward(g)
    y = ops.bilinear_head_fwd(_nhwc(x), cu, (64, 64), align)
    assert y.dtype == torch.float32 and y.shape == yref.shape
    assert (y - yref).abs().max().item() <= 1e-5 * (1 + yref.abs().max().item())
    dx = ops.bilinear_head_bwd(g, (16, 16, 16), align)
    _close(_nchw(dx)[:, :cu], xr.grad, "head bwd")
    assert torch.count_nonzero(dx[..., cu:]) == 0


@pytest.mark.parametrize("geom", [((64, 64), (1, 1)), ((64, 64), (2, 2)), ((64, 64), (3, 3)), ((64, 64), (6, 6)),
                                  ((32, 32), (1, 1)), ((10, 7), (3, 2))])
def test_adaptive_avgpool(geom):
    """AdaptiveAvgPool2d incl. the unequal overlapping bins of 64 -> 3 / 6 (smp PSPModule)."""
    from pdac_pathological_image_segmentation_b200 import ops
    (h, w), (ho, wo) = geom
    x = _rnd(2, 32, h, w, seed=1)
    xr = x.float().requires_grad_(True)
    yref = F.adaptive_avg_pool2d(xr, (ho, wo))
    dy = _rnd(2, 32, ho, wo, seed=2)
    yref.backward(dy.float())
    _close(_nchw(ops.adaptive_avgpool_fwd(_nhwc(x), (ho, wo))), yref, "adaptive pool fwd", abs_=1e-3)
    _close(_nchw(ops.adaptive_avgpool_bwd(_nhwc(dy), (h, w))), xr.grad, "adaptive pool bwd", abs_=1e-4)


def test_layout_roundtrip_and_stem_prep():
    from pdac_pathological_image_segmentation_b200 import ops
    x = torch.randn(2, 3, 32, 32, device=DEV)
    y = ops.nchw_to_nhwc(x, c_pad=16)
    assert torch.equal(y[..., :3].float(), x.to(torch.bfloat16).float().permute(0, 2, 3, 1))
    assert torch.count_nonzero(y[..., 3:]) == 0
    back = ops.nhwc_to_nchw(y, 3)
    assert torch.equal(back, x.to(torch.bfloat16).float())
    # uint8 HWC patches: normalise (train_main.py:73) + geometric augmentation codes
    img = torch.randint(0, 256, (4, 32, 32, 3), dtype=torch.uint8, device=DEV)
    mask = torch.randint(0, 2, (4, 32, 32), dtype=torch.uint8, device=DEV)
    aug = torch.tensor([0, 1, 2, 3], dtype=torch.int32, device=DEV)
    z, mo = ops.stem_prep_u8(img, mask, aug)
    mean = torch.tensor(ops.IMAGENET_MEAN, device=DEV)
    std = torch.tensor(ops.IMAGENET_STD, device=DEV)
    xn = ((img.float() / 255 - mean) / std).permute(0, 3, 1, 2)      # N,3,H,W
    def geo(t, code):                                                # t: (..., H, W)
        if code == 1:
            return t.flip(-1)
        if code == 2:
            return t.flip(-2)
        if code == 3:
            return torch.rot90(t, 1, (-2, -1))
        return t
    xa = torch.stack([geo(xn[i], int(aug[i])) for i in range(4)])
    ma = torch.stack([geo(mask[i], int(aug[i])) for i in range(4)])
    zref = ops.stem_prep_f32(xa.contiguous())
    assert (z.float() - zref.float()).abs().max().item() <= 2.0 ** -7 * 3
    assert torch.equal(mo, ma)
