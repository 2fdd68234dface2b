"""Parity of the tcgen05 implicit-GEMM convolutions (through the C ABI) against torch fp32 on bf16-rounded inputs.

Tolerance: inputs/weights are exactly representable in bf16, accumulation is fp32, so the only error is the final
bf16 rounding of the output (relative 2^-8) plus fp32 summation-order noise.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _rnd(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(torch.bfloat16).to(DEV)


def _nhwc(x_nchw_bf16):
    return x_nchw_bf16.permute(0, 2, 3, 1).contiguous()


def _close(got, ref, what, rel=2.0 ** -7):
    got, ref = got.float(), ref.float()
    err = (got - ref).abs().max().item()
    tol = rel * ref.abs().max().item() + 1e-6
    assert err <= tol, f"{what}: max abs err {err:.4g} > tol {tol:.4g} (ref max {ref.abs().max().item():.4g})"


FWD_CASES = [
    # (N, H, W, [cins], cout, R, stride, pad, dil, relu, bias, addend)
    (2, 16, 16, [64], 64, 3, 1, 1, 1, False, False, False),
    (2, 32, 32, [32], 16, 3, 1, 1, 1, False, False, False),
    (1, 32, 32, [16], 16, 3, 1, 1, 1, True, False, False),
    (2, 16, 16, [128], 128, 3, 1, 1, 1, True, True, False),
    (2, 16, 16, [64, 64], 64, 3, 1, 1, 1, False, True, True),
    (2, 16, 16, [128, 64], 64, 3, 1, 1, 1, False, False, False),
    (2, 16, 16, [64], 256, 1, 1, 0, 1, False, False, False),
    (2, 32, 32, [64], 128, 3, 2, 1, 1, False, False, False),
    (2, 32, 32, [64], 128, 1, 2, 0, 1, False, False, False),
    (2, 16, 16, [64], 64, 3, 1, 2, 2, False, False, False),
    (3, 10, 12, [64], 32, 3, 1, 1, 1, False, False, False),
    (4, 8, 8, [128], 64, 3, 1, 1, 1, False, False, False),
    (2, 2, 2, [128], 32, 1, 1, 0, 1, True, True, False),
    (2, 3, 3, [128], 32, 1, 1, 0, 1, False, False, False),
    (2, 16, 16, [16], 1, 3, 1, 1, 1, False, True, False),
    (2, 16, 16, [304], 256, 1, 1, 0, 1, False, False, False),
    (1, 16, 16, [512], 512, 3, 1, 1, 1, False, False, False),
    (2, 16, 16, [64], 48, 1, 1, 0, 1, False, False, False),
    # halo-tile kernel (conv_halo.cu): many tiles per image, partial tiles, every swizzle width, resident and
    # streamed weights, concatenated sources, several N tiles, dilation, fused addend/bias/ReLU
    (2, 96, 80, [16], 16, 3, 1, 1, 1, True, True, False),
    (1, 128, 128, [32], 16, 3, 1, 1, 1, False, False, True),
    (2, 70, 66, [64], 64, 3, 1, 1, 1, False, False, False),
    (1, 64, 64, [64, 64], 32, 3, 1, 1, 1, True, False, False),
    (1, 64, 64, [32], 128, 3, 1, 1, 1, False, False, True),
    (1, 48, 48, [128], 128, 3, 1, 1, 1, False, True, False),
    (1, 32, 32, [256, 128], 128, 3, 1, 1, 1, True, False, False),
    (1, 32, 32, [64], 192, 3, 1, 1, 1, False, False, False),
    (1, 32, 32, [256], 256, 3, 1, 1, 1, False, False, True),
    (1, 40, 40, [128], 64, 3, 1, 3, 3, False, False, False),
    (1, 33, 47, [512], 256, 3, 1, 2, 2, False, False, False),
    (3, 256, 256, [16], 16, 3, 1, 1, 1, False, False, False),
]


@pytest.mark.parametrize("case", FWD_CASES, ids=[str(c) for c in FWD_CASES])
def test_conv_fwd(case):
    from pdac_pathological_image_segmentation_b200 import ops
    N, H, W, cins, cout, R, stride, pad, dil, relu, use_bias, use_add = case
    xs = [_rnd(N, c, H, W, seed=i + 1) for i, c in enumerate(cins)]
    w = _rnd(cout, sum(cins), R, R, scale=0.05, seed=11)
    bias = (torch.randn(cout, device=DEV) if use_bias else None)
    ref = F.conv2d(torch.cat(xs, 1).float(), w.float(), bias, stride=stride, padding=pad, dilation=dil)
    addend = None
    if use_add:
        addend = _rnd(*ref.shape, seed=21)
        ref = ref + addend.float()
    if relu:
        ref = ref.relu()
    cpad = ops.pad16(cout) if cout % 8 else cout
    wp, _ = ops.pack_conv_weight(w.float().contiguous(), want_dgrad=False, cout_pad=cpad)
    stats = ops.new_stats(cpad, DEV)
    bias_p = None
    if bias is not None:
        bias_p = torch.zeros(cpad, device=DEV)
        bias_p[:cout] = bias
    out = ops.conv2d_fwd([_nhwc(x) for x in xs], wp, cpad, R, R, stride, pad, dil, relu=relu, bias=bias_p,
                         addend=_nhwc(addend) if addend is not None else None, stats=stats)
    ops.check_device_flag()
    got = out.permute(0, 3, 1, 2)[:, :cout]
    _close(got, ref, "conv fwd")
    # BN statistics of the stored values
    o32 = out.float()[..., :cout]
    s_ref = o32.sum((0, 1, 2))
    q_ref = (o32 * o32).sum((0, 1, 2))
    stats = stats.sum(0)
    assert torch.allclose(stats[0, :cout], s_ref, rtol=1e-3, atol=1e-2 * o32.abs().max().item() + 1e-3)
    assert torch.allclose(stats[1, :cout], q_ref, rtol=1e-3, atol=1e-3)


DGRAD_CASES = [
    # (N, H, W, cin, cout, R, stride, pad, dil, addend)
    (2, 16, 16, 64, 64, 3, 1, 1, 1, False),
    (2, 16, 16, 128, 64, 3, 1, 1, 1, True),
    (2, 32, 32, 64, 128, 3, 2, 1, 1, False),
    (2, 32, 32, 64, 128, 3, 2, 1, 1, True),
    (2, 32, 32, 64, 128, 1, 2, 0, 1, False),
    (2, 16, 16, 64, 64, 3, 1, 2, 2, False),
    (2, 16, 16, 32, 16, 3, 1, 1, 1, False),
    (2, 16, 16, 16, 1, 3, 1, 1, 1, False),
    (2, 16, 16, 256, 64, 1, 1, 0, 1, False),
]


@pytest.mark.parametrize("case", DGRAD_CASES, ids=[str(c) for c in DGRAD_CASES])
def test_conv_dgrad(case):
    from pdac_pathological_image_segmentation_b200 import ops
    N, H, W, cin, cout, R, stride, pad, dil, use_add = case
    w = _rnd(cout, cin, R, R, scale=0.05, seed=3)
    ho, wo = ops.conv_out_hw(H, W, R, R, stride, pad, dil)
    dy = _rnd(N, cout, ho, wo, seed=4)
    ref = torch.nn.grad.conv2d_input((N, cin, H, W), w.float(), dy.float(), stride=stride, padding=pad, dilation=dil)
    addend = None
    if use_add:
        addend = _rnd(N, cin, H, W, seed=5)
        ref = ref + addend.float()
    cpad = ops.pad16(cout) if cout % 8 else cout
    _, wpt = ops.pack_conv_weight(w.float().contiguous(), want_fwd=False, stride=stride, pad=pad, cout_pad=cpad)
    dyp = torch.zeros(N, ho, wo, cpad, dtype=torch.bfloat16, device=DEV)
    dyp[..., :cout] = _nhwc(dy)
    dx = ops.conv2d_dgrad(dyp, wpt, cin, (H, W), R, R, stride, pad, dil,
                          addend=_nhwc(addend) if addend is not None else None)
    ops.check_device_flag()
    _close(dx.permute(0, 3, 1, 2), ref, "conv dgrad")


WGRAD_CASES = [
    # (N, H, W, [cins], cout, R, stride, pad, dil)
    (2, 16, 16, [64], 64, 3, 1, 1, 1),
    (2, 32, 32, [32], 16, 3, 1, 1, 1),
    (2, 32, 32, [16], 16, 3, 1, 1, 1),
    (2, 16, 16, [128], 128, 3, 1, 1, 1),
    (2, 16, 16, [256], 256, 3, 1, 1, 1),
    (1, 16, 16, [64], 512, 3, 1, 1, 1),
    (2, 16, 16, [128, 64], 64, 3, 1, 1, 1),
    (2, 32, 32, [64], 128, 3, 2, 1, 1),
    (2, 32, 32, [64], 128, 1, 2, 0, 1),
    (2, 16, 16, [16], 1, 3, 1, 1, 1),
    (3, 10, 12, [64], 32, 3, 1, 1, 1),
    (2, 16, 16, [64], 48, 1, 1, 0, 1),
    (4, 64, 64, [64], 64, 3, 1, 1, 1),
    # halo-tile weight gradient (wgrad_halo.cu): partial tiles, every slab composition (chunk pairs, tap pairs,
    # kernel-row runs for 16/32-channel tensors), several slab groups per CTA grid, dilation
    (2, 40, 50, [16], 16, 3, 1, 1, 1),
    (1, 128, 128, [32], 16, 3, 1, 1, 1),
    (2, 24, 32, [32, 32], 32, 3, 1, 1, 1),
    (1, 32, 32, [64, 64], 32, 3, 1, 1, 1),
    (1, 32, 48, [192], 64, 3, 1, 1, 1),
    (1, 32, 32, [384], 128, 3, 1, 1, 1),
    (2, 16, 16, [128], 128, 3, 1, 2, 2),
    (1, 24, 24, [16], 32, 3, 1, 3, 3),
    (1, 32, 32, [48], 16, 3, 1, 1, 1),
]


@pytest.mark.parametrize("case", WGRAD_CASES, ids=[str(c) for c in WGRAD_CASES])
def test_conv_wgrad(case):
    from pdac_pathological_image_segmentation_b200 import ops
    N, H, W, cins, cout, R, stride, pad, dil = case
    xs = [_rnd(N, c, H, W, seed=i + 1) for i, c in enumerate(cins)]
    ho, wo = ops.conv_out_hw(H, W, R, R, stride, pad, dil)
    dy = _rnd(N, cout, ho, wo, seed=7)
    ref = torch.nn.grad.conv2d_weight(torch.cat(xs, 1).float(), (cout, sum(cins), R, R), dy.float(), stride=stride,
                                      padding=pad, dilation=dil)
    cpad = ops.pad16(cout) if cout % 16 else cout
    dyp = torch.zeros(N, ho, wo, cpad, dtype=torch.bfloat16, device=DEV)
    dyp[..., :cout] = _nhwc(dy)
    dw = ops.conv2d_wgrad([_nhwc(x) for x in xs], dyp, R, R, stride, pad, dil)   # [RS, cpad, Ctot]
    ops.check_device_flag()
    got = dw.view(R, R, cpad, sum(cins)).permute(2, 3, 0, 1)[:cout]
    err = (got - ref).abs().max().item()
    tol = 1e-3 * ref.abs().max().item() + 1e-4
    assert err <= tol, f"wgrad: max abs err {err:.4g} > {tol:.4g}"


def test_convt2x2_all():
    from pdac_pathological_image_segmentation_b200 import ops
    N, H, W, cin, cout = 2, 16, 16, 128, 64
    x = _rnd(N, cin, H, W, seed=1)
    w = _rnd(cin, cout, 2, 2, scale=0.05, seed=2)
    bias = torch.randn(cout, device=DEV)
    ref = F.conv_transpose2d(x.float(), w.float(), bias, stride=2)
    wp_f, wp_t = ops.pack_convt_weight(w.float().contiguous())
    out = ops.convt2x2_fwd(_nhwc(x), wp_f, cout, bias=bias)
    ops.check_device_flag()
    _close(out.permute(0, 3, 1, 2), ref, "convT fwd")
    dy = _rnd(N, cout, 2 * H, 2 * W, seed=3)
    xr = x.float().requires_grad_(True)
    wr = w.float().requires_grad_(True)
    F.conv_transpose2d(xr, wr, None, stride=2).backward(dy.float())
    dx = ops.convt2x2_dgrad(_nhwc(dy), wp_t, cin)
    ops.check_device_flag()
    _close(dx.permute(0, 3, 1, 2), xr.grad, "convT dgrad")
    dw = ops.convt2x2_wgrad(_nhwc(x), _nhwc(dy))      # [4, Cout, Cin]
    ops.check_device_flag()
    got = dw.view(2, 2, cout, cin).permute(3, 2, 0, 1)
    err = (got - wr.grad).abs().max().item()
    assert err <= 1e-3 * wr.grad.abs().max().item() + 1e-4, f"convT wgrad err {err}"


def test_stem_as_4x4_s2d():
    """7x7 stride-2 stem == 4x4 stride-1 conv over the space-to-depth input (layout.cu packing)."""
    from pdac_pathological_image_segmentation_b200 import ops
    N, H, W = 2, 64, 64
    x = _rnd(N, 3, H, W, seed=1)
    w = _rnd(64, 3, 7, 7, scale=0.05, seed=2)
    ref = F.conv2d(x.float(), w.float(), None, stride=2, padding=3)
    z = ops.stem_prep_f32(x.float())
    wp = ops.pack_stem_weight(w.float().contiguous())
    out = ops.stem_conv_fwd(z, wp, 64)
    ops.check_device_flag()
    _close(out.permute(0, 3, 1, 2), ref, "stem fwd")
    dy = _rnd(N, 64, H // 2, W // 2, seed=3)
    ref_dw = torch.nn.grad.conv2d_weight(x.float(), (64, 3, 7, 7), dy.float(), stride=2, padding=3)
    dwp = ops.stem_conv_wgrad(z, _nhwc(dy))
    dw = ops.unpack_stem_wgrad(dwp, 64)
    ops.check_device_flag()
    err = (dw - ref_dw).abs().max().item()
    assert err <= 1e-3 * ref_dw.abs().max().item() + 1e-4, f"stem wgrad err {err}"


# ------------------------------------------------------------------ data gradient with the nearest x2 backward fused
POOL_CASES = [
    # N, H, W, cout (dy channels), c_pool, c_rest, dil
    (2, 64, 64, 16, 32, 0, 1),        # last Unet block: no skip, two CTAs per SM
    (2, 32, 32, 32, 64, 64, 1),       # 64 up + 64 skip in one 128-column tile
    (2, 16, 16, 64, 128, 64, 1),      # pool boundary at a tile boundary (128 | 64)
    (1, 24, 40, 128, 256, 128, 1),    # three N tiles, sizes that are no multiple of the tile
    (2, 8, 8, 256, 512, 256, 1),      # streamed weights
    (1, 36, 28, 16, 16, 16, 1),       # pool boundary inside a 32-column pass (16-channel dy: the two-kernel form's
                                      # data gradient is not tap-stacked, which would change its summation order)
    (1, 20, 12, 16, 24, 8, 1),        # 8-channel granularity
    (1, 32, 32, 32, 32, 32, 2),       # dilated 3x3
    (1, 6, 130, 16, 16, 0, 1),        # wider than every tile, lower than one accumulator
]


@pytest.mark.parametrize("case", POOL_CASES, ids=[str(c) for c in POOL_CASES])
def test_conv_dgrad_pool2_is_bit_identical_to_the_two_kernel_form(case):
    """pdac_conv2d_dgrad_pool2 == pdac_conv2d_dgrad + pdac_upsample_nearest2x_bwd on the upsampled channels (smp
    UnetDecoder: interpolate(nearest, x2) -> cat -> conv), bit for bit; and both match torch's fp32 gradient."""
    from pdac_pathological_image_segmentation_b200 import ops
    N, H, W, cout, c_pool, c_rest, dil = case
    cin = c_pool + c_rest
    w = _rnd(cout, cin, 3, 3, scale=0.05, seed=13)
    dy = _rnd(N, cout, H, W, seed=14)
    _, wpt = ops.pack_conv_weight(w.float().contiguous(), want_fwd=False, stride=1, pad=dil, cout_pad=cout)
    dyp = _nhwc(dy)
    full = ops.conv2d_dgrad(dyp, wpt, cin, (H, W), 3, 3, 1, dil, dil)
    want_pool = ops.upsample_nearest2x_bwd(full[..., :c_pool])
    res = ops.conv2d_dgrad_pool2(dyp, wpt, c_pool, c_rest, 3, 3, dil, dil)
    ops.check_device_flag()
    assert res is not None, "the fused kernel refused an eligible geometry"
    pooled, rest = res
    assert pooled.shape == (N, H // 2, W // 2, c_pool)
    assert torch.equal(pooled.view(torch.int16), want_pool.view(torch.int16)), \
        (pooled.float() - want_pool.float()).abs().max().item()
    if c_rest:
        assert rest.shape == (N, H, W, c_rest)
        assert torch.equal(rest.view(torch.int16), full[..., c_pool:].contiguous().view(torch.int16))
    else:
        assert rest is None
    ref = torch.nn.grad.conv2d_input((N, cin, H, W), w.float(), dy.float(), stride=1, padding=dil, dilation=dil)
    ref_pool = torch.nn.functional.avg_pool2d(ref[:, :c_pool], 2) * 4
    _close(pooled.permute(0, 3, 1, 2), ref_pool, "pooled dgrad")


def test_conv_dgrad_pool2_declines_what_it_cannot_take():
    from pdac_pathological_image_segmentation_b200 import ops
    w = _rnd(16, 16, 3, 3, scale=0.05, seed=1)
    _, wpt = ops.pack_conv_weight(w.float().contiguous(), want_fwd=False, stride=1, pad=1, cout_pad=16)
    assert ops.conv2d_dgrad_pool2(_nhwc(_rnd(1, 16, 7, 8, seed=2)), wpt, 16, 0, 3, 3, 1, 1) is None      # odd height
    assert ops.conv2d_dgrad_pool2(_nhwc(_rnd(1, 16, 8, 8, seed=2)), wpt, 12, 4, 3, 3, 1, 1) is None      # 8-channel groups


# ------------------------------------------------------------------ inference: conv + folded BatchNorm stored x2 upsampled
UP2_CASES = [
    # N, H, W, [cins], cout
    (2, 16, 16, [64], 32),
    (2, 32, 32, [32], 16),        # two CTAs per SM, tap stacking
    (1, 64, 64, [16], 16),
    (2, 16, 16, [128, 64], 64),   # concatenated sources
    (1, 10, 12, [64], 64),        # partial tiles
    (1, 16, 16, [256], 256),      # two N tiles
    (1, 8, 8, [512, 256], 256),   # streamed weights
]


@pytest.mark.parametrize("case", UP2_CASES, ids=[str(c) for c in UP2_CASES])
def test_conv_affine_up2_equals_conv_then_nearest_upsampling(case):
    """pdac_conv2d_fwd_affine_up2 (smp decoder block output in front of F.interpolate(scale_factor=2, 'nearest'), eval
    mode) == pdac_conv2d_fwd_affine followed by pdac_upsample_nearest2x_fwd, bit for bit."""
    from pdac_pathological_image_segmentation_b200 import ops
    N, H, W, cins, cout = case
    cin = sum(cins)
    xs = [_nhwc(_rnd(N, c, H, W, seed=21 + i)) for i, c in enumerate(cins)]
    w = _rnd(cout, cin, 3, 3, scale=0.05, seed=23)
    g = torch.Generator().manual_seed(24)
    ss = torch.stack([0.5 + torch.rand(cout, generator=g), 0.2 * torch.randn(cout, generator=g)]).to(DEV)
    wp, _ = ops.pack_conv_weight(w.float().contiguous(), want_dgrad=False, cout_pad=cout)
    y = ops.conv2d_fwd(xs, wp, cout, 3, 3, 1, 1, 1, relu=True, bias=ss[1], scale=ss[0])
    want = ops.upsample_nearest2x_fwd(y)
    got = ops.conv2d_fwd_affine_up2(xs, wp, cout, 3, 3, 1, 1, True, ss[0], ss[1])
    ops.check_device_flag()
    assert got is not None, "the fused kernel refused an eligible layer"
    assert got.shape == (N, 2 * H, 2 * W, cout)
    assert torch.equal(got.view(torch.int16), want.view(torch.int16))
    ref = F.relu(F.conv2d(torch.cat([x.permute(0, 3, 1, 2) for x in xs], 1).float(), w.float(), padding=1)
                 * ss[0].view(1, -1, 1, 1) + ss[1].view(1, -1, 1, 1))
    _close(got.permute(0, 3, 1, 2), F.interpolate(ref, scale_factor=2, mode="nearest"), "conv + bn + up2")
