"""The full-resolution segmentation head (head.cu: Conv2d(C, classes, 3, padding=1) + bias as direct kernels) against
torch fp32 convolutions on bf16-representable activations and fp32 weights: forward logits, data gradient, weight and
bias gradients (state_dict layouts, accumulated), partial tiles, 16 / 32 channels, 1 / 2 classes; reproducibility."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("shape", [(2, 16, 40, 72), (3, 16, 17, 33), (1, 32, 64, 64), (2, 32, 9, 5), (4, 16, 128, 128)])
@pytest.mark.parametrize("classes", [1, 2])
def test_head3x3_fwd_dgrad_wgrad(shape, classes):
    from pdac_pathological_image_segmentation_b200 import ops
    N, C, H, W = shape
    g = torch.Generator(device=DEV).manual_seed(7)
    x = torch.randn(N, C, H, W, device=DEV, generator=g).to(torch.bfloat16)
    # bf16-representable weights and gradient: the warp-MMA forward rounds the weights, the weight gradient rounds
    # dL/dlogits to bf16 (as the tensor-core convolutions do); products and the fp32 accumulation are then exact
    w = (torch.randn(classes, C, 3, 3, device=DEV, generator=g) * 0.2).to(torch.bfloat16).float()
    b = torch.randn(classes, device=DEV, generator=g)
    go = torch.randn(N, classes, H, W, device=DEV, generator=g).to(torch.bfloat16).float()
    assert ops.head3x3_supported(C, classes)
    xr = x.float().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    br = b.clone().requires_grad_(True)
    ref = F.conv2d(xr, wr, br, 1, 1)
    ref.backward(go)
    xh = x.permute(0, 2, 3, 1).contiguous()
    out = ops.head3x3_fwd(xh, w, b)
    assert out.shape == ref.shape and out.dtype == torch.float32
    assert (out - ref).abs().max().item() <= 1e-4 * max(ref.abs().max().item(), 1.0)       # fp32 arithmetic both sides
    dx = ops.head3x3_dgrad(go, w, (N, H, W, C))
    err = (dx.permute(0, 3, 1, 2).float() - xr.grad).abs()
    assert (err <= 2.0 ** -7 * xr.grad.abs().clamp_min(1.0)).all(), err.max().item()        # one bf16 rounding
    dw = torch.full_like(w, 0.5)                                                           # accumulated into
    db = torch.full_like(b, -1.0)
    ops.head3x3_wgrad(xh, go, dw, db)
    assert torch.allclose(dw - 0.5, wr.grad, rtol=1e-4, atol=1e-4 * wr.grad.abs().max().item())
    assert torch.allclose(db + 1.0, br.grad, rtol=1e-4, atol=1e-4 * br.grad.abs().max().item())
    # reproducible
    dw2 = torch.full_like(w, 0.5)
    db2 = torch.full_like(b, -1.0)
    ops.head3x3_wgrad(xh, go, dw2, db2)
    assert torch.equal(dw, dw2) and torch.equal(db, db2)
    # no bias
    out0 = ops.head3x3_fwd(xh, w, None)
    assert (out0 - (ref - b.view(1, -1, 1, 1))).abs().max().item() <= 1e-4 * max(ref.abs().max().item(), 1.0)
    ops.check_device_flag()


def test_unet_uses_the_direct_head_and_unsupported_shapes_fall_back():
    from pdac_pathological_image_segmentation_b200 import _lib, ops
    from pdac_pathological_image_segmentation_b200.smp import Unet
    assert not ops.head3x3_supported(64, 1) and not ops.head3x3_supported(16, 3)
    torch.manual_seed(0)
    m = Unet(encoder_weights=None, classes=2).to(DEV).eval()
    _lib.trace, _lib.trace_all = [], True
    with torch.no_grad():
        y = m(torch.randn(1, 3, 64, 64, device=DEV))
    names = [t[5] for t in _lib.trace]
    _lib.trace, _lib.trace_all = None, False
    assert y.shape == (1, 2, 64, 64) and any(n.startswith("head fwd") for n in names)
    m3 = Unet(encoder_weights=None, classes=3).to(DEV).eval()          # 3 classes: tensor-core path, channel-padded
    with torch.no_grad():
        assert m3(torch.randn(1, 3, 64, 64, device=DEV)).shape == (1, 3, 64, 64)
