"""Thin torch-tensor wrappers over the C ABI (include/pdac_b200.h). PyTorch is used for device memory and streams
only; every function here launches hand-written sm_100a kernels from libpdac_b200.so and raises if a tensor is not
on a CUDA device (no CPU fallback).

Activations are NHWC bf16 tensors of shape (N, H, W, C) whose last stride is 1 (channel slices are fine).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _lib
from ._lib import PdacConv, PdacTapConv, PdacView, call

BF16 = torch.bfloat16
STAT_ROWS = _lib.PDAC_STAT_ROWS

# Reproducible by default, as the reference requests (train_main.py:36-43 cudnn.deterministic = True): BatchNorm
# statistics, BN / bias gradient sums, the Dice sums and the split-K weight gradients are all added in a fixed order.
# PDAC_NONDETERMINISTIC=1 (or ops.deterministic = False) lets the weight / bias gradients use fp32 atomics instead.
deterministic = os.environ.get("PDAC_NONDETERMINISTIC", "0") != "1"
SCRATCH_BYTES = 64 << 20
_scratch_cache = {}


def scratch(device):
    """The per-device fp32 scratch buffer of the deterministic reductions (partial split-K weight gradients, bias
    partial rows). One buffer per device: its users are ordered by the stream they run on."""
    key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
    t = _scratch_cache.get(key)
    if t is None:
        t = torch.empty(SCRATCH_BYTES // 4, dtype=torch.float32, device=device)
        _scratch_cache[key] = t
    return t


def _scratch_args(device):
    if not deterministic:
        return None, 0
    t = scratch(device)
    return C.c_void_p(t.data_ptr()), t.numel() * 4


def new_stats(c, device):
    """Zeroed [STAT_ROWS, 2, c] table of the conv epilogues' BatchNorm partial sums (one row per CTA)."""
    return torch.zeros((STAT_ROWS, 2, c), dtype=torch.float32, device=device)


def new_bwd_sums(c, device):
    """[1 + STAT_ROWS, 2, c] table of bn_bwd_reduce: row 0 = the totals, the rest = per-block partials."""
    return torch.empty((1 + STAT_ROWS, 2, c), dtype=torch.float32, device=device)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _need_cuda(t, what="tensor"):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"pdac_b200: {what} must be a CUDA tensor (there is no CPU fallback)")


def view(t):
    """PdacView of an (N,H,W,C) bf16 tensor."""
    _need_cuda(t, "activation")
    if t.dtype != BF16 or t.dim() != 4 or (t.shape[3] > 1 and t.stride(3) != 1):
        raise RuntimeError(f"pdac_b200: activation must be NHWC bf16 with unit channel stride, got {t.dtype} "
                           f"{tuple(t.shape)} strides {t.stride()}")
    n, h, w, c = t.shape
    # torch leaves the stride of an extent-1 dimension arbitrary: give the library the canonical one
    sw = t.stride(2) if w > 1 else c
    sh = t.stride(1) if h > 1 else w * sw
    sn = t.stride(0) if n > 1 else h * sh
    return PdacView(t.data_ptr(), n, h, w, c, sn, sh, sw)


def _vp(t):
    return C.byref(view(t)) if t is not None else None


def _ptr(t):
    if t is None:
        return None
    _need_cuda(t)
    return C.c_void_p(t.data_ptr())


def wrows_for(c):
    """Rows per tap of a packed weight matrix whose row dimension has `c` used channels (matches the kernel's N tile)."""
    return ((c + 15) // 16) * 16 if c <= 128 else ((c + 127) // 128) * 128


def pad16(c):
    return ((c + 15) // 16) * 16


def empty_act(n, h, w, c, device):
    return torch.empty((n, h, w, c), dtype=BF16, device=device)


# ------------------------------------------------------------------------------------------------ convolutions
def _conv_struct(R, S, stride, pad, dil, relu, wrows, cout_valid=0):
    return PdacConv(R, S, stride, pad, dil, int(bool(relu)), wrows, cout_valid)


def conv_out_hw(h, w, R, S, stride, pad, dil):
    return ((h + 2 * pad - dil * (R - 1) - 1) // stride + 1, (w + 2 * pad - dil * (S - 1) - 1) // stride + 1)


def _check_stats(stats, cout):
    if stats is not None:
        assert stats.dtype == torch.float32 and stats.is_contiguous() and tuple(stats.shape) == (STAT_ROWS, 2, cout), \
            f"stats must be the zeroed [STAT_ROWS, 2, {cout}] table of ops.new_stats, got {tuple(stats.shape)}"


def _act_bytes(*ts):
    return sum(t.numel() * t.element_size() for t in ts if t is not None)


def conv2d_fwd(srcs, wp, cout, R, S, stride=1, pad=0, dil=1, relu=False, bias=None, addend=None, stats=None,
               out=None, c_real=None, scale=None):
    """out = [relu](conv(cat(srcs), w) + bias + addend); optional BN statistics (fp32 [2, cout], pre-zeroed)."""
    x0 = srcs[0]
    ho, wo = conv_out_hw(x0.shape[1], x0.shape[2], R, S, stride, pad, dil)
    if out is None:
        out = empty_act(x0.shape[0], ho, wo, cout, x0.device)
    ktot = sum(s.shape[3] for s in srcs)
    wrows = wp.shape[0] // (R * S)
    assert wp.dtype == BF16 and wp.is_contiguous() and wp.shape[1] == ktot, (wp.shape, ktot)
    _check_stats(stats, out.shape[3])
    cv = _conv_struct(R, S, stride, pad, dil, relu, wrows)
    views = (PdacView * len(srcs))(*[view(s) for s in srcs])
    if scale is not None:        # out = act(acc * scale + bias + addend): eval-mode BatchNorm folded into the epilogue
        assert stats is None and bias is not None
        call("pdac_conv2d_fwd_affine", C.byref(cv), len(srcs), views, _ptr(wp), _ptr(scale), _ptr(bias), _vp(addend),
             _vp(out), _stream(), flops=2 * x0.shape[0] * ho * wo * (c_real or cout) * ktot * R * S,
             nbytes=_act_bytes(out, addend, *srcs),
             tag=f"fwd+bn {R}x{S}s{stride}d{dil} {x0.shape[0]}x{ho}x{wo} {ktot}->{cout}")
        return out
    call("pdac_conv2d_fwd", C.byref(cv), len(srcs), views, _ptr(wp), _ptr(bias), _vp(addend), _vp(out), _ptr(stats),
         _stream(), flops=2 * x0.shape[0] * ho * wo * (c_real or cout) * ktot * R * S,
         nbytes=_act_bytes(out, addend, *srcs),
         tag=f"fwd {R}x{S}s{stride}d{dil} {x0.shape[0]}x{ho}x{wo} {ktot}->{cout}")
    return out


def conv2d_dgrad(dy, wpt, cin, in_hw, R, S, stride=1, pad=0, dil=1, addend=None, out=None, c_real=None):
    """dx (N, H, W, cin) = data gradient of the conv; `addend` (same shape as dx) is added (may alias `out`)."""
    if out is None:
        out = empty_act(dy.shape[0], in_hw[0], in_hw[1], cin, dy.device)
    wrows = wpt.shape[0] // (R * S)
    assert wpt.dtype == BF16 and wpt.is_contiguous() and wpt.shape[1] == dy.shape[3]
    cv = _conv_struct(R, S, stride, pad, dil, False, wrows)
    call("pdac_conv2d_dgrad", C.byref(cv), _vp(dy), _ptr(wpt), _vp(addend), _vp(out), _stream(),
         flops=2 * dy.shape[0] * dy.shape[1] * dy.shape[2] * (c_real or dy.shape[3]) * cin * R * S,
         nbytes=_act_bytes(dy, out, addend),
         tag=f"dgrad {R}x{S}s{stride}d{dil} {dy.shape[0]}x{dy.shape[1]}x{dy.shape[2]} {dy.shape[3]}->{cin}")
    return out


fuse_up2_bwd = os.environ.get("PDAC_FUSE_UP2_BWD", "1") != "0"
fuse_up2_infer = os.environ.get("PDAC_FUSE_UP2_INFER", "1") != "0"


def conv2d_fwd_affine_up2(srcs, wp, cout, R, S, pad, dil, relu, scale, shift):
    """[relu](conv(cat(srcs)) * scale + shift) stored nearest-neighbour x2 upsampled -> (N, 2H, 2W, cout), or None when
    the fused kernel does not take the layer (the caller then runs conv2d_fwd + upsample_nearest2x_fwd)."""
    x0 = srcs[0]
    ktot = sum(s.shape[3] for s in srcs)
    if not fuse_up2_infer or cout % 8 or (R - 1) * dil != 2 * pad or (S - 1) * dil != 2 * pad:
        return None
    wrows = wp.shape[0] // (R * S)
    assert wp.dtype == BF16 and wp.is_contiguous() and wp.shape[1] == ktot
    out = empty_act(x0.shape[0], 2 * x0.shape[1], 2 * x0.shape[2], cout, x0.device)
    cv = _conv_struct(R, S, 1, pad, dil, relu, wrows)
    views = (PdacView * len(srcs))(*[view(s) for s in srcs])
    handled = C.c_int(0)
    call("pdac_conv2d_fwd_affine_up2", C.byref(cv), len(srcs), views, _ptr(wp), _ptr(scale), _ptr(shift), _vp(out),
         C.byref(handled), _stream(), flops=2 * x0.shape[0] * x0.shape[1] * x0.shape[2] * cout * ktot * R * S,
         nbytes=_act_bytes(out, *srcs), tag=f"fwd+bn+up2 {R}x{S}s1d{dil} {x0.shape[0]}x{x0.shape[1]}x{x0.shape[2]} {ktot}->{cout}")
    return out if handled.value else None


def conv2d_dgrad_pool2(dy, wpt, c_pool, c_rest, R, S, pad=0, dil=1, c_real=None):
    """Data gradient of a stride-1 same-size conv whose first c_pool input channels were a nearest x2 upsampling:
    -> (dx_pooled (N, H/2, W/2, c_pool) = the 2x2-summed gradient, dx_rest (N, H, W, c_rest) or None), or None when the
    fused kernel does not take the geometry (the caller then runs conv2d_dgrad + upsample_nearest2x_bwd)."""
    n, h, w = dy.shape[0], dy.shape[1], dy.shape[2]
    if not fuse_up2_bwd or (h | w) & 1 or c_pool % 8 or c_rest % 8:
        return None
    wrows = wpt.shape[0] // (R * S)
    assert wpt.dtype == BF16 and wpt.is_contiguous() and wpt.shape[1] == dy.shape[3]
    cv = _conv_struct(R, S, 1, pad, dil, False, wrows)
    pooled = empty_act(n, h // 2, w // 2, c_pool, dy.device)
    rest = empty_act(n, h, w, c_rest, dy.device) if c_rest else None
    handled = C.c_int(0)
    call("pdac_conv2d_dgrad_pool2", C.byref(cv), _vp(dy), _ptr(wpt), _vp(pooled), _vp(rest), C.byref(handled), _stream(),
         flops=2 * n * h * w * (c_real or dy.shape[3]) * (c_pool + c_rest) * R * S,
         nbytes=_act_bytes(dy, pooled, rest),
         tag=f"dgrad+pool {R}x{S}s1d{dil} {n}x{h}x{w} {dy.shape[3]}->{c_pool}/4+{c_rest}")
    if not handled.value:
        return None
    return pooled, rest


def conv2d_wgrad(srcs, dy, R, S, stride=1, pad=0, dil=1, out=None, oihw=False, cout_valid=0):
    """dw fp32 [R*S, dy.C, sum(src.C)] (or, with oihw=True, the state_dict layout [cout, sum(src.C), R, S]),
    accumulated into `out` if given (which must be pre-zeroed). cout_valid: real channels of a padded dy."""
    ktot = sum(s.shape[3] for s in srcs)
    if out is None:
        shape = ((cout_valid or dy.shape[3]), ktot, R, S) if oihw else (R * S, dy.shape[3], ktot)
        out = torch.zeros(shape, dtype=torch.float32, device=dy.device)
    cv = _conv_struct(R, S, stride, pad, dil, False, 0, cout_valid)
    views = (PdacView * len(srcs))(*[view(s) for s in srcs])
    sp, sb = _scratch_args(dy.device)
    call("pdac_conv2d_wgrad", C.byref(cv), len(srcs), views, _vp(dy), _ptr(out), int(bool(oihw)), sp, sb, _stream(),
         flops=2 * dy.shape[0] * dy.shape[1] * dy.shape[2] * (cout_valid or dy.shape[3]) * ktot * R * S,
         nbytes=_act_bytes(dy, *srcs),
         tag=f"wgrad {R}x{S}s{stride}d{dil} {dy.shape[0]}x{dy.shape[1]}x{dy.shape[2]} {ktot}->{dy.shape[3]}")
    return out


def convt2x2_fwd(x, wp, cout, bias=None, relu=False, out=None):
    if out is None:
        out = empty_act(x.shape[0], 2 * x.shape[1], 2 * x.shape[2], cout, x.device)
    wrows = wp.shape[0] // 4
    call("pdac_convt2x2_fwd", _vp(x), _ptr(wp), wrows, _ptr(bias), int(bool(relu)), _vp(out), _stream(),
         flops=8 * x.numel() * cout, nbytes=_act_bytes(x, out))
    return out


def convt2x2_dgrad(dy, wpt, cin, out=None):
    if out is None:
        out = empty_act(dy.shape[0], dy.shape[1] // 2, dy.shape[2] // 2, cin, dy.device)
    wrows = wpt.shape[0] // 4
    call("pdac_convt2x2_dgrad", _vp(dy), _ptr(wpt), wrows, _vp(out), _stream(),
         flops=2 * dy.numel() * cin, nbytes=_act_bytes(dy, out))
    return out


def convt2x2_wgrad(x, dy, out=None, iohw=False):
    """dw fp32 [4, Cout, Cin], or with iohw=True the nn.ConvTranspose2d layout [Cin, Cout, 2, 2]."""
    if out is None:
        shape = (x.shape[3], dy.shape[3], 2, 2) if iohw else (4, dy.shape[3], x.shape[3])
        out = torch.zeros(shape, dtype=torch.float32, device=x.device)
    sp, sb = _scratch_args(x.device)
    call("pdac_convt2x2_wgrad", _vp(x), _vp(dy), _ptr(out), int(bool(iohw)), sp, sb, _stream(),
         flops=2 * dy.numel() * x.shape[3], nbytes=_act_bytes(x, dy))
    return out


def _tapconv(taps, n_cat, wrows, relu):
    tc = PdacTapConv()
    tc.n_taps, tc.n_cat, tc.wrows, tc.relu = len(taps), n_cat, wrows, int(bool(relu))
    for i, (src, dh, dw) in enumerate(taps):
        tc.taps[i].src, tc.taps[i].dh, tc.taps[i].dw = src, dh, dw
    return tc


def tapconv_fwd(views, taps, wp, out, n_cat=1, relu=False, bias=None, addend=None, stats=None, flops=0, scale=None):
    """Generalised tap convolution (pdac_tapconv_fwd): taps = [(first view index, dh, dw), ...]. With `scale` the
    epilogue is act(acc * scale + bias + addend) (pdac_tapconv_fwd_affine; no statistics)."""
    tc = _tapconv(taps, n_cat, wp.shape[0] // len(taps), relu)
    vs = (PdacView * len(views))(*[view(v) for v in views])
    _check_stats(stats, out.shape[3])
    if scale is not None:
        assert stats is None and bias is not None
        call("pdac_tapconv_fwd_affine", C.byref(tc), len(views), vs, _ptr(wp), _ptr(scale), _ptr(bias), _vp(addend),
             _vp(out), _stream(), flops=flops, nbytes=_act_bytes(out, addend, *views[:n_cat]))
        return out
    call("pdac_tapconv_fwd", C.byref(tc), len(views), vs, _ptr(wp), _ptr(bias), _vp(addend), _vp(out), _ptr(stats),
         _stream(), flops=flops, nbytes=_act_bytes(out, addend, *views[:n_cat]))
    return out


def tapconv_wgrad(views, taps, dy, dw, tap_stride, a_stride, b_stride, n_cat=1, b_valid=0, flops=0):
    tc = _tapconv(taps, n_cat, 0, False)
    vs = (PdacView * len(views))(*[view(v) for v in views])
    sp, sb = _scratch_args(dy.device)
    call("pdac_tapconv_wgrad", C.byref(tc), len(views), vs, _vp(dy), _ptr(dw), tap_stride, a_stride, b_stride, b_valid,
         sp, sb, _stream(), flops=flops, nbytes=_act_bytes(dy, *views[:n_cat]))
    return dw


# The ResNet stem (7x7, stride 2, pad 3; torchvision resnet.py:197) as a 4x4 stride-1 convolution over the
# space-to-depth input of stem_prep_*: tap (a, b) reads z[oh + a - 2, ow + b - 2].
STEM_TAPS = [(0, a - 2, b - 2) for a in range(4) for b in range(4)]


def stem_conv_fwd(z, wp, cout, stats=None, out=None, scale=None, shift=None, relu=False):
    if out is None:
        out = empty_act(z.shape[0], z.shape[1], z.shape[2], cout, z.device)
    return tapconv_fwd([z], STEM_TAPS, wp, out, stats=stats, scale=scale, bias=shift, relu=relu,
                       flops=2 * out.shape[0] * out.shape[1] * out.shape[2] * cout * 147)


def stem_conv_wgrad(z, dy, out=None):
    """fp32 [16, Cout, 16] in the packed space-to-depth form (see unpack_stem_wgrad)."""
    if out is None:
        out = torch.zeros((16, dy.shape[3], 16), dtype=torch.float32, device=z.device)
    return tapconv_wgrad([z], STEM_TAPS, dy, out, dy.shape[3] * 16, 1, 16,
                         flops=2 * dy.shape[0] * dy.shape[1] * dy.shape[2] * dy.shape[3] * 147)


# ------------------------------------------------------------------------------------------------ packing / layout
def pack_conv_weight(w, want_fwd=True, want_dgrad=True, stride=1, pad=0, cout_pad=None, wp_f=None, wp_t=None):
    """fp32 [O, I, R, S] -> (wp_f bf16 [R*S*wrows_f, I], wp_t bf16 [R*S*wrows_t, O_pad])."""
    _need_cuda(w, "weight")
    w = w.detach()
    assert w.dtype == torch.float32 and w.is_contiguous()
    O, I, R, S = w.shape
    opad = cout_pad or O
    wrows_f, wrows_t = wrows_for(opad), wrows_for(I)
    if want_fwd and wp_f is None:
        wp_f = torch.empty((R * S * wrows_f, I), dtype=BF16, device=w.device)
    if want_dgrad and wp_t is None:
        wp_t = torch.empty((R * S * wrows_t, opad), dtype=BF16, device=w.device)
    call("pdac_pack_conv_weight", _ptr(w), O, I, R, S, wrows_f, I, _ptr(wp_f) if want_fwd else None, wrows_t, opad,
         _ptr(wp_t) if want_dgrad else None, pad if stride == 2 else -1, _stream())
    return wp_f, wp_t


def conv_pack_recipe(w, wp_f, wp_t, stride=1, pad=0, cout_pad=None):
    """Descriptor of one pack_conv_weight call (for pack_conv_weights_batched): same arguments, given buffers."""
    O, I, R, S = w.shape
    opad = cout_pad or O
    return _lib.PdacPackDesc(w.data_ptr(), wp_f.data_ptr() if wp_f is not None else None,
                             wp_t.data_ptr() if wp_t is not None else None, O, I, R, S, wrows_for(opad), I,
                             wrows_for(I), opad, pad if stride == 2 else -1)


def pack_conv_weights_batched(recipes):
    """Re-packs many conv weights into their existing bf16 buffers with one launch per 24 tensors."""
    if not recipes:
        return
    arr = (_lib.PdacPackDesc * len(recipes))(*recipes)
    call("pdac_pack_conv_weights_batched", arr, len(recipes), _stream())


def pack_convt_weight(w, wp_f=None, wp_t=None):
    """nn.ConvTranspose2d weight fp32 [Cin, Cout, 2, 2] -> (forward matrix [4*wrows(Cout), Cin], dgrad matrix
    [4*wrows(Cin), Cout])."""
    _need_cuda(w, "weight")
    w = w.detach()
    cin, cout = w.shape[0], w.shape[1]
    wr_f, wr_t = wrows_for(cout), wrows_for(cin)
    if wp_f is None:
        wp_f = torch.empty((4 * wr_f, cin), dtype=BF16, device=w.device)
    if wp_t is None:
        wp_t = torch.empty((4 * wr_t, cout), dtype=BF16, device=w.device)
    # with O=Cin, I=Cout the "forward" packing [t][o=cin][i=cout] is the dgrad matrix and vice versa
    call("pdac_pack_conv_weight", _ptr(w), cin, cout, 2, 2, wr_t, cout, _ptr(wp_t), wr_f, cin, _ptr(wp_f), -1,
         _stream())
    return wp_f, wp_t


def pack_stem_weight(w, wp=None):
    _need_cuda(w, "weight")
    w = w.detach()
    assert tuple(w.shape[1:]) == (3, 7, 7) and w.is_contiguous()
    O = w.shape[0]
    if wp is None:
        wp = torch.empty((16 * wrows_for(O), 16), dtype=BF16, device=w.device)
    assert wrows_for(O) == O, "stem Cout must be a multiple of 16"
    call("pdac_pack_stem_weight", _ptr(w), O, _ptr(wp), _stream())
    return wp


def unpack_stem_wgrad(dwp, O, out=None):
    if out is None:
        out = torch.empty((O, 3, 7, 7), dtype=torch.float32, device=dwp.device)
    call("pdac_unpack_stem_wgrad", _ptr(dwp), O, _ptr(out), _stream())
    return out


def nchw_to_nhwc(x, c_pad=None, out=None):
    _need_cuda(x, "input")
    x = x.contiguous()
    assert x.dtype == torch.float32
    n, c, h, w = x.shape
    if out is None:
        out = empty_act(n, h, w, c_pad or c, x.device)
    call("pdac_nchw_f32_to_nhwc_bf16", _ptr(x), n, c, h, w, _vp(out), _stream())
    return out


def nhwc_to_nchw(x, c_used=None, out=None):
    c_used = c_used or x.shape[3]
    if out is None:
        out = torch.empty((x.shape[0], c_used, x.shape[1], x.shape[2]), dtype=torch.float32, device=x.device)
    call("pdac_nhwc_bf16_to_nchw_f32", _vp(x), c_used, _ptr(out), _stream())
    return out


def stem_prep_f32(x, out=None):
    _need_cuda(x, "input")
    x = x.contiguous()
    assert x.dtype == torch.float32 and x.shape[1] == 3
    n, _, h, w = x.shape
    if out is None:
        out = empty_act(n, h // 2, w // 2, 16, x.device)
    call("pdac_stem_prep_f32", _ptr(x), n, h, w, _ptr(out), _stream())
    return out


IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def stem_prep_u8(img, mask=None, aug=None, mean=IMAGENET_MEAN, std=IMAGENET_STD, out=None, mask_out=None):
    """uint8 (N,H,W,3) patches (+ optional uint8 (N,H,W) masks, int32 (N,) augmentation codes) -> stem input."""
    _need_cuda(img, "image batch")
    assert img.dtype == torch.uint8 and img.is_contiguous() and img.shape[3] == 3
    n, h, w, _ = img.shape
    if out is None:
        out = empty_act(n, h // 2, w // 2, 16, img.device)
    if mask is not None and mask_out is None:
        mask_out = torch.empty_like(mask)
    m = (C.c_float * 3)(*mean)
    s = (C.c_float * 3)(*std)
    call("pdac_stem_prep_u8", _ptr(img), _ptr(mask), _ptr(aug), n, h, w, m, s, _ptr(out), _ptr(mask_out), _stream())
    return out, mask_out


class Augment:
    """The reference's training augmentation (train_main.py:70-80), drawn and applied on the GPU: ColorJitter ranges,
    the probability that ColorJitter is applied to a patch (albumentations' default p = 0.5), the OneOf group's p."""

    def __init__(self, brightness=0.3, contrast=0.3, saturation=0.3, hue=0.3, p_color=0.5, p_geom=0.3, seed=0):
        self.brightness, self.contrast, self.saturation, self.hue = brightness, contrast, saturation, hue
        self.p_color, self.p_geom, self.seed = p_color, p_geom, int(seed)

    def struct(self):
        return _lib.PdacAugment(self.brightness, self.contrast, self.saturation, self.hue, self.p_color, self.p_geom,
                                self.seed & 0xFFFFFFFFFFFFFFFF)


def stem_prep_u8_augment(img, mask, cfg, step=None, mean=IMAGENET_MEAN, std=IMAGENET_STD, out=None, mask_out=None,
                         params=None):
    """uint8 (N,H,W,3) patches (+ uint8 masks) -> augmented stem input. -> (z, mask_out, params fp32 [N, 8]: the drawn
    per-patch parameters, see include/pdac_b200.h). step: int64 device scalar mixed into the seed."""
    _need_cuda(img, "image batch")
    assert img.dtype == torch.uint8 and img.is_contiguous() and img.shape[3] == 3
    n, h, w, _ = img.shape
    if out is None:
        out = empty_act(n, h // 2, w // 2, 16, img.device)
    if mask is not None and mask_out is None:
        mask_out = torch.empty_like(mask)
    if params is None:
        params = torch.empty((n, 8), dtype=torch.float32, device=img.device)
    m = (C.c_float * 3)(*mean)
    s = (C.c_float * 3)(*std)
    st = cfg.struct()
    call("pdac_stem_prep_u8_augment", _ptr(img), _ptr(mask), n, h, w, m, s, C.byref(st), _ptr(step), _ptr(params),
         _ptr(out), _ptr(mask_out), _stream())
    return out, mask_out, params


# ------------------------------------------------------------------------------------------------ BN / elementwise
def bn_prepare(stats, count, gamma, beta, eps, momentum, train, running_mean, running_var, scale_shift, mean_invstd,
               num_batches_tracked=None):
    """stats: fp32 [rows, 2, C] partial sums (new_stats) or a plain [2, C] vector; None in eval mode."""
    c = scale_shift.shape[-1]
    n_rows = 0 if stats is None else (stats.shape[0] if stats.dim() == 3 else 1)
    call("pdac_bn_prepare", _ptr(stats), n_rows, float(count), _ptr(gamma), _ptr(beta), eps, momentum, int(bool(train)),
         _ptr(running_mean), _ptr(running_var), _ptr(num_batches_tracked), _ptr(scale_shift), _ptr(mean_invstd), c,
         _stream())


def scale_shift_act(x, scale_shift=None, residual=None, relu=False, out=None, up2=False):
    """up2: the result is stored nearest-neighbour x2 upsampled ((N, 2H, 2W, C); the low-resolution tensor is not)."""
    if out is None:
        out = empty_act(x.shape[0], 2 * x.shape[1], 2 * x.shape[2], x.shape[3], x.device) if up2 else torch.empty_like(x)
    call("pdac_scale_shift_act", _vp(x), _ptr(scale_shift), _vp(residual), int(bool(relu)), _vp(out), _stream())
    return out


def bn_bwd_reduce(dy, y, x, mean_invstd, sums=None, scale_shift=None):
    """y: ReLU output for the mask, or None with scale_shift given to recompute the mask from x (y is not read).
    sums: new_bwd_sums table; sums[0] holds (sum dz, sum dz*xhat) afterwards."""
    if sums is None:
        sums = new_bwd_sums(dy.shape[3], dy.device)
    assert sums.shape == (1 + STAT_ROWS, 2, dy.shape[3]) and sums.is_contiguous()
    call("pdac_bn_bwd_reduce", _vp(dy), _vp(y), _vp(x), _ptr(mean_invstd), _ptr(scale_shift), _ptr(sums), _stream())
    return sums


def bn_bwd_apply(dy, y, x, mean_invstd, gamma, sums, count, train, dx=None, dz_out=None, dgamma=None, dbeta=None,
                 scale_shift=None):
    if dx is None:
        dx = torch.empty_like(dy)
    call("pdac_bn_bwd_apply", _vp(dy), _vp(y), _vp(x), _ptr(mean_invstd), _ptr(gamma), _ptr(scale_shift), _ptr(sums),
         float(count),
         int(bool(train)), _vp(dx), _vp(dz_out), _ptr(dgamma), _ptr(dbeta), _stream())
    return dx


def relu_bwd_bias(dy, y=None, dz=None, dbias=None):
    sp = None
    if dbias is not None and deterministic:
        sc = scratch(dy.device)
        assert STAT_ROWS * dy.shape[3] <= sc.numel()
        sp = C.c_void_p(sc.data_ptr())
    call("pdac_relu_bwd_bias", _vp(dy), _vp(y), _vp(dz), _ptr(dbias), sp, _stream())
    return dz


def add(a, b, relu=False, out=None):
    if out is None:
        out = torch.empty_like(a)
    call("pdac_add", _vp(a), _vp(b), int(bool(relu)), _vp(out), _stream())
    return out


# ------------------------------------------------------------------------------------------------ pooling / resize
def maxpool_fwd(x, want_idx=True):
    n, h, w, c = x.shape
    ho, wo = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    y = empty_act(n, ho, wo, c, x.device)
    idx = torch.empty((n, ho, wo, c), dtype=torch.uint8, device=x.device) if want_idx else None
    call("pdac_maxpool3x3s2_fwd", _vp(x), _vp(y), _ptr(idx), _stream())
    return y, idx


def maxpool_bwd(dy, idx, in_hw, addend=None):
    """dx = scatter of dy to the arg-max positions (+ addend, a gradient already accumulated at the same tensor)."""
    dx = empty_act(dy.shape[0], in_hw[0], in_hw[1], dy.shape[3], dy.device)
    call("pdac_maxpool3x3s2_bwd", _vp(dy), _ptr(idx), _vp(addend), _vp(dx), _stream())
    return dx


def upsample_nearest2x_fwd(x, out=None):
    if out is None:
        out = empty_act(x.shape[0], 2 * x.shape[1], 2 * x.shape[2], x.shape[3], x.device)
    call("pdac_upsample_nearest2x_fwd", _vp(x), _vp(out), _stream())
    return out


def upsample_nearest2x_bwd(dy, addend=None, out=None):
    if out is None:
        out = empty_act(dy.shape[0], dy.shape[1] // 2, dy.shape[2] // 2, dy.shape[3], dy.device)
    call("pdac_upsample_nearest2x_bwd", _vp(dy), _vp(addend), _vp(out), _stream())
    return out


def bilinear_fwd(x, out_hw, align_corners, out=None):
    if out is None:
        out = empty_act(x.shape[0], out_hw[0], out_hw[1], x.shape[3], x.device)
    call("pdac_bilinear_fwd", _vp(x), int(bool(align_corners)), _vp(out), _stream())
    return out


def bilinear_bwd(dy, in_hw, align_corners, out=None):
    if out is None:
        out = empty_act(dy.shape[0], in_hw[0], in_hw[1], dy.shape[3], dy.device)
    call("pdac_bilinear_bwd", _vp(dy), int(bool(align_corners)), _vp(out), _stream())
    return out


def bilinear_head_fwd(x, c_used, out_hw, align_corners, out=None):
    if out is None:
        out = torch.empty((x.shape[0], c_used, out_hw[0], out_hw[1]), dtype=torch.float32, device=x.device)
    call("pdac_bilinear_head_fwd", _vp(x), c_used, int(bool(align_corners)), _ptr(out), out_hw[0], out_hw[1], _stream())
    return out


def bilinear_head_bwd(g, in_shape, align_corners, out=None):
    """g: fp32 (N, Cu, Ho, Wo) -> dx bf16 (N, h, w, Cpad) with in_shape = (h, w, Cpad)."""
    _need_cuda(g, "gradient")
    g = g.contiguous()
    n, cu, ho, wo = g.shape
    if out is None:
        out = empty_act(n, in_shape[0], in_shape[1], in_shape[2], g.device)
    call("pdac_bilinear_head_bwd", _ptr(g), cu, ho, wo, int(bool(align_corners)), _vp(out), _stream())
    return out


def adaptive_avgpool_fwd(x, out_hw):
    y = empty_act(x.shape[0], out_hw[0], out_hw[1], x.shape[3], x.device)
    call("pdac_adaptive_avgpool_fwd", _vp(x), _vp(y), _stream())
    return y


def adaptive_avgpool_bwd(dy, in_hw):
    dx = empty_act(dy.shape[0], in_hw[0], in_hw[1], dy.shape[3], dy.device)
    call("pdac_adaptive_avgpool_bwd", _vp(dy), _vp(dx), _stream())
    return dx


# ------------------------------------------------------------------------------------------------ GroupNorm / dropout / depthwise
def groupnorm_fwd(x, groups, gamma, beta, eps, relu, out=None):
    """-> (y, stats fp32 [N, G, 2] kept for backward)."""
    stats = torch.zeros((x.shape[0], groups, 2), dtype=torch.float32, device=x.device)
    call("pdac_groupnorm_stats", _vp(x), groups, _ptr(stats), _scratch_args(x.device)[0], _stream())
    if out is None:
        out = torch.empty_like(x)
    call("pdac_groupnorm_fwd", _vp(x), groups, _ptr(stats), _ptr(gamma), _ptr(beta), eps, int(bool(relu)), _vp(out),
         _stream())
    return out, stats


def groupnorm_bwd(dy, y, x, groups, stats, gamma, eps, dgamma=None, dbeta=None, out=None):
    """dz = dy * (y > 0) when y is given; returns dx, accumulates dgamma / dbeta."""
    gsums = torch.zeros((x.shape[0], groups, 2), dtype=torch.float32, device=x.device)
    call("pdac_groupnorm_bwd_reduce", _vp(dy), _vp(y), _vp(x), groups, _ptr(stats), _ptr(gamma), eps, _ptr(gsums),
         _ptr(dgamma), _ptr(dbeta), _scratch_args(x.device)[0], _stream())
    if out is None:
        out = torch.empty_like(dy)
    call("pdac_groupnorm_bwd_apply", _vp(dy), _vp(y), _vp(x), groups, _ptr(stats), _ptr(gamma), eps, _ptr(gsums),
         _vp(out), _stream())
    return out


def dropout(x, p, seed, per_channel, out=None, step=None):
    """nn.Dropout2d (per_channel) / nn.Dropout with a counter-based hash mask; backward = same call on the gradient.
    step: optional int64 device scalar mixed into the seed at run time (fresh masks under CUDA-graph replay)."""
    if out is None:
        out = torch.empty_like(x)
    call("pdac_dropout", _vp(x), _vp(out), float(p), int(seed) & 0xFFFFFFFFFFFFFFFF, _ptr(step), int(bool(per_channel)),
         _stream())
    return out


def dwconv3x3(x, w, dil=1, flip=False, out=None):
    """Depthwise 3x3 (stride 1, padding = dilation); w = the fp32 (C, 1, 3, 3) state_dict tensor. flip: data gradient."""
    assert w.dtype == torch.float32 and w.is_contiguous() and w.numel() == x.shape[3] * 9
    if out is None:
        out = torch.empty_like(x)
    call("pdac_dwconv3x3", _vp(x), _ptr(w), int(dil), int(bool(flip)), _vp(out), _stream(),
         flops=18 * x.numel(), nbytes=_act_bytes(x, out), tag=f"dw3x3 d{dil} {tuple(x.shape)}")
    return out


def dwconv3x3_wgrad(x, dy, dil, out):
    """out: fp32 (C, 1, 3, 3) gradient slot (accumulated into)."""
    call("pdac_dwconv3x3_wgrad", _vp(x), _vp(dy), int(dil), _ptr(out), _scratch_args(x.device)[0], _stream())
    return out


# ------------------------------------------------------------------------------------------------ segmentation head
def head3x3_supported(channels, classes):
    return bool(_lib.load().pdac_head3x3_supported(int(channels), int(classes)))


def head3x3_fwd(x, weight, bias, out=None):
    """Conv2d(C, classes, 3, padding=1) + bias on bf16 NHWC x with the fp32 state_dict weight -> fp32 NCHW logits."""
    classes = weight.shape[0]
    assert weight.dtype == torch.float32 and weight.is_contiguous() and tuple(weight.shape[1:]) == (x.shape[3], 3, 3)
    if out is None:
        out = torch.empty((x.shape[0], classes, x.shape[1], x.shape[2]), dtype=torch.float32, device=x.device)
    call("pdac_head3x3_fwd", _vp(x), _ptr(weight), _ptr(bias), classes, _ptr(out), _stream(),
         flops=2 * x.shape[0] * x.shape[1] * x.shape[2] * classes * x.shape[3] * 9, nbytes=_act_bytes(x, out),
         tag=f"head fwd {tuple(x.shape)} -> {classes}")
    return out


def head3x3_dgrad(g, weight, in_shape):
    """g fp32 NCHW (N, classes, H, W) -> dx bf16 NHWC (N, H, W, C)."""
    _need_cuda(g, "gradient")
    g = g.contiguous()
    assert g.dtype == torch.float32
    n, h, w, c = in_shape
    dx = empty_act(n, h, w, c, g.device)
    call("pdac_head3x3_dgrad", _ptr(g), _ptr(weight), weight.shape[0], _vp(dx), _stream(),
         flops=2 * n * h * w * weight.shape[0] * c * 9, nbytes=_act_bytes(g, dx), tag=f"head dgrad {tuple(dx.shape)}")
    return dx


def head3x3_wgrad(x, g, dw, dbias=None):
    """Accumulates dw fp32 [classes, C, 3, 3] and dbias fp32 [classes] (state_dict layouts); fixed-order reduction."""
    g = g.contiguous()
    sc = scratch(x.device)
    call("pdac_head3x3_wgrad", _vp(x), _ptr(g), g.shape[1], _ptr(dw), _ptr(dbias), C.c_void_p(sc.data_ptr()),
         sc.numel() * 4, _stream(), flops=2 * x.numel() * g.shape[1] * 9, nbytes=_act_bytes(x, g),
         tag=f"head wgrad {tuple(x.shape)}")
    return dw


# ------------------------------------------------------------------------------------------------ loss / metric
def _tgt(t):
    _need_cuda(t, "target")
    if t.dtype == torch.float32:
        return t.contiguous(), 0
    if t.dtype in (torch.uint8, torch.bool):
        return t.contiguous().view(torch.uint8), 1
    raise RuntimeError(f"pdac_b200: targets must be float32 or uint8, got {t.dtype}")


def dice_fwd(logits, targets, smooth=1e-6):
    """-> (loss fp32 scalar tensor, sums fp64 (the first 3 entries are sum p*t, sum p, sum t) saved for backward)."""
    _need_cuda(logits, "logits")
    logits = logits.contiguous()
    assert logits.dtype == torch.float32
    t, kind = _tgt(targets)
    assert t.numel() == logits.numel(), "DiceLoss: inputs and targets must have the same number of elements"
    sums = torch.empty(3 * (1 + STAT_ROWS), dtype=torch.float64, device=logits.device)   # [0:3] totals + block rows
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    call("pdac_dice_fwd", _ptr(logits), _ptr(t), kind, logits.numel(), _ptr(sums), smooth, _ptr(loss), _stream())
    return loss, sums


def dice_bwd(logits, targets, sums, grad_out, smooth=1e-6):
    logits = logits.contiguous()
    t, kind = _tgt(targets)
    g = torch.empty_like(logits)
    go = grad_out.contiguous().to(torch.float32) if grad_out is not None else None
    call("pdac_dice_bwd", _ptr(logits), _ptr(t), kind, logits.numel(), _ptr(sums), smooth, _ptr(go), _ptr(g), _stream())
    return g


def dice_bwd_nhwc(logits, targets, sums, cpad, smooth=1e-6):
    """1-class Dice gradient (grad_out = 1) as the bf16 NHWC (N, H, W, cpad) tensor the head's backward reads."""
    logits = logits.contiguous()
    n, c, h, w = logits.shape
    assert c == 1 and logits.dtype == torch.float32
    t, kind = _tgt(targets)
    g = torch.empty((n, h, w, cpad), dtype=BF16, device=logits.device)
    call("pdac_dice_bwd_nhwc", _ptr(logits), _ptr(t), kind, logits.numel(), _ptr(sums), smooth, _ptr(g), cpad, _stream())
    return g


def sigmoid_mask(logits, as_uint8=False):
    _need_cuda(logits, "logits")
    logits = logits.contiguous()
    assert logits.dtype == torch.float32
    out = torch.empty(logits.shape, dtype=torch.uint8 if as_uint8 else torch.float32, device=logits.device)
    call("pdac_sigmoid_mask", _ptr(logits), logits.numel(), None if as_uint8 else _ptr(out),
         _ptr(out) if as_uint8 else None, _stream())
    return out


def dice_score(pred, target, smooth=1e-6, pred_is_logits=False):
    _need_cuda(pred, "pred")
    pred = pred.contiguous()
    if pred.dtype == torch.float32:
        kind = 1 if pred_is_logits else 0
    elif pred.dtype in (torch.uint8, torch.bool):
        pred, kind = pred.view(torch.uint8), 2
    else:
        raise RuntimeError(f"pdac_b200: pred must be float32 or uint8, got {pred.dtype}")
    t, tk = _tgt(target)
    assert t.numel() == pred.numel()
    counts = torch.empty(3, dtype=torch.int64, device=pred.device)
    score = torch.empty((), dtype=torch.float32, device=pred.device)
    call("pdac_dice_score", _ptr(pred), kind, _ptr(t), tk, pred.numel(), _ptr(counts), smooth, _ptr(score), _stream())
    return score, counts


def argmax_nchw(logits):
    _need_cuda(logits, "logits")
    logits = logits.contiguous()
    n, c, h, w = logits.shape
    out = torch.empty((n, h, w), dtype=torch.uint8, device=logits.device)
    call("pdac_argmax_nchw", _ptr(logits), n, c, h * w, _ptr(out), _stream())
    return out


def stitch_tiles(tiles, rows, cols, slide, row0=0):
    """tiles uint8 (T, th, tw); rows/cols int32 (T,) grid coordinates; slide uint8 (Hs, Ws) band starting at row0."""
    _need_cuda(tiles, "tiles")
    assert tiles.dtype == torch.uint8 and tiles.is_contiguous() and slide.dtype == torch.uint8 and slide.is_contiguous()
    assert rows.dtype == torch.int32 and cols.dtype == torch.int32
    t, th, tw = tiles.shape
    call("pdac_stitch_tiles", _ptr(tiles), _ptr(rows), _ptr(cols), t, th, tw, _ptr(slide), slide.shape[0],
         slide.shape[1], row0, _stream())
    return slide


def overlay_blend(rgb, mask, color=(255, 0, 0), alpha_256=102, out=None):
    """rgb uint8 (..., H, W, 3), mask uint8 (..., H, W): tumour overlay by integer blend (oracle overlay_blend).
    out=rgb blends in place."""
    _need_cuda(rgb, "rgb")
    assert rgb.dtype == torch.uint8 and mask.dtype == torch.uint8 and rgb.is_contiguous() and mask.is_contiguous()
    assert rgb.shape[-1] == 3 and rgb.shape[:-1] == mask.shape
    if out is None:
        out = torch.empty_like(rgb)
    assert out.dtype == torch.uint8 and out.is_contiguous() and out.shape == rgb.shape
    call("pdac_overlay_blend", _ptr(rgb), _ptr(mask), mask.numel(), int(color[0]), int(color[1]), int(color[2]),
         int(alpha_256), _ptr(out), _stream())
    return out


def adam_step(params, grads, exp_avg, exp_avg_sq, lr, step, beta1=0.9, beta2=0.999, eps=1e-8, grad_scale=1.0):
    call("pdac_adam_step", _ptr(params), _ptr(grads), _ptr(exp_avg), _ptr(exp_avg_sq), params.numel(), lr, beta1, beta2,
         eps, step, grad_scale, _stream())


def adam_step_dev(params, grads, exp_avg, exp_avg_sq, hyper, step):
    """Adam with the step counter (int64 scalar) and hyper-parameters (fp32 [12]: lr, beta1, beta2, eps, grad_scale,
    2 outputs, 1 - beta1, 1 - beta2) on the device: graph-capturable, lr changeable between replays."""
    assert hyper.dtype == torch.float32 and hyper.numel() >= 12 and step.dtype == torch.int64
    call("pdac_adam_step_dev", _ptr(params), _ptr(grads), _ptr(exp_avg), _ptr(exp_avg_sq), params.numel(), _ptr(hyper),
         _ptr(step), _stream())


def check_device_flag():
    """Synchronises the current stream, then raises if a pipeline barrier timed out (tests)."""
    call("pdac_check_device_flag", _stream())


def peek_device_flag():
    """Non-blocking poll of the pipeline error flag (see _lib.peek_device_flag)."""
    _lib.peek_device_flag()
