"""ctypes binding of libpdac_b200.so (the C ABI declared in include/pdac_b200.h).

The product path has no fallback: if the shared library is missing this module raises at import of any op, and every
entry point raises ``RuntimeError(pdac_last_error())`` on a non-zero status.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PDAC_B200_LIB") or os.path.join(_HERE, "libpdac_b200.so")  # override: diagnostic builds

PDAC_MAX_TAPS = 16
PDAC_STAT_ROWS = 592  # include/pdac_b200.h: row capacity of the fixed-order partial-sum tables


class PdacView(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("N", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("C", C.c_int32),
                ("sN", C.c_int64), ("sH", C.c_int64), ("sW", C.c_int64)]


class PdacConv(C.Structure):
    _fields_ = [("R", C.c_int32), ("S", C.c_int32), ("stride", C.c_int32), ("pad", C.c_int32), ("dil", C.c_int32),
                ("relu", C.c_int32), ("wrows", C.c_int32), ("cout_valid", C.c_int32)]


class PdacAugment(C.Structure):
    _fields_ = [("brightness", C.c_float), ("contrast", C.c_float), ("saturation", C.c_float), ("hue", C.c_float),
                ("p_color", C.c_float), ("p_geom", C.c_float), ("seed", C.c_ulonglong)]


class PdacPackDesc(C.Structure):
    _fields_ = [("w", C.c_void_p), ("wp_f", C.c_void_p), ("wp_t", C.c_void_p)] + \
               [(n, C.c_int32) for n in ("O", "I", "R", "S", "wrows_f", "kf", "wrows_t", "kt", "s2_pad")]


class PdacTap(C.Structure):
    _fields_ = [("src", C.c_int32), ("dh", C.c_int32), ("dw", C.c_int32)]


class PdacTapConv(C.Structure):
    _fields_ = [("n_taps", C.c_int32), ("n_cat", C.c_int32), ("wrows", C.c_int32), ("relu", C.c_int32),
                ("taps", PdacTap * PDAC_MAX_TAPS)]


_VP = C.POINTER(PdacView)
_p, _i, _ll, _f, _d = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_double

# name -> argtypes; every function returns int status (except pdac_version / pdac_last_error)
SIGNATURES = {
    "pdac_check_device_flag": [_p],
    "pdac_tapconv_fwd": [C.POINTER(PdacTapConv), _i, _VP, _p, _p, _VP, _VP, _p, _p],
    "pdac_tapconv_wgrad": [C.POINTER(PdacTapConv), _i, _VP, _VP, _p, C.c_int64, C.c_int64, C.c_int64, _i, _p, _ll, _p],
    "pdac_conv2d_fwd": [C.POINTER(PdacConv), _i, _VP, _p, _p, _VP, _VP, _p, _p],
    "pdac_conv2d_fwd_affine": [C.POINTER(PdacConv), _i, _VP, _p, _p, _p, _VP, _VP, _p],
    "pdac_conv2d_fwd_affine_up2": [C.POINTER(PdacConv), _i, _VP, _p, _p, _p, _VP, C.POINTER(C.c_int), _p],
    "pdac_tapconv_fwd_affine": [C.POINTER(PdacTapConv), _i, _VP, _p, _p, _p, _VP, _VP, _p],
    "pdac_conv2d_dgrad": [C.POINTER(PdacConv), _VP, _p, _VP, _VP, _p],
    "pdac_conv2d_dgrad_pool2": [C.POINTER(PdacConv), _VP, _p, _VP, _VP, C.POINTER(C.c_int), _p],
    "pdac_conv2d_wgrad": [C.POINTER(PdacConv), _i, _VP, _VP, _p, _i, _p, _ll, _p],
    "pdac_convt2x2_fwd": [_VP, _p, _i, _p, _i, _VP, _p],
    "pdac_convt2x2_dgrad": [_VP, _p, _i, _VP, _p],
    "pdac_convt2x2_wgrad": [_VP, _VP, _p, _i, _p, _ll, _p],
    "pdac_nchw_f32_to_nhwc_bf16": [_p, _i, _i, _i, _i, _VP, _p],
    "pdac_nhwc_bf16_to_nchw_f32": [_VP, _i, _p, _p],
    "pdac_stem_prep_f32": [_p, _i, _i, _i, _p, _p],
    "pdac_stem_prep_u8": [_p, _p, _p, _i, _i, _i, C.POINTER(_f), C.POINTER(_f), _p, _p, _p],
    "pdac_stem_prep_u8_augment": [_p, _p, _i, _i, _i, C.POINTER(_f), C.POINTER(_f), C.POINTER(PdacAugment), _p, _p, _p, _p,
                                  _p],
    "pdac_pack_conv_weight": [_p, _i, _i, _i, _i, _i, _i, _p, _i, _i, _p, _i, _p],
    "pdac_pack_conv_weights_batched": [C.POINTER(PdacPackDesc), _i, _p],
    "pdac_pack_stem_weight": [_p, _i, _p, _p],
    "pdac_unpack_stem_wgrad": [_p, _i, _p, _p],
    "pdac_bn_prepare": [_p, _i, _d, _p, _p, _f, _f, _i, _p, _p, _p, _p, _p, _i, _p],
    "pdac_scale_shift_act": [_VP, _p, _VP, _i, _VP, _p],
    "pdac_bn_bwd_reduce": [_VP, _VP, _VP, _p, _p, _p, _p],
    "pdac_bn_bwd_apply": [_VP, _VP, _VP, _p, _p, _p, _p, _d, _i, _VP, _VP, _p, _p, _p],
    "pdac_relu_bwd_bias": [_VP, _VP, _VP, _p, _p, _p],
    "pdac_relu_bwd_bias_noalloc": [_VP, _VP, _VP, _p, _p],
    "pdac_maxpool3x3s2_fwd": [_VP, _VP, _p, _p],
    "pdac_maxpool3x3s2_bwd": [_VP, _p, _VP, _VP, _p],
    "pdac_upsample_nearest2x_fwd": [_VP, _VP, _p],
    "pdac_upsample_nearest2x_bwd": [_VP, _VP, _VP, _p],
    "pdac_bilinear_fwd": [_VP, _i, _VP, _p],
    "pdac_bilinear_bwd": [_VP, _i, _VP, _p],
    "pdac_bilinear_head_fwd": [_VP, _i, _i, _p, _i, _i, _p],
    "pdac_bilinear_head_bwd": [_p, _i, _i, _i, _i, _VP, _p],
    "pdac_adaptive_avgpool_fwd": [_VP, _VP, _p],
    "pdac_adaptive_avgpool_bwd": [_VP, _VP, _p],
    "pdac_add": [_VP, _VP, _i, _VP, _p],
    "pdac_groupnorm_stats": [_VP, _i, _p, _p, _p],
    "pdac_groupnorm_fwd": [_VP, _i, _p, _p, _p, _f, _i, _VP, _p],
    "pdac_groupnorm_bwd_reduce": [_VP, _VP, _VP, _i, _p, _p, _f, _p, _p, _p, _p, _p],
    "pdac_groupnorm_bwd_apply": [_VP, _VP, _VP, _i, _p, _p, _f, _p, _VP, _p],
    "pdac_dropout": [_VP, _VP, _f, C.c_ulonglong, _p, _i, _p],
    "pdac_dwconv3x3": [_VP, _p, _i, _i, _VP, _p],
    "pdac_dwconv3x3_wgrad": [_VP, _VP, _i, _p, _p, _p],
    "pdac_head3x3_supported": [_i, _i],
    "pdac_head3x3_fwd": [_VP, _p, _p, _i, _p, _p],
    "pdac_head3x3_dgrad": [_p, _p, _i, _VP, _p],
    "pdac_head3x3_wgrad": [_VP, _p, _i, _p, _p, _p, _ll, _p],
    "pdac_dice_fwd": [_p, _p, _i, _ll, _p, _f, _p, _p],
    "pdac_dice_bwd": [_p, _p, _i, _ll, _p, _f, _p, _p, _p],
    "pdac_dice_bwd_nhwc": [_p, _p, _i, _ll, _p, _f, _p, _i, _p],
    "pdac_sigmoid_mask": [_p, _ll, _p, _p, _p],
    "pdac_dice_score": [_p, _i, _p, _i, _ll, _p, _f, _p, _p],
    "pdac_argmax_nchw": [_p, _i, _i, _ll, _p, _p],
    "pdac_stitch_tiles": [_p, _p, _p, _i, _i, _i, _p, _ll, _ll, _ll, _p],
    "pdac_overlay_blend": [_p, _p, _ll, _i, _i, _i, _i, _p, _p],
    "pdac_adam_step": [_p, _p, _p, _p, _ll, _f, _f, _f, _f, _i, _f, _p],
    "pdac_adam_step_dev": [_p, _p, _p, _p, _ll, _p, _p, _p],
}

_lib = None
launches = 0  # number of C-ABI calls that launch kernels (bench.py reports it as gpu_launches)


def load():
    """Loads the shared library (once). Raises if it has not been built: there is no fallback implementation."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C pdac_pathological_image_segmentation_b200/csrc`). There is no CPU/PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    lib.pdac_version.restype = C.c_int
    lib.pdac_version.argtypes = []
    lib.pdac_last_error.restype = C.c_char_p
    lib.pdac_last_error.argtypes = []
    lib.pdac_last_kernel_id.restype = C.c_int
    lib.pdac_last_kernel_id.argtypes = []
    lib.pdac_peek_device_flag.restype = C.c_int
    lib.pdac_peek_device_flag.argtypes = []
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = C.c_int
        fn.argtypes = args
    _lib = lib
    return lib


KERNEL_NAMES = {1: "conv_tc_kernel", 2: "conv_halo_kernel", 3: "wgrad_tc_kernel", 4: "wgrad_halo_kernel"}
trace_all = False  # with `trace` set: also time the calls that carry no flops / bytes annotation (profiling scripts)
trace = None  # bench.py sets this to a list: (name, start_event, end_event, flops, bytes) per traced call


def call(name, *args, flops=0, nbytes=0, tag=""):
    """Calls a status-returning entry point; raises RuntimeError with the library's message on failure.
    flops / nbytes: algorithmic work of the call, recorded with CUDA events on the current stream when tracing."""
    global launches
    lib = load()
    if trace is not None and (flops or nbytes or trace_all):
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = getattr(lib, name)(*args)
        e1.record()
        kid = lib.pdac_last_kernel_id() if name.startswith(("pdac_conv", "pdac_tapconv")) else 0
        if not tag:  # shape of the first view argument, for per-layer tables of the elementwise passes
            for a in args:
                v = getattr(a, "_obj", None)
                if isinstance(v, PdacView):
                    tag = f"{v.N}x{v.H}x{v.W}x{v.C}"
                    break
        trace.append((KERNEL_NAMES.get(kid, name), e0, e1, flops, nbytes, tag))
    else:
        rc = getattr(lib, name)(*args)
    launches += 1
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {lib.pdac_last_error().decode()}")
    return rc


def peek_device_flag():
    """Polls the pipeline error flag (mapped pinned host memory) WITHOUT synchronising: raises if any kernel that has
    finished so far reported a timed-out TMA / tcgen05 barrier. Called once per step / batch by the product path."""
    lib = load()
    rc = lib.pdac_peek_device_flag()
    if rc != 0:
        raise RuntimeError(f"pdac_b200 device pipeline error ({rc}): {lib.pdac_last_error().decode()}")
