"""CPU-only checks of the drop-in boundary: libpdac_b200.so loads without a GPU, exports every symbol that
include/pdac_b200.h declares (and the ctypes table binds the same set), rejects host pointers / missing devices with
an error instead of falling back, and the Python modules keep the reference's constructor / state_dict contracts."""
import ctypes
import os
import re
import warnings

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pdac_b200.h")


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    from pdac_pathological_image_segmentation_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        g.build()
    return _lib.load()


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pdac_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from pdac_pathological_image_segmentation_b200 import _lib
    names = _declared()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/pdac_b200.h but not exported by libpdac_b200.so"
    bound = set(_lib.SIGNATURES) | {"pdac_version", "pdac_last_error", "pdac_last_kernel_id", "pdac_peek_device_flag"}
    assert bound == set(names), f"ctypes table and header disagree: {bound ^ set(names)}"
    assert lib.pdac_version() >= 100


def test_struct_layouts_match_header():
    from pdac_pathological_image_segmentation_b200._lib import PdacConv, PdacTap, PdacTapConv, PdacView
    assert ctypes.sizeof(PdacView) == 8 + 4 * 4 + 3 * 8
    assert ctypes.sizeof(PdacConv) == 8 * 4
    assert ctypes.sizeof(PdacTap) == 12
    assert ctypes.sizeof(PdacTapConv) == 16 + 16 * 12


def test_no_cpu_fallback_at_the_abi(lib):
    """Host pointers / no device: every compute entry returns an error code and a message (never computes)."""
    from pdac_pathological_image_segmentation_b200._lib import PdacView
    x = torch.zeros(1, 4, 4, 16, dtype=torch.bfloat16)
    v = PdacView(x.data_ptr(), 1, 4, 4, 16, 256, 64, 16)
    rc = lib.pdac_add(ctypes.byref(v), ctypes.byref(v), 0, ctypes.byref(v), None)
    assert rc != 0
    assert len(lib.pdac_last_error()) > 0
    rc = lib.pdac_dice_fwd(None, None, 0, 0, None, ctypes.c_float(1e-6), None, None)
    assert rc == -1 and b"null" in lib.pdac_last_error()
    # argument validation of the overlay entry point happens before any launch
    assert lib.pdac_overlay_blend(None, None, 16, 255, 0, 0, 300, None, None) != 0 and b"alpha" in lib.pdac_last_error()
    assert lib.pdac_overlay_blend(None, None, 16, 256, 0, 0, 102, None, None) != 0 and b"colour" in lib.pdac_last_error()
    assert lib.pdac_overlay_blend(None, None, 16, 255, 0, 0, 102, None, None) != 0 and b"null" in lib.pdac_last_error()
    assert lib.pdac_overlay_blend(None, None, 0, 255, 0, 0, 102, None, None) == 0      # empty band: nothing to do


def test_modules_refuse_cpu_tensors():
    from pdac_pathological_image_segmentation_b200.resunet import ResUNet
    from pdac_pathological_image_segmentation_b200.util import DiceLoss
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m = ResUNet(1, 64)
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(1, 3, 64, 64))
    with pytest.raises(RuntimeError, match="CUDA"):
        DiceLoss()(torch.zeros(4), torch.zeros(4))
    with pytest.raises(RuntimeError, match="container"):
        m.conv1(torch.zeros(1, 512, 4, 4))


def test_state_dict_layouts_match_reference_and_smp_restatement():
    from oracle import smp_oracle as S
    from pdac_pathological_image_segmentation_b200 import smp as P
    from pdac_pathological_image_segmentation_b200.resunet import ResUNet
    gold = torch.load(os.path.join(ROOT, "tests", "golden", "state_keys.pt"), weights_only=False)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for nc in (1, 2):
            m = ResUNet(nc, 512)
            got = [(k, tuple(v.shape), str(v.dtype)) for k, v in m.state_dict().items()]
            want = [(k, (nc,) + tuple(s[1:]) if k.startswith("conv4") else tuple(s), d) for k, s, d in gold]
            assert got == want
    pairs = [("unet_smp", "Unet", {})]
    for extra in (("fpn", "FPN", {}), ("pspnet", "PSPNet", {"psp_use_batchnorm": True}),
                  ("deeplabv3+", "DeepLabV3Plus", {"decoder_atrous_rates": (3, 6, 9)})):
        if hasattr(P, extra[1]):
            pairs.append(extra)
    for oname, pname, kw in pairs:
        ref = S.build(oname)
        ours = getattr(P, pname)(encoder_name="resnet18", encoder_weights=None, in_channels=3, classes=1,
                                 activation=None, **kw)
        assert [(k, tuple(v.shape), v.dtype) for k, v in ours.state_dict().items()] == \
               [(k, tuple(v.shape), v.dtype) for k, v in ref.state_dict().items()], pname
        ours.load_state_dict(ref.state_dict(), strict=True)


def test_reference_constructor_calls_are_accepted():
    """train_worker.py:150-182 / test.py:79-111 keyword sets."""
    from pdac_pathological_image_segmentation_b200 import smp as P
    with pytest.raises(RuntimeError, match="ImageNet"):
        P.Unet(encoder_name="resnet18", encoder_weights="imagenet", in_channels=3, classes=1, activation=None)
    with pytest.raises(ValueError):
        P.Unet(encoder_name="resnet50", encoder_weights=None)
    with pytest.raises(ValueError):
        P.Unet(encoder_name="resnet18", encoder_weights=None, activation="sigmoid")


def test_synthetic_generators_agree():
    from oracle import resunet_oracle as R
    from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
    a, b = R.synthetic_batch(2, 64, seed=5), synthetic_batch(2, 64, seed=5)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def test_reference_checkpoint_file_round_trip(tmp_path):
    """A .pth as train_worker.py:116-125 writes it ({"model", "optimizer", "epoch", "previous_best"}, DDP "module." key
    prefix) loads through the loop of train_worker.py:185-196 / test.py:118-124 into our module, strictly and
    value-for-value, and Adam's state loads onto our parameters."""
    from collections import OrderedDict
    from oracle import resunet_oracle as R
    from pdac_pathological_image_segmentation_b200.resunet import ResUNet
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        ours = ResUNet(1, 64)
    state = R.make_state(1, seed=3)                                   # reference key order / shapes (models/resunet.py)
    opt = torch.optim.Adam(ours.parameters(), lr=1e-4)
    ckpt = {"model": OrderedDict(("module." + k, v.clone()) for k, v in state.items()),
            "optimizer": opt.state_dict(), "epoch": 7, "previous_best": 0.5}
    path = str(tmp_path / "latest.pth")
    torch.save(ckpt, path)
    loaded = torch.load(path, weights_only=False)
    new_state = OrderedDict()
    for n, v in loaded["model"].items():
        new_state[n.replace("module.", "")] = v
    ours.load_state_dict(new_state)                                   # strict, as the reference calls it
    for (k, v), (k2, v2) in zip(ours.state_dict().items(), ckpt["model"].items()):
        assert "module." + k == k2 and torch.equal(v, v2), k
    torch.optim.Adam(ours.parameters(), lr=1e-4).load_state_dict(loaded["optimizer"])
    assert loaded["epoch"] == 7 and loaded["previous_best"] == 0.5


def test_ctypes_signatures_have_the_declared_number_of_arguments():
    """Every prototype in include/pdac_b200.h against the ctypes table: same number of parameters, pointer parameters
    bound as pointers and integer / floating ones as scalars (a wrong count or kind corrupts the call silently)."""
    import ctypes as C
    from pdac_pathological_image_segmentation_b200 import _lib
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    protos = dict(re.findall(r"\bint\s+(pdac_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S))
    checked = 0
    for name, argtypes in _lib.SIGNATURES.items():
        assert name in protos, f"{name} bound in _lib.py but not declared as `int {name}(...)` in the header"
        params = [p.strip() for p in protos[name].replace("\n", " ").split(",") if p.strip() and p.strip() != "void"]
        assert len(params) == len(argtypes), f"{name}: header has {len(params)} parameters, ctypes binds {len(argtypes)}"
        for i, (decl, ct) in enumerate(zip(params, argtypes)):
            is_ptr_decl = "*" in decl
            is_ptr_ct = ct in (C.c_void_p, C.c_char_p) or hasattr(ct, "contents") or ct.__name__.startswith("LP_")
            assert is_ptr_decl == is_ptr_ct, f"{name} argument {i} `{decl}` bound as {ct.__name__}"
            if not is_ptr_decl:
                want_float = decl.split()[0] in ("float", "double")
                assert (ct in (C.c_float, C.c_double)) == want_float, f"{name} argument {i} `{decl}` bound as {ct.__name__}"
                if want_float:
                    assert (ct is C.c_double) == decl.startswith("double"), f"{name} argument {i} `{decl}` bound as {ct.__name__}"
        checked += 1
    assert checked >= 55
