"""pdac_relu_bwd_bias_noalloc: the allocation-free variant of pdac_relu_bwd_bias (include/pdac_b200.h; no counterpart in
the reference — added for callers that cannot cudaMalloc). Called through the C-ABI
with ctypes, checked against the existing entry point with caller scratch on two calls in a row (the second with fewer
channels and other data, so anything left in the shared static array by the first call would show), plus the size limit.
Collected last on purpose (the file name)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _case(n, h, w, c, seed):
    g = torch.Generator().manual_seed(seed)
    dy = torch.randn(n, h, w, c, generator=g).to(torch.bfloat16).to(DEV)
    y = torch.randn(n, h, w, c, generator=g).relu().to(torch.bfloat16).to(DEV)
    return dy, y


def _run_both(lib, ops, dy, y, with_dz, strict_free):
    """-> ((dz, dbias) of pdac_relu_bwd_bias with caller scratch, (dz, dbias) of pdac_relu_bwd_bias_noalloc)."""
    from pdac_pathological_image_segmentation_b200._lib import PDAC_STAT_ROWS
    c = dy.shape[3]
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    scratch = torch.full((PDAC_STAT_ROWS * c,), float("nan"), device=DEV)
    dz_a = torch.empty_like(dy) if with_dz else None
    dz_b = torch.empty_like(dy) if with_dz else None
    db_a = torch.zeros(c, device=DEV)
    db_b = torch.zeros(c, device=DEV)
    vdy, vy = ops.view(dy), ops.view(y)
    vza = C.byref(ops.view(dz_a)) if with_dz else None
    vzb = C.byref(ops.view(dz_b)) if with_dz else None
    torch.cuda.synchronize()
    rc = lib.pdac_relu_bwd_bias(C.byref(vdy), C.byref(vy), vza, C.c_void_p(db_a.data_ptr()),
                                C.c_void_p(scratch.data_ptr()), stream)
    assert rc == 0, lib.pdac_last_error()
    torch.cuda.synchronize()
    torch_before = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved())
    free_before = torch.cuda.mem_get_info()[0]
    rc = lib.pdac_relu_bwd_bias_noalloc(C.byref(vdy), C.byref(vy), vzb, C.c_void_p(db_b.data_ptr()), stream)
    assert rc == 0, lib.pdac_last_error()
    torch.cuda.synchronize()
    assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == torch_before
    if strict_free:     # the driver's view: no device memory taken (the very first call may still load the module lazily)
        assert torch.cuda.mem_get_info()[0] == free_before
    return (dz_a, db_a), (dz_b, db_b)


def test_noalloc_entry_point_matches_the_existing_one_twice_in_a_row():
    from pdac_pathological_image_segmentation_b200 import _lib, ops
    lib = _lib.load()
    # first call: 512 channels (the limit) and enough pixels to use every row of the static table;
    # second call: 16 channels, odd sizes, other data, no dz output
    for i, (shape, with_dz) in enumerate([((4, 96, 96, 512), True), ((3, 37, 53, 16), False), ((2, 64, 64, 64), True)]):
        dy, y = _case(*shape, seed=100 + i)
        (dz_a, db_a), (dz_b, db_b) = _run_both(lib, ops, dy, y, with_dz, strict_free=i > 0)
        # same kernel, same grid, same fixed-order row sum: the two entry points agree bit for bit ...
        assert torch.equal(db_a, db_b), (i, (db_a - db_b).abs().max().item())
        if with_dz:
            assert torch.equal(dz_a.view(torch.int16), dz_b.view(torch.int16)), i
        # ... and both are the ReLU-masked column sum (tolerance of tests/test_elementwise_gpu.py)
        ref = dy.float() * (y > 0)
        want = ref.sum((0, 1, 2))
        assert torch.allclose(db_b, want, rtol=1e-3, atol=1e-2 + 1e-3 * ref.abs().sum((0, 1, 2)).max().item()), i
        if with_dz:
            assert torch.equal(dz_b.float(), ref), i


def test_noalloc_entry_point_accumulates_and_rejects_sizes_beyond_the_limit():
    from pdac_pathological_image_segmentation_b200 import _lib, ops
    lib = _lib.load()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # dbias += sum: a second call on the same dbias doubles it
    dy, y = _case(2, 16, 16, 32, seed=7)
    db = torch.zeros(32, device=DEV)
    for _ in range(2):
        assert lib.pdac_relu_bwd_bias_noalloc(C.byref(ops.view(dy)), C.byref(ops.view(y)), None,
                                              C.c_void_p(db.data_ptr()), stream) == 0, lib.pdac_last_error()
    torch.cuda.synchronize()
    want = 2 * (dy.float() * (y > 0)).sum((0, 1, 2))
    assert torch.allclose(db, want, rtol=1e-3, atol=1e-2)
    # C = 520 > PDAC_NOALLOC_MAX_C = 512: refused with a message that names the limit, nothing launched
    big = torch.zeros(1, 2, 2, 520, dtype=torch.bfloat16, device=DEV)
    dbig = torch.full((520,), 3.0, device=DEV)
    rc = lib.pdac_relu_bwd_bias_noalloc(C.byref(ops.view(big)), None, None, C.c_void_p(dbig.data_ptr()), stream)
    msg = lib.pdac_last_error()
    assert rc != 0 and b"512" in msg and b"520" in msg, (rc, msg)
    torch.cuda.synchronize()
    assert torch.equal(dbig, torch.full_like(dbig, 3.0))
    # missing dbias / dy
    assert lib.pdac_relu_bwd_bias_noalloc(C.byref(ops.view(dy)), None, None, None, stream) != 0
    assert b"required" in lib.pdac_last_error()
    assert lib.pdac_relu_bwd_bias_noalloc(None, None, None, C.c_void_p(db.data_ptr()), stream) != 0
    # shape mismatch between dy and y
    y2 = torch.zeros(2, 16, 8, 32, dtype=torch.bfloat16, device=DEV)
    assert lib.pdac_relu_bwd_bias_noalloc(C.byref(ops.view(dy)), C.byref(ops.view(y2)), None,
                                          C.c_void_p(db.data_ptr()), stream) != 0
    assert b"shape" in lib.pdac_last_error()
