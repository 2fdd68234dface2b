"""Host-side logic of the engine that needs no GPU: the tape's bookkeeping of pooled (half-resolution) gradients of
nearest x2 upsamplings, the per-step pool of BatchNorm partial-sum tables, and the cache of eval-mode BatchNorm affine
tensors. The CUDA ops the tape would call are replaced by torch CPU equivalents of the same arithmetic."""
import pytest
import torch


def _up2_bwd(dy, addend=None, out=None):
    """ops.upsample_nearest2x_bwd on NHWC tensors: (a + b) + (c + d) (+ addend)."""
    s = (dy[:, 0::2, 0::2] + dy[:, 0::2, 1::2]) + (dy[:, 1::2, 0::2] + dy[:, 1::2, 1::2])
    if addend is not None:
        s = s + addend
    if out is not None:
        out.copy_(s)
        return out
    return s


@pytest.fixture
def cpu_ops(monkeypatch):
    from pdac_pathological_image_segmentation_b200 import ops
    monkeypatch.setattr(ops, "upsample_nearest2x_bwd", _up2_bwd)
    monkeypatch.setattr(ops, "add", lambda a, b, relu=False, out=None: out.copy_(a + b) if out is not None else a + b)
    return ops


def _tape():
    from pdac_pathological_image_segmentation_b200.engine import PackCache, Tape
    return Tape(PackCache(), True, None)


def test_pooled_and_full_resolution_gradients_of_an_upsampled_tensor_combine(cpu_ops):
    g = torch.Generator().manual_seed(3)
    y = torch.zeros(1)                       # stands for the upsampled tensor (the key)
    full1, full2 = torch.randn(1, 4, 4, 8, generator=g), torch.randn(1, 4, 4, 8, generator=g)
    pool1, pool2 = torch.randn(1, 2, 2, 8, generator=g), torch.randn(1, 2, 2, 8, generator=g)
    # pooled first, then a full-resolution gradient from a second consumer
    t = _tape()
    t.put_pooled(y, pool1.clone())
    t.put(y, full1.clone())
    got, pooled = t.pop_up2(y)
    assert pooled and torch.allclose(got, pool1 + _up2_bwd(full1))
    assert t.pop_up2(y) == (None, False)
    # full resolution first, then pooled; then another pooled
    t = _tape()
    t.put(y, full1.clone())
    t.put_pooled(y, pool1.clone())
    t.put_pooled(y, pool2.clone())
    got, pooled = t.pop_up2(y)
    assert pooled and torch.allclose(got, _up2_bwd(full1) + pool1 + pool2)
    # only full-resolution gradients: nothing is pooled behind the consumer's back
    t = _tape()
    t.put(y, full1.clone())
    t.put(y, full2.clone())
    got, pooled = t.pop_up2(y)
    assert not pooled and torch.allclose(got, full1 + full2)
    # a tensor that wants no gradient stays without one
    t = _tape()
    t.no_grad_for(y)
    t.put_pooled(y, pool1.clone())
    assert t.pop_up2(y) == (None, False)
    # backward() forgets the pooled marks with the gradients
    t = _tape()
    t.enabled = True
    t.put_pooled(y, pool1.clone())
    t.backward()
    assert not t.pooled and not t.grads


def test_statistics_tables_come_from_one_zeroed_pool_while_it_lasts():
    from pdac_pathological_image_segmentation_b200 import ops
    t = _tape()
    dev = torch.device("cpu")
    a = t.new_stats(16, dev)                 # no pool yet: separate zeroed tables, demand recorded
    assert a.shape == (ops.STAT_ROWS, 2, 16) and not a.any() and t.stats_used == ops.STAT_ROWS * 2 * 16
    t2 = _tape()
    t2.stats_pool = torch.zeros(ops.STAT_ROWS * 2 * (16 + 32), dtype=torch.float32)
    b, c = t2.new_stats(16, dev), t2.new_stats(32, dev)
    assert b.is_contiguous() and c.is_contiguous() and b.shape == (ops.STAT_ROWS, 2, 16) and c.shape == (ops.STAT_ROWS, 2, 32)
    assert b.data_ptr() == t2.stats_pool.data_ptr()
    assert c.data_ptr() == t2.stats_pool.data_ptr() + 4 * ops.STAT_ROWS * 2 * 16       # adjacent, no overlap
    d = t2.new_stats(64, dev)                # pool exhausted (the model grew): falls back to its own table
    assert d.shape == (ops.STAT_ROWS, 2, 64) and not d.any()
    assert d.data_ptr() < t2.stats_pool.data_ptr() or d.data_ptr() >= t2.stats_pool.data_ptr() + 4 * t2.stats_pool.numel()
    assert t2.stats_used == ops.STAT_ROWS * 2 * (16 + 32 + 64)                         # next step's pool covers all three


def test_eval_batchnorm_affine_is_cached_per_parameter_version():
    from pdac_pathological_image_segmentation_b200.containers import BatchNorm2d
    from pdac_pathological_image_segmentation_b200.engine import PackCache
    bn = BatchNorm2d(8)
    calls = []

    def compute(ss, mi):
        calls.append(1)
        inv = torch.rsqrt(bn.running_var + bn.eps)
        ss[0].copy_(bn.weight.detach() * inv)
        ss[1].copy_(bn.bias.detach() - bn.running_mean * ss[0])
        mi[0].copy_(bn.running_mean)
        mi[1].copy_(inv)

    packs = PackCache()
    ss, mi = packs.bn_eval(bn, compute)
    ss_again, _ = packs.bn_eval(bn, compute)
    assert len(calls) == 1 and ss_again is ss                      # second forward: no launch
    with torch.no_grad():
        bn.running_mean.add_(1.0)                                  # a checkpoint load / torch-side update bumps the version
    ss3, _ = packs.bn_eval(bn, compute)
    assert len(calls) == 2 and ss3 is ss                           # recomputed IN PLACE (a captured graph keeps its pointers)
    assert torch.allclose(ss[1], bn.bias.detach() - bn.running_mean * ss[0])
    packs.drop_bn(bn)                                              # a training-mode step updated the statistics by raw pointer
    packs.bn_eval(bn, compute)
    assert len(calls) == 3
