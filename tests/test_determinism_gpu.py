"""Reproducibility (the reference sets cudnn.deterministic = True, train_main.py:36-43): no floating-point atomics on the
training path, so identical inputs give bit-identical statistics, gradients and parameter updates — across launches,
between the fused TrainStep harness and the drop-in module under torch autograd, and between eager execution and the
CUDA-graph replay of the whole step."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _unet(seed=3, train=True):
    from pdac_pathological_image_segmentation_b200.smp import Unet
    torch.manual_seed(seed)
    return Unet(encoder_weights=None).to(DEV).train(train)


def _batch(n, size, seed):
    from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
    img, mask = synthetic_batch(n, size, seed=seed)
    return img.to(DEV), mask.to(DEV)


def test_kernel_reductions_are_bit_reproducible():
    """Conv epilogue statistics, split-K weight gradients (halo and generic kernels), BN backward sums, bias gradient and
    Dice sums: two launches on the same data give identical bits (many CTAs / split-K shares per output)."""
    from pdac_pathological_image_segmentation_b200 import ops
    assert ops.deterministic
    g = torch.Generator(device=DEV).manual_seed(1)
    x = torch.randn(8, 96, 96, 64, device=DEV, generator=g).to(torch.bfloat16)
    dy = torch.randn(8, 96, 96, 64, device=DEV, generator=g).to(torch.bfloat16)
    w = torch.randn(64, 64, 3, 3, device=DEV, generator=g) * 0.05
    wp, _ = ops.pack_conv_weight(w, want_dgrad=False)

    def conv_stats():
        st = ops.new_stats(64, x.device)
        ops.conv2d_fwd([x], wp, 64, 3, 3, 1, 1, 1, stats=st)
        ss, mi = torch.empty(2, 64, device=DEV), torch.empty(2, 64, device=DEV)
        ops.bn_prepare(st, x.shape[0] * 96 * 96, None, None, 1e-5, 0.1, True, None, None, ss, mi)
        return torch.cat([ss.flatten(), mi.flatten()])

    def wgrad3():
        return ops.conv2d_wgrad([x], dy, 3, 3, 1, 1, 1, oihw=True)

    def wgrad_s2():          # stride 2 -> generic split-K kernel
        return ops.conv2d_wgrad([x], dy[:, ::2, ::2].contiguous(), 3, 3, 2, 1, 1, oihw=True)

    def bn_sums():
        mi = torch.stack([torch.zeros(64, device=DEV), torch.ones(64, device=DEV)])
        return ops.bn_bwd_reduce(dy, None, x, mi)[0].clone()

    def dbias():
        db = torch.zeros(64, device=DEV)
        ops.relu_bwd_bias(dy, None, None, db)
        return db

    def dice():
        lg = x.float().reshape(-1)
        loss, sums = ops.dice_fwd(lg, (dy.float().reshape(-1) > 0).to(torch.uint8))
        return torch.cat([loss.reshape(1).double(), sums[:3]])

    for name, fn in (("conv stats", conv_stats), ("wgrad halo", wgrad3), ("wgrad generic", wgrad_s2),
                     ("bn bwd sums", bn_sums), ("bias grad", dbias), ("dice sums", dice)):
        a = fn().clone()
        for _ in range(3):
            assert torch.equal(a, fn()), f"{name}: not bit-reproducible"
    ops.check_device_flag()


@pytest.mark.parametrize("size,batch", [(128, 4), (256, 2)])
def test_identical_train_steps_are_bit_identical(size, batch):
    """Two TrainStep runs from the same seed on the same batch: losses, gradient arenas and updated parameters are
    bit-identical (fp32 atomics made them differ by 0.28-0.38 relative L2)."""
    from pdac_pathological_image_segmentation_b200.trainer import TrainStep
    img, mask = _batch(batch, size, 7)
    runs = []
    for _ in range(2):
        ts = TrainStep(_unet(), lr=1e-3)
        losses = []
        for _ in range(3):
            loss, score = ts.step(img, mask)
            losses.append((loss.item(), score.item()))
        runs.append((losses, ts.arena.flat.clone(), ts.flat_params.clone(), ts.exp_avg_sq.clone()))
    assert runs[0][0] == runs[1][0]
    assert runs[0][1].abs().max().item() > 0
    for a, b in zip(runs[0][1:], runs[1][1:]):
        assert torch.equal(a, b)


def test_fused_upsampling_backward_leaves_every_gradient_bit_identical():
    """The decoder's nearest x2 backward formed inside the consuming conv's data-gradient kernel
    (pdac_conv2d_dgrad_pool2) against the two-kernel form (PDAC_FUSE_UP2_BWD=0): same bf16 roundings, same summation
    order -> the same loss and the same gradient arena bit for bit; and the fused form really ran."""
    from pdac_pathological_image_segmentation_b200 import _lib, ops
    from pdac_pathological_image_segmentation_b200.trainer import TrainStep
    img, mask = _batch(2, 128, 9)
    runs = []
    for fuse in (True, False):
        old = ops.fuse_up2_bwd
        ops.fuse_up2_bwd = fuse
        try:
            ts = TrainStep(_unet(), lr=1e-3)
            _lib.trace, _lib.trace_all = [], True
            loss, _ = ts.step(img, mask)
            names = [t[0] if t[0].startswith("pdac_") else t[5] for t in _lib.trace]
        finally:
            _lib.trace, _lib.trace_all = None, False
            ops.fuse_up2_bwd = old
        tags = [t for t in names if t.startswith("dgrad+pool")]
        n_up2 = sum(1 for t in names if t == "pdac_upsample_nearest2x_bwd")
        runs.append((loss.item(), ts.arena.flat.clone(), len(tags), n_up2))
    assert runs[0][2] == 5 and runs[0][3] == 0, runs[0][2:]       # five decoder blocks, no separate pooling pass left
    assert runs[1][2] == 0 and runs[1][3] == 5, runs[1][2:]
    assert runs[0][0] == runs[1][0]
    assert runs[0][1].abs().max().item() > 0 and torch.equal(runs[0][1], runs[1][1])


def test_trainstep_gradients_equal_autograd_path_bitwise():
    """The fused harness (uint8 patches -> ... -> Adam) and the drop-in module under torch autograd + torch.optim.Adam
    launch the same kernels on the same prepared input: same loss, the same gradients bit for bit (train mode, batch
    statistics), parameters after Adam within float rounding of torch.optim.Adam (train_worker.py:68-80)."""
    from pdac_pathological_image_segmentation_b200 import ops
    from pdac_pathological_image_segmentation_b200.module_base import _NetFn
    from pdac_pathological_image_segmentation_b200.trainer import TrainStep
    from pdac_pathological_image_segmentation_b200.util import DiceLoss
    m1 = _unet()
    m2 = copy.deepcopy(m1)
    img, mask = _batch(2, 128, 7)
    ts = TrainStep(m1, lr=1e-3)
    before = ts.flat_params.clone()
    loss1, score1 = ts.step(img, mask)
    g1 = {k: ts.arena.slot(p).clone() for k, p in m1.named_parameters()}
    z, _ = ops.stem_prep_u8(img, mask, None)
    opt = torch.optim.Adam(m2.parameters(), lr=1e-3)
    out = _NetFn.apply(m2, z, *m2._trainable())
    loss2 = DiceLoss()(torch.squeeze(out), mask.float())
    loss2.backward()
    opt.step()
    assert loss1.item() == loss2.item()
    assert 0.0 <= score1.item() <= 1.0
    for k, p2 in m2.named_parameters():
        assert torch.equal(g1[k], p2.grad), k
    moved = (ts.flat_params - before).abs().max().item()
    assert 0.5e-3 <= moved <= 1.01e-3      # Adam's first step moves weights by ~lr
    for (k, p1), (_, p2) in zip(m1.named_parameters(), m2.named_parameters()):
        frac = ((p1 - p2).abs() > 2e-4).float().mean().item()   # sign flips of ~0 gradients only
        assert frac <= 0.02, (k, frac)


@pytest.mark.parametrize("train", [True, False])
def test_cuda_graph_replay_equals_eager_steps_bitwise(train):
    """TrainStep(graph=True) captures the whole step (prep, forward, loss, backward, Adam with its device-side step
    counter and learning rate) once and replays it. Six steps over three batches, with a learning-rate change on the
    way, leave bit-identical parameters, moments, BatchNorm buffers and losses in the graph and the eager trainer."""
    from pdac_pathological_image_segmentation_b200.trainer import TrainStep
    m1 = _unet(11, train)
    m2 = copy.deepcopy(m1)
    tg, te = TrainStep(m1, lr=1e-4, graph=True), TrainStep(m2, lr=1e-4)
    batches = [_batch(2, 128, 20 + i) for i in range(3)]
    for it in range(6):                         # steps 0-1 eager warm-up, step 2 captures + replays, 3-5 replay
        if it == 4:
            tg.lr = te.lr = 3e-5                # what ReduceLROnPlateau does through optimizer.param_groups
        lg, sg = tg.step(*batches[it % 3])
        le, se = te.step(*batches[it % 3])
        assert lg.item() == le.item() and sg.item() == se.item(), it
    assert tg._g is not None
    assert tg.step_count == te.step_count == 6
    assert torch.equal(tg.flat_params, te.flat_params)
    assert torch.equal(tg.exp_avg, te.exp_avg) and torch.equal(tg.exp_avg_sq, te.exp_avg_sq)
    for (k, a), (_, b) in zip(m1.state_dict().items(), m2.state_dict().items()):
        assert torch.equal(a, b), k
