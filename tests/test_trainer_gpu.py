"""The fast training loop's reference-format state I/O and validation hook:
train_worker.py:113-125 saves {"model", "optimizer", "epoch", "previous_best"}, :186-201 restores it, :240 drives the
learning rate with ReduceLROnPlateau, :16-37 validates with an eval-mode Dice, :203-206 wraps the model in
DistributedDataParallel(find_unused_parameters=True)."""
import copy
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _model(seed=3):
    import warnings
    from pdac_pathological_image_segmentation_b200.resunet import ResUNet
    torch.manual_seed(seed)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return ResUNet(num_classes=1, output_size=64).to(DEV).train()


def _batch(n, size, seed):
    from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
    img, mask = synthetic_batch(n, size, seed=seed)
    return img.to(DEV), mask.to(DEV)


def test_optimizer_state_round_trips_with_torch_adam(tmp_path):
    """Three real optimizer steps on both sides from the same weights: TrainStep's arena moments / step count, written in
    torch.optim.Adam's state_dict layout, equal torch.optim.Adam's own state after the same three steps on the drop-in
    module; the checkpoint file loads into a fresh torch.optim.Adam and into a fresh TrainStep, which then continue
    identically."""
    from pdac_pathological_image_segmentation_b200 import ops
    from pdac_pathological_image_segmentation_b200.module_base import _NetFn
    from pdac_pathological_image_segmentation_b200.trainer import TrainStep
    from pdac_pathological_image_segmentation_b200.util import DiceLoss
    m1 = _model()
    m2 = copy.deepcopy(m1)
    ts = TrainStep(m1, lr=1e-3)
    opt = torch.optim.Adam(m2.parameters(), lr=1e-3)          # the reference's optimizer (train_worker.py:184)
    batches = [_batch(2, 64, 30 + i) for i in range(3)]
    for it, (img, mask) in enumerate(batches):
        ts.step(img, mask)
        z, _ = ops.stem_prep_u8(img, mask, None)
        opt.zero_grad()
        out = _NetFn.apply(m2, z, *m2._trainable())
        DiceLoss()(torch.squeeze(out), mask.float()).backward()
        opt.step()
        sd1, sd2 = ts.optimizer.state_dict(), opt.state_dict()
        assert sd1["param_groups"][0]["params"] == sd2["param_groups"][0]["params"] == list(range(76))
        assert set(sd1["state"]) == set(sd2["state"]) and len(sd1["state"]) == 74    # encoder.fc.* never get state
        for i in sd2["state"]:
            assert float(sd1["state"][i]["step"]) == float(sd2["state"][i]["step"]) == it + 1.0
            for k in ("exp_avg", "exp_avg_sq"):
                a, b = sd1["state"][i][k], sd2["state"][i][k]
                assert a.shape == b.shape and a.dtype == b.dtype
                if it == 0:     # same weights, bit-identical gradients: the moments agree to float rounding
                    assert torch.allclose(a, b, rtol=1e-5, atol=1e-12), (i, k, (a - b).abs().max().item())
                else:           # the two Adam implementations round differently and the random-init network amplifies
                    # rounding noise (tests/test_smp_gpu.py): later steps only have to stay on the same trajectory
                    assert ((a - b).norm() / (b.norm() + 1e-30)).item() <= 0.5, (it, i, k)
    # reference checkpoint file (train_worker.py:116-125) -> the reference's own resume loop (:186-201) with torch Adam
    path = str(tmp_path / "latest.pth")
    torch.save(ts.state_dict(epoch=5, previous_best=0.61), path)
    ck = torch.load(path, map_location=DEV, weights_only=False)
    assert set(ck) == {"model", "optimizer", "epoch", "previous_best"} and all(k.startswith("module.") for k in ck["model"])
    m3 = _model(seed=99)
    m3.load_state_dict({k.replace("module.", ""): v for k, v in ck["model"].items()})
    opt3 = torch.optim.Adam(m3.parameters(), lr=1e-4)
    opt3.load_state_dict(ck["optimizer"])
    for state in opt3.state.values():
        for k, v in state.items():
            if torch.is_tensor(v):
                state[k] = v.cuda(0) if k != "step" else v
    assert opt3.param_groups[0]["lr"] == 1e-3 and len(opt3.state) == 74
    # ... and into a fresh TrainStep: the fourth step of the resumed run equals the fourth step of the original
    m4 = _model(seed=77)
    ts4 = TrainStep(m4, lr=5e-5)
    epoch, best = ts4.load_state_dict(ck)
    assert (epoch, best) == (5, 0.61) and ts4.step_count == 3 and ts4.lr == 1e-3
    assert torch.equal(ts4.flat_params, ts.flat_params) and torch.equal(ts4.exp_avg, ts.exp_avg)
    img, mask = _batch(2, 64, 40)
    la, _ = ts.step(img, mask)
    lb, _ = ts4.step(img, mask)
    assert la.item() == lb.item() and torch.equal(ts4.flat_params, ts.flat_params)
    # a torch.optim.Adam state saved by the reference loads into TrainStep too
    ts5 = TrainStep(_model(seed=55), lr=1e-4)
    ts5.optimizer.load_state_dict(sd2)
    assert ts5.step_count == 3
    off, n, shape = ts5.arena.offsets[id(ts5.params[0])]
    assert torch.equal(ts5.exp_avg[off:off + n].view(shape), sd2["state"][0]["exp_avg"])


def test_reduce_lr_on_plateau_drives_the_fused_optimizer():
    """train_worker.py:240: scheduler = ReduceLROnPlateau(optimizer, 'max', ...); scheduler.step(score)."""
    from pdac_pathological_image_segmentation_b200.trainer import TrainStep
    ts = TrainStep(_model(), lr=1e-3, graph=True)
    sched = torch.optim.lr_scheduler.ReduceLROnPlateau(ts.optimizer, mode="max", factor=0.1, patience=0)
    img, mask = _batch(2, 64, 1)
    for _ in range(4):
        ts.step(img, mask)
    sched.step(0.5)
    sched.step(0.4)                               # no improvement -> lr * 0.1
    assert abs(ts.lr - 1e-4) < 1e-12
    before = ts.flat_params.clone()
    ts.step(img, mask)                            # replayed graph reads the new learning rate from device memory
    torch.cuda.synchronize()
    assert abs(ts.hyper[0].item() - 1e-4) < 1e-10
    moved = (ts.flat_params - before).abs().max().item()
    assert 0 < moved <= 1.2e-4, moved


def test_validate_is_the_eval_mode_dice():
    from oracle import util_oracle as U
    from pdac_pathological_image_segmentation_b200 import ops
    from pdac_pathological_image_segmentation_b200.trainer import TrainStep
    m = _model()
    ts = TrainStep(m, lr=1e-3)
    img, mask = _batch(4, 64, 2)
    ts.step(img, mask)
    score = ts.validate(img, mask)
    assert m.training                              # mode restored
    m.eval()
    from pdac_pathological_image_segmentation_b200.engine import Tape
    with torch.no_grad():
        z, _ = ops.stem_prep_u8(img, mask, None)
        out, _ = m._run(Tape(m._packs, False, None), z)
    pred = torch.ge(torch.squeeze(out).sigmoid(), 0.5).float()
    want = U.dice_score(pred.cpu(), mask.cpu())
    assert abs(score.item() - want.item()) <= 1e-6


def test_loss_curve_tracks_the_fp32_oracle():
    """Loss-curve parity: 60 Adam steps of the reference loop (train_worker.py:68-80) on one fixed batch,
    the fp32 torch oracle of the smp Unet with torch.optim.Adam on one side and the captured-graph TrainStep on the other,
    from the same weights. Individual steps diverge (bf16 activations, a chaotic random-init network: see
    test_gradients_match_fp32_when_well_conditioned), so the test compares what a user of the loop sees: the first loss,
    that both runs learn, the level of the curve half way and at the end, and the eval-mode Dice afterwards."""
    from oracle import resunet_oracle as R
    from oracle import smp_oracle as S
    from oracle import util_oracle as U
    from pdac_pathological_image_segmentation_b200 import smp as P
    from pdac_pathological_image_segmentation_b200.trainer import TrainStep
    torch.manual_seed(17)
    ref = S.build("unet_smp", classes=1, dropout=0.0)
    ours = P.Unet(encoder_name="resnet18", encoder_weights=None, in_channels=3, classes=1, activation=None)
    ours.load_state_dict(ref.state_dict(), strict=True)
    ref, ours = ref.to(DEV).train(), ours.to(DEV).train()
    img, mask = R.synthetic_batch(8, 64, seed=11)
    img = torch.where(mask[..., None].bool(), img // 2 + 128, img // 2).to(torch.uint8)   # a learnable signal: bright = tumour
    x = R.normalize(img).to(DEV)
    img_d, mask_d = img.to(DEV), mask.to(DEV)
    opt = torch.optim.Adam(ref.parameters(), lr=1e-3)
    ts = TrainStep(ours, lr=1e-3, graph=True)
    la, lb = [], []
    for _ in range(60):
        _, loss, _ = U.train_step(ref, opt, x, mask_d)
        la.append(loss.item())
        loss_o, _ = ts.step(img_d, mask_d)
        lb.append(loss_o.item())
    mean = lambda v: sum(v) / len(v)
    assert abs(la[0] - lb[0]) <= 5e-3, (la[0], lb[0])
    assert mean(la[-10:]) <= 0.6 * la[0] and mean(lb[-10:]) <= 0.6 * lb[0], (la[0], la[-10:], lb[0], lb[-10:])
    assert abs(mean(la[20:30]) - mean(lb[20:30])) <= 0.08, (la[20:30], lb[20:30])
    assert abs(mean(la[-10:]) - mean(lb[-10:])) <= 0.05, (la[-10:], lb[-10:])
    ref.eval()
    with torch.no_grad():
        want = U.dice_score(U.sigmoid_mask(torch.squeeze(ref(x))), mask_d.float()).item()
    got = ts.validate(img_d, mask_d).item()
    assert abs(want - got) <= 0.1, (want, got)
    ts.release_graph()


# ------------------------------------------------------------------------------------------------ 2 GPUs (NCCL)
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _ddp_worker(rank, world, port, out):
    import torch.distributed as dist
    from pdac_pathological_image_segmentation_b200.smp import Unet
    from pdac_pathological_image_segmentation_b200.trainer import TrainStep
    from pdac_pathological_image_segmentation_b200.util import DiceLoss
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    tg = None
    try:
        from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
        img, mask = (t.to(dev) for t in synthetic_batch(2, 128, seed=50 + rank))
        res = {}
        # (1) fused harness, overlapped bucketed allreduce: arena == sum over ranks of the local gradients
        torch.manual_seed(3)
        m = Unet(encoder_weights=None).to(dev).train()
        ts = TrainStep(m, lr=1e-3, bucket_mb=8.0)
        ts.forward_backward(img, mask, reduce=False)
        local = ts.arena.flat.clone()
        gathered = [torch.empty_like(local) for _ in range(world)]
        dist.all_gather(gathered, local)
        want = gathered[0] + gathered[1]
        for bn in m.modules():                       # undo the BatchNorm side effects of the local pass
            if hasattr(bn, "num_batches_tracked"):
                bn.num_batches_tracked.zero_()
        ts.forward_backward(img, mask, reduce=True)
        torch.cuda.synchronize()
        res["sum_ok"] = bool(torch.equal(ts.arena.flat, want))
        res["buckets"] = len(ts.reducer.spans)
        res["order"] = list(ts.reducer.launch_order)
        # (2) whole-step graph with the collectives inside: parameters stay identical on both ranks
        tg = TrainStep(copy.deepcopy(m), lr=1e-3, bucket_mb=8.0, graph=True)
        for _ in range(5):
            tg.step(img, mask)
        torch.cuda.synchronize()
        flat = [torch.empty_like(tg.flat_params) for _ in range(world)]
        dist.all_gather(flat, tg.flat_params)
        res["graph_params_equal"] = bool(torch.equal(flat[0], flat[1]))
        res["graph_comm_in_graph"] = bool(tg._comm_in_graph and tg._g is not None)
        # (3) the drop-in module under torch's own DistributedDataParallel, as train_worker.py:203-206 wraps it
        torch.manual_seed(3)
        m2 = Unet(encoder_weights=None).to(dev).train()
        ddp = torch.nn.parallel.DistributedDataParallel(m2, device_ids=[rank], find_unused_parameters=True)
        opt = torch.optim.Adam(ddp.parameters(), lr=1e-3)
        from pdac_pathological_image_segmentation_b200 import ops
        x = torch.empty(2, 3, 128, 128, device=dev).normal_()
        logits = ddp(x)
        DiceLoss()(torch.squeeze(logits), mask.float()).backward()
        g = getattr(m2.segmentation_head, "0").weight.grad
        gs = [torch.empty_like(g) for _ in range(world)]
        dist.all_gather(gs, g.contiguous())
        res["ddp_grads_equal"] = bool(torch.equal(gs[0], gs[1]))      # DDP averaged them
        opt.step()
        res["ddp_finite"] = bool(all(torch.isfinite(p).all() for p in m2.parameters()))
        ops.check_device_flag()
        out[rank] = dict(res)
    finally:
        if tg is not None:
            tg.release_graph()      # also on failure: a live graph with captured collectives blocks the teardown
        torch.cuda.synchronize()
        dist.destroy_process_group()


@pytest.mark.timeout(240)
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (NCCL)")
def test_two_gpu_allreduce_overlap_graph_and_ddp_wrap():
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_ddp_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    for r in (0, 1):
        res = out[r]
        assert res["sum_ok"], res
        # every bucket exactly once; the last-registered parameters' bucket goes first (backward reaches them first)
        assert res["buckets"] >= 3 and sorted(res["order"]) == list(range(res["buckets"])) and res["order"][0] == 0, res
        assert res["graph_params_equal"], res
        assert res["ddp_grads_equal"] and res["ddp_finite"], res
