"""Host-side logic of the data-parallel path on CPU: bucket construction (DDP order / size rules) and the bucketed
gradient reduction over a world_size-2 gloo group, checked against torch's own DistributedDataParallel averaging."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pdac_pathological_image_segmentation_b200.ddp import BucketReducer, make_buckets
from pdac_pathological_image_segmentation_b200.engine import GradArena


def test_buckets_tile_arena_in_reverse_registration_order():
    sizes = [9408, 64, 64, 36864, 64, 64, 2359296, 512, 512, 16, 1]
    params = [torch.nn.Parameter(torch.zeros(n)) for n in sizes]
    arena = GradArena(params, "cpu")
    offs = [arena.offsets[id(p)][0] for p in params]
    assert offs == sorted(offs) and all(o % 64 == 0 for o in offs)
    assert arena.slot(params[-1], padded=True).numel() == 16          # >= 16 floats for padded-bias reductions
    for mb in (0.01, 0.5, 25.0):
        spans = make_buckets(offs, arena.size, mb)
        assert spans[0][1] == arena.size and spans[-1][0] == 0       # last parameters first
        for (lo, hi), (lo2, hi2) in zip(spans, spans[1:]):
            assert hi2 == lo and lo2 < hi2
        cap = int(mb * 1024 * 1024 / 4)
        for lo, hi in spans:
            inside = [o for o in offs if lo <= o < hi]
            assert inside and (hi - lo <= cap or len(inside) == 1)   # only a single oversized tensor may exceed the cap
    assert len(make_buckets(offs, arena.size, 25.0)) == 1
    assert make_buckets([], 0) == []
    # a capped tail: the last span (first registered parameters, complete only when backward ends) is split off small
    spans = make_buckets(offs, arena.size, 25.0, tail_mb=0.5)
    assert len(spans) == 2 and spans[0][1] == arena.size and spans[0][0] == spans[1][1] and spans[1][0] == 0
    assert 0 < (spans[1][1] - spans[1][0]) * 4 <= 0.5 * 1024 * 1024 and spans[1][1] in offs
    assert make_buckets(offs, arena.size, 25.0, tail_mb=1000.0) == make_buckets(offs, arena.size, 25.0)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        net = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 16), torch.nn.ReLU(),
                                  torch.nn.Linear(16, 1))
        unused = torch.nn.Linear(4, 4)                  # like ResUNet's encoder.fc: registered, never used
        params = list(net.parameters())
        x = torch.randn(4, 8, generator=torch.Generator().manual_seed(100 + rank))
        # torch DDP reference (averages gradients over ranks)
        import copy
        ref = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(net))
        ref(x).sum().backward()
        want = [p.grad.clone() for p in ref.module.parameters()]
        # ours: local gradients into the flat arena, bucketed sum, 1/world folded in afterwards
        arena = GradArena(params, "cpu")
        net(x).sum().backward()
        for p in params:
            arena.slot(p).copy_(p.grad)
        offs = [arena.offsets[id(p)][0] for p in params]
        red = BucketReducer(arena.flat, make_buckets(offs, arena.size, bucket_mb=0.0005), None,
                            {id(p): arena.offsets[id(p)][0] for p in params})
        assert red.world == world and len(red.spans) >= 3
        local = arena.flat.clone()
        red.allreduce_sum()
        ok = all(torch.allclose(arena.slot(p) / world, w, rtol=1e-6, atol=1e-7) for p, w in zip(params, want))
        # the overlapped protocol: parameters report in backward order (last registered first); each bucket's
        # all_reduce is launched the moment its last parameter reports, buckets in index order, every bucket once,
        # and the result equals the non-overlapped reduction
        summed = arena.flat.clone()
        arena.flat.copy_(local)
        red.begin()
        sent_after = []
        for p in reversed(params):
            red.ready([id(p)])
            red.ready([id(p)])                       # reporting twice is harmless
            sent_after.append(len(red.launch_order))
        assert red.launch_order == list(range(len(red.spans))) and sent_after[-1] == len(red.spans)
        assert sent_after[0] <= 1 and sent_after == sorted(sent_after)
        assert sent_after[len(params) // 2] >= 1, "no bucket was sent before the last parameter reported: no overlap"
        red.finish()
        ok = ok and torch.equal(arena.flat, summed)
        # parameters that never report are flushed by finish()
        arena.flat.copy_(local)
        red.begin()
        red.ready([id(params[-1])])
        red.finish()
        ok = ok and torch.equal(arena.flat, summed) and sorted(red.launch_order) == list(range(len(red.spans)))
        assert unused.weight.grad is None
        # construction-time broadcast
        t = torch.full((5,), float(rank))
        red.broadcast([t], src=0)
        ok = ok and bool((t == 0).all())
        out[rank] = ok
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_bucketed_allreduce_matches_torch_ddp_world2():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}
