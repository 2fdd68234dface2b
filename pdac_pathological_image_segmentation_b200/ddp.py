"""Gradient reduction of the data-parallel training path (train_worker.py:203-206: DistributedDataParallel over NCCL).
DDP semantics kept: gradients are averaged over ranks; buckets are contiguous ranges of the flat gradient arena built
from the LAST registered parameter backwards (those gradients are complete first in backward), at most `bucket_mb`
each (torch's default 25 MB); parameters that never receive a gradient (ResUNet's encoder.fc, PSPNet's layer3/4) are
simply not in the arena, which replaces find_unused_parameters=True's per-iteration graph walk.
Device-agnostic (flat CPU tensors + gloo in the CPU tests, CUDA + NCCL on the box)."""
from __future__ import annotations

import torch
import torch.distributed as dist


def make_buckets(param_offsets, total, bucket_mb=25.0, elem_bytes=4):
    """param_offsets: arena offset of every parameter in REGISTRATION order (ascending). -> [(lo, hi)] spans in
    reduction order (last parameters first) that tile [0, total)."""
    cap = max(int(bucket_mb * 1024 * 1024 / elem_bytes), 1)
    spans, hi = [], total
    lo = total
    for off in reversed(list(param_offsets)):
        if hi - off > cap and lo < hi:
            spans.append((lo, hi))
            hi = lo
        lo = off
    if lo > 0:
        lo = 0
    if lo < hi:
        spans.append((lo, hi))
    return spans


class BucketReducer:
    def __init__(self, flat: torch.Tensor, spans, process_group=None):
        self.flat, self.spans, self.pg = flat, list(spans), process_group
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.stream = torch.cuda.Stream(device=flat.device) if (flat.is_cuda and self.world > 1) else None

    def allreduce_sum(self, upto=None):
        """Sums every bucket over the ranks (the mean's 1/world is folded into the optimizer's grad_scale). On CUDA the
        collectives run on a side stream that waits for the producing stream and is joined before the optimizer."""
        if self.world == 1:
            return
        if self.stream is None:
            for lo, hi in self.spans:
                dist.all_reduce(self.flat[lo:hi], op=dist.ReduceOp.SUM, group=self.pg)
            return
        cur = torch.cuda.current_stream(self.flat.device)
        self.stream.wait_stream(cur)
        with torch.cuda.stream(self.stream):
            for lo, hi in self.spans:
                dist.all_reduce(self.flat[lo:hi], op=dist.ReduceOp.SUM, group=self.pg)
        cur.wait_stream(self.stream)

    def broadcast(self, tensors, src=0):
        if self.world > 1:
            for t in tensors:
                dist.broadcast(t, src, group=self.pg)
