"""Gradient reduction of the data-parallel training path (train_worker.py:203-206: DistributedDataParallel over NCCL).
DDP semantics kept: gradients are averaged over ranks; buckets are contiguous ranges of the flat gradient arena built
from the LAST registered parameter backwards (those gradients are complete first in backward), at most `bucket_mb`
each (torch's default 25 MB); parameters that never receive a gradient (ResUNet's encoder.fc, PSPNet's layer3/4) are
simply not in the arena, which replaces find_unused_parameters=True's per-iteration graph walk.

Overlap with backward (what torch's reducer does during loss.backward(), train_worker.py:79): the engine's tape reports
after every backward step which parameters it completed (`ready`); when the last parameter of a bucket is complete an
event is recorded on the compute stream and the bucket's all_reduce is enqueued on a side stream that waits for that
event only — the rest of backward keeps running on the compute stream. `finish` joins the side stream before the
optimizer. All of it is stream-ordered, so it can be captured into the train step's CUDA graph.
Device-agnostic (flat CPU tensors + gloo in the CPU tests, CUDA + NCCL on the box)."""
from __future__ import annotations

import torch
import torch.distributed as dist


def make_buckets(param_offsets, total, bucket_mb=25.0, elem_bytes=4, tail_mb=0.0):
    """param_offsets: arena offset of every parameter in REGISTRATION order (ascending). -> [(lo, hi)] spans in
    reduction order (last parameters first) that tile [0, total). tail_mb > 0 caps the LAST span (the first registered
    parameters, whose gradients complete at the very end of backward and whose all_reduce nothing can hide) at that
    size by splitting it at a parameter boundary — the mirror image of torch DDP's small first bucket."""
    offs = list(param_offsets)
    cap = max(int(bucket_mb * 1024 * 1024 / elem_bytes), 1)
    spans, hi = [], total
    lo = total
    for off in reversed(offs):
        if hi - off > cap and lo < hi:
            spans.append((lo, hi))
            hi = lo
        lo = off
    if lo > 0:
        lo = 0
    if lo < hi:
        spans.append((lo, hi))
    if tail_mb > 0 and spans:
        tail = max(int(tail_mb * 1024 * 1024 / elem_bytes), 1)
        lo, hi = spans[-1]
        cut = [o for o in offs if lo < o < hi and o - lo <= tail]
        if hi - lo > tail and cut:
            spans[-1:] = [(max(cut), hi), (lo, max(cut))]
    return spans


class BucketReducer:
    def __init__(self, flat: torch.Tensor, spans, process_group=None, param_offsets=None):
        """param_offsets: {key: arena offset} of every parameter that reports readiness (key = id(parameter))."""
        self.flat, self.spans, self.pg = flat, list(spans), process_group
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.stream = torch.cuda.Stream(device=flat.device) if (flat.is_cuda and self.world > 1) else None
        self.bucket_of, self.members = {}, [0] * len(self.spans)
        for key, off in (param_offsets or {}).items():
            for b, (lo, hi) in enumerate(self.spans):
                if lo <= off < hi:
                    self.bucket_of[key] = b
                    self.members[b] += 1
                    break
        self.record_timeline = False    # profiling: CUDA events around every bucket's all_reduce (scripts/ddp_timeline.py)
        self.timeline = []
        self.begin()
        self.launch_order = []          # bucket indices in the order their all_reduce was enqueued (tests / profiling)

    # -- overlapped protocol: begin() ... ready(keys) per backward step ... finish() ------------------------------------
    def begin(self):
        self._left = list(self.members)
        self._sent = [False] * len(self.spans)
        self._seen = set()
        self.launch_order = []
        self.timeline = []

    def ready(self, keys):
        """The gradient slots of these parameters are complete on the current stream."""
        if self.world == 1:
            return
        for k in keys:
            b = self.bucket_of.get(k)
            if b is None or k in self._seen:
                continue
            self._seen.add(k)
            self._left[b] -= 1
            if self._left[b] == 0 and not self._sent[b]:
                self._launch(b)

    def _launch(self, b):
        lo, hi = self.spans[b]
        self._sent[b] = True
        self.launch_order.append(b)
        if self.stream is None:
            dist.all_reduce(self.flat[lo:hi], op=dist.ReduceOp.SUM, group=self.pg)
            return
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.flat.device))
        self.stream.wait_event(ev)
        with torch.cuda.stream(self.stream):
            if self.record_timeline:
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record(self.stream)
            dist.all_reduce(self.flat[lo:hi], op=dist.ReduceOp.SUM, group=self.pg)
            if self.record_timeline:
                t1.record(self.stream)
                self.timeline.append((b, (hi - lo) * self.flat.element_size(), t0, t1))

    def finish(self):
        """Reduces whatever has not been sent yet (buckets whose parameters never reported) and makes the current
        stream wait for all collectives: after this the arena holds the SUM over ranks (1/world is folded into the
        optimizer's grad_scale)."""
        if self.world == 1:
            return
        for b in range(len(self.spans)):
            if not self._sent[b]:
                self._launch(b)
        if self.stream is not None:
            torch.cuda.current_stream(self.flat.device).wait_stream(self.stream)

    def allreduce_sum(self):
        """Non-overlapped form: every bucket after the producing stream has finished all of backward."""
        self.begin()
        self.finish()

    def broadcast(self, tensors, src=0):
        if self.world > 1:
            for t in tensors:
                dist.broadcast(t, src, group=self.pg)
