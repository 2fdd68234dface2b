"""The reference's hot loop (train_worker.py:67-83: zero_grad / forward / sigmoid-mask / dice_score / DiceLoss /
backward [+ DDP gradient allreduce] / Adam.step) as one stream-ordered sequence of hand-written kernels with no
host synchronisation: uint8 HWC patches are decoded, normalised and augmented on the GPU (dataset.py:24-30,
train_main.py:70-88), parameters / gradients / Adam moments live in flat fp32 arenas, and the gradient arena is
all-reduced over NCCL in reverse-registration-order buckets launched on a side stream while the backward tape is
still running (DistributedDataParallel semantics: mean over ranks, train_worker.py:204)."""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import ops
from .engine import GradArena, Tape


class TrainStep:
    def __init__(self, model, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, process_group=None, bucket_mb=25.0):
        self.model = model
        self.params = model._trainable()
        dev = self.params[0].device
        if dev.type != "cuda":
            raise RuntimeError("TrainStep needs the model on a CUDA device (model.to(gpu) first); there is no CPU path")
        self.device = dev
        self.lr, self.betas, self.eps = lr, betas, eps
        self.step_count = 0
        self.arena = GradArena(self.params, dev)
        # parameters re-homed into one flat arena with the gradient arena's slot layout (zero-padded slots)
        self.flat_params = torch.zeros_like(self.arena.flat)
        for p in self.params:
            off, n, shape = self.arena.offsets[id(p)]
            view = self.flat_params[off:off + n].view(shape)
            view.copy_(p.data)
            p.data = view
            p._pdac_padded_slot = True
        self.exp_avg = torch.zeros_like(self.arena.flat)
        self.exp_avg_sq = torch.zeros_like(self.arena.flat)
        self.world = 1
        self.pg = process_group
        if dist.is_available() and dist.is_initialized():
            self.world = dist.get_world_size(process_group)
        self.comm_stream = torch.cuda.Stream(device=dev) if self.world > 1 else None
        self._buckets = self._make_buckets(bucket_mb)
        model._packs.invalidate()

    # DDP-style buckets: contiguous ranges of the arena, built from the LAST registered parameter backwards (the
    # decoder / head gradients are complete first), each <= bucket_mb
    def _make_buckets(self, bucket_mb):
        cap = int(bucket_mb * 1024 * 1024 / 4)
        spans, hi, lo = [], self.arena.size, self.arena.size
        for p in reversed(self.params):
            off, _, _ = self.arena.offsets[id(p)]
            if hi - off > cap and lo < hi:
                spans.append((lo, hi))
                hi = lo
            lo = off
        if lo < hi:
            spans.append((lo, hi))
        return spans

    def broadcast_parameters(self, src=0):
        """DDP's construction-time parameter / buffer broadcast (train_worker.py:204)."""
        if self.world > 1:
            dist.broadcast(self.flat_params, src, group=self.pg)
            for b in self.model.buffers():
                dist.broadcast(b, src, group=self.pg)

    def _allreduce(self):
        if self.world == 1:
            return
        cur = torch.cuda.current_stream(self.device)
        self.comm_stream.wait_stream(cur)
        with torch.cuda.stream(self.comm_stream):
            for lo, hi in self._buckets:
                dist.all_reduce(self.arena.flat[lo:hi], op=dist.ReduceOp.SUM, group=self.pg)
        cur.wait_stream(self.comm_stream)

    def forward_backward(self, img_u8, mask_u8, aug=None):
        """-> (loss, dice score) as 0-d device tensors. img_u8 (B,H,W,3) uint8, mask_u8 (B,H,W) uint8 {0,1}."""
        model = self.model
        z, mask = ops.stem_prep_u8(img_u8, mask_u8, aug)
        tape = Tape(model._packs, model.training, self.arena)
        logits, feed = model._run(tape, z)
        if logits.shape[1] != 1:
            raise RuntimeError("TrainStep implements the reference's 1-class sigmoid/Dice step (train_worker.py:150-182)")
        score, _ = ops.dice_score(logits, mask, pred_is_logits=True)
        loss, sums = ops.dice_fwd(logits, mask)
        g = ops.dice_bwd(logits, mask, sums, None)
        self.arena.zero()
        feed(g)
        tape.backward()
        return loss, score

    def optimizer_step(self):
        self.step_count += 1
        ops.adam_step(self.flat_params, self.arena.flat, self.exp_avg, self.exp_avg_sq, self.lr, self.step_count,
                      self.betas[0], self.betas[1], self.eps, grad_scale=1.0 / self.world)
        self.model._packs.invalidate()

    def step(self, img_u8, mask_u8, aug=None):
        loss, score = self.forward_backward(img_u8, mask_u8, aug)
        self._allreduce()
        self.optimizer_step()
        return loss, score
