"""The reference's hot loop (train_worker.py:67-83: zero_grad / forward / sigmoid-mask / dice_score / DiceLoss /
backward [+ DDP gradient allreduce] / Adam.step) as one stream-ordered sequence of hand-written kernels with no
host synchronisation: uint8 HWC patches are decoded, normalised and augmented on the GPU (dataset.py:24-30,
train_main.py:70-88), parameters / gradients / Adam moments live in flat fp32 arenas, and the gradient arena is
all-reduced over NCCL in reverse-registration-order buckets launched on a side stream while the backward tape is
still running (DistributedDataParallel semantics: mean over ranks, train_worker.py:204)."""
from __future__ import annotations

import torch

from . import ops
from .ddp import BucketReducer, make_buckets
from .engine import GradArena, Tape


class TrainStep:
    def __init__(self, model, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, process_group=None, bucket_mb=25.0, graph=False):
        """graph=True: after two eager steps per input shape, forward + loss + backward (about 260 kernel launches) are
        captured once into a CUDA graph and replayed; the gradient allreduce and Adam stay outside the graph."""
        self.model = model
        self.use_graph = bool(graph)
        self._g_key = None
        self._host_rng = False
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
        offsets = [self.arena.offsets[id(p)][0] for p in self.params]
        self.reducer = BucketReducer(self.arena.flat, make_buckets(offsets, self.arena.size, bucket_mb), process_group)
        self.world = self.reducer.world
        model._packs.invalidate()

    def broadcast_parameters(self, src=0):
        """DDP's construction-time parameter / buffer broadcast (train_worker.py:204)."""
        self.reducer.broadcast([self.flat_params] + list(self.model.buffers()), src)

    def _allreduce(self):
        self.reducer.allreduce_sum()

    def forward_backward(self, img_u8, mask_u8, aug=None):
        """-> (loss, dice score) as 0-d device tensors. img_u8 (B,H,W,3) uint8, mask_u8 (B,H,W) uint8 {0,1}."""
        model = self.model
        z, mask = ops.stem_prep_u8(img_u8, mask_u8, aug)
        tape = Tape(model._packs, model.training, self.arena)
        logits, feed = model._run(tape, z)
        self._host_rng = bool(getattr(tape, "host_rng", False))
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
        if self.use_graph:
            loss, score = self._graph_forward_backward(img_u8, mask_u8, aug)
        else:
            loss, score = self.forward_backward(img_u8, mask_u8, aug)
        self._allreduce()
        self.optimizer_step()
        return loss, score

    def _graph_forward_backward(self, img_u8, mask_u8, aug):
        """forward_backward through a captured CUDA graph: inputs are copied into static buffers, every kernel argument
        (tensor maps, arena slots, the allocator's activation addresses) is baked at capture time. BatchNorm's
        num_batches_tracked and running statistics advance on the device, so replays keep the module state exact."""
        from . import _lib
        key = (tuple(img_u8.shape), tuple(mask_u8.shape), None if aug is None else tuple(aug.shape), self.model.training)
        if key != self._g_key:
            self._g_key, self._g, self._g_warm = key, None, 0
            self._g_img, self._g_mask = torch.empty_like(img_u8), torch.empty_like(mask_u8)
            self._g_aug = None if aug is None else torch.empty_like(aug)
        self._g_img.copy_(img_u8, non_blocking=True)
        self._g_mask.copy_(mask_u8, non_blocking=True)
        if aug is not None:
            self._g_aug.copy_(aug, non_blocking=True)
        if self._g is None:
            if self._g_warm < 2 or self._host_rng:
                # eager: lazy one-time host initialisation inside the library, allocator warm-up; and for good when the
                # step draws host random numbers (active Dropout: the seed is a launch argument)
                self._g_warm += 1
                return self.forward_backward(self._g_img, self._g_mask, self._g_aug)
            l0 = _lib.launches
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._g_out = self.forward_backward(self._g_img, self._g_mask, self._g_aug)
            self._g_launches = _lib.launches - l0
            _lib.launches = l0
            self._g = g
        self._g.replay()
        _lib.launches += self._g_launches
        return self._g_out[0].clone(), self._g_out[1].clone()
