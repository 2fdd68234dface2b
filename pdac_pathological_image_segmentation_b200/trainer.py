"""The reference's hot loop (train_worker.py:67-83: zero_grad / forward / sigmoid-mask / dice_score / DiceLoss /
backward [+ DDP gradient allreduce] / Adam.step) as one stream-ordered sequence of hand-written kernels with no
host synchronisation: uint8 HWC patches are decoded, normalised and augmented on the GPU (dataset.py:24-30,
train_main.py:70-88), parameters / gradients / Adam moments live in flat fp32 arenas, and the gradient arena is
all-reduced over NCCL in reverse-registration-order buckets, each launched on a side stream as soon as the backward
tape has completed its last parameter (DistributedDataParallel semantics: mean over ranks, train_worker.py:204).

Checkpoints (train_worker.py:113-125, 186-201): `TrainStep.state_dict()` / `load_state_dict()` write and read the
reference's {"model", "optimizer", "epoch", "previous_best"} dictionaries ("module."-prefixed keys, torch.optim.Adam's
state layout), `TrainStep.optimizer` is a torch.optim.Optimizer whose param_groups a ReduceLROnPlateau can drive
(train_worker.py:240), and `TrainStep.validate()` is the eval-mode Dice of train_worker.py:16-37, all-reduced."""
from __future__ import annotations

import os
from collections import OrderedDict

import torch
import torch.distributed as dist

from . import _lib, ops
from .ddp import BucketReducer, make_buckets
from .engine import GradArena, Tape


class ArenaAdam(torch.optim.Optimizer):
    """torch.optim.Adam's interface (param_groups, state_dict layout, step) over TrainStep's fused arena update.
    Parameter indices follow `model.parameters()` order, never-used parameters included (they simply have no state),
    so an `optimizer.state_dict()` saved by the reference (train_worker.py:119) loads, and ours loads into
    torch.optim.Adam(model.parameters())."""

    def __init__(self, owner, all_params, lr, betas, eps):
        defaults = dict(torch.optim.Adam([torch.nn.Parameter(torch.zeros(1))], lr=lr, betas=betas, eps=eps).defaults)
        super().__init__(list(all_params), defaults)
        self._owner = owner

    @torch.no_grad()
    def step(self, closure=None):
        self._owner.optimizer_step()

    def zero_grad(self, set_to_none=True):
        self._owner.arena.zero()

    def state_dict(self):
        ts = self._owner
        step_count = ts.step_count
        state = {}
        if step_count > 0:
            for i, p in enumerate(self.param_groups[0]["params"]):
                slot = ts.arena.offsets.get(id(p))
                if slot is None:
                    continue
                off, n, shape = slot
                state[i] = {"step": torch.tensor(float(step_count)),
                            "exp_avg": ts.exp_avg[off:off + n].view(shape).clone(),
                            "exp_avg_sq": ts.exp_avg_sq[off:off + n].view(shape).clone()}
        group = {k: v for k, v in self.param_groups[0].items() if k != "params"}
        group["params"] = list(range(len(self.param_groups[0]["params"])))
        return {"state": state, "param_groups": [group]}

    @torch.no_grad()
    def load_state_dict(self, sd):
        ts = self._owner
        groups = sd["param_groups"]
        if len(groups) != 1 or len(groups[0]["params"]) != len(self.param_groups[0]["params"]):
            raise ValueError("loaded state dict has a different number of parameter groups / parameters")
        params = self.param_groups[0]["params"]
        for k, v in groups[0].items():
            if k != "params":
                self.param_groups[0][k] = v
        ts.exp_avg.zero_()
        ts.exp_avg_sq.zero_()
        step = 0
        for i, st in sd["state"].items():
            slot = ts.arena.offsets.get(id(params[int(i)]))
            if slot is None:
                continue            # state of a parameter the fast path never trains (e.g. encoder.fc): nothing to map
            off, n, shape = slot
            ts.exp_avg[off:off + n].view(shape).copy_(st["exp_avg"])
            ts.exp_avg_sq[off:off + n].view(shape).copy_(st["exp_avg_sq"])
            step = max(step, int(float(st["step"])))
        ts.step_count = step


class TrainStep:
    def __init__(self, model, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, process_group=None, bucket_mb=25.0, graph=False,
                 overlap=True, augment=None):
        """graph=True: after two eager steps per input shape the WHOLE step (input prep, forward, loss, backward, the
        bucketed NCCL all_reduces on their side stream, Adam) is captured once into a CUDA graph and replayed.
        overlap=False keeps all collectives behind the end of backward (for A/B timing).
        augment: None, True (the reference's ColorJitter + OneOf pipeline, train_main.py:70-80) or an ops.Augment:
        drawn on the device per step from (seed, device step counter) — graph-replayable, reproducible."""
        self.model = model
        self.use_graph = bool(graph)
        self.overlap = bool(overlap) and os.environ.get("PDAC_DDP_OVERLAP", "1") != "0"
        self._g_key = None
        self._host_rng = False
        self._stats_need = 0      # floats of BatchNorm partial-sum tables the last step asked for
        self._comm_in_graph = True
        self.params = model._trainable()
        dev = self.params[0].device
        if dev.type != "cuda":
            raise RuntimeError("TrainStep needs the model on a CUDA device (model.to(gpu) first); there is no CPU path")
        self.device = dev
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
        offsets = {id(p): self.arena.offsets[id(p)][0] for p in self.params}
        # 25 MB buckets as torch DDP, plus a small last bucket: the stem / layer1 gradients are complete only when
        # backward ends, so their all_reduce is the one piece of communication nothing can overlap
        self.reducer = BucketReducer(self.arena.flat,
                                     make_buckets(sorted(offsets.values()), self.arena.size, bucket_mb, tail_mb=1.0),
                                     process_group, offsets)
        self.world = self.reducer.world
        self.optimizer = ArenaAdam(self, model.parameters(), lr, betas, eps)
        # device-side Adam state: step counter and hyper-parameters (the graph reads them, schedulers write lr)
        self.step_dev = torch.zeros((), dtype=torch.int64, device=dev)
        self.hyper = torch.zeros(12, dtype=torch.float32, device=dev)
        self._hyper_host = None
        self._seed = int(torch.randint(0, 2 ** 62, (1,)).item())      # Dropout masks / augmentation draws (torch.manual_seed)
        self.augment = ops.Augment(seed=self._seed) if augment is True else augment
        self.last_aug_params = None
        self._sync_hyper()
        model._packs.invalidate()

    # ------------------------------------------------------------------------------------------------ hyper-parameters
    @property
    def lr(self):
        return self.optimizer.param_groups[0]["lr"]

    @lr.setter
    def lr(self, v):
        self.optimizer.param_groups[0]["lr"] = float(v)

    def _sync_hyper(self):
        g = self.optimizer.param_groups[0]
        b1, b2 = float(g["betas"][0]), float(g["betas"][1])
        host = (float(g["lr"]), b1, b2, float(g["eps"]), 1.0 / self.world, 0.0, 0.0, 1.0 - b1, 1.0 - b2)
        if host != self._hyper_host:
            self.hyper[:9].copy_(torch.tensor(host, dtype=torch.float32), non_blocking=False)
            self._hyper_host = host

    @property
    def step_count(self):
        """Adam's step counter; it lives on the device (reading it synchronises: checkpoints only)."""
        return int(self.step_dev.item())

    @step_count.setter
    def step_count(self, step):
        self.step_dev.fill_(int(step))

    # ------------------------------------------------------------------------------------------------ data parallel
    def broadcast_parameters(self, src=0):
        """DDP's construction-time parameter / buffer broadcast (train_worker.py:204)."""
        self.reducer.broadcast([self.flat_params] + list(self.model.buffers()), src)

    # ------------------------------------------------------------------------------------------------ the step
    def forward_backward(self, img_u8, mask_u8, aug=None, reduce=True):
        """-> (loss, dice score) as 0-d device tensors. img_u8 (B,H,W,3) uint8, mask_u8 (B,H,W) uint8 {0,1}. With
        reduce=True the gradient arena holds the SUM over the ranks when the returned work has run."""
        model = self.model
        if self.augment is not None and aug is None and model.training:
            z, mask, self.last_aug_params = ops.stem_prep_u8_augment(img_u8, mask_u8, self.augment, step=self.step_dev)
        else:
            z, mask = ops.stem_prep_u8(img_u8, mask_u8, aug)
        tape = Tape(model._packs, model.training, self.arena, step_dev=self.step_dev, base_seed=self._seed)
        if self._stats_need:      # one memset for all BatchNorm partial-sum tables of the step
            tape.stats_pool = torch.zeros(self._stats_need, dtype=torch.float32, device=self.device)
        logits, feed = model._run(tape, z)
        self._stats_need = tape.stats_used
        self._host_rng = bool(getattr(tape, "host_rng", False))
        if logits.shape[1] != 1:
            raise RuntimeError("TrainStep implements the reference's 1-class sigmoid/Dice step (train_worker.py:150-182)")
        score, _ = ops.dice_score(logits, mask, pred_is_logits=True)
        loss, sums = ops.dice_fwd(logits, mask)
        cpad = getattr(feed, "nhwc_cpad", None)
        g = ops.dice_bwd_nhwc(logits, mask, sums, cpad) if cpad else ops.dice_bwd(logits, mask, sums, None)
        self.arena.zero()
        feed(g)
        if reduce and self.world > 1:
            red = self.reducer
            red.begin()
            tape.backward((lambda ps: red.ready([id(p) for p in ps])) if self.overlap else None)
            red.finish()
        else:
            tape.backward()
        return loss, score

    def optimizer_step(self):
        ops.adam_step_dev(self.flat_params, self.arena.flat, self.exp_avg, self.exp_avg_sq, self.hyper, self.step_dev)
        self.model._packs.refresh()     # bf16 GEMM copies of every weight, re-packed in place by one batched launch

    def _whole_step(self, img_u8, mask_u8, aug, comm=True):
        out = self.forward_backward(img_u8, mask_u8, aug, reduce=comm)
        if comm:
            self.optimizer_step()
        return out

    def step(self, img_u8, mask_u8, aug=None):
        _lib.peek_device_flag()      # non-blocking: a pipeline time-out reported by an earlier step's kernels raises here
        self._sync_hyper()
        if self.use_graph:
            return self._graph_step(img_u8, mask_u8, aug)
        return self._whole_step(img_u8, mask_u8, aug)

    def _graph_step(self, img_u8, mask_u8, aug):
        """The step through a captured CUDA graph: inputs are copied into static buffers, every kernel argument (tensor
        maps, arena slots, the allocator's activation addresses) is baked at capture time. BatchNorm's
        num_batches_tracked / running statistics, Adam's step counter and the learning rate live on the device, so
        replays keep module and optimizer state exact."""
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
                return self._whole_step(self._g_img, self._g_mask, self._g_aug)
            self._capture()
        self._g.replay()
        _lib.launches += self._g_launches
        if not self._comm_in_graph:
            self.reducer.allreduce_sum()
            self.optimizer_step()
        return self._g_out[0].clone(), self._g_out[1].clone()

    def _capture(self):
        comm = self._comm_in_graph
        l0 = _lib.launches
        g = torch.cuda.CUDAGraph()
        try:
            # NCCL's watchdog thread polls events while we capture: thread-local capture mode keeps that legal
            with torch.cuda.graph(g, capture_error_mode="thread_local" if self.world > 1 else "global"):
                self._g_out = self._whole_step(self._g_img, self._g_mask, self._g_aug, comm=comm)
        except Exception:
            if not (comm and self.world > 1):
                raise
            # collectives could not be captured on this stack: graph = forward + loss + backward, allreduce + Adam eager
            self._comm_in_graph = False
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._g_out = self._whole_step(self._g_img, self._g_mask, self._g_aug, comm=False)
        self._g_launches = _lib.launches - l0
        _lib.launches = l0
        self._g = g

    def release_graph(self):
        """Drops the captured CUDA graph (the next step re-captures). REQUIRED before torch.distributed's
        destroy_process_group() when the graph holds the NCCL all_reduces: NCCL's communicator teardown waits for every
        CUDA graph that captured collectives on it to be destroyed."""
        self._g = None
        self._g_out = None
        self._g_key = None
        torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------------------------------------ validation
    @torch.no_grad()
    def validate(self, img_u8, mask_u8):
        """Eval-mode forward + hard Dice of one batch (train_worker.py:16-37), with the three integer counts summed over
        the ranks so that every rank steps its scheduler on the same score. -> 0-d fp32 device tensor."""
        model = self.model
        was_training = model.training
        model.eval()
        try:
            z, mask = ops.stem_prep_u8(img_u8, mask_u8, None)      # the packed weights are current: refreshed by every step
            logits, _ = model._run(Tape(model._packs, False, None), z)
            _, counts = ops.dice_score(logits, mask, pred_is_logits=True)
        finally:
            model.train(was_training)
        if self.world > 1:
            dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=self.reducer.pg)
        c = counts.to(torch.float32)
        smooth = 1e-6
        return (2.0 * c[0] + smooth) / (c[1] + c[2] + smooth)

    # ------------------------------------------------------------------------------------------------ checkpoints
    def state_dict(self, epoch=0, previous_best=0.0):
        """The dictionary train_worker.py:116-125 saves as latest.pth / best.pth (keys carry DDP's "module." prefix)."""
        model_sd = OrderedDict(("module." + k, v.detach().clone()) for k, v in self.model.state_dict().items())
        return {"model": model_sd, "optimizer": self.optimizer.state_dict(), "epoch": epoch,
                "previous_best": previous_best}

    def load_state_dict(self, ckpt):
        """Restores what train_worker.py:186-201 restores; -> (epoch, previous_best)."""
        sd = OrderedDict((k.replace("module.", ""), v) for k, v in ckpt["model"].items())
        with torch.no_grad():
            own = self.model.state_dict()
            missing = [k for k in own if k not in sd]
            extra = [k for k in sd if k not in own]
            if missing or extra:
                raise RuntimeError(f"checkpoint does not match the model: missing {missing[:3]}, unexpected {extra[:3]}")
            for k, t in own.items():
                t.copy_(sd[k])          # in place: the parameters keep living in the flat arena
        self.model._packs.refresh()     # in place as well: a captured graph keeps pointing at the same packed buffers
        if "optimizer" in ckpt:
            self.optimizer.load_state_dict(ckpt["optimizer"])
        self._sync_hyper()
        return ckpt.get("epoch", 0), ckpt.get("previous_best", 0.0)
