"""Where the gradient all_reduces sit relative to backward (run under torchrun, one rank per GPU):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/ddp_timeline.py
Prints, for one eager smp-Unet train step at batch 64, the start / end of every bucket's all_reduce and the end of
backward, in ms after the start of the step (CUDA events; max over ranks is not needed for a timeline: rank 0 prints),
and the step time with and without overlap."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from pdac_pathological_image_segmentation_b200.smp import Unet
from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
from pdac_pathological_image_segmentation_b200.trainer import TrainStep

rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
img, mask = (t.to(dev) for t in synthetic_batch(B, 512, seed=41 + rank))
for overlap in (True, False):
    torch.manual_seed(41)
    ts = TrainStep(Unet(encoder_weights=None).to(dev).train(), overlap=overlap)
    ts.broadcast_parameters()
    for _ in range(3):
        ts.step(img, mask)
    ts.reducer.record_timeline = True
    e0, e_bwd, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0.record()
    ts.forward_backward(img, mask)          # includes reducer.finish(): the compute stream has joined the collectives
    e_bwd.record()
    ts.optimizer_step()
    e1.record()
    torch.cuda.synchronize()
    tl = list(ts.reducer.timeline)
    ts.reducer.record_timeline = False
    # steady-state step time
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        ts.step(img, mask)
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / 10], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"overlap={overlap}: world {dist.get_world_size()}, batch {B}/GPU, eager step {t.item():.3f} ms (max over ranks); "
              f"forward+backward+join {e0.elapsed_time(e_bwd):.3f} ms, +Adam {e0.elapsed_time(e1):.3f} ms")
        for bkt, nbytes, t0, t1 in tl:
            print(f"   bucket {bkt} ({nbytes / 2**20:.1f} MiB): all_reduce {e0.elapsed_time(t0):8.3f} -> {e0.elapsed_time(t1):8.3f} ms "
                  f"after step start (backward + join ends at {e0.elapsed_time(e_bwd):.3f})", flush=True)
    del ts
dist.destroy_process_group()
