"""CUDA-graph replay vs eager forward_backward of the same TrainStep (eval mode: deterministic forward)."""
import os, sys, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pdac_pathological_image_segmentation_b200.smp import Unet
from pdac_pathological_image_segmentation_b200.synthetic import synthetic_batch
from pdac_pathological_image_segmentation_b200.trainer import TrainStep
torch.manual_seed(11)
m = Unet(encoder_weights=None).cuda().eval()
ts = TrainStep(m, lr=1e-3, graph=True)
batches = [tuple(x.cuda() for x in synthetic_batch(2, 128, seed=20 + i)) for i in range(3)]
names = {id(p): k for k, p in m.named_parameters()}
for it in range(6):
    img, mask = batches[it % 3]
    loss, score = ts._graph_forward_backward(img, mask, None)
    g_graph = ts.arena.flat.clone()
    le, se = ts.forward_backward(img, mask)
    g_eager = ts.arena.flat.clone()
    worst = []
    for p in ts.params:
        off, n, shape = ts.arena.offsets[id(p)]
        a, b = g_graph[off:off + n].double(), g_eager[off:off + n].double()
        worst.append((((a - b).norm() / (b.norm() + 1e-30)).item(), names[id(p)], b.norm().item()))
    worst.sort(reverse=True)
    print(f"it {it} graph={'replay' if ts._g is not None else 'eager'} loss {loss.item():.6f} vs {le.item():.6f}; "
          f"|g| {g_eager.double().norm().item():.3e}; worst: {[(round(w[0], 4), w[1]) for w in worst[:4]]}")
    ts.optimizer_step()
