import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from oracle import resunet_oracle as R
from oracle.emulate import bf16_storage
import test_smp_gpu as T
for size in (64, 128):
    ref, ours = T._pair("unet_smp")
    x, _ = T._inputs(size)
    ref.eval(); ours.eval()
    with torch.no_grad():
        a = ref(x); b = ours(x.cuda()).cpu()
        with bf16_storage(ref): e = ref(R.round_bf16(x))
    print(size, "ours-vs-fp32 max", (a-b).abs().max().item(), "mean", (a-b).abs().mean().item(),
          "| emu-vs-fp32 max", (a-e).abs().max().item(), "mean", (a-e).abs().mean().item(),
          "| ours-vs-emu max", (b-e).abs().max().item(), "mean", (b-e).abs().mean().item(), "| max|ref|", a.abs().max().item())
