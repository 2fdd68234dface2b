import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import test_smp_gpu as T
from oracle.emulate import bf16_storage
from oracle.resunet_oracle import round_bf16
name = sys.argv[1] if len(sys.argv) > 1 else "deeplabv3+"
ref, ours = T._pair(name)
x, mask = T._inputs()
for train in (False, True):
    ref.train(train); ours.train(train)
    with torch.no_grad():
        a = ref(x); b = ours(x.cuda()).cpu()
        with bf16_storage(ref):
            e = ref(round_bf16(x))
    d, c = (b - a), (e - a)
    print(f"train={train} |ref|max {a.abs().max():.3f} ours: max {d.abs().max():.4f} mean {d.abs().mean():.5f} bias {d.mean():.5f} | emu: max {c.abs().max():.4f} mean {c.abs().mean():.5f} bias {c.mean():.5f}")
    print("   slope ours/ref", ((a*b).sum()/(a*a).sum()).item(), "emu/ref", ((a*e).sum()/(a*a).sum()).item())
