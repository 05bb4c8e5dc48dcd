"""GraphedNLLStep on a config-#4 batch of B events (public call), timed with CUDA events; for launch lists.
    python scripts/prof_step.py [B] [reps]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import posteriflow_b200 as P  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = P.LeanNPE(tier="bf16").to(dev)
g = torch.Generator().manual_seed(1)
nsig = (torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), B, replacement=True, generator=g) + 1).to(dev)
params = P.ParamScaler().denormalize(torch.rand(B, 5, 11, generator=g) * 2 - 1).to(dev)
ctx = torch.randn(B, 256, generator=g).to(dev).requires_grad_(True)
step = P.GraphedNLLStep(m)
ps = [p for p in m.parameters() if p.requires_grad]


def one():
    for p in ps:
        p.grad = None
    ctx.grad = None
    return step(params, nsig, ctx)


for _ in range(3):
    one()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    loss = one()
e1.record()
torch.cuda.synchronize()
print(f"B {B}: {e0.elapsed_time(e1) / reps:.3f} ms per step (back to back, no L2 flush), loss {float(loss):.4f}, "
      f"fused glue {bool(step.slot_graphs)} (slot mode {step._use_slots(B, dev)}), kernels per replay {step.last_launches}")
