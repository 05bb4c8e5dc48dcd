"""Where the 0.55 ms between the flow-only training step and the full one go: clip_grad_norm_(5.0), fused AdamW, re-pack.
    python scripts/prof_opt.py [events]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import posteriflow_b200 as P  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = P.LeanNPE(tier="bf16").to(dev)
g = torch.Generator().manual_seed(1)
nsig = (torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), B, replacement=True, generator=g) + 1).to(dev)
params = m.scaler.denormalize((torch.rand(B, 5, 11, generator=g) * 2 - 1).to(dev))
ctx = torch.randn(B, 256, generator=g).to(dev)
step = P.GraphedNLLStep(m)
plist = [p for p in list(m.flow.parameters()) + list(m.rank_embed.parameters()) if p.requires_grad]
opt = torch.optim.AdamW(plist, lr=0.0, fused=True)


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def flow_step():
    for p in plist:
        p.grad = None
    return step(params, nsig, ctx)


flow_step()
print(f"flow step                     {timed(flow_step):.3f} ms")
print(f"clip_grad_norm_(5.0)          {timed(lambda: torch.nn.utils.clip_grad_norm_(plist, 5.0)):.3f} ms  ({len(plist)} tensors)")
print(f"clip on the flat buffer       {timed(lambda: step.flat_grad.mul_(torch.clamp(5.0 / (torch.linalg.vector_norm(step.flat_grad) + 1e-6), max=1.0))):.3f} ms")
print(f"AdamW(fused).step()           {timed(opt.step):.3f} ms")


def repack():
    with torch.no_grad():
        step.flow_params[0].add_(0.0)
    m.flow.handle()


print(f"re-pack                       {timed(repack):.3f} ms")


def full():
    flow_step()
    torch.nn.utils.clip_grad_norm_(plist, 5.0)
    opt.step()


print(f"flow step + clip + AdamW (+ re-pack inside the next step)  {timed(full):.3f} ms")
