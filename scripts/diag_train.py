"""Where the NLL training step's time goes: host issue time vs device time, and a per-kernel table (torch.profiler)."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import posteriflow_b200 as P  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
model = P.LeanNPE(tier=os.environ.get("PF_TIER", "bf16")).to(dev)
Bt, D = 512, 11
g = torch.Generator().manual_seed(100)
nsig = torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), Bt, replacement=True, generator=g) + 1
ev = torch.repeat_interleave(torch.arange(Bt), nsig)
rk = torch.cat([torch.arange(int(k)) for k in nsig])
Rt = int(ev.numel())
theta = P.ParamScaler().denormalize(torch.rand(Rt, D, generator=g) * 2 - 1).to(dev)
tctx = torch.randn(Bt, 256, generator=g).to(dev).requires_grad_(True)
ev, rk = ev.to(dev), rk.to(dev)
params = [p for p in model.parameters() if p.requires_grad]


def step():
    for p in params:
        p.grad = None
    tctx.grad = None
    loss = model.nll(None, theta, rk, context=tctx[ev]).sum() / Rt
    loss.backward()
    return loss


for _ in range(5):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    step()
e1.record()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"rows {Rt}: host issue {1e2 * (t1 - t0):.3f} ms/step, device {e0.elapsed_time(e1) / 10:.3f} ms/step, "
      f"wall {1e2 * (t2 - t0):.3f} ms/step")
from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=30, max_name_column_width=70))
