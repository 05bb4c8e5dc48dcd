"""One fused training step of the flow head (forward-with-saves + backward) on R rows, for launch lists / timing.
    python scripts/prof_fused.py [rows] [reps]        (PF_TRAIN_FUSED=0 for the GEMM-per-launch form)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import posteriflow_b200 as P  # noqa: E402

R = int(sys.argv[1]) if len(sys.argv) > 1 else 7608
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = P.LeanNPE(tier=os.environ.get("PF_TIER", "bf16")).to(dev)
fh = m.flow.handle()
w = [p.detach() for p in m.flow.layer_weights()]
x = (torch.rand(R, 11, device=dev) * 2 - 1)
ctx = torch.randn(R, 288, device=dev)
gz = torch.randn(R, 11, device=dev) / R
gl = -torch.ones(R, device=dev) / R
nr = P.ops.no_row2ctx(x)


def step():
    z, ld, saved = torch.ops.posteriflow.nsf_forward(x, ctx, w, fh.id, 1, True, nr)
    return torch.ops.posteriflow.nsf_backward(saved, ctx, gz, gl, w, fh.id, 1, None)


for _ in range(2):
    step()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
tf = tb = 0.0
for _ in range(reps):
    ev[0].record()
    z, ld, saved = torch.ops.posteriflow.nsf_forward(x, ctx, w, fh.id, 1, True, nr)
    ev[1].record()
    outs = torch.ops.posteriflow.nsf_backward(saved, ctx, gz, gl, w, fh.id, 1, None)
    ev[2].record()
    torch.cuda.synchronize()
    tf += ev[0].elapsed_time(ev[1])
    tb += ev[1].elapsed_time(ev[2])
fh.check()
print(f"rows {R}: forward(save) {tf / reps:.3f} ms, backward {tb / reps:.3f} ms, fused={os.environ.get('PF_TRAIN_FUSED', '1')}")
