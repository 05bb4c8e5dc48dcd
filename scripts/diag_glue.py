"""How much of the slot-mode training step is ATen glue around the two flow ops?  Replays (a) the full GraphedNLLStep
graph and (b) a graph holding ONLY nsf_forward + nsf_backward on static buffers of the same 2 560 slot rows."""
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
plist = step.flow_params


def timed(fn, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def full():
    for p in plist:
        p.grad = None
    return step(params, nsig, ctx)


full()
st = next(iter(step.slot_graphs.values())) if step.slot_graphs else next(iter(step.graphs.values()))
print(f"GraphedNLLStep call (graph + host glue)      {timed(full):.3f} ms")
print(f"its graph replay alone                       {timed(st['graph'].replay):.3f} ms")
R = B * 5 if B in step.slot_graphs else int(nsig.sum())
fh = m.flow.handle()
w = [p.detach() for p in plist]
x = (torch.rand(R, 11, device=dev) * 2 - 1)
c = torch.randn(R, 288, device=dev)
gz = torch.randn(R, 11, device=dev) / R
gl = -torch.ones(R, device=dev) / R
nr = P.ops.no_row2ctx(x)


def ops_only():
    z, ld, saved = torch.ops.posteriflow.nsf_forward(x, c, w, fh.id, 1, True, nr)
    return torch.ops.posteriflow.nsf_backward(saved, c, gz, gl, w, fh.id, 1, None)


s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3):
        ops_only()
torch.cuda.current_stream().wait_stream(s)
gr = torch.cuda.CUDAGraph()
with torch.cuda.graph(gr):
    keep = ops_only()
print(f"graph of nsf_forward + nsf_backward only     {timed(gr.replay):.3f} ms   ({R} rows)")
