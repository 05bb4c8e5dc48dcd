"""Where does the NLL training step's time go?  (config #4 shard: 512 events -> ~950 flow rows per GPU)

    python scripts/prof_train.py [events] [mode]

mode = "eager"  : batch_nll + backward through autograd, as bench.py's training step (default)
       "raw"    : the two custom ops alone (nsf_forward(save) + nsf_backward), static inputs, no autograd
       "graph"  : the two ops captured in ONE CUDA graph and replayed: device time without any host issue cost
       "once"   : one eager step after warm-up and exit (for `ncu --metrics gpu__time_duration.sum`)
Prints ms per step (CUDA events, 20 steps after 5 warm-up steps).
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import posteriflow_b200 as P  # noqa: E402
from posteriflow_b200 import ops  # noqa: E402


def main():
    Bt = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    mode = sys.argv[2] if len(sys.argv) > 2 else "eager"
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model = P.LeanNPE(tier=os.environ.get("PF_TIER", "bf16")).to(dev)
    D = 11
    g = torch.Generator().manual_seed(100)
    nsig = torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), Bt, replacement=True, generator=g) + 1
    params_b = P.ParamScaler().denormalize(torch.rand(Bt, 5, D, generator=g) * 2 - 1).to(dev)
    nsig_dev = nsig.to(dev)
    tctx = torch.randn(Bt, 256, generator=g).to(dev).requires_grad_(True)
    params = [p for p in model.parameters() if p.requires_grad]
    Rt = int(nsig.sum())

    def step_eager():
        for p in params:
            p.grad = None
        tctx.grad = None
        loss = P.batch_nll(model, None, params_b, nsig_dev, context=tctx)
        loss.backward()
        return loss

    # static inputs of the two ops
    flow = model.flow
    fh = flow.handle()
    weights = [w.detach() for w in flow.layer_weights()]
    with torch.no_grad():
        event, rank = P.train.signal_rows(nsig_dev, 5)
        x = model.scaler.normalize(params_b[event, rank, :]).contiguous()
        ctx = model._full_context(tctx.detach()[event], rank).contiguous()
    gz = torch.randn(Rt, D, device=dev)
    gl = -torch.ones(Rt, device=dev)

    def step_raw():
        fh.ensure_packed(weights)
        z, ld, saved = torch.ops.posteriflow.nsf_forward(x, ctx, weights, fh.id, 1, True, P.ops.no_row2ctx(x))
        return torch.ops.posteriflow.nsf_backward(saved, ctx, gz, gl, weights, fh.id, 1)

    def timeit(fn, steps=20, warm=5):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    if mode == "once":
        for _ in range(3):
            step_eager()
        torch.cuda.synchronize()
        l0 = fh.launches
        step_eager()
        torch.cuda.synchronize()
        print(f"rows {Rt}: one eager step, {fh.launches - l0} library launches")
        return
    if mode == "eager":
        print(f"rows {Rt}: eager batch_nll + backward {timeit(step_eager):.3f} ms/step")
    elif mode == "raw":
        print(f"rows {Rt}: raw ops {timeit(step_raw):.3f} ms/step")
    elif mode == "graph":
        for _ in range(3):
            step_raw()
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            out = step_raw()
        torch.cuda.synchronize()
        print(f"rows {Rt}: graph replay {timeit(gr.replay):.3f} ms/step")
        ref = step_raw()
        gr.replay()
        torch.cuda.synchronize()
        err = max(float((a - b).abs().max()) for a, b in zip(out, ref))
        print(f"graph vs eager max |diff| over {len(out)} gradient tensors: {err:.3e}")


if __name__ == "__main__":
    main()
