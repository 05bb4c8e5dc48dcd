"""Fused training step (forward-with-saves / backward chain / weight-gradient kernels) against the GEMM-per-launch form
of the same tier and against the FP32-FMA tier, tensor by tensor.   python scripts/diag_fused.py [rows] [layers]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import nsf_oracle as O  # noqa: E402
import posteriflow_b200 as P  # noqa: E402

R = int(sys.argv[1]) if len(sys.argv) > 1 else 300
L = int(sys.argv[2]) if len(sys.argv) > 2 else 10
stress = (sys.argv[3] == "stress") if len(sys.argv) > 3 else True
dev = torch.device("cuda:0")
spec = O.FlowSpec(num_layers=L)
w = O.init_weights(spec, seed=1, stress=stress)
sd = O.full_state_dict(spec, w)


def make(tier, fused):
    os.environ["PF_TRAIN_FUSED"] = "1" if fused else "0"
    os.environ["PF_TRAIN_TC_MIN_ROWS"] = "0"
    f = P.NSFPosteriorFlow(features=11, context_features=288, hidden_features=256, num_layers=L, num_bins=16,
                           tail_bound=5.0, temperature_scale=1.0, tier=tier)
    f.load_state_dict(sd, strict=True)
    f = f.to(dev)
    f.handle()
    return f


g = torch.Generator().manual_seed(3)
x = (torch.rand(R, 11, generator=g) * 2 - 1).to(dev)
ctx = torch.randn(R, 288, generator=g).to(dev)
gw = torch.linspace(0.5, 1.5, R).to(dev) / R


def run(f):
    f.zero_grad(set_to_none=True)
    xx = x.clone().requires_grad_(True)
    cc = ctx.clone().requires_grad_(True)
    nll = f.compute_psd_aware_nll(xx, cc, None)
    (nll * gw).sum().backward()
    torch.cuda.synchronize()
    f.handle().check()
    out = {"nll": nll.detach(), "gx": xx.grad, "gctx": cc.grad}
    for n, p in f.named_parameters():
        if p.grad is not None and n.startswith("transform."):
            out[n.replace("transform._transforms.", "T").replace(".autoregressive_net", "")] = p.grad
    return out


ref = run(make("fp32", False))
for name, fl in (("bf16 unfused", make("bf16", False)), ("bf16 fused", make("bf16", True))):
    got = run(fl)
    print(f"== {name} vs fp32 tier ({R} rows, {L} layers, stress={stress})")
    worst = []
    for k in ref:
        a, b = got[k].double(), ref[k].double()
        sc = float(b.abs().max()) + 1e-30
        e = float((a - b).abs().max()) / sc
        rn = float((a - b).norm() / (b.norm() + 1e-30))
        worst.append((e, k, rn))
    worst.sort(reverse=True)
    for e, k, rn in worst[:14]:
        print(f"   {k:55s} max|d|/max|ref| {e:9.2e}   rel norm {rn:9.2e}")
    print("   median over tensors of max|d|/max|ref|:", sorted(w_[0] for w_ in worst)[len(worst) // 2])
