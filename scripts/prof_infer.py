"""One sampler / log-prob call at a given shape, for ncu captures.   python scripts/prof_infer.py sample|logprob E n [tier]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import posteriflow_b200 as P  # noqa: E402

what, E, n = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
tier = sys.argv[4] if len(sys.argv) > 4 else "bf16"
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = P.LeanNPE(tier=tier).to(dev)
ctx = m._full_context(torch.randn(E, 256, device=dev), torch.zeros(E, dtype=torch.long, device=dev))
if what == "sample":
    for i in range(2):
        th = P.sample_fused(m.flow, m.scaler, ctx, n, z=None, seed=10 + i)[0]
    torch.cuda.synchronize()
    print("ok sample", tuple(th.shape), bool(torch.isfinite(th).all()))
else:
    y = torch.rand(E * n, 11, device=dev) * 2 - 1
    with torch.no_grad():
        for i in range(2):
            nll = m.flow.compute_psd_aware_nll(y, ctx, None, n_per_ctx=n)
    torch.cuda.synchronize()
    print("ok logprob", tuple(nll.shape), bool(torch.isfinite(nll).all()))
m.flow.handle().check()
