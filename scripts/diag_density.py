"""Density direction (log-prob) throughput of the tensor-core tier: rows/s of flow.compute_psd_aware_nll on
E events x n rows (context broadcast), device-timed.  PF_TC_DENSE=0 selects the stepwise kernel for an A/B."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import posteriflow_b200 as P  # noqa: E402

E = int(sys.argv[1]) if len(sys.argv) > 1 else 512
n = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = P.LeanNPE(tier=os.environ.get("PF_TIER", "bf16")).to(dev)
ctx = m._full_context(torch.randn(E, 256, device=dev), torch.zeros(E, dtype=torch.long, device=dev))
y = (torch.rand(E * n, 11, device=dev) * 2 - 1)
with torch.no_grad():
    for _ in range(2):
        nll = m.flow.compute_psd_aware_nll(y, ctx, None, n_per_ctx=n)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        nll = m.flow.compute_psd_aware_nll(y, ctx, None, n_per_ctx=n)
    e1.record()
    torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
print(f"density: {E} x {n} rows, {ms:.3f} ms, {E * n / ms * 1e3 / 1e6:.2f} M rows/s, dense={os.environ.get('PF_TC_DENSE', '1')}, "
      f"nll mean {float(nll.mean()):.4f} finite {bool(torch.isfinite(nll).all())}")
