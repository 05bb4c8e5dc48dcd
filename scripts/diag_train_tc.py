"""Training path of the tensor-core tier (split-bf16 tcgen05 GEMMs, pf_tc_gemm.cuh) against the FP32 tier:
z / logdet / every gradient on the same inputs, error relative to each tensor's own max |value|.

    python scripts/diag_train_tc.py [rows] [case]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import load_golden, make_flow  # noqa: E402


def run(flow, x, ctx, gw):
    x = x.clone().requires_grad_(True)
    ctx = ctx.clone().requires_grad_(True)
    for p in flow.parameters():
        p.grad = None
    z, ld = flow.forward(x, ctx)
    loss = ((z * z).sum(dim=1) * 0.5 - ld) @ gw
    loss.backward()
    grads = {n: p.grad.clone() for n, p in flow.named_parameters() if p.grad is not None}
    return z.detach(), ld.detach(), x.grad, ctx.grad, grads


def main():
    R = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    case = sys.argv[2] if len(sys.argv) > 2 else "lean_default"
    spec, w, g = load_golden(case)
    gen = torch.Generator().manual_seed(5)
    x = ((torch.rand(R, spec.features, generator=gen) * 2 - 1) * 1.5).cuda()
    ctx = torch.randn(R, spec.context_features, generator=gen).cuda()
    gw = torch.rand(R, generator=gen).cuda()

    def variant(tier, streams):
        os.environ["PF_TRAIN_STREAMS"] = "1" if streams else "0"     # read when the handle is created
        fl = make_flow(spec, w, "cuda", tier=tier)
        out = run(fl, x, ctx, gw)
        torch.cuda.synchronize()
        return out
    a = variant("fp32", False)
    which = sys.argv[3] if len(sys.argv) > 3 else "tc1"
    b = variant("fp32" if which.startswith("f") else "bf16", which.endswith("1"))
    print(f"reference: fp32 tier, one stream; candidate: {which}")

    def rel(u, v):
        return float((u - v).abs().max() / v.abs().max().clamp_min(1e-30))
    print(f"rows {R} case {case}")
    for name, u, v in (("z", b[0], a[0]), ("logdet", b[1], a[1]), ("grad x", b[2], a[2]), ("grad ctx", b[3], a[3])):
        print(f"  {name:10s} max|d|/max|ref| = {rel(u, v):.3e}")
    worst = sorted(((rel(b[4][n], a[4][n]), n) for n in a[4]), reverse=True)
    for e, n in worst[:6]:
        print(f"  grad {n:75s} {e:.3e}")
    print(f"  weight grads: {len(worst)} tensors, worst {worst[0][0]:.3e}, median {worst[len(worst) // 2][0]:.3e}")


if __name__ == "__main__":
    main()
