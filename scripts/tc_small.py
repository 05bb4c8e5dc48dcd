"""Smallest tensor-core-tier invocation (one 128-row tile) -- for compute-sanitizer / ncu on a B200."""
import os
import sys

import torch

os.environ.setdefault("PF_TC_DEBUG", "1")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
from util import load_golden, make_flow  # noqa: E402

spec, w, g = load_golden("lean_default")
tc = make_flow(spec, w, "cuda", tier="bf16")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
gen = torch.Generator().manual_seed(5)
ctx = torch.randn(1, spec.context_features, generator=gen).cuda()
z = torch.randn(n, spec.features, generator=gen).cuda()
x, ld = tc.inverse(z, ctx, n_per_ctx=n)
torch.cuda.synchronize()
print("ok", x.shape, float(x.abs().max()), float(ld.mean()))
