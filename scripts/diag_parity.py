"""Error statistics of the CUDA path against the float64 reference run, next to the float32 reference's
own deviation from float64 (test infrastructure: reads tests/golden, imports oracle/)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
from util import load_golden, make_flow, t  # noqa: E402


def stats(name, got, r32, r64):
    got, r32, r64 = (np.asarray(a, np.float64) for a in (got, r32, r64))
    den = np.maximum(1.0, np.abs(r64))
    e = np.abs(got - r64) / den
    o = np.abs(r32 - r64) / den
    q = lambda a: "med %.2e p90 %.2e p99 %.2e max %.2e" % (np.median(a), np.quantile(a, .9), np.quantile(a, .99), a.max())
    print(f"  {name:16s} ours: {q(e)}\n  {'':16s} ref32: {q(o)}")


for case in sys.argv[1:] or ["lean_default", "lean_stress", "lean_stress_2layer", "head_12x384x24"]:
    spec, w, g = load_golden(case)
    flow = make_flow(spec, w, "cuda")
    print(case)
    z, ctx = t(g["z"], "cuda"), t(g["ctx"], "cuda")
    x, ld = flow.inverse(z, ctx)
    stats("inverse x", x.cpu().numpy(), g["inv_x_f32"], g["inv_x_f64"])
    stats("inverse logdet", ld.cpu().numpy(), g["inv_logdet_f32"], g["inv_logdet_f64"])
    with torch.no_grad():
        zz, fl = flow.forward(t(g["fwd_x"], "cuda"), ctx)
    stats("forward z", zz.cpu().numpy(), g["fwd_z_f32"], g["fwd_z_f64"])
    stats("forward logdet", fl.cpu().numpy(), g["fwd_logdet_f32"], g["fwd_logdet_f64"])
