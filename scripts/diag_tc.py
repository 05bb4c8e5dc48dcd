"""Tensor-core (bf16) tier of the sampler vs the float64 reference run and vs the FP32 tier, plus a quick
throughput reading (test infrastructure: reads tests/golden, imports oracle/).  Run on a B200."""
import os
import sys
import time

import numpy as np
import torch

os.environ.setdefault("PF_TC_DEBUG", "1")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
from util import load_golden, make_flow, t  # noqa: E402
import posteriflow_b200 as P  # noqa: E402


def q(a):
    a = np.asarray(a, np.float64)
    return "med %.2e p90 %.2e p99 %.2e max %.2e" % (np.median(a), np.quantile(a, .9), np.quantile(a, .99), a.max())


for case in [c for c in sys.argv[1:] if not c.startswith("-")] or ["lean_default", "lean_stress"]:
    spec, w, g = load_golden(case)
    f32 = make_flow(spec, w, "cuda", tier="fp32")
    tc = make_flow(spec, w, "cuda", tier="bf16")
    print(case, flush=True)
    z, ctx = t(g["z"], "cuda"), t(g["ctx"], "cuda")
    xa, la = f32.inverse(z, ctx)
    xb, lb = tc.inverse(z, ctx)
    torch.cuda.synchronize()
    r64, l64 = g["inv_x_f64"], g["inv_logdet_f64"]
    den = np.maximum(1.0, np.abs(r64))
    print("  rows", z.shape[0])
    print("  x   fp32 tier vs f64:", q(np.abs(xa.cpu().numpy() - r64) / den))
    print("  x   bf16 tier vs f64:", q(np.abs(xb.cpu().numpy() - r64) / den))
    print("  ld  fp32 tier vs f64:", q(np.abs(la.cpu().numpy() - l64) / np.maximum(1, np.abs(l64))))
    print("  ld  bf16 tier vs f64:", q(np.abs(lb.cpu().numpy() - l64) / np.maximum(1, np.abs(l64))))
    with torch.no_grad():
        za, fa = f32.forward(t(g["fwd_x"], "cuda"), ctx)
        zb, fb = tc.forward(t(g["fwd_x"], "cuda"), ctx)
    z64, f64 = g["fwd_z_f64"], g["fwd_logdet_f64"]
    print("  fwd z  fp32 tier vs f64:", q(np.abs(za.cpu().numpy() - z64) / np.maximum(1, np.abs(z64))))
    print("  fwd z  bf16 tier vs f64:", q(np.abs(zb.cpu().numpy() - z64) / np.maximum(1, np.abs(z64))))
    print("  fwd ld bf16 tier vs f64:", q(np.abs(fb.cpu().numpy() - f64) / np.maximum(1, np.abs(f64))))
    # broadcast form, many tiles, ragged last tile
    gen = torch.Generator().manual_seed(5)
    E, n = 3, 1111
    c2 = torch.randn(E, spec.context_features, generator=gen).cuda()
    z2 = torch.randn(E * n, spec.features, generator=gen).cuda()
    xa, la = f32.inverse(z2, c2, n_per_ctx=n)
    xb, lb = tc.inverse(z2, c2, n_per_ctx=n)
    d = (xa - xb).abs() / xa.abs().clamp_min(1)
    print("  broadcast 3x1111: x bf16 vs fp32 tier:", q(d.cpu().numpy()), " ld:",
          q(((la - lb).abs() / la.abs().clamp_min(1)).cpu().numpy()), flush=True)

if "--time" in sys.argv:
    spec, w, g = load_golden("lean_default")
    sc = P.ParamScaler().to("cuda")
    for tier in ("bf16", "fp32"):
        flow = make_flow(spec, w, "cuda", tier=tier)
        E, n = (148, 5000) if tier == "bf16" else (32, 5000)
        ctx = torch.randn(E, spec.context_features, device="cuda")
        for it in range(3):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            th, ld, rl = P.sample_fused(flow, sc, ctx, n, z=None, seed=it)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            print(f"  {tier}: {E}x{n} samples in {dt*1e3:.1f} ms -> {E*n/dt/1e6:.2f} M samples/s", flush=True)
        y = th.reshape(-1, spec.features)
        yn = sc.normalize(y)
        for it in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            with torch.no_grad():
                nll = flow.compute_psd_aware_nll(yn, ctx, None, n_per_ctx=n)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            print(f"  {tier}: log-prob of {E*n} rows in {dt*1e3:.1f} ms -> {E*n/dt/1e6:.2f} M rows/s", flush=True)
