"""Shared helpers for the parity tests (test infrastructure; may import oracle/)."""
import os

import numpy as np
import torch

import nsf_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    g = dict(np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False))
    s = g["spec"]
    spec = O.FlowSpec(int(s[0]), int(s[1]), int(s[2]), int(s[3]), int(s[4]), float(s[5]), int(s[6]))
    if any(k.startswith("w:") for k in g):
        w = {k[2:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("w:")}
    else:
        w = O.init_weights(spec, seed=int(g["seed"]), stress=bool(g["stress"]))
    assert abs(O.weights_checksum(w) - float(g["weights_checksum"])) <= 1e-9 * abs(float(g["weights_checksum"])), \
        "regenerated weights differ from the ones the golden fixture was made with"
    return spec, w, g


def t(a, device="cpu"):
    return torch.from_numpy(np.ascontiguousarray(a)).to(device)


def make_flow(spec, w, device, tier="fp32"):
    """posteriflow_b200.NSFPosteriorFlow holding the oracle weight set `w` (strict state-dict load)."""
    import posteriflow_b200 as P
    flow = P.NSFPosteriorFlow(features=spec.features, context_features=spec.context_features,
                              hidden_features=spec.hidden_features, num_layers=spec.num_layers,
                              num_bins=spec.num_bins, tail_bound=spec.tail_bound, dropout=0.0,
                              temperature_scale=1.0, use_masked_context=False, tier=tier)
    flow.load_state_dict(O.full_state_dict(spec, w), strict=True)
    return flow.to(device)


def rel_err(a, ref):
    """max |a - ref| / max(1, |ref|) -- the tolerance form of SURVEY.md 8c / BASELINE north_star."""
    a, ref = np.asarray(a, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    return float(np.max(np.abs(a - ref) / np.maximum(1.0, np.abs(ref)))) if a.size else 0.0


ARBITRATED = {}   # name -> largest arbitrated fraction seen in this process (printed by the parity tests)
TOL = 1e-4  # FP32 tier: |d| <= 1e-4 * max(1, |ref|)  (BASELINE.json north_star)


def check_parity(name, got, ref32, ref64=None, tol=TOL, scale=None, median_factor=2.0):
    """Pass if `got` is within `tol` of the float32 reference.  Where it is not and a float64 run of
    the same reference code is available, the float64 run arbitrates: with ill-conditioned (stress)
    weights the float32 reference ITSELF sits further than `tol` from float64, and any other valid
    float32 evaluation order (ours) is a second draw from the same error distribution.  It must then be
    statistically indistinguishable from the reference's own float32 error: median within 2x, maximum
    within 8x (+tol).  `scale` switches from per-element max(1,|ref|) to one global scale (gradients); median_factor=None
    drops the median test for entries whose errors are correlated (norms of one backward pass)."""
    got, ref32 = np.asarray(got, np.float64), np.asarray(ref32, np.float64)
    assert got.shape == ref32.shape, (name, got.shape, ref32.shape)
    if got.size == 0:
        return 0.0
    den = (lambda r: np.maximum(1.0, np.abs(r))) if scale is None else (lambda r: scale)
    e32 = np.abs(got - ref32) / den(ref32)
    # fraction of entries that are NOT within `tol` of the float32 reference, i.e. that needed the float64 arbiter:
    # returned so that callers can assert 0 for the well-conditioned weight sets and report it for the stress sets
    arbitrated = float(np.mean(e32 > tol))
    ARBITRATED[name] = max(ARBITRATED.get(name, 0.0), arbitrated)
    if e32.max() <= tol:
        return 0.0
    assert ref64 is not None, f"{name}: rel err {e32.max():.3e} > {tol} (no float64 arbiter)"
    ref64 = np.asarray(ref64, np.float64)
    e = np.abs(got - ref64) / den(ref64)
    own = np.abs(ref32 - ref64) / den(ref64)
    ok = e.max() <= 8 * own.max() + tol
    if median_factor is not None:
        ok = ok and np.median(e) <= median_factor * np.median(own) + 1e-6
    assert ok, (f"{name}: vs f32 max {e32.max():.3e}; vs f64 median {np.median(e):.3e} max {e.max():.3e}; "
                f"f32 reference's own vs f64 median {np.median(own):.3e} max {own.max():.3e}")
    return arbitrated


def load_quantile_fixture(name):
    """(spec, w, g, ctx [E, C], z [E*n, D]) of a `lean_quantiles_*` fixture; z regenerated from its stored seed."""
    spec, w, g = load_golden(name)
    E, n = int(g["E"]), int(g["n"])
    gen = torch.Generator().manual_seed(int(g["noise_seed"]))
    ctx = torch.randn(E, spec.context_features, generator=gen)
    z = torch.randn(E * n, spec.features, generator=gen)
    assert float((ctx - t(g["ctx"])).abs().max()) == 0.0
    assert abs(float(z.double().abs().sum()) - float(g["z_checksum"])) <= 1e-9 * float(g["z_checksum"])
    return spec, w, g, ctx, z
