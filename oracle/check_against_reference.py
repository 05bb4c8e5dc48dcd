"""Pin the standalone oracle (oracle/nsf_oracle.py) to the reference's OWN code.

TEST INFRASTRUCTURE ONLY; runs in the build container where /root/reference is mounted:

    python oracle/check_against_reference.py

It imports the unmodified reference classes (`ahsd.models.flows.NSFPosteriorFlow`,
`ahsd.models.lean_npe.{LeanNPE, ParamScaler}`) over the restated `oracle/nflows` shim, loads the
same seeded weights into them with a STRICT `load_state_dict` (which also checks the state-dict
key set of SURVEY.md Appendix B), and asserts bit-for-bit float32 equality on CPU for
forward / inverse / NLL / ParamScaler / sample_posterior glue / physical log-prob, for the
LeanNPE shape, the two `experiments/frozen_context_heads.py:159-163` shapes and a tiny shape,
with default-init and stress weights.
"""
from __future__ import annotations

import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("POSTERIFLOW_REF", "/root/reference")
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(REF, "src"))

import logging  # noqa: E402
import warnings  # noqa: E402

import torch  # noqa: E402

warnings.filterwarnings("ignore")
logging.disable(logging.CRITICAL)

import nsf_oracle as O  # noqa: E402


def reference_flow(spec: O.FlowSpec, w):
    from ahsd.models.flows import NSFPosteriorFlow
    flow = NSFPosteriorFlow(features=spec.features, context_features=spec.context_features,
                            hidden_features=spec.hidden_features, num_layers=spec.num_layers,
                            num_bins=spec.num_bins, tail_bound=spec.tail_bound, dropout=0.0,
                            temperature_scale=1.0, use_masked_context=False)
    flow.load_state_dict(O.full_state_dict(spec, w), strict=True)
    return flow.eval()


def make_inputs(spec: O.FlowSpec, R: int, seed: int):
    g = torch.Generator().manual_seed(seed)
    ctx = torch.randn(R, spec.context_features, generator=g)
    z = torch.randn(R, spec.features, generator=g) * 1.7
    # edge cases: beyond the tails, exactly on the tail bound, zeros
    z[0, 0] = spec.tail_bound
    z[0, 1] = -spec.tail_bound
    z[1, 2] = spec.tail_bound + 1.5
    z[1, 3] = -spec.tail_bound - 0.25
    z[2, :] = 0.0
    return z, ctx


def check(spec: O.FlowSpec, seed: int, stress: bool, R: int = 96) -> None:
    w = O.init_weights(spec, seed=seed, stress=stress)
    flow = reference_flow(spec, w)
    z, ctx = make_inputs(spec, R, seed + 100)
    with torch.no_grad():
        x_ref, ld_ref = flow.transform.inverse(z, context=ctx)
        x_ora, ld_ora = O.transform_inverse(spec, w, z, ctx)
        assert torch.equal(x_ref, x_ora) and torch.equal(ld_ref, ld_ora), "transform.inverse differs"
        xc_ref, _ = flow.inverse(z, ctx)
        xc_ora, _ = O.flow_inverse(spec, w, z, ctx)
        assert torch.equal(xc_ref, xc_ora), "NSFPosteriorFlow.inverse differs"
        y = xc_ref
        z_ref, ldf_ref = flow.forward(y, ctx)
        z_ora, ldf_ora = O.flow_forward(spec, w, y, ctx)
        assert torch.equal(z_ref, z_ora) and torch.equal(ldf_ref, ldf_ora), "forward differs"
        nll_ref = flow.compute_psd_aware_nll(y, ctx, torch.zeros_like(y))
        nll_ora = O.flow_nll(spec, w, y, ctx)
        assert torch.equal(nll_ref, nll_ora), "nll differs"
        ls = torch.randn(R, spec.features) * 0.1
        assert torch.equal(flow.compute_psd_aware_nll(y, ctx, ls), O.flow_nll(spec, w, y, ctx, ls))
        # incremental ordering == D-pass ordering (float64: identical up to rounding of unused terms)
        w64 = {k: (v.double() if v.is_floating_point() else v) for k, v in w.items()}
        a, la = O.transform_inverse(spec, w64, z.double(), ctx.double())
        b, lb = O.transform_inverse_incremental(spec, w64, z.double(), ctx.double())
        assert (a - b).abs().max() < 1e-12 and (la - lb).abs().max() < 1e-11, "incremental inverse differs"
        # NaN/Inf context sanitise path
        cbad = ctx.clone()
        cbad[3, 5] = float("nan"); cbad[4, 7] = float("inf"); cbad[5, 9] = float("-inf")
        assert torch.equal(flow.inverse(z, cbad)[0], O.flow_inverse(spec, w, z, cbad)[0])
    print(f"  ok  spec={spec} stress={stress}  max|x|={x_ref.abs().max():.3f} "
          f"logdet range [{ld_ref.min():.2f},{ld_ref.max():.2f}]")


def check_lean_npe() -> None:
    from ahsd.models.lean_npe import LeanNPE, ParamScaler, PARAM_NAMES
    assert PARAM_NAMES == O.PARAM_NAMES
    spec = O.FlowSpec()
    w = O.init_weights(spec, seed=3, stress=True)
    torch.manual_seed(0)
    model = LeanNPE().eval()
    model.flow.load_state_dict(O.full_state_dict(spec, w), strict=True)
    for premerger in (False, True):
        a, b = ParamScaler(premerger=premerger), O.Scaler(premerger=premerger)
        assert torch.equal(a.lo, b.lo) and torch.equal(a.hi, b.hi)
        assert torch.equal(a.log_mask, b.log_mask) and torch.equal(a.circ_mask, b.circ_mask)
        y = torch.randn(64, 11) * 1.5
        assert torch.equal(a.wrap(y), b.wrap(y))
        assert torch.equal(a.denormalize(a.wrap(y)), b.denormalize(b.wrap(y)))
        th = a.denormalize(a.wrap(y))
        assert torch.equal(a.normalize(th), b.normalize(th))
    E, n = 3, 40
    g = torch.Generator().manual_seed(11)
    context = torch.randn(E, 256, generator=g)
    rank = torch.tensor([0, 2, 4])
    with torch.no_grad():
        fc_ref = model._full_context(context, rank)
        fc_ora = O.full_context(context, model.rank_embed.weight, rank)
        assert torch.equal(fc_ref, fc_ora)
        # nll through LeanNPE.nll with a given context
        theta = O.Scaler().denormalize(torch.rand(E, 11) * 2 - 1)
        nll_ref = model.nll(None, theta, rank, context=context)
        nll_ora = O.flow_nll(spec, w, O.Scaler().normalize(theta), fc_ora)
        assert torch.equal(nll_ref, nll_ora)
        # sample_posterior glue: patch the encoder + randn so the reference method itself runs
        z = torch.randn(E * n, 11, generator=g)
        model.encode = lambda strain, asd=None: context
        r0 = torch.zeros(E, dtype=torch.long)
        orig_randn = torch.randn
        torch.randn = lambda *a, **k: z
        try:
            s_ref = model.sample_posterior(torch.zeros(E, 3, 8), rank=0, n_samples=n)
        finally:
            torch.randn = orig_randn
        s_ora = O.sample_posterior(spec, w, O.Scaler(), O.full_context(context, model.rank_embed.weight, r0), z, n)
        assert torch.equal(s_ref, s_ora)
        # physical log-prob (pipeline.py:57-76); ahsd.inference imports heavy deps lazily -> restate call
        try:
            from ahsd.inference.pipeline import _log_prob_physical
            ynorm = O.Scaler().wrap(torch.randn(32, 11))
            fc = fc_ora[:1].expand(32, -1)
            assert torch.equal(_log_prob_physical(model, ynorm, fc), O.log_prob_physical(spec, w, O.Scaler(), ynorm, fc))
            print("  ok  pipeline._log_prob_physical")
        except ImportError as e:  # gwpy/bilby absent: the function body is restated verbatim in the oracle
            print(f"  (pipeline import unavailable here: {e}; _log_prob_physical checked by formula only)")
    print("  ok  LeanNPE glue / ParamScaler / sample_posterior")


def main() -> None:
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    print("reference:", REF)
    check(O.FlowSpec(), seed=0, stress=False)
    check(O.FlowSpec(), seed=1, stress=True)
    check(O.FlowSpec(11, 288, 192, 8, 16, 3.0), seed=2, stress=True, R=48)    # frozen_context_heads 8x192x16
    check(O.FlowSpec(11, 288, 384, 12, 24, 3.0), seed=3, stress=True, R=32)   # frozen_context_heads 12x384x24
    check(O.FlowSpec(5, 12, 32, 3, 8, 2.0), seed=4, stress=True, R=64)        # tiny
    check_lean_npe()
    print("ORACLE PINNED TO REFERENCE-OVER-SHIM: all checks bit-exact")


if __name__ == "__main__":
    main()
