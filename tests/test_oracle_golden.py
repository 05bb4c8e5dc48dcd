"""CPU: the standalone oracle against the committed golden fixtures, which were produced by the
REFERENCE's own NSFPosteriorFlow / LeanNPE classes (oracle/make_golden.py, over the restated nflows
shim).  Bit-exactness was asserted in the build container by oracle/check_against_reference.py; here a
small tolerance absorbs CPU BLAS differences between machines."""
import numpy as np
import pytest
import torch

import nsf_oracle as O
from util import load_golden, rel_err, t

CASES = ["lean_default", "lean_stress_1layer", "head_8x192x16", "tiny_explicit"]


@pytest.mark.parametrize("case", CASES)
def test_oracle_inverse_forward_nll(case):
    spec, w, g = load_golden(case)
    z, ctx = t(g["z"]), t(g["ctx"])
    with torch.no_grad():
        x, ld = O.flow_inverse(spec, w, z, ctx)
        assert rel_err(x.numpy(), g["inv_x_f32"]) <= 2e-5
        assert rel_err(ld.numpy(), g["inv_logdet_f32"]) <= 2e-5
        y = t(g["fwd_x"])
        zz, fl = O.flow_forward(spec, w, y, ctx)
        assert rel_err(zz.numpy(), g["fwd_z_f32"]) <= 2e-5
        assert rel_err(fl.numpy(), g["fwd_logdet_f32"]) <= 2e-5
        assert rel_err(O.flow_nll(spec, w, y, ctx).numpy(), g["nll_f32"]) <= 2e-5
        xb, _ = O.flow_inverse(spec, w, z, t(g["ctx_bad"]))
        assert rel_err(xb.numpy(), g["inv_x_ctx_bad_f32"]) <= 2e-5


def test_oracle_float64_arbiter_tiny():
    spec, w, g = load_golden("tiny_explicit")
    w64 = {k: (v.double() if v.is_floating_point() else v) for k, v in w.items()}
    with torch.no_grad():
        x, ld = O.flow_inverse(spec, w64, t(g["z"]).double(), t(g["ctx"]).double())
    assert rel_err(x.numpy(), g["inv_x_f64"]) <= 1e-12
    assert rel_err(ld.numpy(), g["inv_logdet_f64"]) <= 1e-11


def test_incremental_inverse_equals_nflows_order():
    """The CUDA sampler's evaluation order (one conditioner pass per layer) is the same map as nflows'
    D-pass inverse."""
    spec, w, g = load_golden("tiny_explicit")
    w64 = {k: (v.double() if v.is_floating_point() else v) for k, v in w.items()}
    z, ctx = t(g["z"]).double(), t(g["ctx"]).double()
    with torch.no_grad():
        a, la = O.transform_inverse(spec, w64, z, ctx)
        b, lb = O.transform_inverse_incremental(spec, w64, z, ctx)
    assert float((a - b).abs().max()) < 1e-12 and float((la - lb).abs().max()) < 1e-11


def test_glue_fixture():
    spec, w, g = load_golden("lean_glue")
    E, n = int(g["E"]), int(g["n"])
    sc = O.Scaler()
    with torch.no_grad():
        fc = O.full_context(t(g["context"]), t(g["rank_embed"]), torch.zeros(E, dtype=torch.long))
        s = O.sample_posterior(spec, w, sc, fc, t(g["z"]), n)
        fcr = O.full_context(t(g["context"]), t(g["rank_embed"]), t(g["nll_ranks"]))
        nll = O.flow_nll(spec, w, sc.normalize(t(g["nll_theta"])), fcr)
    nonc = ~sc.circ_mask.numpy()
    assert rel_err(s.numpy()[..., nonc], g["samples_rank0"][..., nonc]) <= 1e-4
    assert rel_err(nll.numpy(), g["nll"]) <= 2e-5


def test_masks_and_counts():
    """SURVEY Appendix A.2 non-zero counts; per-layer and total parameter counts (Appendix B)."""
    spec = O.FlowSpec()
    m = O.made_masks(spec)
    assert int(m["initial"].sum()) == 1396 and int(m["hidden"].sum()) == 36046 and int(m["final"].sum()) == 66740
    w = O.init_weights(spec)
    n = sum(v.numel() for k, v in w.items() if k.endswith((".weight", ".bias")))
    assert n == 6210610 and n // 10 == 621061


def test_quantile_fixture_is_reproduced_by_the_oracle():
    """The posterior-scale fixture (16 000 samples of the reference flow, oracle/make_golden.py:quantile_case): the base
    noise regenerates from its seed, and the standalone oracle reproduces a slice of the stored samples."""
    from util import load_quantile_fixture
    spec, w, g, ctx, z = load_quantile_fixture("lean_quantiles_default")
    n = int(g["n"])
    with torch.no_grad():
        x, ld = O.flow_inverse(spec, w, z[n:n + 192], ctx[1:2].expand(192, -1))
        nll = O.flow_nll(spec, w, t(g["inv_x_f32"])[:192], ctx[0:1].expand(192, -1))
    assert rel_err(x.numpy(), g["inv_x_f32"][n:n + 192]) <= 2e-5
    assert rel_err(ld.numpy(), g["inv_logdet_f32"][n:n + 192]) <= 2e-5
    assert rel_err(nll.numpy(), g["nll_f64"][:192]) <= 1e-3      # float32 evaluation against the stored float64 one
