"""Minimal restatement of the parts of `nflows` (bayesiains/nflows, MIT, last release 0.14)
that PosteriFlow's LeanNPE flow path imports.

TEST INFRASTRUCTURE ONLY.  This package exists so that the reference's own
`ahsd.models.flows` / `ahsd.models.lean_npe` import and run UNMODIFIED on CPU:

    PYTHONPATH=/root/repo/oracle:/root/reference/src python -c "from ahsd.models.lean_npe import LeanNPE"

The real nflows is not vendored in /root/reference, is unpinned there
(environment.yaml:35 `- nflows`) and cannot be installed offline, so its published
algorithm is restated here in plain PyTorch (dtype-generic: also runs in float64).
Import sites in the reference that this shim must satisfy:
  src/ahsd/models/flows.py:11        from nflows import distributions
  src/ahsd/models/flows.py:21-25     Flow, StandardNormal, CompositeTransform,
                                     MaskedPiecewiseRationalQuadraticAutoregressiveTransform,
                                     ReversePermutation
  src/ahsd/models/flows.py:203,258   nflows.transforms.made.MaskedLinear, _get_input_degrees
  src/ahsd/models/flows.py:337-338   AutoregressiveTransform, splines.rational_quadratic defaults

PARITY UNPINNED: the reference ships no test, fixture or checkpoint for this path
(SURVEY.md section 4), and real nflows cannot be run here, so the three recalled
behaviours below are one-line switches.
"""

# --- recalled-behaviour switches (SURVEY.md section 8c "unverifiable-here items") -------------
# (i) nflows' MADE does not expose `.hidden_features`; if it did, the spline width/height
#     logits would be divided by sqrt(hidden_features) (autoregressive.py `_elementwise`).
MADE_EXPOSES_HIDDEN_FEATURES = False
# (ii) MaskedResidualBlock gates its output with a GLU on the context projection
#      (F.glu(cat(temps, context_layer(context)))) rather than adding it.
RESIDUAL_BLOCK_CONTEXT_GLU = True

from . import distributions, flows, transforms, utils  # noqa: E402,F401

__version__ = "0.14+oracle-shim"
