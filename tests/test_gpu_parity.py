"""GPU parity: the CUDA path (through torch.library ops -> C-ABI) against the golden fixtures made by
the reference's own classes (oracle/make_golden.py) and against the standalone oracle on fresh inputs.

Tolerance of the FP32 tier (BASELINE.json north_star): |dtheta|, |dlogq| <= 1e-4 * max(1, |ref|).
The reference run in float64 is the arbiter where float32 evaluations legitimately differ: an error is
accepted if it is within 1e-4 of the float32 reference OR statistically indistinguishable from the float32
reference's own deviation from float64 (util.check_parity: median within 2x, maximum within 8x).
"""
import numpy as np
import pytest
import torch

import nsf_oracle as O
from util import check_parity, load_golden, make_flow, rel_err, t

pytestmark = pytest.mark.gpu
TOL = 1e-4
CASES = ["lean_default", "lean_stress", "lean_stress_1layer", "lean_stress_2layer", "head_8x192x16",
         "head_12x384x24", "tiny_explicit"]


check = check_parity
# Fraction of entries that may need the float64 arbiter, per weight set (measured on the B200, round 2, + margin): a
# regression that pushes entries from "within 1e-4 of the float32 reference" to "arbitrated" must fail, not pass quietly.
# Only lean_default carries well-conditioned (default-init) weights: there the float32 reference itself is within 1e-4
# of float64 except on single entries (measured: 1 of 64 inverse log-dets, 1 of 704 z entries, 3 of 64 forward log-dets --
# the forward inputs sit exactly on spline knots).  Every other fixture -- the two
# head shapes and tiny_explicit included -- is built with the ill-conditioned stress weights (make_golden.py: stress=True),
# where a third of the reference's own float32 entries sit further than 1e-4 from its float64 run.
ARBITRATED_BOUND = {"lean_default": 0.08, "tiny_explicit": 0.10, "head_8x192x16": 0.60, "head_12x384x24": 0.75}


def strict(case, frac, what):
    print(f"[parity] {case}: {what}: fraction of entries that needed the float64 arbiter = {frac:.4f}")
    bound = ARBITRATED_BOUND.get(case)
    if bound is not None:
        assert frac <= bound, f"{case}: {what}: {frac:.4f} of the entries exceed 1e-4 against the float32 reference (bound {bound})"


@pytest.mark.parametrize("case", CASES)
def test_inverse_matches_reference(case):
    spec, w, g = load_golden(case)
    flow = make_flow(spec, w, "cuda")
    z, ctx = t(g["z"], "cuda"), t(g["ctx"], "cuda")
    x, ld = flow.inverse(z, ctx)
    strict(case, check("inverse x", x.cpu().numpy(), g["inv_x_f32"], g["inv_x_f64"]), "inverse x")
    strict(case, check("inverse logdet", ld.cpu().numpy(), g["inv_logdet_f32"], g["inv_logdet_f64"]), "inverse logdet")
    assert float(x.abs().max()) <= 3.0
    # NaN/Inf context is sanitised exactly like flows.py:626-632
    xb, _ = flow.inverse(z, t(g["ctx_bad"], "cuda"))
    # no float64 run of the bad-context case is stored: only rows 3..5 differ from the case above, so the
    # same arbiter is used for all other rows and the three sanitised rows are checked on their own
    xb, ref_b = xb.cpu().numpy(), g["inv_x_ctx_bad_f32"]
    keep = np.ones(len(xb), bool)
    keep[3:6] = False
    check("inverse x (bad ctx, untouched rows)", xb[keep], ref_b[keep], g["inv_x_f64"][keep])
    assert np.isfinite(xb).all()
    assert np.median(np.abs(xb[3:6] - ref_b[3:6])) <= 1e-5 and np.abs(xb[3:6] - ref_b[3:6]).max() <= 2e-2


@pytest.mark.parametrize("case", CASES)
def test_forward_and_nll_match_reference(case):
    spec, w, g = load_golden(case)
    flow = make_flow(spec, w, "cuda")
    x, ctx = t(g["fwd_x"], "cuda"), t(g["ctx"], "cuda")
    with torch.no_grad():
        z, ld = flow.forward(x, ctx)
        nll = flow.compute_psd_aware_nll(x, ctx, torch.zeros_like(x))
        nll_none = flow.compute_psd_aware_nll(x, ctx, None)
    strict(case, check("forward z", z.cpu().numpy(), g["fwd_z_f32"], g["fwd_z_f64"]), "forward z")
    strict(case, check("forward logdet", ld.cpu().numpy(), g["fwd_logdet_f32"], g["fwd_logdet_f64"]), "forward logdet")
    strict(case, check("nll", nll.cpu().numpy(), g["nll_f32"], g["nll_f64"]), "nll")
    assert torch.equal(nll, nll_none)


@pytest.mark.parametrize("tier", ["fp32", "bf16"])
@pytest.mark.parametrize("case", ["lean_default", "lean_stress", "lean_stress_1layer", "tiny_explicit"])
def test_backward_matches_reference_autograd(case, tier, monkeypatch):
    """tier="bf16": the training path of the tensor-core tier (split-bf16 tcgen05 GEMMs, pf_tc_gemm.cuh: three MMAs per
    product keep ~16 significand bits), forced on for these small batches (PF_TRAIN_TC_MIN_ROWS=0; by default calls
    under 512 rows stay on the FP32-FMA kernels).  Stated tolerance of that path, float64-arbitrated like the FP32
    one: gradients of x / context within 2e-3 of the gradient's scale, weight-gradient norms within 1e-3, sampled
    weight-gradient entries within 5e-2 of the tensor's rms entry (measured 2.8e-2: the fixture puts inputs exactly ON
    spline knots, where a 1e-5 change of a knot moves a row to the neighbouring bin -- the flow is C1 there, the
    parameter gradients are not), median error within 16x of the float32 reference's own (measured 5-7x)."""
    mf, tol_g, tol_n, tol_s = (2.0, TOL, TOL, TOL) if tier == "fp32" else (16.0, 2e-3, 1e-3, 5e-2)
    if tier == "bf16":
        monkeypatch.setenv("PF_TRAIN_TC_MIN_ROWS", "0")      # read when the handle is created
    if tier == "bf16" and case == "tiny_explicit":
        pytest.skip("the tensor-core tier covers the LeanNPE / head shapes only (tests/test_gpu_tc.py)")
    spec, w, g = load_golden(case)
    flow = make_flow(spec, w, "cuda", tier=tier)
    x = t(g["fwd_x"], "cuda").requires_grad_(True)
    ctx = t(g["ctx"], "cuda").requires_grad_(True)
    gw = t(g["grad_upstream"], "cuda").float()
    loss = (flow.compute_psd_aware_nll(x, ctx, torch.zeros_like(x)) * gw).sum()
    loss.backward()
    def gcheck(name, got, r32, r64):  # relative to the gradient's own scale (max |ref|)
        check_parity(name, got, r32, r64, tol=tol_g, scale=max(1e-6, float(np.abs(r64).max())), median_factor=mf)
    gcheck("grad x", x.grad.cpu().numpy(), g["grad_x_f32"], g["grad_x_f64"])
    gcheck("grad ctx", ctx.grad.cpu().numpy(), g["grad_ctx_f32"], g["grad_ctx_f64"])
    names = [str(n) for n in g["grad_w_names"]]
    params = dict(flow.named_parameters())
    got_norm, got_s, ref_s32, ref_s64 = [], [], [], []
    for i, n in enumerate(names):
        p = params[n]
        assert p.grad is not None, n
        gr = p.grad.reshape(-1).double().cpu()
        got_norm.append(float(gr.norm()))
        idx = torch.arange(0, gr.numel(), max(1, gr.numel() // 64))[:64]
        sc = max(1e-9, float(g["grad_w_norms_f64"][i]) / np.sqrt(gr.numel()))  # rms entry of this tensor
        pad = np.zeros(64)
        pad[: idx.numel()] = gr[idx].numpy() / sc
        got_s.append(pad)
        ref_s32.append(g["grad_w_samples_f32"][i] / sc)
        ref_s64.append(g["grad_w_samples_f64"][i] / sc)
    n64 = g["grad_w_norms_f64"]
    check_parity("weight-grad norms", np.array(got_norm) / n64, g["grad_w_norms_f32"] / n64, np.ones_like(n64),
                 tol=tol_n, median_factor=None)
    # sampled entries, each tensor normalised by its own rms gradient entry
    check_parity("weight-grad samples", np.stack(got_s), np.stack(ref_s32), np.stack(ref_s64),
                 tol=tol_s, scale=max(1.0, float(np.abs(np.stack(ref_s64)).max())), median_factor=mf)
    # masked entries receive exactly zero gradient, like d(W*mask)/dW
    net = flow.transform._transforms[1].autoregressive_net
    for lin in (net.initial_layer, net.final_layer, net.blocks[0].linear_layers[0], net.blocks[1].linear_layers[1]):
        assert float((lin.weight.grad * (1 - lin.mask)).abs().max()) == 0.0
