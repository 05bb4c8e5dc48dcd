"""Host logic of the callers either side of the hot path, on CPU:
`signal_rows` reproduces the row sets of the reference's per-rank loop (train_lean_npe.py:114-124) and
`save_posterior` writes the files of PosteriorResult.save (result.py:242-252) byte for byte."""
import numpy as np
import torch

from posteriflow_b200 import PARAM_NAMES
from posteriflow_b200.result import load_posterior, posterior_arrays, save_posterior
from posteriflow_b200.train import batch_nll, batch_nll_reference_order, signal_rows


def _reference_rows(nsig):
    ev, rk = [], []
    for r in range(int(nsig.max().item())):          # train_lean_npe.py:116-122
        mask = nsig > r
        if not mask.any():
            continue
        idx = mask.nonzero(as_tuple=True)[0]
        ev.append(idx)
        rk.append(torch.full((idx.numel(),), r, dtype=torch.long))
    return torch.cat(ev), torch.cat(rk)


def test_signal_rows_are_the_reference_loop_rows():
    g = torch.Generator().manual_seed(3)
    for B in (1, 7, 512):
        nsig = torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), B, replacement=True, generator=g) + 1
        ev, rk = signal_rows(nsig, 5)
        rev, rrk = _reference_rows(nsig)
        assert torch.equal(ev, rev) and torch.equal(rk, rrk)
        assert ev.numel() == int(nsig.sum())
    # ragged / degenerate inputs: no signals at all, a rank level nobody reaches, nsig above max_signals
    ev, rk = signal_rows(torch.zeros(4, dtype=torch.long), 5)
    assert ev.numel() == 0 and rk.numel() == 0
    ev, rk = signal_rows(torch.tensor([1, 3, 1]), 5)
    assert ev.tolist() == [0, 1, 2, 1, 1] and rk.tolist() == [0, 0, 0, 1, 2]
    ev, rk = signal_rows(torch.tensor([9]), 5)       # params hold max_signals slots: ranks stop there
    assert rk.tolist() == [0, 1, 2, 3, 4]


class _ToyModel:
    """nll(params, rank, context) with the call signature of LeanNPE.nll; counts calls."""
    max_signals = 5

    def __init__(self):
        self.calls = 0

    def nll(self, strain, params_phys, rank, context=None, asd_bands=None):
        self.calls += 1
        return (params_phys ** 2).sum(1) + context.sum(1) * (rank.float() + 1.0)


def test_batch_nll_is_one_call_and_equals_the_loop():
    g = torch.Generator().manual_seed(0)
    B = 64
    nsig = torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), B, replacement=True, generator=g) + 1
    params = torch.randn(B, 5, 11, generator=g, dtype=torch.float64)
    ctx = torch.randn(B, 8, generator=g, dtype=torch.float64, requires_grad=True)
    m = _ToyModel()
    fused = batch_nll(m, None, params, nsig, context=ctx)
    assert m.calls == 1
    (gf,) = torch.autograd.grad(fused, ctx)
    m2 = _ToyModel()
    loop = batch_nll_reference_order(m2, None, params, nsig, context=ctx)
    assert m2.calls == int(nsig.max())
    (gl,) = torch.autograd.grad(loop, ctx)
    assert abs(float(fused) - float(loop)) <= 1e-12 * abs(float(loop))
    assert torch.allclose(gf, gl, rtol=1e-12, atol=1e-14)
    empty = batch_nll(m, None, params, torch.zeros(B, dtype=torch.long), context=ctx)
    assert float(empty) == 0.0 and empty.requires_grad


def _reference_save(out, samples, log_prob, param_names):
    """result.py:242-252 restated (np.save x2 + np.savetxt with the header line)."""
    np.save(out / "posterior_samples.npy", samples)
    np.save(out / "posterior_log_prob.npy", log_prob)
    header = ",".join(param_names + ["log_prob"])
    np.savetxt(out / "posterior_samples.csv", np.column_stack([samples, log_prob]), delimiter=",", header=header,
               comments="")


def test_save_posterior_writes_the_reference_files(tmp_path):
    g = torch.Generator().manual_seed(1)
    theta = torch.randn(3, 40, 11, generator=g)               # device-side layout [E, n, D] float32
    logq = torch.randn(120, generator=g)
    ours = save_posterior(tmp_path / "ours", theta, logq)
    ref = tmp_path / "ref"
    ref.mkdir()
    # infer() hands PosteriorResult float64 host arrays (pipeline.py:163-165)
    _reference_save(ref, theta.reshape(-1, 11).numpy().astype(np.float64), logq.numpy().astype(np.float64),
                    list(PARAM_NAMES))
    for name in ("posterior_samples.npy", "posterior_log_prob.npy", "posterior_samples.csv"):
        assert (ours / name).read_bytes() == (ref / name).read_bytes(), name
    s, lp, names = load_posterior(ours)
    assert s.dtype == np.float64 and s.shape == (120, 11) and lp.shape == (120,)
    assert names == list(PARAM_NAMES)
    assert (ours / "posterior_samples.csv").read_text().splitlines()[0] == \
        "mass_1,mass_2,luminosity_distance,ra,dec,theta_jn,psi,phase,geocent_time,a1,a2,log_prob"


def test_save_posterior_matches_the_reference_class_itself(tmp_path):
    """The same three files written by the reference's OWN `PosteriorResult.save` (result.py:242-252), imported
    unmodified over the nflows shim.  Build-container only: /root/reference does not exist on the GPU box."""
    import os
    import sys
    import pytest
    ref_src = os.path.join(os.environ.get("POSTERIFLOW_REF", "/root/reference"), "src")
    if not os.path.isdir(ref_src):
        pytest.skip("reference tree not present (GPU box)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for pth in (os.path.join(root, "oracle"), ref_src):
        if pth not in sys.path:
            sys.path.insert(0, pth)
    import logging
    logging.disable(logging.CRITICAL)
    try:
        # the file itself, unmodified (`import ahsd.inference` would pull in the bilby-based pipeline, absent here)
        import importlib.util
        sp = importlib.util.spec_from_file_location("ahsd_inference_result",
                                                    os.path.join(ref_src, "ahsd", "inference", "result.py"))
        mod = importlib.util.module_from_spec(sp)
        sys.modules["ahsd_inference_result"] = mod
        sp.loader.exec_module(mod)
        PosteriorResult = mod.PosteriorResult
    finally:
        logging.disable(logging.NOTSET)
    g = torch.Generator().manual_seed(5)
    theta = torch.randn(2, 64, 11, generator=g).abs() + 0.5        # positive: summary() takes logs of nothing, but be safe
    logq = torch.randn(128, generator=g)
    ours = save_posterior(tmp_path / "ours", theta, logq)
    res = PosteriorResult(samples=theta.reshape(-1, 11).numpy().astype(np.float64),
                          log_prob=logq.numpy().astype(np.float64), param_names=list(PARAM_NAMES), trigger_gps=None)
    ref = res.save(str(tmp_path / "ref"), plots=False)
    for name in ("posterior_samples.npy", "posterior_log_prob.npy", "posterior_samples.csv"):
        assert (ours / name).read_bytes() == (ref / name).read_bytes(), name


def test_posterior_arrays_shapes_and_errors():
    s, lp, rm = posterior_arrays(np.zeros((2, 5, 11), np.float32), np.zeros(10, np.float32),
                                 np.array([0, 1] * 5, np.uint8))
    assert s.shape == (10, 11) and lp.shape == (10,) and rm.dtype == bool and rm.sum() == 5
    s, lp, rm = posterior_arrays(np.zeros((0, 11)), np.zeros(0))          # empty posterior
    assert s.shape == (0, 11) and lp.shape == (0,) and rm is None
    try:
        posterior_arrays(np.zeros((4, 11)), np.zeros(3))
    except ValueError:
        pass
    else:
        raise AssertionError("length mismatch must raise")
