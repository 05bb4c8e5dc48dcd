"""N3: on-device posterior summaries (pf_posterior_summary) against the numpy restatement of the reference's
host-side reports (oracle/summary_oracle.py <- overlap_benchmark.py:67-89, validate_checkpoint.py:102-128).
Order statistics are exact, so medians / lower medians / SBC ranks / railing fractions must match bit for bit;
interpolated quantiles to 1e-6 relative (numpy interpolates in float32, the kernel in double)."""
import numpy as np
import pytest
import torch

import summary_oracle as SO

pytestmark = pytest.mark.gpu
QTOL = 1e-6


def _check(samples, truth=None, railed=None, q=(0.05, 0.25, 0.5, 0.75, 0.95), f32_check=True):
    import posteriflow_b200 as P
    th = torch.from_numpy(samples).cuda()
    tr = torch.from_numpy(truth).cuda() if truth is not None else None
    rl = torch.from_numpy(railed).cuda() if railed is not None else None
    got = P.posterior_summary(th, truth=tr, railed=rl, q=q)
    ref = SO.summaries(samples, truth, railed, q=q)
    gq, rq = got["quantiles"].cpu().numpy().astype(np.float64), ref["quantiles"].astype(np.float64)
    assert gq.shape == rq.shape
    # arbiter: the same numpy formula evaluated in float64 -- the kernel must be its correctly rounded float32 value
    r64 = np.stack([np.quantile(samples.astype(np.float64), lev, axis=1) for lev in q], axis=1)
    assert np.all(np.abs(gq - r64) <= 1.2e-7 * np.maximum(np.abs(r64), 1e-30)), np.abs(gq - r64).max()
    if f32_check:   # numpy's float32 evaluation (what the reference scripts get): its own lerp rounding, <= 1e-6
        assert np.all(np.abs(gq - rq) <= QTOL * np.maximum(1.0, np.abs(rq))), np.abs(gq - rq).max()
    gm = got["median"].cpu().numpy().astype(np.float64)
    m64 = np.median(samples.astype(np.float64), axis=1)
    assert np.all(np.abs(gm - m64) <= 1.2e-7 * np.maximum(np.abs(m64), 1e-30))
    assert np.array_equal(got["lower_median"].cpu().numpy(), ref["lower_median"])
    if truth is not None:
        assert np.array_equal(got["sbc_rank"].cpu().numpy(), ref["sbc_rank"].astype(np.float32))
    if railed is not None:
        assert np.array_equal(got["rail_frac"].cpu().numpy(), ref["rail_frac"].astype(np.float32))
    return got


def test_summary_matches_numpy_on_posterior_like_samples():
    rng = np.random.default_rng(0)
    E, n, D = 7, 5000, 11
    scale = np.array([30, 20, 800, 2, .5, 1, 1, 2, .01, .3, .3], np.float32)
    shift = np.array([40, 25, 900, 3, 0, 1.5, 1.5, 3, 0, .5, .5], np.float32)
    samples = (rng.standard_normal((E, n, D)).astype(np.float32) * scale + shift).astype(np.float32)
    truth = (rng.standard_normal((E, D)).astype(np.float32) * scale + shift).astype(np.float32)
    railed = (rng.random(E * n) < 0.03).astype(np.uint8)
    _check(samples, truth, railed)


def test_summary_edge_cases():
    rng = np.random.default_rng(1)
    # n = 1, 2, 3 and sizes around the warp / block granularity; ties; negative values; zeros of both signs
    for n in (1, 2, 3, 31, 32, 33, 511, 512, 513, 1000):
        s = rng.standard_normal((3, n, 11)).astype(np.float32)
        _check(s, rng.standard_normal((3, 11)).astype(np.float32))
    ties = rng.integers(-3, 4, size=(4, 777, 11)).astype(np.float32)
    ties[0, :, 0] = 2.5                      # a constant column
    ties[1, ::2, 1] = -0.0                   # signed zeros compare equal; both are exact order statistics of value 0
    _check(ties, np.zeros((4, 11), np.float32), q=(0.0, 0.1, 0.5, 0.9, 1.0))
    wide = (rng.standard_normal((2, 4096, 11)) * np.exp(rng.uniform(-20, 20, size=(2, 4096, 11)))).astype(np.float32)
    # 40 e-folds of dynamic range: every radix pass discriminates.  numpy's float32 lerp between order statistics
    # that far apart is itself off by several ulp, so only the float64 arbiter applies
    _check(wide, q=(0.001, 0.5, 0.999), f32_check=False)
    _check(rng.standard_normal((1, 64, 3)).astype(np.float32))      # D != 11


def test_summary_large_single_event_and_uncached_path():
    rng = np.random.default_rng(2)
    # BASELINE config #2: 1 event x 50 000 samples (keys cached in shared memory) and a size beyond the cache
    for n in (50000, 70001):
        s = (rng.standard_normal((1, n, 11)) * 3 + 1).astype(np.float32)
        _check(s, rng.standard_normal((1, 11)).astype(np.float32), (rng.random(n) < 0.5).astype(np.uint8))


def test_calib_block_matches_reference_report():
    import posteriflow_b200 as P
    rng = np.random.default_rng(3)
    E, K = 256, 512
    sc = P.ParamScaler()
    truth = sc.denormalize(torch.from_numpy(rng.uniform(-0.9, 0.9, size=(E, 11)).astype(np.float32))).numpy()
    y = np.clip(rng.standard_normal((E, K, 11)).astype(np.float32) * 0.2 +
                sc.normalize(torch.from_numpy(truth)).numpy()[:, None, :] +
                rng.standard_normal((E, 1, 11)).astype(np.float32) * 0.2, -1, 1)
    samples = sc.denormalize(torch.from_numpy(y)).numpy()
    got = P.calib_block(torch.from_numpy(samples).cuda(), torch.from_numpy(truth).cuda())
    ref = SO.calib_block(samples, truth)
    assert got["n"] == ref["n"]
    for tag in ("cov50", "cov90", "sbc_ks_p"):
        for p in SO.PARAM_NAMES:
            assert abs(got[tag][p] - ref[tag][p]) <= 1e-3 + 1e-9, (tag, p, got[tag][p], ref[tag][p])
    for k in ("cov50_headline_mean", "cov90_headline_mean", "mc_frac_err_median", "dl_frac_err_median",
              "tc_abs_err_median_s"):
        assert abs(got[k] - ref[k]) <= 1e-3 + 1e-9, (k, got[k], ref[k])


def test_summary_rejects_bad_arguments():
    import posteriflow_b200 as P
    th = torch.zeros(2, 8, 11, device="cuda")
    with pytest.raises(ValueError):
        P.posterior_summary(th, q=[0.1 * i for i in range(1, 10)])
    with pytest.raises(Exception):
        P.posterior_summary(th, q=(1.5,))
    with pytest.raises(ValueError):
        P.posterior_summary(th, truth=torch.zeros(3, 11, device="cuda"))
    with pytest.raises(RuntimeError):
        P.posterior_summary(torch.zeros(2, 8, 11))      # CPU tensor: no fallback
