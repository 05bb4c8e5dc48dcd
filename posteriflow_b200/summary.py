"""N3 of SURVEY.md 8(f): posterior summaries computed where the samples are.

The reference copies every posterior sample to the host and summarises there with numpy
(/root/reference/scripts/overlap_benchmark.py:67-89 `calib_block`, scripts/validate_checkpoint.py:102-128,
experiments/train_lean_npe.py:160-167).  `posterior_summary` produces the same per-event statistics from the device
tensor `sample_fused` returns, in one kernel (`pf_posterior_summary`); `calib_block` then reduces the [E, D]-sized
results to the reference's report (coverage, error medians, SBC KS p-values) -- a few kilobytes cross to the host.
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import torch
from torch import Tensor

from .lean import PARAM_NAMES

DEFAULT_Q = (0.05, 0.25, 0.5, 0.75, 0.95)
HEADLINE = ["mass_1", "mass_2", "luminosity_distance", "ra", "dec", "theta_jn", "geocent_time"]  # overlap_benchmark.py:37


def posterior_summary(theta: Tensor, truth: Optional[Tensor] = None, railed: Optional[Tensor] = None,
                      q: Sequence[float] = DEFAULT_Q) -> Dict[str, Tensor]:
    """theta [E, n, D] (device) -> {"q": levels, "quantiles" [E, Q, D] (np.quantile 'linear'), "median" [E, D]
    (np.median), "lower_median" [E, D] (torch.median), "sbc_rank" [E, D] ((samples < truth).mean, if truth given),
    "rail_frac" [E] (if railed given)}."""
    q = [float(v) for v in q]
    qq = q if 0.5 in q else q + [0.5]
    if len(qq) > 8:
        raise ValueError("at most 8 quantile levels per call")
    quant, lower, below, rail = torch.ops.posteriflow.posterior_summary(theta, qq, truth, railed)
    out = {"q": q, "quantiles": quant[:, :len(q), :], "median": quant[:, qq.index(0.5), :], "lower_median": lower}
    if truth is not None:
        out["sbc_rank"] = below
    if railed is not None:
        out["rail_frac"] = rail
    return out


def _mchirp(m1, m2):
    return (m1 * m2) ** 0.6 / (m1 + m2) ** 0.2


def calib_block(theta: Tensor, truth: Tensor, param_names: Sequence[str] = PARAM_NAMES) -> Dict:
    """The reference's calibration report (overlap_benchmark.py:67-89) from device samples [n_ev, K, D] and
    truth [n_ev, D]; same keys and rounding.  `sbc_ks_p` needs scipy on the host (as the reference does)."""
    import numpy as np
    s = posterior_summary(theta, truth=truth, q=(0.05, 0.25, 0.5, 0.75, 0.95))
    qt, med = s["quantiles"], s["median"]
    names = list(param_names)
    blk: Dict = {"n": int(truth.shape[0])}
    for tag, lo, hi in (("cov50", qt[:, 1], qt[:, 3]), ("cov90", qt[:, 0], qt[:, 4])):
        inside = ((truth >= lo) & (truth <= hi)).double().mean(dim=0).cpu().numpy()
        blk[tag] = {p: round(float(inside[j]), 3) for j, p in enumerate(names)}
        blk[f"{tag}_headline_mean"] = round(float(np.mean([inside[names.index(p)] for p in HEADLINE if p in names])), 3)
    med64, tr64 = med.double(), truth.double()
    j1, j2, jd, jt = (names.index(k) for k in ("mass_1", "mass_2", "luminosity_distance", "geocent_time"))
    mc_t, mc_p = _mchirp(tr64[:, j1], tr64[:, j2]), _mchirp(med64[:, j1], med64[:, j2])
    blk["mc_frac_err_median"] = round(float(torch.quantile((mc_p - mc_t).abs() / mc_t, 0.5)), 4)
    blk["dl_frac_err_median"] = round(float(torch.quantile((med64[:, jd] - tr64[:, jd]).abs() / tr64[:, jd], 0.5)), 4)
    blk["tc_abs_err_median_s"] = round(float(torch.quantile((med64[:, jt] - tr64[:, jt]).abs(), 0.5)), 4)
    ranks = s["sbc_rank"].cpu().numpy()
    import scipy.stats as sps
    blk["sbc_ks_p"] = {p: round(float(sps.kstest(ranks[:, j], "uniform").pvalue), 4) for j, p in enumerate(names)}
    return blk
