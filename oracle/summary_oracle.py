"""TEST INFRASTRUCTURE ONLY (may be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg).

CPU restatement, in numpy, of the reference's host-side posterior summaries:
  /root/reference/scripts/overlap_benchmark.py:67-89   calib_block (median, quantile coverage, SBC ranks, KS p)
  /root/reference/scripts/validate_checkpoint.py:102-128  coverage / ranks / railing_frac
  /root/reference/experiments/train_lean_npe.py:160-167   torch .median(dim=1).values / .quantile(q, dim=1)
Parity unpinned: the reference holds no golden vector for these reports; the arithmetic is numpy's own.
"""
import numpy as np

PARAM_NAMES = ["mass_1", "mass_2", "luminosity_distance", "ra", "dec", "theta_jn", "psi", "phase",
               "geocent_time", "a1", "a2"]
HEADLINE = ["mass_1", "mass_2", "luminosity_distance", "ra", "dec", "theta_jn", "geocent_time"]
JM1, JM2, JD, JT = 0, 1, 2, 8


def summaries(samples, truth=None, railed=None, q=(0.05, 0.25, 0.5, 0.75, 0.95)):
    """samples [E, n, D] float32 -> dict of numpy arrays with the reference's formulas."""
    out = {"quantiles": np.stack([np.quantile(samples, lev, axis=1) for lev in q], axis=1),   # overlap_benchmark.py:72-73
           "median": np.median(samples, axis=1),                                              # :69
           "lower_median": np.sort(samples, axis=1)[:, (samples.shape[1] - 1) // 2, :]}       # torch.median(dim).values
    if truth is not None:
        out["sbc_rank"] = (samples < truth[:, None, :]).mean(axis=1)                          # :86
    if railed is not None:
        out["rail_frac"] = railed.reshape(samples.shape[0], -1).astype(np.float64).mean(axis=1)
    return out


def mchirp(m1, m2):
    return (m1 * m2) ** 0.6 / (m1 + m2) ** 0.2


def calib_block(samples, truth):
    """overlap_benchmark.py:67-89, line for line (scipy's kstest for the SBC p-values)."""
    import scipy.stats as sps
    med = np.median(samples, axis=1)
    blk = {"n": int(len(truth))}
    for lev, tag in ((0.50, "cov50"), (0.90, "cov90")):
        lo = np.quantile(samples, 0.5 - lev / 2, axis=1)
        hi = np.quantile(samples, 0.5 + lev / 2, axis=1)
        inside = (truth >= lo) & (truth <= hi)
        blk[tag] = {p: round(float(inside[:, j].mean()), 3) for j, p in enumerate(PARAM_NAMES)}
        blk[f"{tag}_headline_mean"] = round(float(np.mean(
            [inside[:, PARAM_NAMES.index(p)].mean() for p in HEADLINE])), 3)
    mc_t = mchirp(truth[:, JM1], truth[:, JM2])
    mc_p = mchirp(med[:, JM1], med[:, JM2])
    blk["mc_frac_err_median"] = round(float(np.median(np.abs(mc_p - mc_t) / mc_t)), 4)
    blk["dl_frac_err_median"] = round(float(np.median(np.abs(med[:, JD] - truth[:, JD]) / truth[:, JD])), 4)
    blk["tc_abs_err_median_s"] = round(float(np.median(np.abs(med[:, JT] - truth[:, JT]))), 4)
    ranks = (samples < truth[:, None, :]).mean(axis=1)
    blk["sbc_ks_p"] = {p: round(float(sps.kstest(ranks[:, j], "uniform").pvalue), 4)
                       for j, p in enumerate(PARAM_NAMES)}
    return blk
