"""N4 of SURVEY.md 8(f): the data format on the output side of the hot path.

`PosteriorResult.save` (/root/reference/src/ahsd/inference/result.py:242-252) persists a posterior as
`posterior_samples.npy` (float64 [N, 11]), `posterior_log_prob.npy` (float64 [N]) and `posterior_samples.csv`
(`np.savetxt`, header = the parameter names in PARAM_NAMES order + "log_prob", no comment prefix), which is what the
reference's downstream tools (bilby bridge, plots, pandas / R readers) load.  `infer()` builds those arrays as float64
host copies of the device tensors (pipeline.py:163-186).  This module writes the same three files, byte for byte,
from the device tensors `posterior_chunk` / `LeanNPE.sample_with_log_prob` return.
"""
from __future__ import annotations

from pathlib import Path
from typing import Optional, Sequence, Union

import numpy as np
import torch

from .lean import PARAM_NAMES

ArrayLike = Union[np.ndarray, torch.Tensor]


def _host_f64(a: ArrayLike) -> np.ndarray:
    if isinstance(a, torch.Tensor):
        a = a.detach().to("cpu").numpy()
    return np.ascontiguousarray(a, dtype=np.float64)


def posterior_arrays(theta: ArrayLike, logq: ArrayLike, railed: Optional[ArrayLike] = None):
    """(samples float64 [N, D], log_prob float64 [N], rail_mask bool [N] | None) in infer()'s conventions
    (pipeline.py:163-165, 181-183): event-major flattening of theta [E, n, D] / [N, D]."""
    s = _host_f64(theta)
    s = s.reshape(-1, s.shape[-1])
    lp = _host_f64(logq).reshape(-1)
    if lp.shape[0] != s.shape[0]:
        raise ValueError(f"{s.shape[0]} samples but {lp.shape[0]} log-probabilities")
    rm = None
    if railed is not None:
        if isinstance(railed, torch.Tensor):
            railed = railed.detach().to("cpu").numpy()
        rm = np.asarray(railed).reshape(-1).astype(bool)
    return s, lp, rm


def save_posterior(outdir: Union[str, Path], theta: ArrayLike, logq: ArrayLike,
                   param_names: Sequence[str] = PARAM_NAMES) -> Path:
    """Write posterior_samples.npy / posterior_log_prob.npy / posterior_samples.csv exactly as
    PosteriorResult.save does (result.py:242-252).  Returns the output directory."""
    samples, log_prob, _ = posterior_arrays(theta, logq)
    if samples.shape[1] != len(param_names):
        raise ValueError(f"samples have {samples.shape[1]} columns, {len(param_names)} parameter names given")
    out = Path(outdir)
    out.mkdir(parents=True, exist_ok=True)
    np.save(out / "posterior_samples.npy", samples)
    np.save(out / "posterior_log_prob.npy", log_prob)
    header = ",".join(list(param_names) + ["log_prob"])
    np.savetxt(out / "posterior_samples.csv", np.column_stack([samples, log_prob]), delimiter=",", header=header,
               comments="")
    return out


def load_posterior(outdir: Union[str, Path]):
    """(samples [N, D] float64, log_prob [N] float64, param_names) from a directory `save_posterior` (or the
    reference's PosteriorResult.save) wrote."""
    out = Path(outdir)
    samples = np.load(out / "posterior_samples.npy")
    log_prob = np.load(out / "posterior_log_prob.npy")
    with open(out / "posterior_samples.csv") as fh:
        names = fh.readline().strip().split(",")[:-1]
    return samples, log_prob, names
