"""posteriflow_b200 -- B200-native (sm_100a) posterior-flow hot path of PosteriFlow's LeanNPE.

Importing the package registers the `posteriflow::*` torch.library ops.  The CUDA library is loaded
on first use and there is no CPU fallback (see _cabi.load_library).
"""
from . import _cabi  # noqa: F401
from . import ops  # noqa: F401
from .flow import FLOW_NORM_BOUND, NSFPosteriorFlow  # noqa: F401
from .lean import (LeanNPE, ParamScaler, PARAM_NAMES, accelerate, posterior_chunk, sample_fused,  # noqa: F401
                   sample_ranks, symmetrized_log_q)
from .result import load_posterior, posterior_arrays, save_posterior  # noqa: F401
from .summary import calib_block, posterior_summary  # noqa: F401
from .train import GraphedNLLStep, batch_nll, signal_rows  # noqa: F401

__all__ = ["NSFPosteriorFlow", "LeanNPE", "ParamScaler", "PARAM_NAMES", "FLOW_NORM_BOUND", "accelerate",
           "sample_fused", "sample_ranks", "symmetrized_log_q", "posterior_chunk", "ops", "batch_nll", "signal_rows", "GraphedNLLStep", "save_posterior", "load_posterior",
           "posterior_arrays", "posterior_summary", "calib_block"]
