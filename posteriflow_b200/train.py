"""N1 of SURVEY.md 8(f): the training-side caller of the hot path, `batch_nll`
(/root/reference/experiments/train_lean_npe.py:108-127), as ONE flow call for all signal ranks.

The reference loops over ranks r = 0..max(nsig)-1 and calls `model.nll` once per rank on the rows
`nonzero(nsig > r)`: up to five flow forward passes (and five autograd graphs) per step plus a host sync
per rank (`mask.any()`, `nsig.max().item()`).  Here the (event, rank) rows of all ranks are built at once,
in the reference's own rank-major order, and go through the flow as a single [R, 11] batch with the rank as
a tensor -- the rank embedding gather already takes one.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import Tensor

from .dist import global_mean_loss


def signal_rows(nsig: Tensor, max_signals: int) -> Tuple[Tensor, Tensor]:
    """(event index [R], rank [R]) of every present signal, rank-major then event-ascending: exactly the
    concatenation over r of `nonzero(nsig > r)` that train_lean_npe.py:117-124 iterates over."""
    r = torch.arange(max_signals, device=nsig.device).unsqueeze(1)
    rank, event = (nsig.unsqueeze(0) > r).nonzero(as_tuple=True)
    return event, rank


def batch_nll(model, strain: Optional[Tensor], params: Tensor, nsig: Tensor, asd_bands: Optional[Tensor] = None,
              context: Optional[Tensor] = None, distributed: bool = False) -> Tensor:
    """Mean per-signal NLL of a batch of events (train_lean_npe.py:108-127), one flow call.

    strain [B, 3, T] (or None with `context` [B, context_dim] given), params [B, max_signals, 11] physical,
    nsig [B] int.  `distributed=True` divides by the GLOBAL row count, so that the SUM all-reduce of
    `dist.allreduce_gradients` yields the gradient of the global mean (the reference divides by the local
    count and lets DDP average).  An empty batch returns a zero that still carries the graph of `context`."""
    if context is None:
        context = model.encode(strain, asd_bands)
    event, rank = signal_rows(nsig.to(context.device), int(model.max_signals))
    count = event.numel()
    if count == 0:
        return context.sum() * 0.0
    nll = model.nll(None, params[event, rank, :], rank, context=context[event])
    total = nll.sum()
    if distributed:
        return global_mean_loss(total, torch.tensor(float(count), device=total.device))
    return total / count


def batch_nll_reference_order(model, strain, params, nsig, asd_bands=None, context=None) -> Tensor:
    """The reference's own loop (train_lean_npe.py:114-127), kept for the parity test of `batch_nll`."""
    if context is None:
        context = model.encode(strain, asd_bands)
    count, losses = 0, []
    for r in range(int(nsig.max().item())):
        idx = (nsig > r).nonzero(as_tuple=True)[0]
        if idx.numel() == 0:
            continue
        rank = torch.full((idx.numel(),), r, dtype=torch.long, device=context.device)
        losses.append(model.nll(None, params[idx, r, :], rank, context=context[idx]).sum())
        count += idx.numel()
    return torch.stack(losses).sum() / count
