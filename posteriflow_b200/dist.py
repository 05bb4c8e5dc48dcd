"""Multi-GPU plumbing of the hot path: one process per GPU over torch.distributed.

Sampling / log-prob shard by rows with NO data-path collective: every (event, rank, sample) row is
independent given the (replicated, 25 MB) weights, so events are partitioned contiguously across ranks --
or, when there are fewer events than ranks (BASELINE config #2: 1 event x 50k samples), each event's samples
are split into contiguous blocks.  Training is data-parallel: the only exchange is one gradient all-reduce
per step (NCCL over NVLink on the GPUs; the same code runs over gloo on CPU in the tests), with the loss
normaliser made globally consistent (train_lean_npe.py:127 divides by the number of flow rows).
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced [begin, end) of `n` items for `rank` (first n % world ranks get one extra)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} / world {world}")
    q, r = divmod(n, world)
    b = rank * q + min(rank, r)
    return b, b + q + (1 if rank < r else 0)


def shard_sampling(n_events: int, n_samples: int, rank: int, world: int) -> Tuple[Tuple[int, int], Tuple[int, int]]:
    """((event_begin, event_end), (sample_begin, sample_end)) this rank draws.

    Events >= ranks: whole events per rank.  Fewer events than ranks: every rank keeps all events and takes a
    contiguous block of each event's samples (the context row is simply replicated: 1.2 KB)."""
    if n_events >= world:
        return shard_range(n_events, rank, world), (0, n_samples)
    return (0, n_events), shard_range(n_samples, rank, world)


def _world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_seed(seed: int, rank: int) -> int:
    """Philox key of a rank's shard: the in-kernel counter is the LOCAL row index, so the key carries the shard --
    distinct keys give statistically independent, non-overlapping streams (no row can repeat across ranks)."""
    return (int(seed) * 1000003 + int(rank)) & (2 ** 63 - 1)


def sample_posterior_sharded(model, context: torch.Tensor, rank_index: int, n_samples: int, seed: int = 0,
                             gather: bool = False, rank_world: Optional[Tuple[int, int]] = None) -> torch.Tensor:
    """This rank's share of `LeanNPE.sample_posterior` for `context` [E, context_dim] (same on every rank).

    Returns theta [E_local, n_local, D]; `gather=True` all-gathers the shards (equal shard sizes required)
    into the full [E, n, D] on every rank -- an optional convenience, not part of the hot path."""
    r, w = rank_world if rank_world is not None else _world()   # (rank, world) override: single-process tests
    (e0, e1), (s0, s1) = shard_sampling(context.shape[0], n_samples, r, w)
    ctx = context[e0:e1]
    from .lean import sample_fused
    rk = torch.full((ctx.shape[0],), rank_index, dtype=torch.long, device=ctx.device)
    full = model._full_context(ctx, rk)
    theta = sample_fused(model.flow, model.scaler, full, s1 - s0, z=None, seed=shard_seed(seed, r))[0]
    if not gather or w == 1 or rank_world is not None:
        return theta
    parts = [torch.empty_like(theta) for _ in range(w)]
    dist.all_gather(parts, theta.contiguous())
    return torch.cat(parts, dim=0 if context.shape[0] >= w else 1)


def allreduce_gradients(params: Iterable[torch.nn.Parameter], bucket_bytes: int = 64 << 20) -> int:
    """Sum `.grad` of `params` over all ranks, in place, in flat buckets (one collective per bucket; 36 MB of
    fp32 gradients = one bucket at the default).  Returns the number of collectives issued."""
    r, w = _world()
    if w == 1:
        return 0
    # A rank whose batch produced no gradient for a parameter (no signal in its shard) contributes zeros: the list of
    # collectives and their sizes must be identical on every rank.
    grads: List[torch.Tensor] = []
    for p in params:
        if not getattr(p, "requires_grad", True):
            continue
        if p.grad is None:
            p.grad = torch.zeros_like(p)
        grads.append(p.grad)
    n_coll, i = 0, 0
    while i < len(grads):
        j, size = i, 0
        while j < len(grads) and (j == i or size + grads[j].numel() * grads[j].element_size() <= bucket_bytes):
            size += grads[j].numel() * grads[j].element_size()
            j += 1
        flat = torch.cat([g.reshape(-1) for g in grads[i:j]])
        dist.all_reduce(flat)
        o = 0
        for g in grads[i:j]:
            g.copy_(flat[o:o + g.numel()].view_as(g))
            o += g.numel()
        n_coll += 1
        i = j
    return n_coll


def global_mean_loss(nll_sum: torch.Tensor, count: torch.Tensor) -> torch.Tensor:
    """Loss whose gradient, after `allreduce_gradients` (a SUM), equals the gradient of the global mean NLL:
    local sum / global count.  (`batch_nll` divides by the local count: train_lean_npe.py:127.)"""
    r, w = _world()
    total = count.detach().clone().to(nll_sum.dtype)
    if w > 1:
        dist.all_reduce(total)
    return nll_sum / total.clamp_min(1.0)
