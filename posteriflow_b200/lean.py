"""LeanNPE glue over the B200 flow: ParamScaler, rank embedding, fused sample_posterior / nll.

Mirrors /root/reference/src/ahsd/models/lean_npe.py:48-114 (ParamScaler) and :255-338 (LeanNPE) for
the part of the model on the hot path.  The strain encoder (lean_npe.py:131-252,
coherent_encoder.py) is NOT part of this package: it stays the reference's PyTorch module and is
passed in (`LeanNPE(encoder=...)`), or an existing reference model is converted in place with
`accelerate(model)`.
"""
from __future__ import annotations

import math
import types
from typing import List, Optional

import torch
from torch import Tensor, nn

from . import _cabi as A
from .flow import NSFPosteriorFlow

PARAM_NAMES = ["mass_1", "mass_2", "luminosity_distance", "ra", "dec", "theta_jn", "psi", "phase",
               "geocent_time", "a1", "a2"]  # lean_npe.py:41-46


class ParamScaler:
    """Fixed invertible map physical <-> [-1, 1] (lean_npe.py:48-114); same attributes and methods."""

    RANGES = {
        "mass_1": (1.0, 105.0, True), "mass_2": (1.0, 105.0, True),
        "luminosity_distance": (40.0, 2200.0, True),
        "ra": (0.0, 2 * math.pi, False), "dec": (-math.pi / 2, math.pi / 2, False),
        "theta_jn": (0.0, math.pi, False), "psi": (0.0, math.pi, False),
        "phase": (0.0, 2 * math.pi, False), "geocent_time": (-1.6, 1.6, False),
        "a1": (0.0, 1.0, False), "a2": (0.0, 1.0, False),
    }
    CIRCULAR = ("ra", "phase", "psi")

    def __init__(self, param_names: List[str] = PARAM_NAMES, premerger: bool = False):
        self.param_names, self.premerger = param_names, premerger
        rows = []
        for p in param_names:
            lo, hi, is_log = self.RANGES[p]
            if premerger and p == "geocent_time":
                lo, hi = -1.6, 5.2
            rows.append((math.log(lo), math.log(hi), True) if is_log else (lo, hi, False))
        self.lo = torch.tensor([r[0] for r in rows], dtype=torch.float32)
        self.hi = torch.tensor([r[1] for r in rows], dtype=torch.float32)
        self.log_mask = torch.tensor([r[2] for r in rows], dtype=torch.bool)
        self.circ_mask = torch.tensor([p in self.CIRCULAR for p in param_names], dtype=torch.bool)

    def to(self, device):
        for k in ("lo", "hi", "log_mask", "circ_mask"):
            setattr(self, k, getattr(self, k).to(device))
        return self

    def wrap(self, y: Tensor) -> Tensor:
        return torch.where(self.circ_mask, torch.remainder(y + 1.0, 2.0) - 1.0, y.clamp(-1.0, 1.0))

    def normalize(self, x: Tensor) -> Tensor:
        x = torch.where(self.log_mask, torch.log(x.clamp_min(1e-6)), x)
        return (2 * (x - self.lo) / (self.hi - self.lo) - 1).clamp(-1.0, 1.0)

    def denormalize(self, y: Tensor) -> Tensor:
        x = (y.clamp(-1.0, 1.0) + 1) / 2 * (self.hi - self.lo) + self.lo
        return torch.where(self.log_mask, torch.exp(x), x)


SAMPLE_FLAGS = A.PF_FLAG_SANITIZE_CTX | A.PF_FLAG_CLAMP | A.PF_FLAG_WRAP | A.PF_FLAG_DENORMALIZE


def sample_fused(flow: NSFPosteriorFlow, scaler: ParamScaler, ctx_full: Tensor, n_samples: int,
                 z: Optional[Tensor] = None, seed: int = 0, flags: int = SAMPLE_FLAGS):
    """lean_npe.py:326-332 in one C-ABI call: (theta [E, n, D] physical, logdet [E*n], railed [E*n]).
    The context is NOT repeated: row e of `ctx_full` serves samples [e*n, (e+1)*n)."""
    fh = flow.handle()
    if not flow._ar_identity and (flags & (A.PF_FLAG_WRAP | A.PF_FLAG_DENORMALIZE)):
        raise NotImplementedError("the fused sampler applies ParamScaler per autoregressive POSITION in-kernel; with a "
                                  "non-identity set_autoregressive_order call flow.inverse and the scaler separately")
    with torch.no_grad():
        return torch.ops.posteriflow.nsf_sample_fused(ctx_full, z, int(seed), int(n_samples), fh.id, int(flags),
                                                      scaler.lo, scaler.hi, scaler.log_mask, scaler.circ_mask)


def posterior_chunk(flow: NSFPosteriorFlow, scaler: ParamScaler, ctx_full: Tensor, n_samples: int,
                    z: Optional[Tensor] = None, seed: int = 0, sort_masses: bool = True):
    """One iteration of the reference's infer() sampling loop (inference/pipeline.py:168-186) for E events
    at once, everything on the device: base noise -> inverse flow -> wrap -> rail mask -> flow NLL of the wrapped
    samples -> physical log-density (ParamScaler Jacobian) -> denormalise -> m1 >= m2.
    Returns (theta [E, n, D] physical, logq [E*n] physical, railed [E*n] uint8).  Four kernels after the two
    context projections: sampler, density direction, base log-prob, physical epilogue."""
    wrap_flags = A.PF_FLAG_SANITIZE_CTX | A.PF_FLAG_CLAMP | A.PF_FLAG_WRAP
    y, _, railed = sample_fused(flow, scaler, ctx_full, n_samples, z=z, seed=seed, flags=wrap_flags)
    E, n, D = y.shape
    with torch.no_grad():
        nll = flow.compute_psd_aware_nll(y.reshape(E * n, D), ctx_full, None, n_per_ctx=n)
        theta, logq = torch.ops.posteriflow.nsf_physical(y.reshape(E * n, D), nll, bool(sort_masses), scaler.lo,
                                                         scaler.hi, scaler.log_mask, scaler.circ_mask)
    return theta.reshape(E, n, D), logq, railed


def sample_ranks(model, context: Tensor, max_rank: int, n_samples: int, z: Optional[Tensor] = None, seed: int = 0,
                 sort_masses: bool = True) -> Tensor:
    """scripts/overlap_benchmark.py:40-64 `sample_ranks` after the encoder: posterior draws for ranks
    0..max_rank-1 of each event, [N, max_rank, K, 11] physical with m1 >= m2 -- as ONE launch of the fused sampler
    instead of one flow call per rank: the N * max_rank (event, rank) contexts are laid out event-major / rank-minor
    and each serves its K consecutive flow rows (BASELINE config #5).  `z` ([N*max_rank*K, 11], same order) fixes
    the base noise; otherwise it is drawn in-kernel."""
    N = context.shape[0]
    ev = torch.arange(N, device=context.device).repeat_interleave(max_rank)
    rk = torch.arange(max_rank, device=context.device).repeat(N)
    ctx = model._full_context(context.index_select(0, ev), rk)            # [N * max_rank, C]
    flags = SAMPLE_FLAGS | (A.PF_FLAG_SORT_MASSES if sort_masses else 0)
    theta, _, _ = sample_fused(model.flow, model.scaler, ctx, n_samples, z=z, seed=seed, flags=flags)
    return theta.reshape(N, max_rank, n_samples, theta.shape[-1])


def symmetrized_log_q(model, theta: Tensor, ctx_full: Tensor, j1: int = 0, j2: int = 1) -> Tensor:
    """inference/importance.py:69-95 on the device: log[q(m1, m2, ..) + q(m2, m1, ..)] in physical units for
    theta [N, 11] of ONE event (`ctx_full` [1, C] = `_full_context(encoder(strain), rank)`).  The reference runs two
    flow passes per batch of 2048; here both mass orderings go through the flow as ONE call of 2N rows that share the
    event's context projection, followed by the physical-density epilogue and a logaddexp."""
    sc = model.scaler
    N, D = theta.shape
    sw = theta.clone()
    sw[:, [j1, j2]] = theta[:, [j2, j1]]
    y = sc.normalize(torch.cat([theta, sw], dim=0))
    with torch.no_grad():
        nll = model.flow.compute_psd_aware_nll(y, ctx_full[:1], None, n_per_ctx=2 * N)
        _, logq = torch.ops.posteriflow.nsf_physical(y, nll, False, sc.lo, sc.hi, sc.log_mask, sc.circ_mask)
    return torch.logaddexp(logq[:N], logq[N:])


class LeanNPE(nn.Module):
    """encoder (caller's PyTorch module) + rank embedding + B200 flow (lean_npe.py:255-338).

    `encoder(strain[, asd_bands]) -> [B, context_dim]`.  `encoder=None` builds a flow head that must be
    given `context=` explicitly (cached-context training, frozen_context_heads.py)."""

    def __init__(self, param_names: List[str] = PARAM_NAMES, context_dim: int = 256, rank_dim: int = 32,
                 max_signals: int = 5, flow_layers: int = 10, flow_hidden: int = 256, flow_bins: int = 16,
                 premerger: bool = False, psd_cond: bool = False, encoder: Optional[nn.Module] = None,
                 tier: str = "fp32"):
        super().__init__()
        self.param_names, self.max_signals, self.context_dim = param_names, max_signals, context_dim
        self.psd_cond = psd_cond
        self.scaler = ParamScaler(param_names, premerger=premerger)
        self.encoder = encoder
        self.rank_embed = nn.Embedding(max_signals, rank_dim)
        self.flow = NSFPosteriorFlow(features=len(param_names), context_features=context_dim + rank_dim,
                                     hidden_features=flow_hidden, num_layers=flow_layers, num_bins=flow_bins,
                                     tail_bound=5.0, dropout=0.0, temperature_scale=1.0, use_masked_context=False,
                                     tier=tier)
        self.flow.temperature.requires_grad_(False)
        self.sample_seed = 0

    def _full_context(self, context: Tensor, rank: Tensor) -> Tensor:
        return torch.cat([context, self.rank_embed(rank)], dim=1)

    def encode(self, strain: Tensor, asd_bands: Optional[Tensor] = None) -> Tensor:
        if self.encoder is None:
            raise RuntimeError("this LeanNPE was built without an encoder: pass context= explicitly")
        return self.encoder(strain, asd_bands) if self.psd_cond else self.encoder(strain)

    def nll(self, strain: Optional[Tensor], params_phys: Tensor, rank: Tensor, context: Optional[Tensor] = None,
            asd_bands: Optional[Tensor] = None) -> Tensor:
        if context is None:
            context = self.encode(strain, asd_bands)
        ctx = self._full_context(context, rank)
        y = self.scaler.normalize(params_phys)
        return self.flow.compute_psd_aware_nll(y, ctx, None)

    @torch.no_grad()
    def sample_posterior(self, strain: Optional[Tensor], rank: int = 0, n_samples: int = 256,
                         asd_bands: Optional[Tensor] = None, context: Optional[Tensor] = None,
                         z: Optional[Tensor] = None) -> Tensor:
        """[B, n_samples, D] physical samples, event-major.  `z` ([B*n, D]) fixes the base noise (parity);
        otherwise it is drawn in-kernel (Philox keyed by an internal counter seeded from torch's RNG)."""
        if context is None:
            context = self.encode(strain, asd_bands)
        B = context.shape[0]
        r = torch.full((B,), rank, dtype=torch.long, device=context.device)
        ctx = self._full_context(context, r)
        if z is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        else:
            seed = 0
        theta, _, _ = sample_fused(self.flow, self.scaler, ctx, n_samples, z=z, seed=seed)
        return theta

    @torch.no_grad()
    def sample_posterior_host(self, context: Tensor, rank: int = 0, n_samples: int = 256, out: Optional[Tensor] = None,
                              chunk_events: Optional[int] = None, seed: Optional[int] = None) -> Tensor:
        """`sample_posterior(None, rank, n_samples, context=context)` for a consumer on the HOST (the reference's callers
        move every batch of draws off the device: `.cpu().numpy()`, pipeline.py:179, overlap_benchmark.py:59):
        `context` [E, context_dim] may live in (pinned) host memory, the result is theta [E, n, D] in pinned host memory
        (`out`, allocated if None).  Events go through the sampler in chunks and chunk i's device-to-host copy runs on a
        side stream under chunk i + 1's sampling, so only the last chunk's copy (1/8 of the output by default) is not
        hidden.  Chunk i draws its noise in-kernel under the Philox key `dist.shard_seed(seed, i)` (disjoint streams).
        The caller's stream is made to wait for the copies: synchronize it (or the device) before reading `out`."""
        from .dist import shard_seed
        dev = self.rank_embed.weight.device
        E, D = int(context.shape[0]), self.flow.features
        if out is None:
            out = torch.empty(E, n_samples, D, dtype=torch.float32, pin_memory=True)
        if tuple(out.shape) != (E, n_samples, D) or out.dtype != torch.float32 or out.is_cuda:
            raise ValueError(f"out must be a host float32 tensor of shape {(E, n_samples, D)}")
        if E == 0:
            return out
        chunk = int(chunk_events) if chunk_events else max(1, (E + 7) // 8)
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        cur = torch.cuda.current_stream(dev)
        copy = getattr(self, "_copy_stream", None)
        if copy is None or copy.device != dev:
            copy = torch.cuda.Stream(device=dev)
            object.__setattr__(self, "_copy_stream", copy)
        c_dev = context.to(dev, non_blocking=True)
        r = torch.full((E,), rank, dtype=torch.long, device=dev)
        ctx = self._full_context(c_dev, r)
        copy.wait_stream(cur)             # `out` may still be read by earlier work of the caller's stream
        for i, e0 in enumerate(range(0, E, chunk)):
            e1 = min(E, e0 + chunk)
            th = sample_fused(self.flow, self.scaler, ctx[e0:e1], n_samples, z=None, seed=shard_seed(seed, i))[0]
            done = torch.cuda.Event()
            done.record(cur)
            copy.wait_event(done)
            with torch.cuda.stream(copy):
                out[e0:e1].copy_(th, non_blocking=True)
            th.record_stream(copy)
        cur.wait_stream(copy)
        return out

    @torch.no_grad()
    def sample_with_log_prob(self, strain: Optional[Tensor], rank: int = 0, n_samples: int = 256,
                             asd_bands: Optional[Tensor] = None, context: Optional[Tensor] = None,
                             z: Optional[Tensor] = None, sort_masses: bool = True):
        """The body of infer()'s loop (pipeline.py:161-186) as one call: (theta, log q physical, railed)."""
        if context is None:
            context = self.encode(strain, asd_bands)
        r = torch.full((context.shape[0],), rank, dtype=torch.long, device=context.device)
        ctx = self._full_context(context, r)
        seed = 0 if z is not None else int(torch.randint(0, 2 ** 62, (1,)).item())
        return posterior_chunk(self.flow, self.scaler, ctx, n_samples, z=z, seed=seed, sort_masses=sort_masses)

    def _apply(self, fn, *args, **kwargs):
        # .to() / .cuda() / .cpu(): the ParamScaler tensors are plain attributes (lean_npe.py:87-98), they follow the
        # parameters (the reference moves them in its own `to`, lean_npe.py:334-338)
        out = super()._apply(fn, *args, **kwargs)
        self.scaler.to(next(self.parameters()).device)
        return out


def accelerate(model: nn.Module, tier: str = "fp32") -> nn.Module:
    """Convert a *reference* `ahsd.models.lean_npe.LeanNPE` in place: its nflows-backed `flow` is
    replaced by the B200 flow holding the same weights (strict state-dict load) and `sample_posterior`
    is rebound to the fused, context-broadcasting path.  Every other attribute (encoder, rank_embed,
    scaler, nll, encode) is untouched, so all reference call sites keep working."""
    old = model.flow
    dev = next(old.parameters()).device
    new = NSFPosteriorFlow(features=old.features, context_features=old.context_features,
                           hidden_features=old.transform._transforms[1].autoregressive_net.final_layer.weight.shape[1],
                           num_layers=old.num_layers,
                           num_bins=(old.transform._transforms[1].autoregressive_net.final_layer.weight.shape[0]
                                     // old.features + 1) // 3,
                           tail_bound=float(old.tail_bounds_list[0]), dropout=0.0,
                           temperature_scale=float(old.temperature.detach()), use_masked_context=False, tier=tier)
    new.load_state_dict(old.state_dict(), strict=True)
    new.temperature.requires_grad_(old.temperature.requires_grad)
    model.flow = new.to(dev)

    @torch.no_grad()
    def sample_posterior(self, strain, rank: int = 0, n_samples: int = 256, asd_bands=None, z=None):
        context = self.encode(strain, asd_bands)
        B = context.shape[0]
        r = torch.full((B,), rank, dtype=torch.long, device=context.device)
        ctx = self._full_context(context, r)
        seed = 0 if z is not None else int(torch.randint(0, 2 ** 62, (1,)).item())
        return sample_fused(self.flow, self.scaler, ctx, n_samples, z=z, seed=seed)[0]

    model.sample_posterior = types.MethodType(sample_posterior, model)
    return model
