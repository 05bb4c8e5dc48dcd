"""Drop-in `NSFPosteriorFlow` for PosteriFlow's LeanNPE, routed through the sm_100a C-ABI.

Mirrors the reference class at /root/reference/src/ahsd/models/flows.py:363-939 for the branch
LeanNPE uses (`use_masked_context=False`, flows.py:508-526): same constructor arguments, same
`forward / inverse / compute_psd_aware_nll` signatures and return conventions, same state-dict
key set so `load_state_dict(..., strict=True)` of a reference checkpoint
works (pipeline.py:46).  The parameter containers below hold tensors only -- no arithmetic runs in
PyTorch: every call goes through `torch.ops.posteriflow.*` (ops.py) into libposteriflow_b200.so.
There is no CPU / eager fallback.
"""
from __future__ import annotations

import math
from typing import List, Optional, Tuple

import torch
from torch import Tensor, nn

from . import _cabi as A
from . import ops

FLOW_NORM_BOUND = 3.0  # /root/reference/src/ahsd/models/parameter_scalers.py:27


def _hidden_degrees(D: int, H: int) -> Tensor:
    return torch.arange(H) % max(1, D - 1) + min(1, D - 1)


class _Lin(nn.Module):
    """weight/bias holder with nn.Linear's default init; `zero_init` = nflows' U(-1e-3, 1e-3)."""

    def __init__(self, in_f: int, out_f: int, zero_init: bool = False):
        super().__init__()
        bound = 1e-3 if zero_init else 1.0 / math.sqrt(in_f)
        self.weight = nn.Parameter(torch.empty(out_f, in_f).uniform_(-bound, bound))
        self.bias = nn.Parameter(torch.empty(out_f).uniform_(-bound, bound))


class _MaskedLin(_Lin):
    """Holder for an nflows MaskedLinear: adds the `mask` and `degrees` buffers (Appendix A.2)."""

    def __init__(self, in_deg: Tensor, out_deg: Tensor, is_output: bool, zero_init: bool = False):
        super().__init__(in_deg.numel(), out_deg.numel(), zero_init)
        mask = (out_deg[:, None] > in_deg[None, :]) if is_output else (out_deg[:, None] >= in_deg[None, :])
        self.register_buffer("mask", mask.float())
        self.register_buffer("degrees", out_deg.clone())


class _Block(nn.Module):
    def __init__(self, deg_h: Tensor, C: int):
        super().__init__()
        H = deg_h.numel()
        self.context_layer = _Lin(C, H)
        self.linear_layers = nn.ModuleList([_MaskedLin(deg_h, deg_h, False), _MaskedLin(deg_h, deg_h, False, True)])


class _MADE(nn.Module):
    def __init__(self, D: int, C: int, H: int, mult: int, num_blocks: int = 2):
        super().__init__()
        deg_in = torch.arange(1, D + 1)
        deg_h = _hidden_degrees(D, H)
        deg_out = deg_in.repeat_interleave(mult)
        self.initial_layer = _MaskedLin(deg_in, deg_h, False)
        self.context_layer = _Lin(C, H)
        self.blocks = nn.ModuleList([_Block(deg_h, C) for _ in range(num_blocks)])
        self.final_layer = _MaskedLin(deg_h, deg_out, True)


class _MAF(nn.Module):
    def __init__(self, D, C, H, mult):
        super().__init__()
        self.autoregressive_net = _MADE(D, C, H, mult)


class _Reverse(nn.Module):
    def __init__(self, D: int):
        super().__init__()
        self.register_buffer("_permutation", torch.arange(D - 1, -1, -1))


class _Composite(nn.Module):
    def __init__(self, transforms: List[nn.Module]):
        super().__init__()
        self._transforms = nn.ModuleList(transforms)


class _FlowAlias(nn.Module):
    """nflows `Flow(transform, base)` registers the same CompositeTransform again as `_transform`."""

    def __init__(self, transform: nn.Module):
        super().__init__()
        self._transform = transform


class NSFPosteriorFlow(nn.Module):
    """Conditional masked-autoregressive RQ-spline flow, B200-native (flows.py:363-548 replaced).

    Extra, non-reference keyword: `tier` ("fp32" parity tier | "bf16" tensor-core tier | "bf16x3" split-bf16
    tensor-core tier at the FP32 tolerance for the density direction / training; it samples on the FP32-FMA kernels).
    Extra, non-reference argument on `inverse` / `forward`: `n_per_ctx` -- context row e serves flow
    rows [e*n, (e+1)*n) so the context is projected once per event instead of being repeated
    (lean_npe.py:328).  With n_per_ctx=None the reference's dense [R, C] form is used; a context that
    is an `expand`ed single row (pipeline.py:171) is recognised by its zero stride and broadcast.
    """

    def __init__(self, features: int, context_features: int = 0, hidden_features: int = 256, num_layers: int = 12,
                 num_bins: int = 16, tail_bound: float = FLOW_NORM_BOUND, max_overlaps: int = 6, dropout: float = 0.0,
                 temperature_scale: float = 1.5, use_masked_context: Optional[bool] = None, tier: str = "fp32",
                 **kwargs):
        super().__init__()
        if isinstance(tail_bound, dict):
            raise NotImplementedError("per-parameter tail bounds belong to the masked-context branch "
                                      "(flows.py:463-507), which LeanNPE does not use")
        if use_masked_context:
            raise NotImplementedError("use_masked_context=True (flows.py:112-360) is off the LeanNPE hot path")
        if context_features <= 0:
            raise NotImplementedError("the hot path is the conditional flow (context_features > 0)")
        if dropout != 0.0:
            raise NotImplementedError("LeanNPE builds the flow with dropout=0.0 (lean_npe.py:293)")
        self.features, self.context_features = features, context_features
        self.hidden_features, self.num_layers, self.num_bins = hidden_features, num_layers, num_bins
        self.tail_bound = float(tail_bound)
        self.max_overlaps = max_overlaps
        self.use_masked_context = False
        self.tier = {"fp32": A.PF_TIER_FP32, "bf16": A.PF_TIER_BF16, "bf16x3": A.PF_TIER_BF16X3}[tier]
        self.temperature = nn.Parameter(torch.tensor(temperature_scale, dtype=torch.float32))
        self.temperature_scale_init = temperature_scale
        mult = 3 * num_bins - 1
        transforms: List[nn.Module] = []
        for _ in range(num_layers):
            transforms.append(_Reverse(features))
            transforms.append(_MAF(features, context_features, hidden_features, mult))
        self.transform = _Composite(transforms)
        self.flow = _FlowAlias(self.transform)
        self.register_buffer("_ar_perm", torch.arange(features, dtype=torch.long))
        self.register_buffer("_ar_inv_perm", torch.arange(features, dtype=torch.long))
        self._handle: Optional[ops.FlowHandle] = None
        self._weights: Optional[List[Tensor]] = None
        self._ar_identity = True

    # -- plumbing ----------------------------------------------------------------------------------
    def layer_weights(self) -> List[Tensor]:
        """Flat [L * 18] parameter list in ops.LAYER_KEYS order.  The list of Parameter OBJECTS is cached: `.to()`,
        `load_state_dict` and optimizers all write into the same objects, and a moved / rewritten storage is caught
        by the (data_ptr, _version) key of FlowHandle.ensure_packed -- no per-call walk over the module tree."""
        if self._weights is None:
            out: List[Tensor] = []
            for l in range(self.num_layers):
                net = self.transform._transforms[2 * l + 1].autoregressive_net
                sd = dict(net.named_parameters())
                out.extend(sd[k] for k in ops.LAYER_KEYS)
            self._weights = out
        return self._weights

    def _apply(self, fn, *args, **kwargs):   # .to() / .cuda() / .double(): parameter objects may be replaced
        out = super()._apply(fn, *args, **kwargs)
        self._weights = None
        return out

    def handle(self) -> ops.FlowHandle:
        w = self.layer_weights()
        dev = w[0].device
        if dev.type != "cuda":
            raise RuntimeError("posteriflow_b200.NSFPosteriorFlow runs on CUDA (sm_100a) only; move the module "
                               "to a CUDA device. There is no CPU fallback.")
        if self._handle is None or self._handle.device != dev:
            self._handle = ops.FlowHandle(self.features, self.context_features, self.hidden_features,
                                          self.num_layers, self.num_bins, self.tail_bound, FLOW_NORM_BOUND,
                                          self.tier, dev)
        fp = getattr(self, "_flat_param", None)
        if fp is not None and fp._version != self._flat_seen:
            # GraphedNLLStep.flatten_parameters(): the optimizer writes the flat buffer the parameters alias -- their own
            # version counters do not move, so the pack key is bumped here
            self._flat_seen = fp._version
            self._handle.pack_key = None
        self._handle.ensure_packed(w)
        return self._handle

    @property
    def gpu_launches(self) -> int:
        return 0 if self._handle is None else self._handle.launches

    def _ctx_form(self, context: Tensor, R: int, n_per_ctx: Optional[int]) -> Tuple[Tensor, int]:
        if context is None:
            raise ValueError("the conditional flow needs a context")
        if n_per_ctx is not None:
            return context, int(n_per_ctx)
        if context.dim() == 2 and context.shape[0] == R and R > 1 and context.stride(0) == 0:
            return context[:1], R  # expanded single event (pipeline.py:171)
        if context.shape[0] != R:
            raise ValueError(f"context has {context.shape[0]} rows, inputs have {R}")
        return context, 1

    # -- reference surface -------------------------------------------------------------------------
    def forward(self, x: Tensor, context: Tensor = None, n_per_ctx: Optional[int] = None,
                row2ctx: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
        """flows.py:610-618: x (params) -> (z, log|det dz/dx|); differentiable wrt x, context, weights.
        `row2ctx` ([R] int64): explicit context row per flow row (several ranks / repeated events in one call)."""
        fh = self.handle()
        if not self._ar_identity:
            x = x.index_select(1, self._ar_perm)               # flows.py:612
        if row2ctx is None:
            context, n = self._ctx_form(context, x.shape[0], n_per_ctx)
        else:
            n = 1
        w = self.layer_weights()
        need = torch.is_grad_enabled() and (x.requires_grad or context.requires_grad or any(p.requires_grad for p in w))
        z, logdet, _ = torch.ops.posteriflow.nsf_forward(x, context, w, fh.id, n, bool(need),
                                                         row2ctx if row2ctx is not None else ops.no_row2ctx(x))
        return z, logdet

    def inverse(self, z: Tensor, context: Tensor = None, n_overlaps: Optional[Tensor] = None,
                n_per_ctx: Optional[int] = None, row2ctx: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
        """flows.py:620-655: z -> (x clamped to +-3, log_det), context sanitised in-kernel."""
        fh = self.handle()
        if row2ctx is None:
            context, n = self._ctx_form(context, z.shape[0], n_per_ctx)
        else:
            n = 1
        flags = A.PF_FLAG_SANITIZE_CTX | A.PF_FLAG_CLAMP
        with torch.no_grad():
            x, logdet = torch.ops.posteriflow.nsf_inverse(z, context, fh.id, n, flags, None, None, None, None, row2ctx)
            if not self._ar_identity:                          # flows.py:648 (nan_to_num / clamp are per element:
                x = x.index_select(1, self._ar_inv_perm)       #  they commute with the column gather)
        return x, logdet

    def compute_psd_aware_nll(self, x: Tensor, context: Tensor, log_sigma_psd: Optional[Tensor] = None,
                              n_per_ctx: Optional[int] = None, row2ctx: Optional[Tensor] = None) -> Tensor:
        """flows.py:727-779 with the PSDScaledNormal base (flows.py:73-85)."""
        fh = self.handle()
        if not self._ar_identity:
            x = x.index_select(1, self._ar_perm)               # through self.forward, flows.py:752
        if row2ctx is None:
            ctx, n = self._ctx_form(context, x.shape[0], n_per_ctx)
        else:
            ctx, n = context, 1
        w = self.layer_weights()
        need = torch.is_grad_enabled() and (x.requires_grad or ctx.requires_grad or any(p.requires_grad for p in w)
                                            or (log_sigma_psd is not None and log_sigma_psd.requires_grad))
        if not need:
            return torch.ops.posteriflow.nsf_nll(x, ctx, log_sigma_psd, fh.id, n, row2ctx)
        z, logdet, _ = torch.ops.posteriflow.nsf_forward(x, ctx, w, fh.id, n, True,
                                                         row2ctx if row2ctx is not None else ops.no_row2ctx(x))
        const = self.features * math.log(2 * math.pi)
        if log_sigma_psd is None:
            logp = -0.5 * ((z * z).sum(dim=1) + const)
        else:
            logp = -0.5 * (((z / torch.exp(log_sigma_psd)) ** 2).sum(dim=1) + 2 * log_sigma_psd.sum(dim=1) + const)
        return -(logp + logdet)

    def set_autoregressive_order(self, order) -> None:
        """flows.py:550-586: `order[i]` = original feature index placed at autoregressive position i.  A fixed column
        gather in front of forward() and its inverse behind inverse(); the kernels are untouched.  (The per-position
        tail-bound re-patching of flows.py:577-584 belongs to the per-parameter-tail-bound branch, not built here.)"""
        order = [int(i) for i in order]
        if sorted(order) != list(range(self.features)):
            raise ValueError(f"order must be a permutation of range({self.features}), got {order}")
        perm = torch.tensor(order, dtype=torch.long, device=self._ar_perm.device)
        self._ar_perm = perm
        self._ar_inv_perm = torch.argsort(perm)
        self._ar_identity = order == list(range(self.features))

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)
        # a checkpoint written after set_autoregressive_order carries the permutation in its buffers
        self._ar_identity = bool((self._ar_perm.cpu() == torch.arange(self.features)).all())
