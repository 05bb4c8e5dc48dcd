"""Standalone CPU oracle for PosteriFlow's LeanNPE posterior-flow hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under `posteriflow_b200/` may import this file; only
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs do.

It restates, in plain PyTorch on CPU (dtype-generic: float32 or float64), the arithmetic that
the reference reaches through the third-party `nflows` package (unpinned, not vendored; last
release 0.14) at these call sites:

  /root/reference/src/ahsd/models/flows.py:508-529   stack = [ReversePermutation, MAF-RQS] x L
  /root/reference/src/ahsd/models/flows.py:610-618   NSFPosteriorFlow.forward
  /root/reference/src/ahsd/models/flows.py:620-655   NSFPosteriorFlow.inverse (sanitise, clamp +-3)
  /root/reference/src/ahsd/models/flows.py:56-85     PSDScaledNormal.log_prob
  /root/reference/src/ahsd/models/flows.py:727-779   compute_psd_aware_nll
  /root/reference/src/ahsd/models/lean_npe.py:48-114 ParamScaler
  /root/reference/src/ahsd/models/lean_npe.py:299-332 _full_context / nll / sample_posterior
  /root/reference/src/ahsd/inference/pipeline.py:57-76 _log_prob_physical

It does not import the reference or the `oracle/nflows` shim, so it runs on the GPU box where
/root/reference does not exist.  `oracle/check_against_reference.py` (run in the build
container) asserts that this file and the reference's own classes executed over the shim agree
bit-for-bit in float32 on CPU, and `oracle/make_golden.py` stores reference outputs as fixtures.

PARITY UNPINNED: the reference has no test / golden vector / checkpoint for this path and the
real nflows cannot be installed here (SURVEY.md sections 0.3-0.4, 4, 8c), so the oracle is
pinned to the reference's own code run over a *restated* nflows, not to upstream nflows itself.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

FLOW_NORM_BOUND = 3.0  # /root/reference/src/ahsd/models/parameter_scalers.py:27
MIN_BIN_WIDTH = 1e-3   # nflows rational_quadratic.DEFAULT_MIN_BIN_WIDTH
MIN_BIN_HEIGHT = 1e-3
MIN_DERIVATIVE = 1e-3

# recalled-behaviour switches, mirrored from oracle/nflows/__init__.py
MADE_EXPOSES_HIDDEN_FEATURES = False
RESIDUAL_BLOCK_CONTEXT_GLU = True


@dataclass(frozen=True)
class FlowSpec:
    """Shape of one NSFPosteriorFlow.  Defaults = LeanNPE (lean_npe.py:285-295)."""
    features: int = 11
    context_features: int = 288
    hidden_features: int = 256
    num_layers: int = 10
    num_bins: int = 16
    tail_bound: float = 5.0
    num_blocks: int = 2

    @property
    def multiplier(self) -> int:
        return 3 * self.num_bins - 1

    @property
    def out_features(self) -> int:
        return self.features * self.multiplier


# ----------------------------------------------------------------------------------------------
# degrees / masks
# ----------------------------------------------------------------------------------------------
def input_degrees(D: int) -> torch.Tensor:
    return torch.arange(1, D + 1)


def hidden_degrees(D: int, H: int) -> torch.Tensor:
    return torch.arange(H) % max(1, D - 1) + min(1, D - 1)


def output_degrees(D: int, mult: int) -> torch.Tensor:
    return input_degrees(D).repeat_interleave(mult)


def made_masks(spec: FlowSpec) -> Dict[str, torch.Tensor]:
    D, H = spec.features, spec.hidden_features
    din, dh, dout = input_degrees(D), hidden_degrees(D, H), output_degrees(D, spec.multiplier)
    return {
        "initial": (dh[:, None] >= din[None, :]).float(),
        "hidden": (dh[:, None] >= dh[None, :]).float(),
        "final": (dout[:, None] > dh[None, :]).float(),
        "deg_in": din, "deg_hidden": dh, "deg_out": dout,
    }


def _prefix(layer: int) -> str:
    return f"transform._transforms.{2 * layer + 1}.autoregressive_net."


def _linear_init(out_f: int, in_f: int, gen: torch.Generator) -> Tuple[torch.Tensor, torch.Tensor]:
    """Same distribution as torch.nn.Linear's default init (U(+-1/sqrt(fan_in)) for both)."""
    bound = 1.0 / math.sqrt(in_f)
    w = (torch.rand(out_f, in_f, generator=gen) * 2 - 1) * bound
    b = (torch.rand(out_f, generator=gen) * 2 - 1) * bound
    return w, b


def init_weights(spec: FlowSpec, seed: int = 0, stress: bool = False) -> Dict[str, torch.Tensor]:
    """Random-init flow weights under nflows' state-dict key names,
    relative to `LeanNPE.flow.`; loadable into the reference module with
    `flow.load_state_dict(..., strict=True)` (see `full_state_dict`).

    stress=True gives the (normally ~zero-initialised) block outputs a plain Linear init and scales
    the final layer (w x2, b x4) so that the conditioner really depends on x/context and all 16
    bins and both tails get exercised.  Stronger scales make the 10-layer map so ill-conditioned
    that float32 and float64 evaluations of the *reference itself* differ by >1e-1.
    Deterministic given (spec, seed, stress) on the CPU generator."""
    gen = torch.Generator(device="cpu").manual_seed(int(seed))
    D, C, H = spec.features, spec.context_features, spec.hidden_features
    masks = made_masks(spec)
    sd: Dict[str, torch.Tensor] = {}
    for l in range(spec.num_layers):
        p = _prefix(l)
        w, b = _linear_init(H, D, gen)
        sd[p + "initial_layer.weight"], sd[p + "initial_layer.bias"] = w, b
        sd[p + "initial_layer.mask"] = masks["initial"].clone()
        sd[p + "initial_layer.degrees"] = masks["deg_hidden"].clone()
        w, b = _linear_init(H, C, gen)
        sd[p + "context_layer.weight"], sd[p + "context_layer.bias"] = w, b
        for k in range(spec.num_blocks):
            q = p + f"blocks.{k}."
            w, b = _linear_init(H, C, gen)
            sd[q + "context_layer.weight"], sd[q + "context_layer.bias"] = w, b
            w, b = _linear_init(H, H, gen)
            sd[q + "linear_layers.0.weight"], sd[q + "linear_layers.0.bias"] = w, b
            if stress:
                w, b = _linear_init(H, H, gen)
            else:  # nflows zero_initialization: U(-1e-3, 1e-3)
                w = (torch.rand(H, H, generator=gen) * 2 - 1) * 1e-3
                b = (torch.rand(H, generator=gen) * 2 - 1) * 1e-3
            sd[q + "linear_layers.1.weight"], sd[q + "linear_layers.1.bias"] = w, b
            for j in (0, 1):
                sd[q + f"linear_layers.{j}.mask"] = masks["hidden"].clone()
                sd[q + f"linear_layers.{j}.degrees"] = masks["deg_hidden"].clone()
        w, b = _linear_init(spec.out_features, H, gen)
        if stress:
            w, b = w * 2.0, b * 4.0
        sd[p + "final_layer.weight"], sd[p + "final_layer.bias"] = w, b
        sd[p + "final_layer.mask"] = masks["final"].clone()
        sd[p + "final_layer.degrees"] = masks["deg_out"].clone()
    return sd


def full_state_dict(spec: FlowSpec, w: Dict[str, torch.Tensor], temperature: float = 1.0) -> Dict[str, torch.Tensor]:
    """Expand `init_weights` output to the complete `NSFPosteriorFlow.state_dict()` key set
    (SURVEY.md Appendix B): permutation buffers, temperature, and the duplicate `flow._transform.`
    aliases created by `Flow(self.transform, self.base_dist)` (flows.py:529-532)."""
    D = spec.features
    sd = {"temperature": torch.tensor(float(temperature)),
          "_ar_perm": torch.arange(D), "_ar_inv_perm": torch.arange(D)}
    for l in range(spec.num_layers):
        sd[f"transform._transforms.{2 * l}._permutation"] = torch.arange(D - 1, -1, -1)
    sd.update(w)
    for k in list(sd.keys()):
        if k.startswith("transform."):
            sd["flow._" + k] = sd[k]
    return sd


def weights_checksum(w: Dict[str, torch.Tensor]) -> float:
    """Order-independent fingerprint of a weight set (float64 sum of |w| * (1 + index hash))."""
    tot = 0.0
    for k in sorted(w.keys()):
        if k.endswith(".weight") or k.endswith(".bias"):
            t = w[k].double().reshape(-1)
            tot += float((t.abs() * (1.0 + (torch.arange(t.numel(), dtype=torch.float64) % 7))).sum())
    return tot


# ----------------------------------------------------------------------------------------------
# MADE conditioner (Appendix A.3) and RQ spline (Appendix A.5-A.7)
# ----------------------------------------------------------------------------------------------
def made_forward(spec: FlowSpec, w: Dict[str, torch.Tensor], layer: int, x: torch.Tensor,
                 ctx: Optional[torch.Tensor]) -> torch.Tensor:
    p = _prefix(layer)
    t = F.linear(x, w[p + "initial_layer.weight"] * w[p + "initial_layer.mask"], w[p + "initial_layer.bias"])
    if ctx is not None:
        t = t + F.relu(F.linear(ctx, w[p + "context_layer.weight"], w[p + "context_layer.bias"]))
    for k in range(spec.num_blocks):
        q = p + f"blocks.{k}."
        u = F.relu(t)
        u = F.linear(u, w[q + "linear_layers.0.weight"] * w[q + "linear_layers.0.mask"], w[q + "linear_layers.0.bias"])
        u = F.relu(u)
        u = F.linear(u, w[q + "linear_layers.1.weight"] * w[q + "linear_layers.1.mask"], w[q + "linear_layers.1.bias"])
        if ctx is not None:
            g = F.linear(ctx, w[q + "context_layer.weight"], w[q + "context_layer.bias"])
            if RESIDUAL_BLOCK_CONTEXT_GLU:
                u = F.glu(torch.cat((u, g), dim=1), dim=1)
            else:
                u = u + g
        t = t + u
    return F.linear(t, w[p + "final_layer.weight"] * w[p + "final_layer.mask"], w[p + "final_layer.bias"])


def _knots(unnorm: torch.Tensor, lo: float, hi: float, min_size: float) -> Tuple[torch.Tensor, torch.Tensor]:
    K = unnorm.shape[-1]
    s = F.softmax(unnorm, dim=-1)
    s = min_size + (1 - min_size * K) * s
    cum = torch.cumsum(s, dim=-1)
    cum = F.pad(cum, pad=(1, 0), mode="constant", value=0.0)
    cum = (hi - lo) * cum + lo
    cum[..., 0] = lo
    cum[..., -1] = hi
    return cum, cum[..., 1:] - cum[..., :-1]


def rq_spline(x: torch.Tensor, uw: torch.Tensor, uh: torch.Tensor, ud: torch.Tensor, inverse: bool,
              B: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """Elementwise monotone RQ spline on [-B, B] for points known to be inside the interval.
    `ud` already carries the two boundary constants (K+1 entries)."""
    cw, wdt = _knots(uw, -B, B, MIN_BIN_WIDTH)
    d = MIN_DERIVATIVE + F.softplus(ud)
    ch, hgt = _knots(uh, -B, B, MIN_BIN_HEIGHT)

    knots = ch if inverse else cw
    knots[..., -1] += 1e-6  # nflows searchsorted eps, applied in place to the last knot
    k = (torch.sum(x[..., None] >= knots, dim=-1) - 1)[..., None]

    x_k = cw.gather(-1, k)[..., 0]
    w_k = wdt.gather(-1, k)[..., 0]
    y_k = ch.gather(-1, k)[..., 0]
    delta = hgt / wdt
    s_k = delta.gather(-1, k)[..., 0]
    d_k = d.gather(-1, k)[..., 0]
    d_k1 = d[..., 1:].gather(-1, k)[..., 0]
    h_k = hgt.gather(-1, k)[..., 0]

    if inverse:
        dy = x - y_k
        t = d_k + d_k1 - 2 * s_k
        a = dy * t + h_k * (s_k - d_k)
        b = h_k * d_k - dy * t
        c = -s_k * dy
        disc = b.pow(2) - 4 * a * c
        assert (disc >= 0).all()
        root = (2 * c) / (-b - torch.sqrt(disc))
        out = root * w_k + x_k
        tt = root * (1 - root)
        den = s_k + t * tt
        num = s_k.pow(2) * (d_k1 * root.pow(2) + 2 * s_k * tt + d_k * (1 - root).pow(2))
        return out, -(torch.log(num) - 2 * torch.log(den))
    theta = (x - x_k) / w_k
    tt = theta * (1 - theta)
    num = h_k * (s_k * theta.pow(2) + d_k * tt)
    den = s_k + (d_k + d_k1 - 2 * s_k) * tt
    out = y_k + num / den
    dnum = s_k.pow(2) * (d_k1 * theta.pow(2) + 2 * s_k * tt + d_k * (1 - theta).pow(2))
    return out, torch.log(dnum) - 2 * torch.log(den)


def spline_elementwise(spec: FlowSpec, x: torch.Tensor, params: torch.Tensor, inverse: bool
                       ) -> Tuple[torch.Tensor, torch.Tensor]:
    """MaskedPiecewiseRationalQuadraticAutoregressiveTransform._elementwise with tails='linear'."""
    R, D, K, B = x.shape[0], spec.features, spec.num_bins, spec.tail_bound
    tp = params.view(R, D, spec.multiplier)
    uw, uh, ud = tp[..., :K], tp[..., K:2 * K], tp[..., 2 * K:]
    if MADE_EXPOSES_HIDDEN_FEATURES:
        uw = uw / math.sqrt(spec.hidden_features)
        uh = uh / math.sqrt(spec.hidden_features)
    inside = (x >= -B) & (x <= B)
    out = torch.zeros_like(x)
    lad = torch.zeros_like(x)
    ud = F.pad(ud, pad=(1, 1))
    const = math.log(math.exp(1 - MIN_DERIVATIVE) - 1)
    ud[..., 0] = const
    ud[..., -1] = const
    out[~inside] = x[~inside]
    if torch.any(inside):
        out[inside], lad[inside] = rq_spline(x[inside], uw[inside, :], uh[inside, :], ud[inside, :], inverse, B)
    return out, lad.sum(dim=1)


# ----------------------------------------------------------------------------------------------
# flow = [Reverse, MAF] x L  (Appendix A.1, A.4)
# ----------------------------------------------------------------------------------------------
def transform_forward(spec: FlowSpec, w, x: torch.Tensor, ctx: Optional[torch.Tensor]
                      ) -> Tuple[torch.Tensor, torch.Tensor]:
    rev = torch.arange(spec.features - 1, -1, -1)
    total = x.new_zeros(x.shape[0])
    for l in range(spec.num_layers):
        x = torch.index_select(x, 1, rev)
        params = made_forward(spec, w, l, x, ctx)
        x, lad = spline_elementwise(spec, x, params, inverse=False)
        total = total + lad
    return x, total


def transform_inverse(spec: FlowSpec, w, z: torch.Tensor, ctx: Optional[torch.Tensor]
                      ) -> Tuple[torch.Tensor, torch.Tensor]:
    """nflows' generic autoregressive inverse: D full conditioner passes per layer."""
    rev = torch.arange(spec.features - 1, -1, -1)
    total = z.new_zeros(z.shape[0])
    for l in reversed(range(spec.num_layers)):
        out = torch.zeros_like(z)
        lad = None
        for _ in range(spec.features):
            params = made_forward(spec, w, l, out, ctx)
            out, lad = spline_elementwise(spec, z, params, inverse=True)
        total = total + lad
        z = torch.index_select(out, 1, rev)
    return z, total


def transform_inverse_incremental(spec: FlowSpec, w, z: torch.Tensor, ctx: Optional[torch.Tensor]
                                  ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Same map as `transform_inverse`, computed the way the CUDA sampler does it: per layer,
    dimension d's spline parameters are evaluated once, after x_1..x_{d-1} are final (one
    conditioner pass of useful work instead of D).  Used only to show the two orders agree."""
    rev = torch.arange(spec.features - 1, -1, -1)
    total = z.new_zeros(z.shape[0])
    D, M = spec.features, spec.multiplier
    for l in reversed(range(spec.num_layers)):
        out = torch.zeros_like(z)
        lad_sum = z.new_zeros(z.shape[0])
        for d in range(D):
            params = made_forward(spec, w, l, out, ctx)   # only columns [d*M,(d+1)*M) are used
            sub = FlowSpec(1, spec.context_features, spec.hidden_features, 1, spec.num_bins,
                           spec.tail_bound, spec.num_blocks)
            o, lad = spline_elementwise(sub, z[:, d:d + 1], params[:, d * M:(d + 1) * M].contiguous(), True)
            out = out.clone()
            out[:, d] = o[:, 0]
            lad_sum = lad_sum + lad
        total = total + lad_sum
        z = torch.index_select(out, 1, rev)
    return z, total


def sanitize_context(ctx: torch.Tensor) -> torch.Tensor:
    """flows.py:626-632: only if any NaN/Inf is present, nan->0, +inf->1e-3, -inf->-1e-3."""
    if torch.isnan(ctx).any() or torch.isinf(ctx).any():
        ctx = torch.nan_to_num(ctx, nan=0.0, posinf=1e-3, neginf=-1e-3)
    return ctx


def flow_forward(spec: FlowSpec, w, x: torch.Tensor, ctx: torch.Tensor):
    """NSFPosteriorFlow.forward (identity _ar_perm, use_masked_context=False)."""
    return transform_forward(spec, w, x, ctx)


def flow_inverse(spec: FlowSpec, w, z: torch.Tensor, ctx: torch.Tensor):
    """NSFPosteriorFlow.inverse: sanitise ctx, invert, nan_to_num, clamp to +-FLOW_NORM_BOUND."""
    ctx = sanitize_context(ctx)
    try:
        x, lad = transform_inverse(spec, w, z, ctx)
    except AssertionError:  # flows.py:638-642 whole-batch identity fallback
        x, lad = z, z.new_zeros(z.shape[0])
    if torch.isnan(x).any() or torch.isinf(x).any():
        x = torch.nan_to_num(x, nan=0.0, posinf=1.0, neginf=-1.0)
    return torch.clamp(x, -FLOW_NORM_BOUND, FLOW_NORM_BOUND), lad


def base_log_prob(z: torch.Tensor, log_sigma: torch.Tensor) -> torch.Tensor:
    """PSDScaledNormal.log_prob (flows.py:56-85)."""
    sigma = torch.exp(log_sigma)
    quad = ((z / sigma) ** 2).sum(dim=1)
    var = 2 * log_sigma.sum(dim=1)
    const = z.shape[1] * math.log(2 * math.pi)
    return -0.5 * (quad + var + const)


def flow_nll(spec: FlowSpec, w, x: torch.Tensor, ctx: torch.Tensor,
             log_sigma: Optional[torch.Tensor] = None) -> torch.Tensor:
    """NSFPosteriorFlow.compute_psd_aware_nll (flows.py:727-779)."""
    if log_sigma is None:
        log_sigma = torch.zeros_like(x)
    z, lad = flow_forward(spec, w, x, ctx)
    return -(base_log_prob(z, log_sigma) + lad)


# ----------------------------------------------------------------------------------------------
# ParamScaler + LeanNPE glue (lean_npe.py)
# ----------------------------------------------------------------------------------------------
PARAM_NAMES = ["mass_1", "mass_2", "luminosity_distance", "ra", "dec", "theta_jn", "psi", "phase",
               "geocent_time", "a1", "a2"]
_RANGES = {
    "mass_1": (1.0, 105.0, True), "mass_2": (1.0, 105.0, True),
    "luminosity_distance": (40.0, 2200.0, True),
    "ra": (0.0, 2 * math.pi, False), "dec": (-math.pi / 2, math.pi / 2, False),
    "theta_jn": (0.0, math.pi, False), "psi": (0.0, math.pi, False),
    "phase": (0.0, 2 * math.pi, False), "geocent_time": (-1.6, 1.6, False),
    "a1": (0.0, 1.0, False), "a2": (0.0, 1.0, False),
}
_CIRCULAR = ("ra", "phase", "psi")


class Scaler:
    """lean_npe.py:48-114 ParamScaler restated."""

    def __init__(self, premerger: bool = False, dtype=torch.float32):
        lo, hi, lg = [], [], []
        for p in PARAM_NAMES:
            l, h, g = _RANGES[p]
            if p == "geocent_time" and premerger:
                l, h = -1.6, 5.2
            lo.append(math.log(l) if g else l)
            hi.append(math.log(h) if g else h)
            lg.append(g)
        self.lo = torch.tensor(lo, dtype=torch.float32).to(dtype)
        self.hi = torch.tensor(hi, dtype=torch.float32).to(dtype)
        self.log_mask = torch.tensor(lg, dtype=torch.bool)
        self.circ_mask = torch.tensor([p in _CIRCULAR for p in PARAM_NAMES], dtype=torch.bool)

    def wrap(self, y):
        wrapped = torch.remainder(y + 1.0, 2.0) - 1.0
        return torch.where(self.circ_mask, wrapped, y.clamp(-1.0, 1.0))

    def normalize(self, x):
        x = torch.where(self.log_mask, torch.log(x.clamp_min(1e-6)), x)
        return (2 * (x - self.lo) / (self.hi - self.lo) - 1).clamp(-1.0, 1.0)

    def denormalize(self, y):
        x = (y.clamp(-1.0, 1.0) + 1) / 2 * (self.hi - self.lo) + self.lo
        return torch.where(self.log_mask, torch.exp(x), x)


def full_context(context: torch.Tensor, rank_embed_weight: torch.Tensor, rank: torch.Tensor) -> torch.Tensor:
    """LeanNPE._full_context (lean_npe.py:299-300)."""
    return torch.cat([context, F.embedding(rank, rank_embed_weight)], dim=1)


def sample_posterior(spec: FlowSpec, w, scaler: Scaler, ctx_full: torch.Tensor, z: torch.Tensor,
                     n_samples: int, chunk: Optional[int] = None) -> torch.Tensor:
    """LeanNPE.sample_posterior (lean_npe.py:318-332) after the encoder, with the base noise z
    [E*n, D] given (event-major).  Returns physical samples [E, n, D]."""
    E = ctx_full.shape[0]
    rep = ctx_full.unsqueeze(1).expand(E, n_samples, ctx_full.shape[1]).reshape(E * n_samples, -1)
    if chunk is None:
        y, _ = flow_inverse(spec, w, z, rep)
    else:
        ys = [flow_inverse(spec, w, z[i:i + chunk], rep[i:i + chunk])[0] for i in range(0, z.shape[0], chunk)]
        y = torch.cat(ys)
    y = scaler.wrap(y).reshape(E, n_samples, -1)
    return scaler.denormalize(y)


def log_prob_physical(spec: FlowSpec, w, scaler: Scaler, y_norm: torch.Tensor, ctx_full: torch.Tensor
                      ) -> torch.Tensor:
    """pipeline.py:57-76 _log_prob_physical."""
    neg = flow_nll(spec, w, y_norm, ctx_full)
    theta = scaler.denormalize(y_norm)
    span = scaler.hi - scaler.lo
    jac = (math.log(2.0) - torch.log(span)).sum()
    log_theta = torch.where(scaler.log_mask, torch.log(theta.clamp_min(1e-6)), torch.zeros_like(theta))
    return -neg + jac - log_theta.sum(dim=1)


def infer_chunk(spec: FlowSpec, w, scaler: Scaler, ctx_full_row: torch.Tensor, z: torch.Tensor):
    """One iteration of infer()'s sampling loop (pipeline.py:168-180): inverse, wrap, rail mask,
    denormalise, physical log-prob.  ctx_full_row: [1, C]."""
    k = z.shape[0]
    fc = ctx_full_row.expand(k, -1)
    y, _ = flow_inverse(spec, w, z, fc)
    y = scaler.wrap(y)
    railed = ((y.abs() > 0.999) & ~scaler.circ_mask).any(dim=1)
    return scaler.denormalize(y), log_prob_physical(spec, w, scaler, y, fc), railed
