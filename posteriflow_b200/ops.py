"""torch.library custom ops `posteriflow::*` over the C-ABI.

PyTorch is plumbing here: it owns device memory and the current CUDA stream; every op body is one
call into libposteriflow_b200.so.  There is no eager / CPU fallback: CPU tensors raise.
"""
from __future__ import annotations

import ctypes as C
import weakref
from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch import Tensor

from . import _cabi as A

# per-layer tensor order of the flat `weights` list handed to the ops (nflows state-dict names,
# relative to `transform._transforms.{2l+1}.autoregressive_net.`)
LAYER_KEYS = [
    "initial_layer.weight", "initial_layer.bias",
    "context_layer.weight", "context_layer.bias",
    "blocks.0.context_layer.weight", "blocks.0.context_layer.bias",
    "blocks.1.context_layer.weight", "blocks.1.context_layer.bias",
    "blocks.0.linear_layers.0.weight", "blocks.0.linear_layers.0.bias",
    "blocks.0.linear_layers.1.weight", "blocks.0.linear_layers.1.bias",
    "blocks.1.linear_layers.0.weight", "blocks.1.linear_layers.0.bias",
    "blocks.1.linear_layers.1.weight", "blocks.1.linear_layers.1.bias",
    "final_layer.weight", "final_layer.bias",
]
NW = len(LAYER_KEYS)

# id -> handle, WEAK: the owning NSFPosteriorFlow holds the only strong reference, so dropping the module (or moving
# it to another device) lets weakref.finalize release the arena, weight images, side streams and events
_HANDLES: "weakref.WeakValueDictionary[int, FlowHandle]" = weakref.WeakValueDictionary()
_NEXT_ID = [1]


def _ptr(t: Optional[Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _require_cuda(*ts: Tensor) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("posteriflow_b200 ops run on CUDA (sm_100a) only; there is no CPU fallback")


def _f32c(t: Tensor) -> Tensor:
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


class FlowHandle:
    """Owns one `pf_flow*` (packed weights of one NSFPosteriorFlow on one device)."""

    def __init__(self, features: int, context_features: int, hidden_features: int, num_layers: int,
                 num_bins: int, tail_bound: float, clamp_bound: float, tier: int, device: torch.device):
        self.lib = A.load_library()
        if device.type != "cuda":
            raise RuntimeError("posteriflow_b200 needs a CUDA device; there is no CPU fallback")
        self.device = device
        self.index = device.index if device.index is not None else torch.cuda.current_device()
        self.shape = (features, context_features, hidden_features, num_layers, num_bins)
        self.tier = tier
        desc = A.FlowDesc(features, context_features, hidden_features, num_layers, num_bins, 2,
                          float(tail_bound), float(clamp_bound), tier, self.index)
        h = C.c_void_p()
        A.check(self.lib.pf_flow_create(C.byref(desc), C.byref(h)), "pf_flow_create")
        self.h = h
        self.pack_key = None
        self.pack_ptrs = None
        self.id = _NEXT_ID[0]
        _NEXT_ID[0] += 1
        _HANDLES[self.id] = self
        self.launches = 0
        weakref.finalize(self, FlowHandle._destroy, self.lib, h, self.id)

    @staticmethod
    def _destroy(lib, h, hid):
        _HANDLES.pop(hid, None)
        try:
            lib.pf_flow_destroy(h)
        except Exception:
            pass

    def check(self) -> None:
        """Synchronise the current stream and raise if a tensor-core kernel of this handle hit a barrier timeout
        (pf_flow_check; such a kernel also writes NaN into the rows it abandons)."""
        with torch.cuda.device(self.index):
            A.check(self.lib.pf_flow_check(self.h, _stream()), "pf_flow_check")

    # -- packing ---------------------------------------------------------------------------------
    def ensure_packed(self, weights: Sequence[Tensor]) -> None:
        """Re-pack iff a weight tensor moved or was written in place since the last pack (optimizer step,
        load_state_dict).  The key is two integers per tensor; `weights` is the module's cached parameter list."""
        key = [w._version for w in weights]
        ptrs = [w.data_ptr() for w in weights]
        if key == self.pack_key and ptrs == self.pack_ptrs:
            return
        self.pack(weights)
        self.pack_key, self.pack_ptrs = key, ptrs

    def pack(self, weights: Sequence[Tensor], masks: Optional[Sequence[Dict[str, Tensor]]] = None) -> None:
        L = self.shape[3]
        if len(weights) != NW * L:
            raise ValueError(f"expected {NW * L} weight tensors, got {len(weights)}")
        keep = [_f32c(w.detach()) for w in weights]
        _require_cuda(*keep)
        arr = (A.LayerPtrs * L)()
        for l in range(L):
            w = keep[l * NW:(l + 1) * NW]
            p = arr[l]
            p.initial_w, p.initial_b, p.ctx_w, p.ctx_b = _ptr(w[0]), _ptr(w[1]), _ptr(w[2]), _ptr(w[3])
            p.blk_ctx_w[0], p.blk_ctx_b[0], p.blk_ctx_w[1], p.blk_ctx_b[1] = (_ptr(w[4]), _ptr(w[5]), _ptr(w[6]),
                                                                              _ptr(w[7]))
            for k in range(2):
                for j in range(2):
                    p.blk_lin_w[k][j] = _ptr(w[8 + 4 * k + 2 * j])
                    p.blk_lin_b[k][j] = _ptr(w[9 + 4 * k + 2 * j])
            p.final_w, p.final_b = _ptr(w[16]), _ptr(w[17])
            if masks is not None:
                m = masks[l]
                mk = {k: _f32c(v) for k, v in m.items()}
                keep.extend(mk.values())
                p.initial_mask = _ptr(mk.get("initial"))
                p.final_mask = _ptr(mk.get("final"))
                for k in range(2):
                    for j in range(2):
                        p.blk_lin_mask[k][j] = _ptr(mk.get(f"hidden{k}{j}"))
        with torch.cuda.device(self.index):
            A.check(self.lib.pf_flow_pack(self.h, arr, _stream()), "pf_flow_pack")
        self._note_launches()

    def _note_launches(self) -> None:
        self.launches += int(self.lib.pf_flow_last_launch_count(self.h))

    @property
    def fused_training(self) -> bool:
        """forward(save) / backward run as the fused persistent kernels rather than GEMM per launch"""
        return bool(self.lib.pf_flow_fused_training(self.h))

    def hidden_perm(self) -> List[int]:
        buf = (C.c_int32 * self.shape[2])()
        A.check(self.lib.pf_flow_hidden_perm(self.h, buf), "pf_flow_hidden_perm")
        return list(buf)

    def workspace(self, op: int, R: int, E: int) -> Tensor:
        n = int(self.lib.pf_flow_workspace_bytes(self.h, op, R, E))
        return torch.empty(max(n, 256), dtype=torch.uint8, device=self.device)

    def saved_bytes(self, R: int, E: int) -> int:
        return int(self.lib.pf_flow_saved_bytes(self.h, R, E))


def get_handle(hid: int) -> FlowHandle:
    try:
        return _HANDLES[hid]
    except KeyError:
        raise RuntimeError(f"posteriflow_b200: unknown flow handle {hid}") from None


def _rows(x: Tensor, ctx: Tensor, n_per_ctx: int, row2ctx: Optional[Tensor] = None) -> Tuple[int, int]:
    R, E = x.shape[0], ctx.shape[0]
    if row2ctx is not None:
        if row2ctx.dtype != torch.int64 or row2ctx.dim() != 1 or row2ctx.shape[0] != R or not row2ctx.is_cuda:
            raise ValueError(f"row2ctx must be a CUDA int64 tensor with one context index per flow row ({R})")
        return R, E
    if n_per_ctx < 1 or (R > 0 and (R - 1) // n_per_ctx >= E):
        raise ValueError(f"context rows ({E}) do not cover {R} flow rows at {n_per_ctx} rows per context")
    return R, E


def _r2c(row2ctx: Optional[Tensor]) -> Optional[Tensor]:
    return None if row2ctx is None else row2ctx.contiguous()


# ------------------------------------------------------------------------------------------------
# density direction
# ------------------------------------------------------------------------------------------------
@torch.library.custom_op("posteriflow::nsf_forward", mutates_args=())
def nsf_forward(x: Tensor, ctx: Tensor, weights: Sequence[Tensor], handle: int, n_per_ctx: int,
                save: bool, row2ctx: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """NSFPosteriorFlow.forward (flows.py:610-618): (z [R,D], logdet [R], saved activations).
    `row2ctx` ([R] int64) addresses the context row of every flow row explicitly instead of `row // n_per_ctx`;
    an EMPTY tensor means "not given" (this op is differentiable: its inputs keep a fixed tensor structure)."""
    _require_cuda(x, ctx)
    fh = get_handle(handle)
    x, ctx, row2ctx = _f32c(x), _f32c(ctx), (_r2c(row2ctx) if row2ctx.numel() else None)
    R, E = _rows(x, ctx, n_per_ctx, row2ctx)
    z = torch.empty_like(x)
    logdet = torch.empty(R, dtype=torch.float32, device=x.device)
    saved = torch.empty(fh.saved_bytes(R, E) if save else 0, dtype=torch.uint8, device=x.device)
    ws = fh.workspace(A.PF_OP_FORWARD, R, E)
    with torch.cuda.device(fh.index):
        A.check(fh.lib.pf_flow_forward(fh.h, x.data_ptr(), ctx.data_ptr(), _ptr(row2ctx), n_per_ctx, R, E, z.data_ptr(),
                                       logdet.data_ptr(), None, None, 0, saved.data_ptr() if save else None,
                                       ws.data_ptr(), ws.numel(), _stream()), "pf_flow_forward")
    fh._note_launches()
    return z, logdet, saved


@nsf_forward.register_fake
def _(x, ctx, weights, handle, n_per_ctx, save, row2ctx):
    return torch.empty_like(x), x.new_empty(x.shape[0]), x.new_empty(0, dtype=torch.uint8)


def no_row2ctx(like: Tensor) -> Tensor:
    return torch.empty(0, dtype=torch.int64, device=like.device)


@torch.library.custom_op("posteriflow::nsf_backward", mutates_args=())
def nsf_backward(saved: Tensor, ctx: Tensor, gz: Tensor, glogdet: Tensor, weights: Sequence[Tensor], handle: int,
                 n_per_ctx: int, row2ctx: Optional[Tensor] = None) -> List[Tensor]:
    """Backward of nsf_forward: [gx, gctx, *gweights] (weights in the nflows layout, masked entries zero)."""
    _require_cuda(saved, ctx, gz, glogdet)
    fh = get_handle(handle)
    ctx, gz, glogdet, row2ctx = _f32c(ctx), _f32c(gz), _f32c(glogdet), _r2c(row2ctx)
    R, E = _rows(gz, ctx, n_per_ctx, row2ctx)
    L = fh.shape[3]
    gx = torch.empty_like(gz)
    gctx = torch.empty_like(ctx)
    gws = [torch.empty(w.shape, dtype=torch.float32, device=gz.device) for w in weights]
    torch._foreach_zero_([gx, gctx] + gws)   # the wgrad kernels accumulate with atomics: one multi-tensor fill
    arr = (A.LayerGrads * L)()
    for l in range(L):
        g = gws[l * NW:(l + 1) * NW]
        p = arr[l]
        p.initial_w, p.initial_b, p.ctx_w, p.ctx_b = _ptr(g[0]), _ptr(g[1]), _ptr(g[2]), _ptr(g[3])
        p.blk_ctx_w[0], p.blk_ctx_b[0], p.blk_ctx_w[1], p.blk_ctx_b[1] = _ptr(g[4]), _ptr(g[5]), _ptr(g[6]), _ptr(g[7])
        for k in range(2):
            for j in range(2):
                p.blk_lin_w[k][j] = _ptr(g[8 + 4 * k + 2 * j])
                p.blk_lin_b[k][j] = _ptr(g[9 + 4 * k + 2 * j])
        p.final_w, p.final_b = _ptr(g[16]), _ptr(g[17])
    ws = fh.workspace(A.PF_OP_BACKWARD, R, E)
    with torch.cuda.device(fh.index):
        A.check(fh.lib.pf_flow_backward(fh.h, saved.data_ptr(), ctx.data_ptr(), _ptr(row2ctx), n_per_ctx, R, E, gz.data_ptr(),
                                        glogdet.data_ptr(), gx.data_ptr(), gctx.data_ptr(), arr, ws.data_ptr(),
                                        ws.numel(), _stream()), "pf_flow_backward")
    fh._note_launches()
    return [gx, gctx] + gws


def backward_into(fh: "FlowHandle", saved: Tensor, ctx: Tensor, gz: Tensor, glogdet: Tensor, grad_views: Sequence[Tensor],
                  gctx: Tensor) -> None:
    """pf_flow_backward with the weight gradients ACCUMULATED straight into `grad_views` (the per-parameter views of
    GraphedNLLStep's flat buffer, zeroed by the caller) and d ctx into `gctx` (zeroed here): no allocation, no copy of
    25 MB of gradients afterwards.  Plain function (called inside a CUDA-graph capture, not through autograd)."""
    R, E = int(gz.shape[0]), int(ctx.shape[0])
    L = fh.shape[3]
    arr = (A.LayerGrads * L)()
    for l in range(L):
        g = grad_views[l * NW:(l + 1) * NW]
        p = arr[l]
        p.initial_w, p.initial_b, p.ctx_w, p.ctx_b = _ptr(g[0]), _ptr(g[1]), _ptr(g[2]), _ptr(g[3])
        p.blk_ctx_w[0], p.blk_ctx_b[0], p.blk_ctx_w[1], p.blk_ctx_b[1] = _ptr(g[4]), _ptr(g[5]), _ptr(g[6]), _ptr(g[7])
        for k in range(2):
            for j in range(2):
                p.blk_lin_w[k][j] = _ptr(g[8 + 4 * k + 2 * j])
                p.blk_lin_b[k][j] = _ptr(g[9 + 4 * k + 2 * j])
        p.final_w, p.final_b = _ptr(g[16]), _ptr(g[17])
    ws = fh.workspace(A.PF_OP_BACKWARD, R, E)
    gctx.zero_()
    with torch.cuda.device(fh.index):
        A.check(fh.lib.pf_flow_backward(fh.h, saved.data_ptr(), ctx.data_ptr(), None, 1, R, E, gz.data_ptr(),
                                        glogdet.data_ptr(), None, gctx.data_ptr(), arr, ws.data_ptr(), ws.numel(), _stream()),
                "pf_flow_backward")
    fh._note_launches()


@nsf_backward.register_fake
def _(saved, ctx, gz, glogdet, weights, handle, n_per_ctx, row2ctx=None):
    return [torch.empty_like(gz), torch.empty_like(ctx)] + [torch.empty_like(w) for w in weights]


def _fwd_setup(ctx, inputs, output):
    x, c, weights, handle, n_per_ctx, save, row2ctx = inputs
    _, _, saved = output
    if not save:
        raise RuntimeError("nsf_forward was called with save=False but a gradient is required")
    ctx.save_for_backward(saved, c, row2ctx, *weights)
    ctx.handle, ctx.n_per_ctx = handle, n_per_ctx


def _fwd_backward(ctx, gz, glogdet, gsaved):
    saved, c, row2ctx, *weights = ctx.saved_tensors
    outs = torch.ops.posteriflow.nsf_backward(saved, c, gz, glogdet, weights, ctx.handle, ctx.n_per_ctx,
                                              row2ctx if row2ctx.numel() else None)
    return outs[0], outs[1], list(outs[2:]), None, None, None, None


nsf_forward.register_autograd(_fwd_backward, setup_context=_fwd_setup)


@torch.library.custom_op("posteriflow::nsf_nll", mutates_args=())
def nsf_nll(x: Tensor, ctx: Tensor, log_sigma: Optional[Tensor], handle: int, n_per_ctx: int,
            row2ctx: Optional[Tensor] = None) -> Tensor:
    """compute_psd_aware_nll (flows.py:727-779) fused, inference only (no autograd)."""
    _require_cuda(x, ctx)
    fh = get_handle(handle)
    x, ctx, row2ctx = _f32c(x), _f32c(ctx), _r2c(row2ctx)
    ls = _f32c(log_sigma) if log_sigma is not None else None
    R, E = _rows(x, ctx, n_per_ctx, row2ctx)
    nll = torch.empty(R, dtype=torch.float32, device=x.device)
    ws = fh.workspace(A.PF_OP_FORWARD, R, E)
    with torch.cuda.device(fh.index):
        A.check(fh.lib.pf_flow_forward(fh.h, x.data_ptr(), ctx.data_ptr(), _ptr(row2ctx), n_per_ctx, R, E, None, None,
                                       nll.data_ptr(), _ptr(ls), 0, None, ws.data_ptr(), ws.numel(), _stream()),
                "pf_flow_forward(nll)")
    fh._note_launches()
    return nll


@nsf_nll.register_fake
def _(x, ctx, log_sigma, handle, n_per_ctx, row2ctx=None):
    return x.new_empty(x.shape[0])


# ------------------------------------------------------------------------------------------------
# sampling direction
# ------------------------------------------------------------------------------------------------
def _scaler_struct(lo, hi, log_mask, circ_mask):
    if lo is None:
        return None, []
    keep = [_f32c(lo), _f32c(hi), log_mask.to(torch.uint8).contiguous(), circ_mask.to(torch.uint8).contiguous()]
    _require_cuda(*keep)
    return A.Scaler(*[t.data_ptr() for t in keep]), keep


@torch.library.custom_op("posteriflow::nsf_inverse", mutates_args=())
def nsf_inverse(z: Tensor, ctx: Tensor, handle: int, n_per_ctx: int, flags: int, lo: Optional[Tensor],
                hi: Optional[Tensor], log_mask: Optional[Tensor], circ_mask: Optional[Tensor],
                row2ctx: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """NSFPosteriorFlow.inverse (flows.py:620-655), optionally fused with ParamScaler.wrap/denormalize."""
    _require_cuda(z, ctx)
    fh = get_handle(handle)
    z, ctx, row2ctx = _f32c(z), _f32c(ctx), _r2c(row2ctx)
    R, E = _rows(z, ctx, n_per_ctx, row2ctx)
    x = torch.empty_like(z)
    logdet = torch.empty(R, dtype=torch.float32, device=z.device)
    sc, keep = _scaler_struct(lo, hi, log_mask, circ_mask)
    ws = fh.workspace(A.PF_OP_INVERSE, R, E)
    with torch.cuda.device(fh.index):
        A.check(fh.lib.pf_flow_inverse(fh.h, z.data_ptr(), ctx.data_ptr(), _ptr(row2ctx), n_per_ctx, R, E, x.data_ptr(),
                                       logdet.data_ptr(), flags, C.byref(sc) if sc is not None else None,
                                       ws.data_ptr(), ws.numel(), _stream()), "pf_flow_inverse")
    fh._note_launches()
    return x, logdet


@nsf_inverse.register_fake
def _(z, ctx, handle, n_per_ctx, flags, lo, hi, log_mask, circ_mask, row2ctx=None):
    return torch.empty_like(z), z.new_empty(z.shape[0])


@torch.library.custom_op("posteriflow::nsf_sample_fused", mutates_args=())
def nsf_sample_fused(ctx: Tensor, z: Optional[Tensor], seed: int, n: int, handle: int, flags: int, lo: Tensor,
                     hi: Tensor, log_mask: Tensor, circ_mask: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """LeanNPE.sample_posterior after the encoder (lean_npe.py:326-332): theta [E, n, D], logdet [E*n],
    railed [E*n] (pipeline.py:177).  z=None draws the base noise in-kernel (Philox keyed by `seed`)."""
    _require_cuda(ctx, z)
    fh = get_handle(handle)
    ctx = _f32c(ctx)
    E, D = ctx.shape[0], fh.shape[0]
    R = E * n
    if z is not None:
        z = _f32c(z)
        if z.shape[0] != R:
            raise ValueError(f"z has {z.shape[0]} rows, expected {R}")
    theta = torch.empty(E, n, D, dtype=torch.float32, device=ctx.device)
    logdet = torch.empty(R, dtype=torch.float32, device=ctx.device)
    railed = torch.empty(R, dtype=torch.uint8, device=ctx.device)
    sc, keep = _scaler_struct(lo, hi, log_mask, circ_mask)
    ws = fh.workspace(A.PF_OP_SAMPLE, R, E)
    with torch.cuda.device(fh.index):
        A.check(fh.lib.pf_sample_fused(fh.h, ctx.data_ptr(), C.byref(sc), _ptr(z), seed & (2 ** 64 - 1), E, n,
                                       theta.data_ptr(), logdet.data_ptr(), railed.data_ptr(), flags, ws.data_ptr(),
                                       ws.numel(), _stream()), "pf_sample_fused")
    fh._note_launches()
    return theta, logdet, railed


@nsf_sample_fused.register_fake
def _(ctx, z, seed, n, handle, flags, lo, hi, log_mask, circ_mask):
    E = ctx.shape[0]
    return (ctx.new_empty(E, n, lo.shape[0]), ctx.new_empty(E * n), ctx.new_empty(E * n, dtype=torch.uint8))


@torch.library.custom_op("posteriflow::nsf_physical", mutates_args=())
def nsf_physical(y: Tensor, nll: Tensor, sort_masses: bool, lo: Tensor, hi: Tensor, log_mask: Tensor,
                 circ_mask: Tensor) -> Tuple[Tensor, Tensor]:
    """Tail of infer()'s loop (pipeline.py:57-76, 180, 184-186): (theta physical [R, D], log q physical [R])."""
    _require_cuda(y, nll)
    y, nll = _f32c(y), _f32c(nll)
    R, D = y.shape
    theta = torch.empty_like(y)
    logq = torch.empty(R, dtype=torch.float32, device=y.device)
    sc, keep = _scaler_struct(lo, hi, log_mask, circ_mask)
    lib = A.load_library()
    idx = y.device.index if y.device.index is not None else torch.cuda.current_device()
    with torch.cuda.device(idx):
        A.check(lib.pf_physical_logq(y.data_ptr(), nll.data_ptr(), C.byref(sc), D, R, theta.data_ptr(), logq.data_ptr(),
                                     A.PF_FLAG_SORT_MASSES if sort_masses else 0, idx, _stream()), "pf_physical_logq")
    return theta, logq


@nsf_physical.register_fake
def _(y, nll, sort_masses, lo, hi, log_mask, circ_mask):
    return torch.empty_like(y), y.new_empty(y.shape[0])


@torch.library.custom_op("posteriflow::posterior_summary", mutates_args=())
def posterior_summary_op(theta: Tensor, q: Sequence[float], truth: Optional[Tensor], railed: Optional[Tensor]
                         ) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Per-event summaries of theta [E, n, D] on the device (overlap_benchmark.py:67-89, validate_checkpoint.py:102-128):
    (quantiles [E, Q, D], lower median [E, D], fraction of samples below truth [E, D], railing fraction [E])."""
    _require_cuda(theta, truth, railed)
    theta = _f32c(theta)
    if theta.dim() != 3:
        raise ValueError("theta must be [E, n, D]")
    E, n, D = theta.shape
    Q = len(q)
    dev = theta.device
    quant = torch.empty(E, Q, D, dtype=torch.float32, device=dev)
    lower = torch.empty(E, D, dtype=torch.float32, device=dev)
    below = torch.zeros(E, D, dtype=torch.float32, device=dev)
    rail = torch.zeros(E, dtype=torch.float32, device=dev)
    tr = _f32c(truth) if truth is not None else None
    rl = railed.to(torch.uint8).contiguous() if railed is not None else None
    if tr is not None and tuple(tr.shape) != (E, D):
        raise ValueError(f"truth must be [{E}, {D}]")
    if rl is not None and rl.numel() != E * n:
        raise ValueError(f"railed must have {E * n} entries")
    lib = A.load_library()
    qa = (C.c_double * max(Q, 1))(*[float(v) for v in q])
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    with torch.cuda.device(idx):
        A.check(lib.pf_posterior_summary(theta.data_ptr(), E, n, D, qa, Q, _ptr(tr), _ptr(rl),
                                         quant.data_ptr() if Q else None, lower.data_ptr(),
                                         below.data_ptr() if tr is not None else None,
                                         rail.data_ptr() if rl is not None else None, idx, _stream()),
                "pf_posterior_summary")
    return quant, lower, below, rail


@posterior_summary_op.register_fake
def _(theta, q, truth, railed):
    E, n, D = theta.shape
    return theta.new_empty(E, len(q), D), theta.new_empty(E, D), theta.new_empty(E, D), theta.new_empty(E)
