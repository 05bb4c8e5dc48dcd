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
        # a rank without signals still takes part in the count all-reduce (and, through dist.allreduce_gradients'
        # zero-fill of missing .grad tensors, in the gradient exchange): no rank may skip a collective
        zero = context.sum() * 0.0
        if distributed:
            return global_mean_loss(zero, torch.tensor(0.0, device=zero.device))
        return zero
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


class GraphedNLLStep:
    """`batch_nll` + backward of the flow head as ONE CUDA-graph replay per step.

    The flow's forward-with-saved-activations and backward are ~320 dependent kernel launches per step
    (DESIGN.md 5.6); issued eagerly through autograd the host needs longer to launch them than the device to run
    them.  Here the two custom ops (`posteriflow::nsf_forward(save=True)`, `posteriflow::nsf_backward`) and the
    NLL arithmetic between them are captured once per row-count bucket and replayed:

        step = GraphedNLLStep(model)
        loss = step(params, nsig, context)      # .grad of the flow weights, rank embedding and `context` are set

    Rows are padded to a multiple of `bucket` (weight 0: a padded row contributes exactly zero to every
    gradient), so a handful of graphs covers the step-to-step variation of the number of signals.  The part of
    the graph of `context` that lies outside the flow (row gather, rank embedding, the caller's encoder) stays
    with autograd: the graph hands back d loss / d full-context and `Tensor.backward` carries it on.
    Same value and gradients as `batch_nll(...).backward()` (tests/test_gpu_glue.py).

    **Slot mode** (`slots=True`; the default whenever all B * max_signals (event, rank) slots of a batch fit one wave of
    128-row tiles, i.e. B * max_signals <= 128 * #SMs): the fused training kernels keep one tile per SM from the first
    layer to the last, so a step costs one tile's latency whether 8 or 148 SMs are busy.  The step then runs over EVERY
    slot of the batch with weight [rank < nsig] / count instead of gathering the present signals first: all shapes are
    static, so the row gather, the scaler, the rank embedding, their backward and the signal count move INTO the
    captured graph -- no `nonzero` (a host sync), no ATen glue around the replay, one graph for every batch of that
    size.  Absent slots carry weight zero and x = 0 (their parameters are never read): they add
    exactly nothing to the loss and to every gradient.

    **Compact rows form** (`compact=True`; the default for batches beyond one wave, and for EVERY batch whose `nsig` is
    passed as a host tensor -- the row count then needs no device sync): only the present pairs become flow
    rows, but WHICH pair a row is gets worked out on the device inside the graph (prefix sum of nsig, `pf_rows_prepare` /
    `pf_rows_ctx_grad` of csrc/pf_glue.cu), so the ATen gather around the replay goes away here too; the host reads the
    row count only, to choose the bucket.  `compact=False` keeps the gather with ATen (the form the first paragraph
    describes), which the tests hold against it."""

    def __init__(self, model, bucket: int = 128, distributed: bool = False, slots: Optional[bool] = None,
                 compact: bool = True):
        self.model, self.bucket, self.distributed = model, int(bucket), distributed
        self.slots = slots
        self.compact = compact          # rows mode through pf_rows_prepare / pf_rows_ctx_grad (False: the ATen gather path)
        self.slot_graphs = {}
        self.graphs = {}
        self.flow_params = model.flow.layer_weights()
        self._gbuf = None
        self.flat_grad = None
        self.flat_param = None
        self.last_launches = 0

    def _capture(self, cap: int):
        m = self.model
        dev = self.flow_params[0].device
        D, C = m.flow.features, m.flow.context_features
        st = {"x": torch.zeros(cap, D, device=dev), "ctx": torch.zeros(cap, C, device=dev),
              "w": torch.zeros(cap, device=dev)}
        fh = m.flow.handle()
        weights = [p.detach() for p in self.flow_params]
        from .ops import no_row2ctx
        no_r2c = no_row2ctx(st["x"])

        self._ensure_flat()

        # loss + its first backward step as one library kernel, weight gradients accumulated straight into the views of the
        # flat buffer that the data-parallel all-reduce exchanges (csrc/pf_glue.cu, ops.backward_into): no ATen glue and no
        # 25 MB gradient copy inside the replay
        from . import _cabi as A_
        from .ops import _stream, backward_into
        lib, idx = fh.lib, fh.index
        loss_buf = torch.zeros(1, device=dev)
        gzbuf = torch.zeros(cap, D, device=dev)
        glbuf = torch.zeros(cap, device=dev)
        gcbuf = torch.zeros(cap, C, device=dev)

        def body():
            z, ld, saved = torch.ops.posteriflow.nsf_forward(st["x"], st["ctx"], weights, fh.id, 1, True, no_r2c)
            with torch.cuda.device(idx):
                A_.check(lib.pf_slot_loss(z.data_ptr(), ld.data_ptr(), st["w"].data_ptr(), cap, D, loss_buf.data_ptr(),
                                          gzbuf.data_ptr(), glbuf.data_ptr(), idx, _stream()), "pf_slot_loss")
            self.flat_step.zero_()
            backward_into(fh, saved, st["ctx"], gzbuf, glbuf, self._step_views, gcbuf)
            return loss_buf.reshape(()), gcbuf, (loss_buf, gzbuf, glbuf)

        s = torch.cuda.Stream(device=dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s):
            for _ in range(2):      # warm-up outside capture (lazy attribute setting, allocator)
                body()
        torch.cuda.current_stream(dev).wait_stream(s)
        g = torch.cuda.CUDAGraph()
        l0 = fh.launches
        with torch.cuda.graph(g):
            st["loss"], st["gctx"], st["gw"] = body()
        st["graph"] = g
        st["launches"] = fh.launches - l0 + 1      # library kernels inside one replay (+ the loss kernel)
        return st

    # -- slot mode ---------------------------------------------------------------------------------------------------
    def _use_slots(self, B: int, dev) -> bool:
        if self.slots is not None:
            return bool(self.slots)
        if not self.model.flow.handle().fused_training:      # GEMM-per-launch kernels: more rows = more time
            return False
        # one wave of tiles: the chain kernels give every 128-row tile a PAIR of SMs (csrc/pf_tc_pair.cuh)
        n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
        return B * int(self.model.max_signals) <= 128 * (n_sm // 2)

    def _capture_slots(self, B: int, ctx_requires_grad: bool, compact_cap: Optional[int] = None):
        """`compact_cap` None: one flow row per (event, rank) slot.  An integer: the COMPACT form -- only the present pairs
        become rows (event-major, found on the device from a prefix sum of nsig), padded to `compact_cap` rows with weight 0;
        the graph is then keyed on (B, compact_cap) and the host reads only the row count, to choose the bucket."""
        m = self.model
        dev = self.flow_params[0].device
        D, C, S = m.flow.features, m.flow.context_features, int(m.max_signals)
        R = B * S
        cap = (R + 127) // 128 * 128 if compact_cap is None else int(compact_cap)
        st = {"params": torch.zeros(B, S, D, device=dev), "nsig": torch.zeros(B, dtype=torch.long, device=dev),
              "context": torch.zeros(B, m.context_dim, device=dev), "inv_count": torch.zeros((), device=dev),
              "use_inv": torch.zeros((), device=dev)}
        fh = m.flow.handle()
        weights = [p.detach() for p in self.flow_params]
        from .ops import no_row2ctx
        no_r2c = no_row2ctx(st["params"])
        self._ensure_flat()
        sc = m.scaler
        xbuf = torch.zeros(cap, D, device=dev)
        cbuf = torch.zeros(cap, C, device=dev)
        wbuf = torch.zeros(cap, device=dev)
        emb = m.rank_embed.weight

        # the glue around the two flow calls as three kernels of the library (csrc/pf_glue.cu) instead of ~30 ATen launches
        # (0.16 ms of a 1.10 ms step): slot preparation, loss + its first backward step, context / embedding gradients;
        # the weight gradients are accumulated straight into the flat buffer's views (no 25 MB copy afterwards)
        import ctypes as C_
        from . import _cabi as A_
        from .ops import _scaler_struct, _stream, backward_into
        lib = fh.lib
        sc_struct, sc_keep = _scaler_struct(sc.lo, sc.hi, sc.log_mask, sc.circ_mask)
        zbuf_loss = torch.zeros(1, device=dev)
        gzbuf = torch.zeros(cap, D, device=dev)
        glbuf = torch.zeros(cap, device=dev)
        gcbuf = torch.zeros(cap, C, device=dev)
        gcontext = torch.zeros(B, m.context_dim, device=dev)
        gemb = torch.zeros(S, C - m.context_dim, device=dev)
        emb_c = emb.detach()
        assert emb_c.is_contiguous() and emb_c.shape[0] >= S and emb_c.shape[1] == C - m.context_dim
        idx = fh.index
        offs = torch.zeros(B + 1, dtype=torch.int32, device=dev)

        def body():
            with torch.cuda.device(idx):
                if compact_cap is None:
                    A_.check(lib.pf_slot_prepare(st["params"].data_ptr(), st["nsig"].data_ptr(), st["context"].data_ptr(),
                                                 emb_c.data_ptr(), C_.byref(sc_struct), st["inv_count"].data_ptr(),
                                                 st["use_inv"].data_ptr(), B, S, D, m.context_dim, C, xbuf.data_ptr(),
                                                 cbuf.data_ptr(), wbuf.data_ptr(), idx, _stream()), "pf_slot_prepare")
                else:
                    A_.check(lib.pf_rows_prepare(st["params"].data_ptr(), st["nsig"].data_ptr(), st["context"].data_ptr(),
                                                 emb_c.data_ptr(), C_.byref(sc_struct), st["inv_count"].data_ptr(),
                                                 st["use_inv"].data_ptr(), B, S, D, m.context_dim, C, cap, offs.data_ptr(),
                                                 xbuf.data_ptr(), cbuf.data_ptr(), wbuf.data_ptr(), idx, _stream()),
                             "pf_rows_prepare")
            z, ld, saved = torch.ops.posteriflow.nsf_forward(xbuf, cbuf, weights, fh.id, 1, True, no_r2c)
            with torch.cuda.device(idx):
                A_.check(lib.pf_slot_loss(z.data_ptr(), ld.data_ptr(), wbuf.data_ptr(), cap, D, zbuf_loss.data_ptr(),
                                          gzbuf.data_ptr(), glbuf.data_ptr(), idx, _stream()), "pf_slot_loss")
            self.flat_step.zero_()
            backward_into(fh, saved, cbuf, gzbuf, glbuf, self._step_views, gcbuf)
            with torch.cuda.device(idx):
                if compact_cap is None:
                    A_.check(lib.pf_slot_ctx_grad(gcbuf.data_ptr(), B, S, m.context_dim, C, gcontext.data_ptr(),
                                                  gemb.data_ptr(), idx, _stream()), "pf_slot_ctx_grad")
                else:
                    A_.check(lib.pf_rows_ctx_grad(gcbuf.data_ptr(), offs.data_ptr(), B, S, m.context_dim, C,
                                                  gcontext.data_ptr(), gemb.data_ptr(), idx, _stream()), "pf_rows_ctx_grad")
            return zbuf_loss.reshape(()), gcontext, gemb

        s = torch.cuda.Stream(device=dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s):
            for _ in range(2):
                body()
        torch.cuda.current_stream(dev).wait_stream(s)
        g = torch.cuda.CUDAGraph()
        l0 = fh.launches
        with torch.cuda.graph(g):
            st["loss"], st["gcontext"], st["gemb"] = body()
        # everything the captured kernels read or write must outlive this frame: the graph holds raw addresses
        st["keep"] = (xbuf, cbuf, wbuf, no_r2c, weights, emb, emb_c, sc_struct, sc_keep, zbuf_loss, gzbuf, glbuf, gcbuf,
                      gcontext, gemb, offs)
        st["graph"] = g
        # + the glue kernels (not calls on the flow handle): prepare (scan + prepare when compact), loss, context gradients
        st["launches"] = fh.launches - l0 + (3 if compact_cap is None else 4)
        return st

    def _call_slots(self, params: Tensor, nsig: Tensor, context: Tensor, compact_cap: Optional[int] = None) -> Tensor:
        m = self.model
        B = int(context.shape[0])
        key = B if compact_cap is None else (B, int(compact_cap))
        st = self.slot_graphs.get(key)
        if st is None:
            st = self.slot_graphs[key] = self._capture_slots(B, context.requires_grad, compact_cap)
        with torch.no_grad():
            st["params"].copy_(params[:, :int(m.max_signals), :])
            st["nsig"].copy_(nsig)
            st["context"].copy_(context)
            if self.distributed:
                from .dist import _world
                import torch.distributed as dist
                if _world()[1] > 1:
                    count = nsig.clamp(max=int(m.max_signals)).sum().to(torch.float32)
                    dist.all_reduce(count)
                    st["inv_count"].copy_(1.0 / count.clamp_min(1.0))
                    st["use_inv"].fill_(1.0)
        st["graph"].replay()
        self.last_launches = st["launches"]
        self._assign_flow_grads()
        w = m.rank_embed.weight
        if w.requires_grad:
            if w.grad is None:
                w.grad = st["gemb"].clone()
            else:
                w.grad.add_(st["gemb"])
        if context.requires_grad:
            context.backward(st["gcontext"])
        return st["loss"].detach()

    def _ensure_packed(self):
        """re-pack if an optimizer step changed the weights (the cached parameter list skips NSFPosteriorFlow.handle()'s
        per-call walk over the module tree: ~0.3 ms of host time per step)"""
        flow = self.model.flow
        fh = flow._handle
        if fh is None or fh.device != self.flow_params[0].device:
            return flow.handle()
        fp = self.flat_param
        if fp is not None and fp._version != flow._flat_seen:      # flat mode: the optimizer wrote the flat buffer
            flow._flat_seen = fp._version
            fh.pack_key = None
        fh.ensure_packed(self.flow_params)
        return fh

    def _assign_flow_grads(self):
        if self.flat_param is not None:
            # flat-parameter mode: every step is a fresh gradient; the optimizer owns `flat_param` only
            self.flat_grad.copy_(self.flat_step)
            self.flat_param.grad = self.flat_grad
            if self.flow_params[0].grad is not self._gbuf[0]:
                for p, v in zip(self.flow_params, self._gbuf):
                    p.grad = v
            return
        fresh = [i for i, p in enumerate(self.flow_params) if p.requires_grad and p.grad is None]
        if len(fresh) == len(self.flow_params):
            # the usual step (zero_grad(set_to_none=True) before it): ONE flat copy, then .grad = the views
            self.flat_grad.copy_(self.flat_step)
            for p, v in zip(self.flow_params, self._gbuf):
                p.grad = v
        else:
            accum = [i for i, p in enumerate(self.flow_params) if p.requires_grad and p.grad is not None]
            if fresh:
                torch._foreach_copy_([self._gbuf[i] for i in fresh], [self._step_views[i] for i in fresh])
                for i in fresh:
                    self.flow_params[i].grad = self._gbuf[i]
            if accum:
                torch._foreach_add_([self.flow_params[i].grad for i in accum], [self._step_views[i] for i in accum])

    def _ensure_flat(self):
        """Views of ONE flat buffer, one per flow parameter: the data-parallel all-reduce is a single collective.
        `flat_step` (same layout) receives each replay's gradients inside the graph; `flat_grad` holds what `.grad` points
        at (equal to `flat_step` after a fresh step, the running sum under gradient accumulation)."""
        if self._gbuf is None:
            dev = self.flow_params[0].device
            n = sum(p.numel() for p in self.flow_params)
            self.flat_grad = torch.zeros(n, dtype=torch.float32, device=dev)
            self.flat_step = torch.zeros(n, dtype=torch.float32, device=dev)
            self._gbuf, self._step_views, o = [], [], 0
            for p in self.flow_params:
                self._gbuf.append(self.flat_grad[o:o + p.numel()].view(p.shape))
                self._step_views.append(self.flat_step[o:o + p.numel()].view(p.shape))
                o += p.numel()

    def __call__(self, params: Tensor, nsig: Tensor, context: Tensor) -> Tensor:
        m = self.model
        if context.is_cuda:
            self._ensure_packed()
            if self.compact and not nsig.is_cuda and self.slots is not True and self.model.flow.handle().fused_training:
                # `nsig` still on the HOST (where a data loader leaves it): the row count costs no device sync, so the
                # compact form is free -- and it gives the three streaming kernels (context projection, weight gradients,
                # d ctx) ~1.9 B rows instead of the max_signals x B slots of slot mode (8 tiles instead of 20 at B = 512)
                R = int(nsig.clamp(0, int(m.max_signals)).sum())
                if R > 0:
                    cap = (R + self.bucket - 1) // self.bucket * self.bucket
                    return self._call_slots(params, nsig.to(context.device, non_blocking=True), context, compact_cap=cap)
            if self._use_slots(int(context.shape[0]), context.device):
                return self._call_slots(params, nsig.to(context.device), context)
            if self.compact:
                # rows mode, compact form: the host learns only HOW MANY rows there are (to choose the bucket); which
                # (event, rank) pair each row is gets worked out on the device inside the graph
                nsig_d = nsig.to(context.device)
                R = int(nsig_d.clamp(0, int(m.max_signals)).sum())
                if R > 0:
                    cap = (R + self.bucket - 1) // self.bucket * self.bucket
                    return self._call_slots(params, nsig_d, context, compact_cap=cap)
        event, rank = signal_rows(nsig.to(context.device), int(m.max_signals))
        R = int(event.numel())
        if R == 0:
            # No signal on this rank: contribute zero gradients, but issue the same collectives as every other rank
            # (the count all-reduce here, the flat gradient all-reduce in allreduce()).
            if self.distributed:
                from .dist import _world
                import torch.distributed as dist
                if _world()[1] > 1:
                    dist.all_reduce(torch.zeros((), device=context.device))
            self._ensure_flat()
            fresh = [i for i, p in enumerate(self.flow_params) if p.requires_grad and p.grad is None]
            if fresh:
                torch._foreach_zero_([self._gbuf[i] for i in fresh])
                for i in fresh:
                    self.flow_params[i].grad = self._gbuf[i]
            return context.sum() * 0.0
        # re-pack if an optimizer step changed the weights (the cached parameter list skips NSFPosteriorFlow.handle()'s
        # per-call walk over the module tree: ~0.3 ms of host time per step)
        self._ensure_packed()
        cap = (R + self.bucket - 1) // self.bucket * self.bucket
        st = self.graphs.get(cap)
        if st is None:
            st = self.graphs[cap] = self._capture(cap)
        # LeanNPE._full_context(context[event], rank), spelled with index_select: same values, and the backward is one
        # index_add kernel per operand instead of the sort-based index_put / embedding backward (~25 small kernels)
        ctx_full = torch.cat([context.index_select(0, event), m.rank_embed.weight.index_select(0, rank)], dim=1)
        inv_count = 1.0 / R                                      # local mean (train_lean_npe.py:127) ...
        if self.distributed:                                      # ... or the global one (dist.global_mean_loss)
            from .dist import _world
            import torch.distributed as dist
            if _world()[1] > 1:
                count = torch.tensor(float(R), device=context.device)
                dist.all_reduce(count)
                inv_count = 1.0 / count
        with torch.no_grad():
            st["x"][:R].copy_(m.scaler.normalize(params[event, rank, :]))
            st["ctx"][:R].copy_(ctx_full)
            st["w"].zero_()
            st["w"][:R] = inv_count
        st["graph"].replay()
        self.last_launches = st["launches"]
        # the graph's gradient tensors are overwritten by the next replay: hand out copies (two multi-tensor kernels)
        self._assign_flow_grads()
        if ctx_full.requires_grad:
            ctx_full.backward(st["gctx"][:R])
        return st["loss"].detach()

    def flatten_parameters(self) -> torch.nn.Parameter:
        """Re-home every flow parameter as a view of ONE flat fp32 buffer (the layout of `flat_grad`) and return that
        buffer as a Parameter whose `.grad` is the flat gradient buffer.  An optimizer built over
        `[step.flatten_parameters(), *other_params]` and `step.clip_grad_norm_()` then touch one tensor instead of 180:
        the reference's `clip_grad_norm_(model.parameters(), 5.0)` + `AdamW.step()` cost 0.54 + 0.47 ms of host-bound
        multi-tensor work per step (scripts/prof_opt.py) next to a 1.26 ms flow step; on the flat buffer they are
        0.04 + ~0.05 ms.  AdamW is elementwise, so the update is the same numbers (one weight decay for all tensors, as
        `train_lean_npe.py` builds it).  The module's own parameters stay valid (they alias the buffer: state_dict,
        sampling and `.grad` per tensor all keep working); every step is a fresh gradient (no accumulation across calls
        in this mode).  Call after the model is on its device."""
        if self.flat_param is None:
            self._ensure_flat()
            flat = torch.empty_like(self.flat_grad)
            o = 0
            with torch.no_grad():
                for p in self.flow_params:
                    v = flat[o:o + p.numel()].view(p.shape)
                    v.copy_(p.data)
                    p.data = v
                    o += p.numel()
            self.graphs.clear()              # graphs captured before the move hold the old parameter tensors
            self.slot_graphs.clear()
            self.flat_param = torch.nn.Parameter(flat, requires_grad=True)
            self.flat_param.grad = self.flat_grad
            self._flat_seen = self.flat_param._version
            # NSFPosteriorFlow.handle() re-packs when the flat buffer is written.  object.__setattr__: a plain attribute,
            # NOT a registered parameter of the module (state_dict / named_parameters stay the reference's)
            object.__setattr__(self.model.flow, "_flat_param", self.flat_param)
            object.__setattr__(self.model.flow, "_flat_seen", self.flat_param._version)
        return self.flat_param

    def clip_grad_norm_(self, max_norm: float, extra_params=()) -> Tensor:
        """`torch.nn.utils.clip_grad_norm_` over the flow's flat gradient buffer (+ `extra_params`): same total norm and
        the same scaling, two small kernels for the flow instead of a 180-tensor multi-tensor pass."""
        sq = torch.linalg.vector_norm(self.flat_grad).square()
        extra = [p.grad for p in extra_params if p.grad is not None]
        for gde in extra:
            sq = sq + torch.linalg.vector_norm(gde).square()
        total = sq.sqrt()
        coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
        self.flat_grad.mul_(coef)
        for gde in extra:
            gde.mul_(coef)
        return total

    def allreduce(self, extra_params=()) -> int:
        """Data-parallel gradient exchange after a step: ONE all-reduce (sum) of the flat flow-gradient
        buffer (25 MB fp32) -- no flatten / unflatten copies -- plus one bucket for `extra_params` (rank embedding,
        encoder).  Falls back to the bucketed path if the .grad tensors are not this object's views (e.g. after
        gradient accumulation into foreign tensors).  Returns the number of collectives issued."""
        import torch.distributed as dist
        from .dist import _world, allreduce_gradients
        if _world()[1] == 1:
            return 0
        mine = self._gbuf is not None and all(
            (not p.requires_grad) or p.grad is self._gbuf[i] for i, p in enumerate(self.flow_params))
        if not mine:
            return allreduce_gradients(list(self.flow_params) + list(extra_params))
        dist.all_reduce(self.flat_grad)
        return 1 + allreduce_gradients(list(extra_params))
