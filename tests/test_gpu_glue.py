"""GPU parity of the LeanNPE glue (ParamScaler epilogue, rank embedding, context broadcast, rail mask)
and size-independent properties of the flow at larger sizes."""
import math
import numpy as np
import pytest
import torch

import nsf_oracle as O
from util import check_parity, load_golden, make_flow, rel_err, t

pytestmark = pytest.mark.gpu


def lean_model(w, g, device="cuda", tier="fp32"):
    import posteriflow_b200 as P
    m = P.LeanNPE(tier=tier)
    m.flow.load_state_dict(O.full_state_dict(O.FlowSpec(), w), strict=True)
    with torch.no_grad():
        m.rank_embed.weight.copy_(t(g["rank_embed"]))
    return m.to(device)


def circ_close(a, b, period, tol):
    d = np.abs(a - b)
    return np.minimum(d, np.abs(period - d)) <= tol


def test_sample_posterior_matches_reference():
    spec, w, g = load_golden("lean_glue")
    m = lean_model(w, g)
    E, n = int(g["E"]), int(g["n"])
    ctx = t(g["context"], "cuda")
    z = t(g["z"], "cuda")
    sc, sc64 = O.Scaler(), O.Scaler(dtype=torch.float64)
    w64 = {k: (v.double() if v.is_floating_point() else v) for k, v in w.items()}
    span = (sc.hi - sc.lo).numpy()
    for rank in (0, 2):
        s = m.sample_posterior(None, rank=rank, n_samples=n, context=ctx, z=z).cpu().numpy()
        ref = g[f"samples_rank{rank}"]  # the reference's own LeanNPE.sample_posterior, float32
        assert s.shape == ref.shape == (E, n, 11)
        with torch.no_grad():  # float64 arbiter: the same (reference-pinned) code evaluated in double
            fc = O.full_context(t(g["context"]).double(), t(g["rank_embed"]).double(),
                                torch.full((E,), rank, dtype=torch.long))
            r64 = O.sample_posterior(spec, w64, sc64, fc, t(g["z"]).double(), n).numpy()
        for j in range(11):
            a, b, c = s[..., j], ref[..., j], r64[..., j]
            if bool(sc.circ_mask[j]):  # compared modulo the period: fold onto the reference
                a = a - span[j] * np.round((a - c) / span[j])
                b = b - span[j] * np.round((b - c) / span[j])
            check_parity(f"rank {rank} dim {j}", a, b, c)


def test_lean_nll_matches_reference():
    spec, w, g = load_golden("lean_glue")
    m = lean_model(w, g)
    nll = m.nll(None, t(g["nll_theta"], "cuda"), t(g["nll_ranks"], "cuda"), context=t(g["context"], "cuda"))
    assert rel_err(nll.detach().cpu().numpy(), g["nll"]) <= 1e-4


def test_infer_chunk_rail_mask_and_logq():
    """pipeline.py:168-180: inverse -> wrap -> rail mask -> denormalize; physical log-prob of the samples."""
    import posteriflow_b200 as P
    from posteriflow_b200 import _cabi as A
    spec, w, g = load_golden("lean_glue")
    m = lean_model(w, g)
    n = int(g["n"])
    ctx = m._full_context(t(g["context"], "cuda")[:1], torch.zeros(1, dtype=torch.long, device="cuda"))
    z = t(g["z"], "cuda")[:n]
    theta, _, railed = P.sample_fused(m.flow, m.scaler, ctx, n, z=z)
    ref = g["infer_theta"]
    nonc = ~O.Scaler().circ_mask.numpy()
    assert np.median(np.abs(theta[0].cpu().numpy()[:, nonc] - ref[:, nonc]) / np.maximum(1, np.abs(ref[:, nonc]))) <= 1e-5
    assert abs(float(railed.float().mean()) - float(g["infer_railed"].mean())) <= 1.0 / n + 1e-6
    # normalized samples (wrap only) -> nll -> physical log-prob
    y, _, _ = P.sample_fused(m.flow, m.scaler, ctx, n, z=z,
                             flags=A.PF_FLAG_SANITIZE_CTX | A.PF_FLAG_CLAMP | A.PF_FLAG_WRAP)
    y = y[0]
    neg = m.flow.compute_psd_aware_nll(y, ctx, None, n_per_ctx=n)
    sc = m.scaler
    th = sc.denormalize(y)
    logq = -neg + (np.log(2.0) - torch.log(sc.hi - sc.lo)).sum() - torch.where(
        sc.log_mask, torch.log(th.clamp_min(1e-6)), torch.zeros_like(th)).sum(dim=1)
    # circular wrap may flip a sample by one period at the boundary; compare where y agrees
    yref = O.Scaler().normalize(torch.from_numpy(ref))
    same = (y.cpu() - yref).abs().max(dim=1).values < 1e-3
    assert same.float().mean() > 0.9
    assert rel_err(logq.detach().cpu().numpy()[same.numpy()], g["infer_logq"][same.numpy()]) <= 2e-3


@pytest.mark.parametrize("tier", ["fp32", "bf16"])
def test_posterior_chunk_is_the_infer_loop_body(tier):
    """posterior_chunk = one iteration of infer()'s loop (pipeline.py:168-186) on the device, against the golden
    theta / log q / rail mask the reference code produced (fp32 tier: 1e-5 median on theta, 2e-3 on log q where
    the wrap agrees; tensor-core tier: its stated 5e-3 / 5e-2)."""
    import posteriflow_b200 as P
    spec, w, g = load_golden("lean_glue")
    m = lean_model(w, g, tier=tier)
    n = int(g["n"])
    ctx = m._full_context(t(g["context"], "cuda")[:1], torch.zeros(1, dtype=torch.long, device="cuda"))
    z = t(g["z"], "cuda")[:n]
    theta, logq, railed = P.posterior_chunk(m.flow, m.scaler, ctx, n, z=z, sort_masses=False)
    ref = g["infer_theta"]
    nonc = ~O.Scaler().circ_mask.numpy()
    tol_t, tol_q = (1e-5, 2e-3) if tier == "fp32" else (5e-3, 5e-2)
    err = np.abs(theta[0].cpu().numpy()[:, nonc] - ref[:, nonc]) / np.maximum(1, np.abs(ref[:, nonc]))
    assert np.median(err) <= tol_t
    assert abs(float(railed.float().mean()) - float(g["infer_railed"].mean())) <= (1.0 if tier == "fp32" else 3.0) / n + 1e-6
    yref = O.Scaler().normalize(torch.from_numpy(ref))
    ygot = O.Scaler().normalize(theta[0].cpu())
    if tier == "fp32":
        same = (ygot - yref).abs().max(dim=1).values < 1e-3
        assert same.float().mean() > 0.9
    else:   # bf16 operands: rows whose non-circular dims stayed within 0.1 (normalised units) of the reference
        same = (ygot - yref)[:, torch.from_numpy(nonc)].abs().max(dim=1).values < 0.1
        assert same.float().mean() > 0.6
    eq = np.abs(logq.cpu().numpy() - g["infer_logq"]) / np.maximum(1, np.abs(g["infer_logq"]))
    assert (np.max(eq[same.numpy()]) if tier == "fp32" else np.median(eq[same.numpy()])) <= tol_q
    # mass ordering (pipeline.py:184-186) leaves everything else untouched
    th2, lq2, _ = P.posterior_chunk(m.flow, m.scaler, ctx, n, z=z, sort_masses=True)
    assert torch.equal(lq2, logq) and bool((th2[..., 0] >= th2[..., 1]).all())
    assert torch.equal(th2[..., 2:], theta[..., 2:])
    assert torch.equal(torch.maximum(theta[..., 0], theta[..., 1]), th2[..., 0])


def test_broadcast_equals_dense_context():
    spec, w, g = load_golden("lean_stress")
    flow = make_flow(spec, w, "cuda")
    gen = torch.Generator().manual_seed(3)
    E, n = 5, 37
    ctx = torch.randn(E, spec.context_features, generator=gen).cuda()
    z = (torch.randn(E * n, spec.features, generator=gen) * 1.5).cuda()
    dense = ctx.unsqueeze(1).expand(E, n, -1).reshape(E * n, -1)
    xa, la = flow.inverse(z, dense)
    xb, lb = flow.inverse(z, ctx, n_per_ctx=n)
    assert torch.allclose(xa, xb, atol=1e-6) and torch.allclose(la, lb, atol=1e-5)
    with torch.no_grad():
        za, fa = flow.forward(xa, dense)
        zb, fb = flow.forward(xa, ctx, n_per_ctx=n)
    assert torch.allclose(za, zb, atol=1e-6) and torch.allclose(fa, fb, atol=1e-5)
    # expanded single-event view (pipeline.py:171) is recognised and broadcast
    one = ctx[:1].expand(n, -1)
    xc, _ = flow.inverse(z[:n], one)
    assert torch.allclose(xc, xa[:n], atol=1e-6)


@pytest.mark.parametrize("R", [1, 127, 20000, 50000])
def test_round_trip_inverse_then_forward(R):
    """forward(inverse(z)) == z and logdets cancel, at sizes up to BASELINE config #2 (1 event x 50k)."""
    spec, w, g = load_golden("lean_default")
    flow = make_flow(spec, w, "cuda")
    from posteriflow_b200 import _cabi as A
    gen = torch.Generator().manual_seed(R)
    ctx = torch.randn(1, spec.context_features, generator=gen).cuda()
    z = torch.randn(R, spec.features, generator=gen).cuda()
    fh = flow.handle()
    x, ld = torch.ops.posteriflow.nsf_inverse(z, ctx, fh.id, R, 0, None, None, None, None)  # no clamp
    with torch.no_grad():
        z2, ld2 = flow.forward(x, ctx, n_per_ctx=R)
    # the inverse is ill-conditioned where a bin is steep, so the maximum over 11*R draws grows with R;
    # the bulk must close to float32 round-off
    # (log-dets are sums of 110 terms of magnitude O(1): compare relative to max(1, |logdet|))
    dz, dl = (z2 - z).abs().flatten(), (ld + ld2).abs() / ld.abs().clamp_min(1.0)
    assert float(dz.median()) <= 1e-5 and float(dl.median()) <= 5e-5
    if R >= 1000:
        assert float(torch.quantile(dz[:1000000], 0.999)) <= 2e-4
    assert float(dz.max()) <= 2e-2 and float(dl.max()) <= 2e-2


def test_empty_and_error_paths():
    spec, w, g = load_golden("tiny_explicit")
    flow = make_flow(spec, w, "cuda")
    z = torch.empty(0, spec.features, device="cuda")
    ctx = torch.empty(0, spec.context_features, device="cuda")
    x, ld = flow.inverse(z, ctx)
    assert x.shape == (0, spec.features) and ld.shape == (0,)
    with torch.no_grad():
        zz, l2 = flow.forward(z, ctx)
    assert zz.shape == (0, spec.features)
    with pytest.raises(Exception):
        flow.inverse(torch.zeros(4, spec.features, device="cuda"), torch.zeros(3, spec.context_features, device="cuda"))
    with pytest.raises(RuntimeError):
        flow.inverse(torch.zeros(4, spec.features), torch.zeros(4, spec.context_features, device="cuda"))


def test_philox_noise_is_standard_normal_and_reproducible():
    import posteriflow_b200 as P
    from posteriflow_b200 import _cabi as A
    spec, w, g = load_golden("lean_default")
    flow = make_flow(spec, w, "cuda")
    sc = P.ParamScaler().to("cuda")
    ctx = torch.zeros(2, spec.context_features, device="cuda")
    a = P.sample_fused(flow, sc, ctx, 4096, z=None, seed=11)[0]
    b = P.sample_fused(flow, sc, ctx, 4096, z=None, seed=11)[0]
    c = P.sample_fused(flow, sc, ctx, 4096, z=None, seed=12)[0]
    assert torch.equal(a, b) and not torch.equal(a, c)
    # push the raw flow output back through the density direction: z must be ~N(0, I)
    y = P.sample_fused(flow, sc, ctx, 4096, z=None, seed=11, flags=0)[0].reshape(-1, spec.features)
    with torch.no_grad():
        z, _ = flow.forward(y, ctx, n_per_ctx=4096)
    assert abs(float(z.mean())) < 0.02 and abs(float(z.std()) - 1.0) < 0.02
    assert abs(float((z ** 4).mean()) - 3.0) < 0.2


def test_batch_nll_fused_matches_loop_and_oracle():
    """N1: one flow call for all signal ranks == the reference's per-rank loop (train_lean_npe.py:108-127)."""
    import posteriflow_b200 as P
    from posteriflow_b200.train import batch_nll_reference_order
    spec, w, g = load_golden("lean_glue")
    m = lean_model(w, g)
    gen = torch.Generator().manual_seed(11)
    B = 48
    nsig = torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), B, replacement=True, generator=gen) + 1
    sc = O.Scaler()
    params = sc.denormalize(torch.rand(B, 5, 11, generator=gen) * 2 - 1)
    ctx = torch.randn(B, 256, generator=gen)
    c1 = ctx.cuda().requires_grad_(True)
    fused = P.batch_nll(m, None, params.cuda(), nsig.cuda(), context=c1)
    l0 = m.flow.gpu_launches
    fused.backward()
    c2 = ctx.cuda().requires_grad_(True)
    loop = batch_nll_reference_order(m, None, params.cuda(), nsig.cuda(), context=c2)
    loop.backward()
    assert m.flow.gpu_launches > l0
    # oracle: the reference loop on the CPU
    tot, cnt = 0.0, 0
    with torch.no_grad():
        for r in range(int(nsig.max())):
            idx = (nsig > r).nonzero(as_tuple=True)[0]
            fc = O.full_context(ctx[idx], t(g["rank_embed"]), torch.full((idx.numel(),), r, dtype=torch.long))
            tot += float(O.flow_nll(spec, w, sc.normalize(params[idx, r, :]), fc).double().sum())
            cnt += idx.numel()
    ref = tot / cnt
    assert abs(float(fused) - ref) <= 1e-4 * max(1.0, abs(ref)), (float(fused), ref)
    assert abs(float(loop) - ref) <= 1e-4 * max(1.0, abs(ref))
    scale = float(c2.grad.abs().max())
    assert float((c1.grad - c2.grad).abs().max()) <= 1e-4 * scale


@pytest.mark.parametrize("tier", ["fp32", "bf16", "bf16-rows", "bf16-rows-aten"])
def test_graphed_nll_step_matches_batch_nll(tier, monkeypatch):
    """GraphedNLLStep (one CUDA-graph replay per step, rows padded to a bucket with zero weight) == batch_nll +
    backward through autograd: same loss, same gradient of the context, the rank embedding and every flow weight;
    replaying the cached graph on a second batch with a different number of rows in the same bucket stays exact,
    and a second call without clearing .grad accumulates."""
    import posteriflow_b200 as P
    # "bf16": the fused training kernels in SLOT mode (every (event, rank) slot with weight [rank < nsig] / count, gather /
    # scaler / embedding and their backward inside the graph); "bf16-rows": the same kernels over the gathered rows
    # (compact form: pf_rows_prepare / pf_rows_ctx_grad inside the graph); "bf16-rows-aten": the gather done by ATen outside it
    slots = None if not tier.startswith("bf16-rows") else False
    compact = tier != "bf16-rows-aten"
    tier = "bf16" if tier.startswith("bf16-rows") else tier
    if tier == "bf16":       # capture the split-bf16 tcgen05 GEMMs too (by default they serve calls of >= 512 rows)
        monkeypatch.setenv("PF_TRAIN_TC_MIN_ROWS", "0")
    spec, w, g = load_golden("lean_glue")
    m = lean_model(w, g, tier=tier)
    sc = O.Scaler()
    step = P.GraphedNLLStep(m, bucket=64, slots=slots, compact=compact)
    assert step._use_slots(40, torch.device("cuda")) == (tier == "bf16" and slots is None)
    gen = torch.Generator().manual_seed(12)
    tol = 1e-4 if tier == "fp32" else 5e-4
    for trial in range(3):
        B = 40
        nsig = torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), B, replacement=True, generator=gen) + 1
        params = sc.denormalize(torch.rand(B, 5, 11, generator=gen) * 2 - 1).cuda()
        ctx = torch.randn(B, 256, generator=gen)
        m.zero_grad(set_to_none=True)
        c1 = ctx.cuda().requires_grad_(True)
        ref = P.batch_nll(m, None, params, nsig.cuda(), context=c1)
        ref.backward()
        ref_g = {n: p.grad.clone() for n, p in m.named_parameters() if p.grad is not None}
        m.zero_grad(set_to_none=True)
        c2 = ctx.cuda().requires_grad_(True)
        loss = step(params, nsig.cuda(), c2)
        assert abs(float(loss) - float(ref)) <= tol * max(1.0, abs(float(ref)))
        assert float((c2.grad - c1.grad).abs().max()) <= tol * float(c1.grad.abs().max())
        got_g = {n: p.grad for n, p in m.named_parameters() if p.grad is not None}
        assert set(got_g) == set(ref_g)
        for n in ref_g:
            assert float((got_g[n] - ref_g[n]).abs().max()) <= tol * max(1e-12, float(ref_g[n].abs().max())), n
        if trial == 2:      # accumulate: a second call on the same batch doubles the flow-weight gradients
            c3 = ctx.cuda().requires_grad_(True)
            step(params, nsig.cuda(), c3)
            n0 = "flow.transform._transforms.1.autoregressive_net.final_layer.weight"
            assert float((m.get_parameter(n0).grad - 2 * ref_g[n0]).abs().max()) <= 2 * tol * float(ref_g[n0].abs().max())
    assert len(step.graphs) + len(step.slot_graphs) <= 2
    assert bool(step.graphs) == (not compact)


@pytest.mark.parametrize("B", [300, 2500])
def test_compact_rows_form_with_empty_events_and_a_long_prefix_sum(B):
    """The compact rows form (prefix sum of nsig -> (event, rank) of every row on the device) against the ATen gather path on
    a batch with events of 0 signals, of more than max_signals (clamped), and B beyond one element per scan thread."""
    import posteriflow_b200 as P
    spec, w, g = load_golden("lean_glue")
    m = lean_model(w, g, tier="bf16")
    sc = O.Scaler()
    gen = torch.Generator().manual_seed(B)
    nsig = torch.randint(0, 8, (B,), generator=gen)               # 0 .. 7 with max_signals = 5
    nsig[0], nsig[-1] = 0, 0
    params = sc.denormalize(torch.rand(B, 5, 11, generator=gen) * 2 - 1).cuda()
    ctx = torch.randn(B, 256, generator=gen)
    out = {}
    for compact in (False, True):
        step = P.GraphedNLLStep(m, slots=False, compact=compact)
        m.zero_grad(set_to_none=True)
        c = ctx.cuda().requires_grad_(True)
        loss = step(params, nsig.cuda(), c)
        out[compact] = (float(loss), c.grad.clone(), m.rank_embed.weight.grad.clone(),
                        [p.grad.clone() for p in m.flow.layer_weights()])
    a, b = out[False], out[True]
    assert abs(a[0] - b[0]) <= 1e-5 * max(1.0, abs(a[0]))
    assert float((a[1] - b[1]).abs().max()) <= 1e-5 * float(a[1].abs().max())
    assert float((a[1][nsig == 0]).abs().max()) == 0.0 and float((b[1][nsig.cuda() == 0]).abs().max()) == 0.0
    assert float((a[2] - b[2]).abs().max()) <= 1e-5 * float(a[2].abs().max())
    for ga, gb in zip(a[3], b[3]):      # row order differs (rank-major vs event-major): atomics / tile sums in another order
        assert float((ga - gb).abs().max()) <= 2e-5 * max(1e-12, float(ga.abs().max()))


@pytest.mark.parametrize("B,S", [(1, 1), (7, 5), (1030, 5), (300, 3)])
def test_glue_entry_points_against_torch(B, S):
    """pf_slot_prepare / pf_rows_prepare / pf_slot_loss / pf_slot_ctx_grad / pf_rows_ctx_grad called through the C-ABI on raw
    buffers, against the torch expressions they replace (train_lean_npe.py:108-127, lean_npe.py:100-109, 296-304): slot and
    compact layouts, events without a signal, nsig beyond S, padding rows, both normalisations (local count / external)."""
    import ctypes as C_
    from posteriflow_b200 import _cabi as A
    from posteriflow_b200.ops import _scaler_struct
    lib = A.load_library()
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(B * 10 + S)
    sc = O.Scaler()
    D, cd, Cc = 11, 256, 288
    nsig = torch.randint(0, S + 3, (B,), generator=g)
    if B > 2:
        nsig[0] = 0
    if int(nsig.clamp(max=S).sum()) == 0:
        nsig[-1] = 1
    n = nsig.clamp(max=S)
    params = sc.denormalize(torch.rand(B, S, D, generator=g) * 2 - 1)
    ctx = torch.randn(B, cd, generator=g)
    emb = torch.randn(S + 1, Cc - cd, generator=g)               # one spare row, as LeanNPE's embedding may have
    pd, nd, cdv, ed = params.to(dev), nsig.to(dev), ctx.to(dev), emb.to(dev)
    sstruct, keep = _scaler_struct(sc.lo.to(dev), sc.hi.to(dev), sc.log_mask.to(dev), sc.circ_mask.to(dev))
    inv_ext = torch.tensor(0.125, device=dev)
    xn = sc.normalize(params)                                   # [B, S, D]
    R = int(n.sum())
    for use_ext in (0.0, 1.0):
        use = torch.tensor(use_ext, device=dev)
        inv = 0.125 if use_ext else 1.0 / max(1, R)
        # ---- slot layout ----
        x = torch.full((B * S, D), 7.0, device=dev)
        cf = torch.full((B * S, Cc), 7.0, device=dev)
        w = torch.full((B * S,), 7.0, device=dev)
        A.check(lib.pf_slot_prepare(pd.data_ptr(), nd.data_ptr(), cdv.data_ptr(), ed.data_ptr(), C_.byref(sstruct),
                                    inv_ext.data_ptr(), use.data_ptr(), B, S, D, cd, Cc, x.data_ptr(), cf.data_ptr(),
                                    w.data_ptr(), 0, None), "pf_slot_prepare")
        present = (torch.arange(S)[None, :] < n[:, None])        # [B, S]
        assert torch.allclose(x.cpu().view(B, S, D), xn * present[..., None], atol=2e-6)
        assert torch.equal(cf.cpu().view(B, S, Cc)[..., :cd], ctx[:, None, :].expand(B, S, cd))
        assert torch.equal(cf.cpu().view(B, S, Cc)[..., cd:], emb[None, :S, :].expand(B, S, Cc - cd))
        assert torch.allclose(w.cpu().view(B, S), present.float() * inv, rtol=1e-6, atol=0)
        # ---- compact layout, padded to a bucket ----
        cap = (R + 127) // 128 * 128
        offs = torch.zeros(B + 1, dtype=torch.int32, device=dev)
        x2 = torch.full((cap, D), 7.0, device=dev)
        cf2 = torch.zeros(cap, Cc, device=dev)
        w2 = torch.full((cap,), 7.0, device=dev)
        A.check(lib.pf_rows_prepare(pd.data_ptr(), nd.data_ptr(), cdv.data_ptr(), ed.data_ptr(), C_.byref(sstruct),
                                    inv_ext.data_ptr(), use.data_ptr(), B, S, D, cd, Cc, cap, offs.data_ptr(), x2.data_ptr(),
                                    cf2.data_ptr(), w2.data_ptr(), 0, None), "pf_rows_prepare")
        ref_offs = torch.cat([torch.zeros(1, dtype=torch.long), n.cumsum(0)])
        assert torch.equal(offs.cpu().long(), ref_offs)
        ev = torch.repeat_interleave(torch.arange(B), n)
        rk = torch.cat([torch.arange(int(k)) for k in n]) if R else torch.zeros(0, dtype=torch.long)
        assert torch.allclose(x2.cpu()[:R], xn[ev, rk], atol=2e-6) and float(x2[R:].abs().max() if cap > R else 0) == 0.0
        assert torch.equal(cf2.cpu()[:R], torch.cat([ctx[ev], emb[rk]], dim=1))
        assert torch.allclose(w2.cpu()[:R], torch.full((R,), inv), rtol=1e-6, atol=0)
        assert cap == R or float(w2[R:].abs().max()) == 0.0
        # ---- loss + first backward step on the compact rows ----
        z = torch.randn(cap, D, generator=g).to(dev)
        ld = torch.randn(cap, generator=g).to(dev)
        loss = torch.zeros(1, device=dev)
        gz = torch.empty(cap, D, device=dev)
        gl = torch.empty(cap, device=dev)
        A.check(lib.pf_slot_loss(z.data_ptr(), ld.data_ptr(), w2.data_ptr(), cap, D, loss.data_ptr(), gz.data_ptr(),
                                 gl.data_ptr(), 0, None), "pf_slot_loss")
        nll = 0.5 * (z.double() ** 2).sum(1) + 0.5 * D * math.log(2 * math.pi) - ld.double()
        ref_loss = float((nll * w2.double()).sum())
        assert abs(float(loss) - ref_loss) <= 1e-5 * max(1.0, abs(ref_loss))
        assert torch.allclose(gz, z * w2[:, None]) and torch.equal(gl, -w2)
        # ---- context / embedding gradients: compact and slot ----
        gfull = torch.randn(cap, Cc, generator=g).to(dev)
        gfull[R:] = 0
        gc = torch.full((B, cd), 7.0, device=dev)
        ge = torch.full((S, Cc - cd), 7.0, device=dev)
        A.check(lib.pf_rows_ctx_grad(gfull.data_ptr(), offs.data_ptr(), B, S, cd, Cc, gc.data_ptr(), ge.data_ptr(), 0, None),
                "pf_rows_ctx_grad")
        ref_gc = torch.zeros(B, cd, dtype=torch.float64).index_add_(0, ev, gfull.cpu()[:R, :cd].double())
        ref_ge = torch.zeros(S, Cc - cd, dtype=torch.float64).index_add_(0, rk, gfull.cpu()[:R, cd:].double())
        assert torch.allclose(gc.cpu().double(), ref_gc, atol=1e-5) and torch.allclose(ge.cpu().double(), ref_ge, atol=2e-4)
        gslot = torch.randn(B * S, Cc, generator=g).to(dev)
        A.check(lib.pf_slot_ctx_grad(gslot.data_ptr(), B, S, cd, Cc, gc.data_ptr(), ge.data_ptr(), 0, None), "pf_slot_ctx_grad")
        assert torch.allclose(gc.cpu().double(), gslot.cpu().double().view(B, S, Cc)[..., :cd].sum(1), atol=1e-5)
        assert torch.allclose(ge.cpu().double(), gslot.cpu().double().view(B, S, Cc)[..., cd:].sum(0), atol=2e-4)
    torch.cuda.synchronize()


def test_host_nsig_takes_the_compact_form_and_matches_slot_mode():
    """`nsig` left on the host (data loader): GraphedNLLStep takes the compact rows form without a device sync -- same loss
    and gradients as the slot-mode step the device tensor gets."""
    import posteriflow_b200 as P
    spec, w, g = load_golden("lean_glue")
    m = lean_model(w, g, tier="bf16")
    sc = O.Scaler()
    gen = torch.Generator().manual_seed(21)
    B = 200
    nsig = torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), B, replacement=True, generator=gen) + 1
    nsig[3] = 0
    params = sc.denormalize(torch.rand(B, 5, 11, generator=gen) * 2 - 1).cuda()
    ctx = torch.randn(B, 256, generator=gen)
    out = {}
    for where in ("device", "host"):
        step = P.GraphedNLLStep(m)
        m.zero_grad(set_to_none=True)
        c = ctx.cuda().requires_grad_(True)
        loss = step(params, nsig.cuda() if where == "device" else nsig.pin_memory(), c)
        keys = list(step.slot_graphs)
        assert len(keys) == 1 and isinstance(keys[0], tuple) == (where == "host")      # (B, cap) = compact, B = slots
        out[where] = (float(loss), c.grad.clone(), m.rank_embed.weight.grad.clone(),
                      [p.grad.clone() for p in m.flow.layer_weights()])
    a, b = out["device"], out["host"]
    assert abs(a[0] - b[0]) <= 1e-5 * max(1.0, abs(a[0]))
    assert float((a[1] - b[1]).abs().max()) <= 1e-5 * float(a[1].abs().max())
    assert float((a[2] - b[2]).abs().max()) <= 1e-5 * float(a[2].abs().max())
    for ga, gb in zip(a[3], b[3]):
        assert float((ga - gb).abs().max()) <= 2e-5 * max(1e-12, float(ga.abs().max()))
