#!/usr/bin/env python
"""bench.py -- headline benchmark of the LeanNPE posterior-flow hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--tier fp32|bf16]

Metric (BASELINE.json): NSF posterior samples/s.  One step = one pass of the fused sampler
(lean_npe.py:326-332: context broadcast -> base noise -> inverse flow -> wrap -> denormalize) over BASELINE
config #3 -- 4096 synthetic events x 5000 samples, rank 0 -- with the events sharded contiguously over the
N GPUs (4096 / N events per GPU: STRONG scaling of the fixed problem; at N = 1 the whole 2.048e7-row problem
runs on one GPU), no data-path collective.  Weights are random-init (seeded), contexts synthetic.  The other
BASELINE configs are reported in the same line: "configs" (#2 one event x 50 000 samples as sample blocks, #5
rank-conditioned sampling through one launch, on a stated sub-sample) and "train" (#4: NLL training step of
the flow head, per-GPU shard and un-sharded, with and without clip + AdamW + re-pack).
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

TOTAL_EVENTS = 4096                 # BASELINE config #3
SAMPLES_PER_EVENT = 5000
CFG5_EVENTS_PER_GPU = 160           # config #5 sub-sample: 160 of the 1250 events a GPU holds at 8 GPUs (x 3 ranks x 10^4)
TRAIN_EVENTS_PER_GPU = 512          # config #4: batch 4096 over 8 GPUs
F_X = 7946240                       # FLOP per flow row, dense single conditioner pass x 10 layers
F_CTX = 4423680                     # FLOP per distinct context row


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.th.join(timeout=2)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit())
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        mx = max(int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower() == "active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": mx, "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's own algorithm (PyTorch eager over nflows semantics) as restated by oracle/
# ------------------------------------------------------------------------------------------------
def cpu_sampler(rows_per_call=4096):
    """Returns (fn(n_rows) -> seconds, cores).  One call = flow.inverse on `n_rows` samples of one event
    in chunks of 4096 (pipeline.py:105), + wrap + denormalize, exactly the reference's op sequence."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import nsf_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = O.FlowSpec()
    w = O.init_weights(spec, seed=0)
    sc = O.Scaler()
    g = torch.Generator().manual_seed(1)
    ctx = torch.randn(1, spec.context_features, generator=g)

    def run(n_rows):
        z = torch.randn(n_rows, spec.features, generator=g)
        t0 = time.perf_counter()
        with torch.no_grad():
            O.sample_posterior(spec, w, sc, ctx, z, n_rows, chunk=rows_per_call)
        return time.perf_counter() - t0
    return run, cores


def cpu_measure(budget_s=12.0):
    run, cores = cpu_sampler()
    run(256)                       # warm-up (thread pool, allocator)
    t = run(1024)
    n = int(max(1024, min(65536, 1024 * budget_s / max(t, 1e-3))))
    n = n // 1024 * 1024
    dt = run(n)
    return {"value": n / dt, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": f"{n} samples of 1 event (chunks of 4096), LeanNPE shape, fp32, oracle/nsf_oracle.py "
                      f"(reference op sequence: 11 conditioner passes per layer), {dt:.1f} s"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    run, cores = cpu_sampler()
    run(256)
    t = run(512)
    per_step = int(max(512, min(16384, 512 * (60.0 / max(1, args.steps + args.warmup)) / max(t, 1e-3)))) // 512 * 512
    for _ in range(args.warmup):
        run(per_step)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run(per_step)
    dt = time.perf_counter() - t0
    v = per_step * args.steps / dt
    sample = (f"{per_step} samples/step of 1 event on {cores} host threads; oracle port of the reference's PyTorch "
              f"path (nflows is not installable here, SURVEY.md 0.3)")
    print(json.dumps({
        "impl": "reference", "metric": "posterior_samples_per_sec", "value": v, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(workload_config(args.gpus), sample="SUB-SAMPLE, EXTRAPOLATED: the CPU arm runs " + sample +
                       "; its samples/s is quoted for the config above assuming linear scaling in the row count"),
        "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def workload_config(n_gpus):
    ev = TOTAL_EVENTS // n_gpus
    return {"workload": f"BASELINE config #3: {TOTAL_EVENTS} events x {SAMPLES_PER_EVENT} samples, rank 0 "
                        f"({ev} events = {ev * SAMPLES_PER_EVENT} flow rows per GPU), LeanNPE flow 10x256x16 bins, "
                        "11-D, 288-D context",
            "events_per_gpu": ev, "samples_per_event": SAMPLES_PER_EVENT,
            "parallelism": f"events sharded over {n_gpus} GPU(s), no data-path collective",
            "l2": "L2 flushed (256 MB write) between timed steps; every step also streams its own outputs "
                  f"({ev * SAMPLES_PER_EVENT * 49 / 1e6:.0f} MB per GPU) through L2; the 5.5 MB bf16 weight image is "
                  "re-read from L2 by design"}


def ncu_traffic(rows_per_launch):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE tc_sampler_kernel launch at exactly this launch shape, from
    a committed `ncu --set full` summary (profiles/r02_sampler_traffic.json: {"<rows per launch>": bytes}), else None:
    a number captured at another shape is not this launch's traffic."""
    p = os.path.join(ROOT, "profiles", "r02_sampler_traffic.json")
    if os.path.exists(p):
        v = json.load(open(p)).get(str(int(rows_per_launch)))
        if v is not None:
            return float(v["dram_bytes"]), v.get("note", "")
    return None, "no ncu capture at this launch shape is committed"


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch.distributed as dist
    import posteriflow_b200 as P
    from posteriflow_b200 import _cabi as A
    from posteriflow_b200.dist import allreduce_gradients
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        tt = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    torch.manual_seed(0)
    model = P.LeanNPE(tier=args.tier).to(dev)      # random-init weights, identical on every rank
    flow, scaler = model.flow, model.scaler
    E = args.events if args.events is not None else TOTAL_EVENTS // world
    n, D = args.samples, 11
    R = E * n
    g = torch.Generator().manual_seed(100 + rank)
    ctx_host = torch.randn(E, 256, generator=g).pin_memory()   # stands in for encoder output (encoder stays PyTorch)
    ctx_dev = ctx_host.to(dev)
    rank_idx = torch.zeros(E, dtype=torch.long, device=dev)
    with torch.no_grad():
        ctx_full = model._full_context(ctx_dev, rank_idx)
    theta_host = torch.empty(E, n, D, dtype=torch.float32).pin_memory()
    fh = flow.handle()

    def step_device(i):
        return P.sample_fused(flow, scaler, ctx_full, n, z=None, seed=1000 + i)

    def step_e2e(i):
        # pinned host context in, pinned host theta out: chunk i's device-to-host copy runs under chunk i + 1's sampling
        model.sample_posterior_host(ctx_host, rank=0, n_samples=n, out=theta_host)

    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def timed(fn, steps, warmup):
        """Device time of exactly `steps` steps: one CUDA-event pair per step on the launching stream, L2
        flushed (256 MB write) between steps outside the pairs, barrier + synchronize on both sides."""
        for i in range(warmup):
            fn(i)
        barrier()
        l0 = fh.launches
        pairs = []
        for i in range(steps):
            flush_buf.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn(warmup + i)
            e1.record()
            pairs.append((e0, e1))
        barrier()
        return max_over_ranks(sum(a.elapsed_time(b) for a, b in pairs)), fh.launches - l0

    with ClockSampler(local) as clk:
        ms, launches = timed(step_device, args.steps, args.warmup)
    clocks = clk.summary()
    ms_e2e, _ = timed(step_e2e, args.steps, max(1, args.warmup // 2))

    # log-prob (density direction) of the same shard, the fused infer() loop body (N2) and the on-device summaries (N3)
    wrap_flags = A.PF_FLAG_SANITIZE_CTX | A.PF_FLAG_CLAMP | A.PF_FLAG_WRAP
    with torch.no_grad():
        y_norm = P.sample_fused(flow, scaler, ctx_full, n, z=None, seed=7, flags=wrap_flags)[0].reshape(R, D)
        theta_dev, _, railed_dev = P.sample_fused(flow, scaler, ctx_full, n, z=None, seed=7)
    truth_dev = theta_dev[:, 0, :].contiguous()

    def step_logprob(i):
        with torch.no_grad():
            return flow.compute_psd_aware_nll(y_norm, ctx_full, None, n_per_ctx=n)

    def step_infer(i):
        return P.posterior_chunk(flow, scaler, ctx_full, n, z=None, seed=2000 + i)

    def step_summary(i):
        return P.posterior_summary(theta_dev, truth=truth_dev, railed=railed_dev)

    ms_lp, launches_lp = timed(step_logprob, args.steps, args.warmup)
    ms_inf, launches_inf = timed(step_infer, args.steps, args.warmup)
    ms_sum, _ = timed(step_summary, args.steps, args.warmup)

    # ---- the tier that meets the FP32 tolerance (north_star: |d theta|, |d log q| <= 1e-4 relative) ----
    # tier="bf16x3": log-prob on the split-operand dense tcgen05 kernel; its sampling direction runs on the FP32-FMA
    # kernels, timed on a bounded sub-sample
    fp32acc = None
    if args.tier == "bf16":
        m3 = P.LeanNPE(tier="bf16x3").to(dev)
        m3.load_state_dict(model.state_dict())
        E3 = min(E, 512)
        R3 = E3 * n
        y3, c3 = y_norm[:R3], ctx_full[:E3]
        Es = min(E, 16)

        def step_lp3(i):
            with torch.no_grad():
                return m3.flow.compute_psd_aware_nll(y3, c3, None, n_per_ctx=n)

        def step_s3(i):
            return P.sample_fused(m3.flow, scaler, ctx_full[:Es], n, z=None, seed=5000 + i)

        fh3 = m3.flow.handle()
        ms_lp3, _ = timed(step_lp3, args.steps, args.warmup)
        ms_s3, _ = timed(step_s3, max(2, args.steps // 2), 3)
        lp3_rows = R3 * args.steps / (ms_lp3 * 1e-3)
        fp32acc = {"tier": "bf16x3 (split fp16 operands: 3 MMAs per product, exact fp32 spline); parity at 1e-4 in "
                           "tests/test_gpu_round2.py::test_split_tier_density_direction_meets_the_fp32_tolerance",
                   "log_prob": {"value": lp3_rows * world, "unit": "rows/s", "rows_per_gpu": R3,
                                "ms_per_step": ms_lp3 / args.steps,
                                "roofline": {"bound": "tensor", "achieved": lp3_rows * (F_X + F_CTX / n) / 1e12,
                                             "peak": float(peaks()[0]["bf16_tflops_sustained"]) / 3.0, "unit": "TFLOP/s",
                                             "frac": lp3_rows * (F_X + F_CTX / n) / 1e12 / (float(peaks()[0]["bf16_tflops_sustained"]) / 3.0),
                                             "frac_burst": lp3_rows * (F_X + F_CTX / n) / 1e12 / (float(peaks()[0]["bf16_tflops"]) / 3.0),
                                             "note": "peak = one third of the bf16 peak: three tensor instructions per product"}},
                   "sampler": {"value": Es * n * world * max(2, args.steps // 2) / (ms_s3 * 1e-3), "unit": "samples/s",
                               "sample": f"{Es} events x {n} samples per GPU (sub-sample)",
                               "kernels": "FP32-FMA tier kernels (launch-bound): the stepwise tensor-core sampler has no "
                                          "split-operand variant -- ten activation arrays of a 128-row tile do not fit on chip"}}
        del m3, fh3

    # ---- BASELINE config #2: ONE event x 50 000 samples, sample blocks over the ranks (dist.shard_sampling), per-rank
    # disjoint Philox keys; at 8 GPUs a 49-tile launch per GPU -- the latency regime of single-event serving ----
    from posteriflow_b200 import dist as PD
    c2_ctx = torch.randn(1, 256, generator=torch.Generator().manual_seed(5)).to(dev)     # same event on every rank
    (_, _), (s0, s1) = PD.shard_sampling(1, 50000, rank, world)

    def step_cfg2(i):
        return PD.sample_posterior_sharded(model, c2_ctx, 0, 50000, seed=3000 + i, rank_world=(rank, world))

    ms_c2, launches_c2 = timed(step_cfg2, args.steps, args.warmup)
    # ---- BASELINE config #5: rank-conditioned sampling, events x ranks 0-2 x 10^4 samples through ONE launch
    # (P.sample_ranks); bounded sub-sample of a GPU's 1250-event shard, throughput is row-count invariant ----
    c5_ev = args.cfg5_events
    c5_ctx = torch.randn(c5_ev, 256, generator=g).to(dev)

    def step_cfg5(i):
        return P.sample_ranks(model, c5_ctx, 3, 10000, seed=4000 + i)

    ms_c5, launches_c5 = timed(step_cfg5, max(2, args.steps // 2), 3)
    c5_steps = max(2, args.steps // 2)
    c5_rows = c5_ev * 3 * 10000

    # NLL training step of the flow head (config #4 shard): forward + backward (+ gradient all-reduce)
    Bt = args.train_events
    nsig = torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), Bt, replacement=True, generator=g) + 1
    ev = torch.repeat_interleave(torch.arange(Bt), nsig)
    rk = torch.cat([torch.arange(int(k)) for k in nsig])
    Rt = int(ev.numel())
    sc_cpu = P.ParamScaler()
    params_b = sc_cpu.denormalize(torch.rand(Bt, 5, D, generator=g) * 2 - 1).to(dev)   # [B, max_signals, 11] physical
    nsig_dev = nsig.to(dev)
    nsig_pin = nsig.clone().pin_memory()
    tctx = torch.randn(Bt, 256, generator=g).to(dev).requires_grad_(True)
    ev, rk = ev.to(dev), rk.to(dev)
    params = [p for p in model.parameters() if p.requires_grad]

    gstep = P.GraphedNLLStep(model, distributed=(world > 1))
    flow_ids = {id(p) for p in gstep.flow_params}
    other_params = [p for p in params if id(p) not in flow_ids]      # rank embedding
    # AdamW with lr = 0: the full optimizer arithmetic runs and every weight tensor is written in place (its version
    # counter moves, so the NEXT step re-packs exactly as after a real update), but the benchmarked weights -- and with
    # them every other number of this line -- stay those of the seeded random init.
    opt = torch.optim.AdamW(params, lr=0.0, weight_decay=0.0, fused=True)
    counted = {"kernels": 0}

    def step_train(i):
        """public training API: batch_nll forward + backward as one CUDA-graph replay (+ gradient all-reduce)"""
        for p in params:
            p.grad = None
        tctx.grad = None
        loss = gstep(params_b, nsig_pin, tctx)        # nsig as the data loader leaves it (pinned host): compact rows form
        counted["kernels"] += gstep.last_launches
        if world > 1:
            gstep.allreduce(extra_params=other_params)
        return loss

    def step_train_opt(i):
        """what train_lean_npe.py:357-369 does around the flow: the step above + clip_grad_norm_(5.0) + AdamW.step();
        the re-pack the updated weights need is paid by the next step's gstep() call, inside the timed region"""
        loss = step_train(i)
        torch.nn.utils.clip_grad_norm_(params, 5.0)
        opt.step()
        return loss

    def step_train_eager(i):
        for p in params:
            p.grad = None
        tctx.grad = None
        loss = P.batch_nll(model, None, params_b, nsig_dev, context=tctx)   # N1: one flow call for all ranks
        loss.backward()
        if world > 1:
            allreduce_gradients(params)
        return loss

    counted["kernels"] = 0
    ms_train, _ = timed(step_train, args.steps, args.warmup)
    launches_train = counted["kernels"] * args.steps // (args.steps + args.warmup)   # counted per call, timed calls only
    kernels_shard = int(gstep.last_launches)        # library kernels of one replay of THIS step (later steps overwrite it)
    l0 = fh.launches
    ms_train_opt, pack_launches = timed(step_train_opt, args.steps, args.warmup)
    ms_train_eager, _ = timed(step_train_eager, args.steps, args.warmup)

    def step_repack(i):
        with torch.no_grad():
            gstep.flow_params[0].add_(0.0)          # bump one version counter: the next handle() call re-packs
        flow.handle()

    ms_pack, launches_pack = timed(step_repack, args.steps, args.warmup)

    # config #4 un-sharded (4096 events on ONE GPU): reported at N = 1 only, next to the per-GPU shard above
    full = None
    if world == 1 and args.train_events_full > 0:
        Bf = args.train_events_full
        nsig_f = torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), Bf, replacement=True, generator=g) + 1
        params_f = sc_cpu.denormalize(torch.rand(Bf, 5, D, generator=g) * 2 - 1).to(dev)
        nsig_f_dev = nsig_f.to(dev)
        fctx = torch.randn(Bf, 256, generator=g).to(dev).requires_grad_(True)

        def step_train_full(i):
            for p in params:
                p.grad = None
            fctx.grad = None
            return gstep(params_f, nsig_f_dev, fctx)

        def step_train_full_opt(i):
            loss = step_train_full(i)
            torch.nn.utils.clip_grad_norm_(params, 5.0)
            opt.step()
            return loss

        ms_full, _ = timed(step_train_full, args.steps, args.warmup)
        ms_full_opt, _ = timed(step_train_full_opt, args.steps, args.warmup)
        Rf = int(nsig_f.sum())
        tf_full = 3 * (F_X + F_CTX) * Rf * args.steps / (ms_full * 1e-3) / 1e12
        full = {"events": Bf, "flow_rows": Rf, "ms_per_step": ms_full / args.steps,
                "steps_per_sec": args.steps / (ms_full * 1e-3), "kernels_per_step": int(gstep.last_launches),
                "e2e_ms": ms_full_opt / args.steps,
                "roofline_frac": tf_full / float(peaks()[0]["bf16_tflops_sustained"]),
                "roofline_frac_burst": tf_full / float(peaks()[0]["bf16_tflops"]),
                "kernels": "split-bf16 tcgen05 GEMMs (calls of >= 512 rows)" if args.tier == "bf16" else "FP32-FMA GEMMs"}

    # the same full step with the flow's parameters re-homed in ONE flat buffer (GraphedNLLStep.flatten_parameters):
    # clip + AdamW touch two tensors instead of 181.  Measured last: it moves the parameters' storage.
    flat = gstep.flatten_parameters()
    opt_flat = torch.optim.AdamW([flat] + other_params, lr=0.0, weight_decay=0.0, fused=True)

    def step_train_flat(i):
        loss = step_train(i)
        gstep.clip_grad_norm_(5.0, other_params)
        opt_flat.step()
        return loss

    ms_train_flat, _ = timed(step_train_flat, args.steps, args.warmup)
    ms_full_flat = None
    if full is not None:
        def step_train_full_flat(i):
            loss = step_train_full(i)
            gstep.clip_grad_norm_(5.0, other_params)
            opt_flat.step()
            return loss
        ms_full_flat, _ = timed(step_train_full_flat, args.steps, args.warmup)
        full["e2e_flat_ms"] = ms_full_flat / args.steps

    pk, pk_kind = peaks()
    rows_s = R * world * args.steps / (ms * 1e-3)
    flop_row = F_X + F_CTX / n
    ach = rows_s / world * flop_row / 1e12
    peak = float(pk["bf16_tflops_sustained"])
    peak_burst = float(pk["bf16_tflops"])
    train_flop = 3 * (F_X + F_CTX) * Rt
    ach_t = train_flop * args.steps / (ms_train * 1e-3) / 1e12
    traffic, traffic_note = ncu_traffic(R)
    lp_tf = R * args.steps / (ms_lp * 1e-3) * flop_row / 1e12
    c2_rows_s = 50000 * args.steps / (ms_c2 * 1e-3)
    c5_rows_s = c5_rows * world * c5_steps / (ms_c5 * 1e-3)
    out = {
        "metric": "posterior_samples_per_sec", "value": rows_s, "unit": "samples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "strong" if args.events is None else "weak",
        "vs_baseline": None, "dtype": "f32" if args.tier == "fp32" else "bf16 operands, f32 accumulate+spline",
        "data": "synthetic (random-init weights, N(0,1) contexts standing in for encoder output, in-kernel Philox base noise)",
        "config": workload_config(world) if args.events is None else
        {"workload": f"{E} events x {n} samples per GPU (non-default)", "parallelism": f"events sharded over {world} GPU(s)"},
        "tier": args.tier + (" (sampler + log-prob: tcgen05 bf16 operands, fp32 accumulate + fp32 spline; training step: split-bf16 tcgen05 GEMMs, fp32 accumulate; FP32-FMA GEMMs for calls under 512 flow rows)"
                             if args.tier == "bf16" else ""),
        "e2e": {"value": R * world * args.steps / (ms_e2e * 1e-3), "unit": "samples/s",
                "h2d_bytes_per_step": int(ctx_host.numel() * 4), "d2h_bytes_per_step": int(theta_host.numel() * 4),
                "ms_per_step": ms_e2e / args.steps,
                "api": "posteriflow_b200.LeanNPE.sample_posterior_host(pinned host context) -> pinned host theta (8 event chunks, copies on a side stream)"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                     "peak_burst": peak_burst, "frac_burst": ach / peak_burst,
                     "traffic": traffic, "traffic_note": traffic_note,
                     "peak_source": f"{pk_kind} bf16_tflops_sustained (frac) / bf16_tflops (frac_burst): the step is "
                                    "long enough for the sustained figure, but no run so far reached the power cap",
                     "flop_per_row": flop_row,
                     "note": "whole sampler step per GPU (dense single-pass conditioner FLOPs, SURVEY 8d); "
                             "hbm: %.1f GB/s of algorithmic output (49 B/row)" % (rows_s / world * 49 / 1e9)},
        "log_prob": {"metric": "log_prob_rows_per_sec", "value": R * world * args.steps / (ms_lp * 1e-3), "unit": "rows/s",
                     "ms_per_step": ms_lp / args.steps, "gpu_launches": int(launches_lp),
                     "api": "NSFPosteriorFlow.compute_psd_aware_nll(y [R,11], context [E,288], n_per_ctx=n), no gradient",
                     "roofline": {"bound": "tensor", "achieved": lp_tf, "peak": peak, "unit": "TFLOP/s",
                                  "frac": lp_tf / peak, "frac_burst": lp_tf / peak_burst, "flop_per_row": flop_row}},
        "fp32_accuracy": fp32acc,
        "infer_loop": {"metric": "samples_with_log_prob_per_sec", "value": R * world * args.steps / (ms_inf * 1e-3),
                       "unit": "samples/s", "ms_per_step": ms_inf / args.steps, "gpu_launches": int(launches_inf),
                       "api": "posteriflow_b200.posterior_chunk (pipeline.py:168-186 on the device: sampler -> wrap -> rail "
                              "mask -> density -> physical log-prob -> m1>=m2)"},
        "summary": {"metric": "posterior_summary_ms", "ms_per_step": ms_sum / args.steps,
                    "api": "posteriflow_b200.posterior_summary(theta [E,n,11], truth, railed): 5 quantiles + medians + SBC "
                           "ranks + railing fraction per event",
                    "roofline": {"bound": "hbm", "achieved": R * D * 4 * args.steps / (ms_sum * 1e-3) / 1e9,
                                 "peak": float(pk["hbm_gbs"]), "unit": "GB/s",
                                 "frac": R * D * 4 * args.steps / (ms_sum * 1e-3) / 1e9 / float(pk["hbm_gbs"]),
                                 "bytes_per_row": D * 4}},
        "configs": {
            "cfg1": "CPU-runnable smoke config of the reference (train a few steps -> sample): correctness only, its GPU "
                    "analogue is tests/test_gpu_round2.py::test_three_adamw_steps_then_every_direction_matches_the_oracle",
            "cfg2": {"workload": "1 event x 50 000 samples, rank 0, sample blocks over the GPUs (dist.shard_sampling), "
                                 "disjoint Philox keys per rank", "rows_per_gpu": int(s1 - s0),
                     "tiles_per_gpu": int((s1 - s0 + 127) // 128), "ms_per_call": ms_c2 / args.steps,
                     "value": c2_rows_s, "unit": "samples/s", "gpu_launches": int(launches_c2),
                     "roofline_frac": c2_rows_s / world * (F_X + F_CTX / (s1 - s0)) / 1e12 / peak,
                     "note": "latency regime: one 128-row tile per SM at most, the call is bound by one tile's 660-stage chain"},
            "cfg3": "the headline of this line",
            "cfg4": "train (per-GPU shard of the 4096-event batch) and train.full_batch_1gpu",
            "cfg5": {"workload": f"rank-conditioned overlap sampling through ONE launch (posteriflow_b200.sample_ranks): "
                                 f"{c5_ev} events x ranks 0-2 x 10 000 samples per GPU -- SUB-SAMPLE of the "
                                 f"{10000 // max(world, 8)}-event shard a GPU holds at 8 GPUs; throughput is row-count invariant",
                     "rows_per_gpu": c5_rows, "ms_per_step": ms_c5 / c5_steps, "value": c5_rows_s, "unit": "samples/s",
                     "gpu_launches": int(launches_c5),
                     "roofline_frac": c5_rows_s / world * (F_X + F_CTX / 10000) / 1e12 / peak,
                     "full_config_seconds_extrapolated": 3e8 / max(c5_rows_s, 1.0)}},
        "train": {"metric": "nll_train_steps_per_sec", "value": args.steps / (ms_train * 1e-3), "unit": "steps/s",
                  "events_per_gpu": Bt, "flow_rows_per_gpu": Rt, "ms_per_step": ms_train / args.steps,
                  "gpu_launches": int(launches_train), "kernels_per_step": kernels_shard,
                  "e2e_ms": ms_train_opt / args.steps,
                  "e2e_includes": "flow-only step + torch.nn.utils.clip_grad_norm_(5.0) + fused AdamW.step() over all flow "
                                  "+ rank-embedding parameters + the re-pack of the updated weights (train_lean_npe.py:363-368); "
                                  "lr = 0 keeps the benchmarked weights fixed while every tensor is rewritten in place",
                  "e2e_flat_ms": ms_train_flat / args.steps,
                  "e2e_flat_includes": "the same step through GraphedNLLStep.flatten_parameters(): flow parameters as views "
                                       "of one flat buffer, GraphedNLLStep.clip_grad_norm_(5.0) + fused AdamW over [flat, rank "
                                       "embedding] + the re-pack (same arithmetic as e2e_ms; tests/test_gpu_round2.py)",
                  "repack_ms": ms_pack / args.steps, "repack_launches": int(launches_pack // max(1, args.steps)),
                  "ms_per_step_eager_autograd": ms_train_eager / args.steps,
                  "full_batch_1gpu": full,
                  "api": "posteriflow_b200.GraphedNLLStep (batch_nll forward + backward captured in one CUDA graph per "
                         "row bucket)",
                  "includes": "batch_nll (one flow call for all signal ranks; nsig handed over as a pinned HOST tensor, as a data loader leaves it, so the compact rows form needs no device sync for the row count): row gather, scaler.normalize, rank embedding, flow "
                              "forward+backward (all weights, context), "
                              + ("NCCL gradient all-reduce" if world > 1 else "no all-reduce at 1 GPU")
                              + "; EXCLUDED: the strain encoder (north_star leaves it in PyTorch; synthetic N(0,1) contexts of "
                                "its output shape [B,256] stand in for encoder(strain [B,3,16384])) and, in ms_per_step, the "
                                "optimizer (see e2e_ms)",
                  "roofline": {"bound": "tensor", "achieved": ach_t, "peak": peak, "unit": "TFLOP/s",
                               "frac": ach_t / peak, "frac_burst": ach_t / peak_burst, "flop_per_row": 3 * (F_X + F_CTX)}},
    }
    if rank == 0:
        if not args.no_cpu and world == 1:
            out["cpu_baseline"] = cpu_measure()
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--tier", default=os.environ.get("PF_TIER", "bf16"), choices=["fp32", "bf16"])
    ap.add_argument("--events", type=int, default=None, help="events PER GPU (default: 4096 / N, strong scaling)")
    ap.add_argument("--cfg5-events", type=int, default=CFG5_EVENTS_PER_GPU)
    ap.add_argument("--samples", type=int, default=SAMPLES_PER_EVENT)
    ap.add_argument("--train-events", type=int, default=TRAIN_EVENTS_PER_GPU)
    ap.add_argument("--train-events-full", type=int, default=4096, help="config #4 un-sharded, timed at N=1 (0 = skip)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
