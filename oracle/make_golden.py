"""Generate the committed golden fixtures under tests/golden/ from the REFERENCE's own classes.

TEST INFRASTRUCTURE ONLY; runs in the build container (needs /root/reference):

    python oracle/make_golden.py

Every output tensor below is produced by the unmodified reference
(`ahsd.models.flows.NSFPosteriorFlow`, `ahsd.models.lean_npe.LeanNPE`) executed over the restated
`oracle/nflows` shim -- float32 as the reference runs it, plus a float64 run of the same modules
as arbiter.  Weights are NOT stored for the big shapes (25 MB): they are regenerated on the test
machine from `nsf_oracle.init_weights(spec, seed, stress)` (CPU generator, deterministic) and
verified against the stored checksum; the tiny case stores its weights explicitly.
"""
from __future__ import annotations

import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("POSTERIFLOW_REF", "/root/reference")
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(REF, "src"))

import logging  # noqa: E402
import warnings  # noqa: E402

import numpy as np  # noqa: E402
import torch  # noqa: E402

warnings.filterwarnings("ignore")
logging.disable(logging.CRITICAL)

import nsf_oracle as O  # noqa: E402
from check_against_reference import reference_flow, make_inputs  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def spec_array(spec: O.FlowSpec) -> np.ndarray:
    return np.array([spec.features, spec.context_features, spec.hidden_features, spec.num_layers,
                     spec.num_bins, spec.tail_bound, spec.num_blocks], dtype=np.float64)


def flow_case(name: str, spec: O.FlowSpec, seed: int, stress: bool, R: int, store_weights: bool = False,
              grads: bool = True) -> None:
    w = O.init_weights(spec, seed=seed, stress=stress)
    flow32 = reference_flow(spec, w)
    flow64 = reference_flow(spec, w).double()
    z, ctx = make_inputs(spec, R, seed + 100)
    out = {"spec": spec_array(spec), "seed": np.int64(seed), "stress": np.bool_(stress),
           "weights_checksum": np.float64(O.weights_checksum(w)),
           "z": z.numpy(), "ctx": ctx.numpy()}
    with torch.no_grad():
        x32, il32 = flow32.inverse(z, ctx)                       # clamped +-3 (flows.py:654)
        xu32, _ = flow32.transform.inverse(z, context=ctx)       # unclamped
        x64, il64 = flow64.inverse(z.double(), ctx.double())
        xu64, _ = flow64.transform.inverse(z.double(), context=ctx.double())
        out.update(inv_x_f32=x32.numpy(), inv_logdet_f32=il32.numpy(), inv_x_unclamped_f32=xu32.numpy(),
                   inv_x_f64=x64.numpy(), inv_logdet_f64=il64.numpy(), inv_x_unclamped_f64=xu64.numpy())
        # forward / nll evaluated at a fixed, well-defined input: the float32 clamped samples
        y = x32
        f32, fl32 = flow32.forward(y, ctx)
        f64, fl64 = flow64.forward(y.double(), ctx.double())
        n32 = flow32.compute_psd_aware_nll(y, ctx, torch.zeros_like(y))
        n64 = flow64.compute_psd_aware_nll(y.double(), ctx.double(), torch.zeros_like(y).double())
        out.update(fwd_x=y.numpy(), fwd_z_f32=f32.numpy(), fwd_logdet_f32=fl32.numpy(), nll_f32=n32.numpy(),
                   fwd_z_f64=f64.numpy(), fwd_logdet_f64=fl64.numpy(), nll_f64=n64.numpy())
        # NaN/Inf context (flows.py:626-632)
        cbad = ctx.clone()
        cbad[3, 5] = float("nan"); cbad[4, 7] = float("inf"); cbad[5, 9] = float("-inf")
        out["ctx_bad"] = cbad.numpy()
        out["inv_x_ctx_bad_f32"] = flow32.inverse(z, cbad)[0].numpy()
    if grads:
        # autograd of sum(nll) through the reference modules, float64 (arbiter) and float32
        for tag, flow, dt in (("f64", flow64, torch.float64), ("f32", flow32, torch.float32)):
            flow.zero_grad(set_to_none=True)
            yy = y.to(dt).clone().requires_grad_(True)
            cc = ctx.to(dt).clone().requires_grad_(True)
            gw = torch.linspace(0.5, 1.5, R, dtype=dt)           # non-uniform upstream gradient
            loss = (flow.compute_psd_aware_nll(yy, cc, torch.zeros_like(yy)) * gw).sum()
            loss.backward()
            out[f"grad_x_{tag}"] = yy.grad.numpy()
            out[f"grad_ctx_{tag}"] = cc.grad.numpy()
            out["grad_upstream"] = gw.double().numpy()
            # per-tensor gradients: layer 0, a middle layer and the last layer in full would be MBs;
            # store L2 norms for every tensor and a strided sample of each
            names, norms, samples = [], [], []
            for k, p in flow.transform.named_parameters():
                g = p.grad.reshape(-1)
                names.append("transform." + k)
                norms.append(float(g.double().norm()))
                idx = torch.arange(0, g.numel(), max(1, g.numel() // 64))[:64]
                s = torch.zeros(64, dtype=torch.float64)
                s[: idx.numel()] = g[idx].double()
                samples.append(s.numpy())
            out[f"grad_w_names"] = np.array(names)
            out[f"grad_w_norms_{tag}"] = np.array(norms)
            out[f"grad_w_samples_{tag}"] = np.stack(samples)
    if store_weights:
        for k, v in w.items():
            out["w:" + k] = v.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(f"wrote {name}.npz  R={R}  |x|max={float(np.abs(out['inv_x_unclamped_f64']).max()):.3f} "
          f"f32-vs-f64 max|dx|={float(np.abs(out['inv_x_f32'] - out['inv_x_f64']).max()):.2e}")


def glue_case() -> None:
    """LeanNPE.sample_posterior / nll / infer()-chunk glue: ParamScaler wrap/denormalize/normalize,
    rank embedding, context broadcast, rail mask, physical log-prob."""
    from ahsd.models.lean_npe import LeanNPE
    spec = O.FlowSpec()
    seed = 7
    w = O.init_weights(spec, seed=seed, stress=True)
    torch.manual_seed(0)
    model = LeanNPE().eval()
    model.flow.load_state_dict(O.full_state_dict(spec, w), strict=True)
    g = torch.Generator().manual_seed(21)
    E, n = 4, 24
    context = torch.randn(E, 256, generator=g)
    rank_w = torch.randn(5, 32, generator=g)
    with torch.no_grad():
        model.rank_embed.weight.copy_(rank_w)
    z = torch.randn(E * n, 11, generator=g) * 1.6
    out = {"spec": spec_array(spec), "seed": np.int64(seed), "stress": np.bool_(True),
           "weights_checksum": np.float64(O.weights_checksum(w)),
           "context": context.numpy(), "rank_embed": rank_w.numpy(), "z": z.numpy(),
           "E": np.int64(E), "n": np.int64(n)}
    model.encode = lambda strain, asd=None: context
    orig = torch.randn
    with torch.no_grad():
        for rank in (0, 2):
            torch.randn = lambda *a, **k: z
            try:
                s = model.sample_posterior(torch.zeros(E, 3, 8), rank=rank, n_samples=n)
            finally:
                torch.randn = orig
            out[f"samples_rank{rank}"] = s.numpy()
        ranks = torch.tensor([0, 1, 3, 4])
        theta = out["samples_rank0"][:, 0, :]
        nll = model.nll(None, torch.from_numpy(theta), ranks, context=context)
        out["nll_ranks"] = ranks.numpy()
        out["nll_theta"] = theta
        out["nll"] = nll.numpy()
        # infer() chunk (pipeline.py:168-180) restated by the oracle over the reference-pinned pieces
        sc = O.Scaler()
        fc = O.full_context(context[:1], rank_w, torch.zeros(1, dtype=torch.long))
        th, logq, railed = O.infer_chunk(spec, w, sc, fc, z[:n])
        out.update(infer_theta=th.numpy(), infer_logq=logq.numpy(), infer_railed=railed.numpy())
    np.savez_compressed(os.path.join(OUT, "lean_glue.npz"), **out)
    print("wrote lean_glue.npz")


def quantile_case(name: str, seed: int, stress: bool, E: int = 2, n: int = 8000) -> None:
    """Posterior-scale fixture for the bf16 tier's quantile / KS / log q checks: E events x n samples
    drawn by the REFERENCE flow (float32, as it runs) from a seeded base noise, plus the reference's own float64
    log-det and NLL of exactly those samples.  z is regenerated on the test machine from the stored seed (CPU
    generator, deterministic); only the outputs are stored."""
    spec = O.FlowSpec()
    w = O.init_weights(spec, seed=seed, stress=stress)
    flow32 = reference_flow(spec, w)
    flow64 = reference_flow(spec, w).double()
    g = torch.Generator().manual_seed(1000 + seed)
    ctx = torch.randn(E, spec.context_features, generator=g)
    z = torch.randn(E * n, spec.features, generator=g)
    cr = ctx.repeat_interleave(n, dim=0)
    with torch.no_grad():
        xs, lds = [], []
        for i in range(0, E * n, 4096):
            x, ld = flow32.inverse(z[i:i + 4096], cr[i:i + 4096])
            xs.append(x); lds.append(ld)
        x32, ild32 = torch.cat(xs), torch.cat(lds)
        f64, n64 = [], []
        for i in range(0, E * n, 4096):
            xx, cc = x32[i:i + 4096].double(), cr[i:i + 4096].double()
            f64.append(flow64.forward(xx, cc)[1])
            n64.append(flow64.compute_psd_aware_nll(xx, cc, torch.zeros_like(xx)))
    out = {"spec": spec_array(spec), "seed": np.int64(seed), "stress": np.bool_(stress),
           "weights_checksum": np.float64(O.weights_checksum(w)), "E": np.int64(E), "n": np.int64(n),
           "noise_seed": np.int64(1000 + seed), "z_checksum": np.float64(float(z.double().abs().sum())),
           "ctx": ctx.numpy(), "inv_x_f32": x32.numpy(), "inv_logdet_f32": ild32.numpy(),
           "fwd_logdet_f64": torch.cat(f64).numpy(), "nll_f64": torch.cat(n64).numpy()}
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(f"wrote {name}.npz  E={E} n={n}  clamped fraction {float((x32.abs() >= 3).float().mean()):.4f}")


def main() -> None:
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    if "--quantiles-only" in sys.argv:
        quantile_case("lean_quantiles_default", seed=0, stress=False)
        quantile_case("lean_quantiles_stress", seed=1, stress=True)
        return
    flow_case("lean_default", O.FlowSpec(), seed=0, stress=False, R=64)
    flow_case("lean_stress", O.FlowSpec(), seed=1, stress=True, R=64)
    flow_case("lean_stress_1layer", O.FlowSpec(num_layers=1), seed=5, stress=True, R=64)
    flow_case("lean_stress_2layer", O.FlowSpec(num_layers=2), seed=6, stress=True, R=64)
    flow_case("head_8x192x16", O.FlowSpec(11, 288, 192, 8, 16, 3.0), seed=2, stress=True, R=32, grads=False)
    flow_case("head_12x384x24", O.FlowSpec(11, 288, 384, 12, 24, 3.0), seed=3, stress=True, R=32, grads=False)
    flow_case("tiny_explicit", O.FlowSpec(5, 12, 32, 3, 8, 2.0), seed=4, stress=True, R=48, store_weights=True)
    glue_case()
    quantile_case("lean_quantiles_default", seed=0, stress=False)
    quantile_case("lean_quantiles_stress", seed=1, stress=True)


if __name__ == "__main__":
    main()
