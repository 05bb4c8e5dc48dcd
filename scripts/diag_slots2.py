import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
os.environ["PF_TRAIN_TC_MIN_ROWS"] = "0"
import nsf_oracle as O
import posteriflow_b200 as P
from util import load_golden, t
spec, w, g = load_golden("lean_glue")
m = P.LeanNPE(tier="bf16")
m.flow.load_state_dict(O.full_state_dict(O.FlowSpec(), w), strict=True)
with torch.no_grad():
    m.rank_embed.weight.copy_(t(g["rank_embed"]))
m = m.to("cuda")
sc = O.Scaler()
step = P.GraphedNLLStep(m, bucket=64)
print("use slots", step._use_slots(40, torch.device("cuda")))
gen = torch.Generator().manual_seed(12)
for trial in range(3):
    B = 40
    nsig = torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), B, replacement=True, generator=gen) + 1
    params = sc.denormalize(torch.rand(B, 5, 11, generator=gen) * 2 - 1).cuda()
    ctx = torch.randn(B, 256, generator=gen)
    m.zero_grad(set_to_none=True)
    c1 = ctx.cuda().requires_grad_(True)
    ref = P.batch_nll(m, None, params, nsig.cuda(), context=c1)
    print(trial, "ref", float(ref.detach()))
    ref.backward()
    m.zero_grad(set_to_none=True)
    c2 = ctx.cuda().requires_grad_(True)
    loss = step(params, nsig.cuda(), c2)
    torch.cuda.synchronize()
    st = list(step.slot_graphs.values())[0]
    print(trial, "loss", float(loss), "ref", float(ref.detach()), "nsig sum", int(nsig.sum()), "st nsig sum", int(st["nsig"].sum()),
          "finite params", bool(torch.isfinite(st["params"]).all()))
    # per-row nll of all slots through the eager op for inspection
    with torch.no_grad():
        S = 5
        p = st["params"]
        y = m.scaler.normalize(p).reshape(B * S, 11)
        cf = torch.cat([st["context"].repeat_interleave(S, 0), m.rank_embed.weight.detach().repeat(B, 1)], 1)
        nll = m.flow.compute_psd_aware_nll(y, cf, None)
        present = (torch.arange(S, device="cuda")[None] < st["nsig"][:, None]).reshape(-1)
        print("   eager all-slot nll: min", float(nll.min()), "max", float(nll.max()), "finite", bool(torch.isfinite(nll).all()),
              "mean present", float(nll[present].mean()))
