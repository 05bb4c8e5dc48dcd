import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import nsf_oracle as O
import posteriflow_b200 as P
torch.manual_seed(0)
m = P.LeanNPE(tier="bf16").to("cuda")
sc = O.Scaler()
gen = torch.Generator().manual_seed(12)
B = 40
nsig = torch.multinomial(torch.tensor([.55, .20, .12, .08, .05]), B, replacement=True, generator=gen) + 1
params = sc.denormalize(torch.rand(B, 5, 11, generator=gen) * 2 - 1).cuda()
ctx = torch.randn(B, 256, generator=gen)
c1 = ctx.cuda().requires_grad_(True)
ref = P.batch_nll(m, None, params, nsig.cuda(), context=c1)
print("ref before backward", float(ref))
ref.backward()
print("ref after backward", float(ref), "gctx", float(c1.grad.abs().max()))
m.zero_grad(set_to_none=True)
for slots in (False, True):
    step = P.GraphedNLLStep(m, bucket=64, slots=slots)
    c2 = ctx.cuda().requires_grad_(True)
    m.zero_grad(set_to_none=True)
    loss = step(params, nsig.cuda(), c2)
    torch.cuda.synchronize()
    print("slots", slots, "loss", float(loss), "ref now", float(ref), "gctx diff", float((c2.grad - c1.grad).abs().max()),
          "launches", step.last_launches)
    st = (step.slot_graphs if slots else step.graphs)
    st = list(st.values())[0]
    if slots:
        print("   gemb", float(st["gemb"].abs().max()), "gcontext", float(st["gcontext"].abs().max()))
m.flow.handle().check()
