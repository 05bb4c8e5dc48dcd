"""Two GPUs, NCCL: the data-parallel NLL step in every form of `GraphedNLLStep` (slot mode, compact rows, ATen gather) gives
the gradients of the GLOBAL mean -- the same numbers as one process stepping over the concatenated batch -- including a
rank whose shard holds no signal.  Skipped on a one-GPU box (the driver's test box); run with `gpurun --gpus 2`."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _batch(B, seed, empty=False):
    import nsf_oracle as O
    g = torch.Generator().manual_seed(seed)
    nsig = torch.randint(0, 6, (B,), generator=g)
    if empty:
        nsig.zero_()
    params = O.Scaler().denormalize(torch.rand(B, 5, 11, generator=g) * 2 - 1)
    ctx = torch.randn(B, 256, generator=g)
    return params, nsig, ctx


def _model(dev):
    import nsf_oracle as O
    import posteriflow_b200 as P
    torch.manual_seed(3)
    m = P.LeanNPE(tier="bf16")
    spec = O.FlowSpec()
    m.flow.load_state_dict(O.full_state_dict(spec, O.init_weights(spec, seed=5, stress=False)), strict=True)
    return m.to(dev)


FORMS = {"slots": dict(slots=True), "compact": dict(slots=False, compact=True), "aten": dict(slots=False, compact=False)}


def _worker(rank, world, port, B, empty_rank, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import posteriflow_b200 as P
        dev = torch.device("cuda", rank)
        m = _model(dev)
        params, nsig, ctx = _batch(B, 100 + rank, empty=(rank == empty_rank))
        res = {}
        for name, kw in FORMS.items():
            step = P.GraphedNLLStep(m, distributed=True, **kw)
            m.zero_grad(set_to_none=True)
            c = ctx.to(dev).requires_grad_(True)
            loss = step(params.to(dev), nsig.to(dev), c)
            step.allreduce([m.rank_embed.weight])
            tot = loss.detach().clone().float().reshape(())
            dist.all_reduce(tot)          # the loss of a rank is its share of the global mean: sum them
            res[name] = (float(tot), [p.grad.detach().cpu().clone() for p in m.flow.layer_weights()][:6],
                         m.rank_embed.weight.grad.detach().cpu().clone(),
                         None if c.grad is None else c.grad.detach().cpu().clone())
        out[rank] = res
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("empty_rank", [-1, 1])
def test_two_gpu_step_equals_the_single_process_step_over_the_whole_batch(empty_rank):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import posteriflow_b200 as P
    world, B, port = 2, 300, _free_port()
    with mp.Manager() as mg:
        out = mg.dict()
        mp.spawn(_worker, args=(world, port, B, empty_rank, out), nprocs=world, join=True)
        res = dict(out)
    # single process, both shards concatenated
    dev = torch.device("cuda", 0)
    m = _model(dev)
    parts = [_batch(B, 100 + r, empty=(r == empty_rank)) for r in range(world)]
    params = torch.cat([p[0] for p in parts]).to(dev)
    nsig = torch.cat([p[1] for p in parts]).to(dev)
    ctx = torch.cat([p[2] for p in parts]).to(dev).requires_grad_(True)
    step = P.GraphedNLLStep(m, slots=False, compact=False)
    loss = float(step(params, nsig, ctx))
    ref_w = [p.grad.detach().cpu() for p in m.flow.layer_weights()][:6]
    ref_e = m.rank_embed.weight.grad.detach().cpu()
    ref_c = ctx.grad.detach().cpu()
    for r in range(world):
        for name in FORMS:
            tot, gw, ge, gc = res[r][name]
            assert abs(tot - loss) <= 1e-5 * max(1.0, abs(loss)), (r, name, tot, loss)
            for a, b in zip(gw, ref_w):
                assert float((a - b).abs().max()) <= 5e-5 * max(1e-12, float(b.abs().max())), (r, name)
            assert float((ge - ref_e).abs().max()) <= 5e-5 * float(ref_e.abs().max()), (r, name)
            if gc is not None and r != empty_rank:
                assert float((gc - ref_c[r * B:(r + 1) * B]).abs().max()) <= 5e-5 * float(ref_c.abs().max()), (r, name)
