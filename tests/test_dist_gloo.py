"""Host-side multi-GPU logic (posteriflow_b200/dist.py) over gloo, world_size 2, on CPU: shard arithmetic,
bucketed gradient all-reduce and the globally consistent loss normaliser.  No CUDA involved."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from posteriflow_b200 import dist as D


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 8, 4096, 10000):
        for w in (1, 2, 3, 8):
            spans = [D.shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_shard_sampling_switches_to_sample_blocks_when_events_are_few():
    # BASELINE config #3: 4096 events over 8 GPUs -> 512 whole events each
    assert D.shard_sampling(4096, 5000, 3, 8) == ((1536, 2048), (0, 5000))
    # BASELINE config #2: 1 event x 50k samples over 8 GPUs -> 6250 samples each of the same event
    assert D.shard_sampling(1, 50000, 7, 8) == ((0, 1), (43750, 50000))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        lin = torch.nn.Linear(5, 3)                      # identical weights on both ranks
        x = torch.arange(20, dtype=torch.float32).reshape(4, 5) / 10 + rank
        rows = 4 if rank == 0 else 2                     # ranks hold different numbers of flow rows
        nll_sum = (lin(x[:rows]) ** 2).sum()
        loss = D.global_mean_loss(nll_sum, torch.tensor(float(rows)))
        loss.backward()
        n = D.allreduce_gradients(lin.parameters(), bucket_bytes=16)   # tiny buckets: several collectives
        out[rank] = (lin.weight.grad.clone(), lin.bias.grad.clone(), n)
    finally:
        dist.destroy_process_group()


def test_gradient_allreduce_matches_single_process_mean():
    world, port = 2, _free_port()
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
        res = dict(out)
    torch.manual_seed(0)
    lin = torch.nn.Linear(5, 3)
    x0 = torch.arange(20, dtype=torch.float32).reshape(4, 5) / 10
    total = ((lin(x0[:4]) ** 2).sum() + (lin((x0 + 1)[:2]) ** 2).sum()) / 6.0
    total.backward()
    for r in range(world):
        gw, gb, n = res[r]
        assert torch.allclose(gw, lin.weight.grad, atol=1e-5) and torch.allclose(gb, lin.bias.grad, atol=1e-6)
        assert n >= 2


def _flat_worker(rank, world, port, out):
    """GraphedNLLStep.allreduce on CPU tensors: the graph replay itself needs a GPU, the exchange does not."""
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import posteriflow_b200 as P
        torch.manual_seed(0)
        m = P.LeanNPE(flow_layers=1, flow_hidden=32, flow_bins=8)
        step = P.GraphedNLLStep(m, distributed=True)
        ps = step.flow_params
        step.flat_grad = torch.empty(sum(p.numel() for p in ps))
        step._gbuf, o = [], 0
        for i, p in enumerate(ps):
            v = step.flat_grad[o:o + p.numel()].view_as(p)
            v.fill_(float(rank + 1) * (i + 1))
            p.grad = v
            step._gbuf.append(v)
            o += p.numel()
        m.rank_embed.weight.grad = torch.full_like(m.rank_embed.weight, float(rank + 1))
        n1 = step.allreduce(extra_params=[m.rank_embed.weight])
        a = (float(ps[0].grad.flatten()[0]), float(ps[-1].grad.flatten()[-1]), float(m.rank_embed.weight.grad[0, 0]))
        # foreign .grad tensors: falls back to the bucketed path, same sums
        for i, p in enumerate(ps):
            p.grad = torch.full_like(p, float(rank + 1) * (i + 1))
        n2 = step.allreduce()
        b = (float(ps[0].grad.flatten()[0]), float(ps[-1].grad.flatten()[-1]))
        out[rank] = (n1, a, n2, b, len(ps))
    finally:
        dist.destroy_process_group()


def test_graphed_step_allreduce_is_one_collective_on_the_flat_buffer():
    world, port = 2, _free_port()
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_flat_worker, args=(world, port, out), nprocs=world, join=True)
        res = dict(out)
    for r in range(world):
        n1, a, n2, b, nps = res[r]
        assert n1 == 2                     # flat flow gradients + one bucket for the rank embedding
        assert a == (3.0, 3.0 * nps, 3.0)  # (1 + 2) * (i + 1)
        assert n2 >= 1 and b == (3.0, 3.0 * nps)


class _ToyNLL(torch.nn.Module):
    """LeanNPE.nll's call signature over one Linear layer (CPU)."""
    max_signals = 5

    def __init__(self):
        super().__init__()
        self.lin = torch.nn.Linear(4, 1)

    def nll(self, strain, params_phys, rank, context=None, asd_bands=None):
        return self.lin(context).squeeze(1) ** 2 + (params_phys ** 2).sum(1) * (rank.float() + 1.0)


def _empty_rank_worker(rank, world, port, out):
    """Rank 1's shard holds no signal at all: it must still issue the count all-reduce and the gradient all-reduce
    (with zeros), or rank 0 hangs / the collectives mismatch."""
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world, timeout=__import__("datetime").timedelta(seconds=60))
    try:
        from posteriflow_b200.train import batch_nll
        import posteriflow_b200 as P
        torch.manual_seed(0)
        m = _ToyNLL()
        g = torch.Generator().manual_seed(1)
        params = torch.randn(6, 5, 11, generator=g)
        ctx = torch.randn(6, 4, generator=g).requires_grad_(True)   # encoder output: carries the graph on every rank
        nsig = torch.tensor([1, 2, 1, 3, 1, 1]) if rank == 0 else torch.zeros(6, dtype=torch.long)
        loss = batch_nll(m, None, params, nsig, context=ctx, distributed=True)
        loss.backward()                                   # rank 1: graph through `context` only, lin gets no grad
        n = D.allreduce_gradients(m.parameters())
        # the graphed step's empty-rank path: count all-reduce + zero flat gradients + the single flat all-reduce
        lean = P.LeanNPE(flow_layers=1, flow_hidden=32, flow_bins=8)
        step = P.GraphedNLLStep(lean, distributed=True)
        if rank == 1:
            l2 = step(params, nsig, torch.zeros(6, 256))
            assert float(l2) == 0.0
        else:                                             # what a populated rank does around its graph replay
            dist.all_reduce(torch.tensor(9.0))
            step._ensure_flat()
            step.flat_grad.fill_(2.0)
            for p, v in zip(step.flow_params, step._gbuf):
                p.grad = v
        n2 = step.allreduce()
        out[rank] = (float(loss), m.lin.weight.grad.clone(), n, n2, float(step.flat_grad.min()), float(step.flat_grad.max()))
    finally:
        dist.destroy_process_group()


def test_rank_without_signals_joins_every_collective():
    world, port = 2, _free_port()
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_empty_rank_worker, args=(world, port, out), nprocs=world, join=True)
        res = dict(out)
    torch.manual_seed(0)
    m = _ToyNLL()
    g = torch.Generator().manual_seed(1)
    params = torch.randn(6, 5, 11, generator=g)
    ctx = torch.randn(6, 4, generator=g)
    from posteriflow_b200.train import batch_nll
    ref = batch_nll(m, None, params, torch.tensor([1, 2, 1, 3, 1, 1]), context=ctx)     # global mean = rank 0's mean
    ref.backward()
    assert abs(res[0][0] - float(ref)) <= 1e-6 and res[1][0] == 0.0
    for r in range(world):
        assert torch.allclose(res[r][1], m.lin.weight.grad, atol=1e-6)
        assert res[r][2] == 1 and res[r][3] == 1
        assert res[r][4] == 2.0 and res[r][5] == 2.0      # 2 (rank 0) + 0 (empty rank)
