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
