"""Time the training step's gradient all-reduce in isolation (one flat fp32 buffer of the flow's 6.2 M gradients).
    torchrun --nproc-per-node N scripts/prof_allreduce.py      (NCCL_ALGO / NCCL_PROTO from the environment)"""
import os

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", 0))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl")
n = 6_210_610
buf = torch.randn(n, device="cuda")
for _ in range(10):
    dist.all_reduce(buf)
torch.cuda.synchronize()
dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ts = []
for rep in range(5):
    e0.record()
    for _ in range(20):
        dist.all_reduce(buf)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) / 20)
t = torch.tensor(min(ts), device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    w = dist.get_world_size()
    print(f"all_reduce {n * 4 / 1e6:.1f} MB on {w} GPUs: {float(t):.3f} ms  (bus bw {2 * (w - 1) / w * n * 4 / float(t) / 1e6:.0f} GB/s)  "
          f"NCCL_ALGO={os.environ.get('NCCL_ALGO', '-')} NCCL_PROTO={os.environ.get('NCCL_PROTO', '-')}", flush=True)
dist.destroy_process_group()
