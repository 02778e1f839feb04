"""Data-parallel training on CPU ranks with the gradients in shared memory.

  python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 examples/cpu_data_parallel.py

Every gradient is a view into the communicator's registered region; bucket all-reduces (1/N fused) run on a worker
thread while backward continues. The reference's use case (allreduce_over_mpi/mpi_mod.hpp:1167-1221) as a training loop.
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flextree as ft  # noqa: E402

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
torch.manual_seed(0)
model = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.ReLU(), torch.nn.Linear(256, 1))
comm = ft.HostCommunicator(window_bytes=ft.HostDataParallel.bytes_needed(model))
dp = ft.HostDataParallel(model, comm, bucket_mb=0.05)
opt = torch.optim.SGD(model.parameters(), lr=0.05)
g = torch.Generator().manual_seed(rank)
w_true = torch.randn(64, 1, generator=torch.Generator().manual_seed(99))
for step in range(30):
    x = torch.randn(32, 64, generator=g)
    loss = torch.nn.functional.mse_loss(dp(x), x @ w_true)
    loss.backward()
    dp.synchronize()
    opt.step()
    dp.zero_grad()
    if rank == 0 and step % 10 == 0:
        print(f"step {step}: loss {loss.item():.4f}")
dp.close()
comm.close()
if rank == 0:
    print("done:", len(dp.buckets), "buckets,", world, "ranks")
dist.destroy_process_group()
