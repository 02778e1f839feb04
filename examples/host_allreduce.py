#!/usr/bin/env python
"""Allreduce of HOST-resident data (the reference's calling convention: MPI_Allreduce on malloc'ed vectors,
allreduce_over_mpi/benchmark.cpp:119-159) through the GPUs:

  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 examples/host_allreduce.py

Communicator.allreduce_host cuts the message into chunks and pipelines H2D copy, in-place NVLink allreduce and D2H copy
over three streams; the tensors stay on the host, the result is there when the call returns."""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flextree as ft  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))   # plumbing only (handle exchange)
comm = ft.init()
n = 1 << 24
x = torch.full((n,), float(rank + 1)).pin_memory()       # pinned: full PCIe speed (pageable memory works, slower)
out = torch.empty(n).pin_memory()
comm.allreduce_host(x, outs=out)                          # warm-up: allocates the symmetric chunk slots (collective)
t0 = time.perf_counter()
comm.allreduce_host(x, outs=out, scale=1.0 / world)       # fused average
dt = time.perf_counter() - t0
want = sum(range(1, world + 1)) / world
assert torch.all(out == want), (out[:4], want)
comm.raise_if_failed()
if rank == 0:
    print(f"host allreduce of {n * 4 >> 20} MiB on {world} GPUs: {dt * 1e3:.2f} ms, result {out[0].item()} (want {want})")
ft.shutdown()
dist.destroy_process_group()
