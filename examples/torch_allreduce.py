"""torchrun --nproc-per-node N --master-addr 127.0.0.1 examples/torch_allreduce.py
allreduce(tensor), the zero-copy path, the halves and the DDP hook in ~30 lines."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flextree as ft

local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))   # plumbing only
comm = ft.init()
world, rank = comm.size, comm.rank

x = torch.full((1 << 20,), float(rank + 1), device="cuda")
ft.allreduce(x)                                   # in place, SUM
assert x[0].item() == world * (world + 1) / 2

g = torch.randn(1 << 22, device="cuda")
ft.allreduce(g, scale=1.0 / world, wire_dtype=torch.bfloat16)   # fused average, 16-bit wire

buf = comm.symm_empty(1 << 24)                    # symmetric memory: zero-copy, NVLS-capable
buf.normal_()
comm.allreduce(buf, scale=1.0 / world)
print(rank, comm.last_call())

part = torch.empty((1 << 20) // world, device="cuda")
comm.reduce_scatter(x, part)                      # the two halves of the algorithm
comm.all_gather(part, x)

model = torch.nn.Linear(1024, 1024).cuda()
ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
ddp.register_comm_hook(ft.FlexTreeHookState(comm), ft.ddp_comm_hook)
ddp(torch.randn(8, 1024, device="cuda")).sum().backward()
torch.cuda.synchronize()
assert comm.error() == 0
ft.shutdown()
dist.destroy_process_group()
