#!/usr/bin/env python
"""CPU data-parallel training step: ft.HostDataParallel (gradients in the registered shared-memory region, bucket
collectives on a worker thread) vs torch DDP over gloo, same model, same data, same machine.

  python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 bench/cpu_ddp.py [--hidden 2048 --layers 6]
"""
import argparse
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flextree as ft  # noqa: E402


def make_model(hidden, layers):
    torch.manual_seed(0)
    mods = []
    for _ in range(layers):
        mods += [torch.nn.Linear(hidden, hidden), torch.nn.ReLU()]
    return torch.nn.Sequential(*mods, torch.nn.Linear(hidden, 16))


def run(step_fn, steps, warmup):
    for _ in range(warmup):
        step_fn()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        step_fn()
    dt = torch.tensor([(time.perf_counter() - t0) / steps], dtype=torch.float64)
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return dt.item()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hidden", type=int, default=2048)
    ap.add_argument("--layers", type=int, default=6)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.set_num_threads(1)
    g = torch.Generator().manual_seed(rank)
    x = torch.randn(args.batch, args.hidden, generator=g)
    y = torch.randint(0, 16, (args.batch,), generator=g)
    out = {"world": world, "params": None}

    model = make_model(args.hidden, args.layers)
    out["params"] = sum(p.numel() for p in model.parameters())
    opt = torch.optim.SGD(model.parameters(), lr=0.01)

    def local_step():  # no communication: the floor
        opt.zero_grad()
        torch.nn.functional.cross_entropy(model(x), y).backward()
        opt.step()
    out["no_comm_ms"] = run(local_step, args.steps, args.warmup) * 1e3

    ddp = torch.nn.parallel.DistributedDataParallel(make_model(args.hidden, args.layers), bucket_cap_mb=25)
    dopt = torch.optim.SGD(ddp.parameters(), lr=0.01)

    def ddp_step():
        dopt.zero_grad()
        torch.nn.functional.cross_entropy(ddp(x), y).backward()
        dopt.step()
    out["gloo_ddp_ms"] = run(ddp_step, args.steps, args.warmup) * 1e3
    del ddp

    m2 = make_model(args.hidden, args.layers)
    comm = ft.HostCommunicator(window_bytes=ft.HostDataParallel.bytes_needed(m2))
    dp = ft.HostDataParallel(m2, comm, bucket_mb=8)
    fopt = torch.optim.SGD(m2.parameters(), lr=0.01)

    def ft_step():
        torch.nn.functional.cross_entropy(dp(x), y).backward()
        dp.synchronize()
        fopt.step()
        dp.zero_grad()
    out["flextree_host_dp_ms"] = run(ft_step, args.steps, args.warmup) * 1e3
    dp.close()
    comm.close()
    if rank == 0:
        out["comm_ms"] = {"gloo_ddp": out["gloo_ddp_ms"] - out["no_comm_ms"], "flextree": out["flextree_host_dp_ms"] - out["no_comm_ms"]}
        print("CPU_DDP " + json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
