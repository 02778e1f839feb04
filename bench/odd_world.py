#!/usr/bin/env python
"""World sizes with no useful factorization (7, 5, 3 GPUs): flat tree vs "+1" tree (one lonely rank folded in) vs "-1" tree
(one phantom rank hosted by rank 0), the candidates the reference's chooseWidth lists (cost_model/ChooseWidth.h:14-31).

  python -m torch.distributed.run --nproc-per-node 7 --master-addr 127.0.0.1 bench/odd_world.py --out gpurun_out/odd_world.jsonl

fp32 SUM on private buffers (lonely and phantom trees run there), device time (CUDA events, max over ranks), clock record
per row, every row checked against an NCCL fp64 sum."""
import argparse
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flextree as ft  # noqa: E402
from flextree.utils.clocks import ClockSampler  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/odd_world.jsonl")
    ap.add_argument("--sizes", default="1048576,16777216,67108864,268435456")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = ft.init()
    sizes = [int(x) for x in args.sizes.split(",")]
    maxcount = max(sizes) // 4
    comm.reserve(max(sizes))
    gen = torch.Generator(device=dev).manual_seed(23 + rank)
    x = torch.randn(maxcount, device=dev, generator=gen)
    buf = torch.empty_like(x)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    fout = open(args.out, "a") if rank == 0 else None
    comm.set_algo("stages")
    trees = [("flat", [world])]
    trees += [(f"{'*'.join(map(str, w))}+1", w) for w in ft.enumerate_widths(world - 1) if len(w) > 1 or world - 1 <= 4]
    trees += [(f"{'*'.join(map(str, w))}-1", w) for w in ft.enumerate_widths(world + 1) if len(w) > 1]

    def timed(fn, nbytes):
        reps = 20 if nbytes <= (64 << 20) else 8
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        t = torch.tensor([a.elapsed_time(b) / reps], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item() * 1e3, (t0, t1)

    for nbytes in sizes:
        count = nbytes // 4
        xs, os_ = x[:count], buf[:count]
        want = xs.double().clone()
        dist.all_reduce(want)
        us, win = timed(lambda: dist.all_reduce(os_), nbytes)
        if rank == 0:
            rec = {"impl": "nccl", "n": world, "bytes": nbytes, "us": round(us, 2), "clocks": sampler.window(win[0], win[1])}
            fout.write(json.dumps(rec) + "\n")
            print(json.dumps(rec), flush=True)
        for name, widths in trees:
            comm.set_widths(widths)
            comm.allreduce(xs, out=os_)
            torch.cuda.synchronize()
            err = (os_.double() - want).abs().max().item()
            ok = comm.error() == 0 and err <= 1e-4 * max(1.0, want.abs().max().item())
            us, win = timed(lambda: comm.allreduce(xs, out=os_), nbytes)
            if rank == 0:
                rec = {"impl": "flextree", "tree": name, "widths": widths, "n": world, "bytes": nbytes, "us": round(us, 2), "ok": bool(ok),
                       "busbw_GBps": round(nbytes / us / 1e3 * 2 * (world - 1) / world, 2), "clocks": sampler.window(win[0], win[1])}
                fout.write(json.dumps(rec) + "\n")
                fout.flush()
                print(json.dumps(rec), flush=True)
        comm.set_widths(None)
    dist.barrier()
    ft.shutdown()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
