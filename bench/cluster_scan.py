#!/usr/bin/env python
"""Cluster (CTA pair per tile, DSMEM combine) vs one CTA per tile for the TMA-staged P2P stages.

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/cluster_scan.py --out gpurun_out/clusters.jsonl

Every width vector of N, 16 / 64 / 256 MiB fp32 SUM in place on registered buffers, cluster off / on, plus private
buffers at 64 MiB. Device time (CUDA events around back-to-back calls, max over ranks), clock record per row,
correctness of every row against an NCCL fp64 sum."""
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
    ap.add_argument("--out", default="gpurun_out/clusters.jsonl")
    ap.add_argument("--sizes", default="16777216,67108864,268435456")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = ft.init()
    sizes = [int(x) for x in args.sizes.split(",")]
    maxcount = max(sizes) // 4
    symm = comm.symm_empty(maxcount, torch.float32)
    priv = torch.empty(maxcount, device=dev)
    comm.reserve(max(sizes))
    gen = torch.Generator(device=dev).manual_seed(5 + rank)
    pristine = torch.randn(maxcount, device=dev, generator=gen)
    want = pristine.double().clone()
    dist.all_reduce(want)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    fout = open(args.out, "a") if rank == 0 else None
    comm.set_algo("stages")
    comm.set_tma(True)

    def timed(fn, nbytes):
        reps = 20 if nbytes <= (64 << 20) else 10
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
        for widths in ft.enumerate_widths(world):
            comm.set_widths(widths)
            for buf, mode in ((symm, "registered"), (priv, "staging")):
                if mode == "staging" and nbytes != (64 << 20):
                    continue
                for cluster in (0, 1):
                    comm.set_tuning("cluster", cluster)
                    view = buf[:count]
                    view.copy_(pristine[:count])
                    comm.allreduce(view)
                    torch.cuda.synchronize()
                    err = (view.double() - want[:count]).abs().max().item()
                    ok = comm.error() == 0 and err <= 1e-4 * max(1.0, want[:count].abs().max().item())
                    us, win = timed(lambda: comm.allreduce(view), nbytes)
                    info = comm.last_call()
                    if rank == 0:
                        rec = {"n": world, "bytes": nbytes, "widths": widths, "mode": info["mode"], "cluster": cluster, "ctas": info["ctas"],
                               "tile_bytes": info["tile_bytes"], "us": round(us, 2), "ok": bool(ok),
                               "busbw_GBps": round(nbytes / us / 1e3 * 2 * (world - 1) / world, 2), "clocks": sampler.window(win[0], win[1])}
                        fout.write(json.dumps(rec) + "\n")
                        fout.flush()
                        print(json.dumps(rec), flush=True)
    comm.set_tuning("cluster", 0)
    dist.barrier()
    ft.shutdown()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
