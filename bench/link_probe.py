#!/usr/bin/env python
"""Measured link model inputs: raw NVLink rates between the GPUs of the box (ft_link_probe, kernels/link_probe.cu).

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/link_probe.py --out gpurun_out/link_probe.jsonl

Patterns: one GPU alone (the other GPUs idle: one link direction carries data) and all GPUs at once (both directions
loaded), reads (pull) vs writes (push), 16-byte register accesses vs TMA bulk copies, 1 .. N-1 concurrent peers,
several CTA counts. GB/s = payload bytes this rank moved / device time (CUDA events, max over the active ranks).
Every row carries a clock record."""
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
    ap.add_argument("--bytes-per-peer", type=int, default=32 << 20)
    ap.add_argument("--out", default="gpurun_out/link_probe.jsonl")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = ft.init()
    bpp = args.bytes_per_peer
    comm.reserve(bpp * world)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    fout = open(args.out, "a") if rank == 0 else None
    reps = 5

    def measure(mode, npeers, ctas, active):
        me = rank in active
        for _ in range(2):
            if me:
                comm.link_probe(mode, npeers, bpp, ctas)
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record()
        if me:
            for _ in range(reps):
                comm.link_probe(mode, npeers, bpp, ctas)
        b.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        t = torch.tensor([a.elapsed_time(b) / reps if me else 0.0], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item()
        if rank == 0:
            rec = {"n": world, "mode": mode, "npeers": npeers, "ctas": ctas, "active_ranks": len(active), "bytes_per_peer": bpp,
                   "ms": round(ms, 4), "GBps_per_gpu": round(npeers * bpp / ms / 1e6, 1), "clocks": sampler.window(t0, t1)}
            fout.write(json.dumps(rec) + "\n")
            fout.flush()
            print(json.dumps(rec), flush=True)

    everyone = list(range(world))
    peers_list = sorted({1, min(2, world - 1), min(4, world - 1), world - 1})
    for mode in ("ldg_pull", "stg_push", "tma_pull", "tma_push"):
        for ctas in (16, 32, 64, 148):
            measure(mode, world - 1, ctas, [0])          # one GPU alone
            measure(mode, world - 1, ctas, everyone)     # all at once
        for npeers in peers_list[:-1]:
            measure(mode, npeers, 64, [0])
            measure(mode, npeers, 64, everyone)
    dist.barrier()
    ft.shutdown()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
