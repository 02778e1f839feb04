#!/usr/bin/env python
"""NVLS on private buffers: scan of the role-split pipeline shape (tile size, CTAs, reducers, copy depth).
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/private_scan.py --out gpurun_out/private_scan.jsonl"""
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
    ap.add_argument("--out", default="gpurun_out/private_scan.jsonl")
    ap.add_argument("--sizes", default="16777216,67108864,268435456")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = ft.init()
    sizes = [int(x) for x in args.sizes.split(",")]
    maxcount = max(sizes) // 4
    comm.reserve(max(sizes))
    x = torch.randn(maxcount, device=dev)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    fout = open(args.out, "a") if rank == 0 else None

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

    def emit(rec, win):
        if rank == 0:
            rec["clocks"] = sampler.window(win[0], win[1])
            fout.write(json.dumps(rec) + "\n")
            fout.flush()
            print(json.dumps(rec), flush=True)

    for nbytes in sizes:
        v = x[: nbytes // 4]
        us, win = timed(lambda: dist.all_reduce(v), nbytes)
        emit({"impl": "nccl", "n": world, "bytes": nbytes, "us": round(us, 2)}, win)
        comm.set_algo("stages")
        us, win = timed(lambda: comm.allreduce(v), nbytes)
        emit({"impl": "flextree", "algo": "stages", "n": world, "bytes": nbytes, "us": round(us, 2), "ctas": comm.last_call()["ctas"]}, win)
        comm.set_algo("nvls")
        os.environ["FT_NVLS_SPLIT_BYTES"] = "4194304"
        for div, tmax in ((64, 256 << 10), (128, 128 << 10), (256, 64 << 10), (256, 32 << 10)):
            for ctas in (64, 96, 132):
                for red in (8, 16):
                    for deep in (0, 8):
                        os.environ["FT_NVLS_PRIV_TILE_DIV"] = str(div)
                        os.environ["FT_NVLS_PRIV_TILE_MAX"] = str(tmax)
                        os.environ["FT_NVLS_PRIV_CTAS"] = str(ctas)
                        os.environ["FT_NVLS_REDUCERS"] = str(red)
                        comm.set_tuning("nvls_flags", deep)   # also invalidates the plan cache
                        us, win = timed(lambda: comm.allreduce(v), nbytes)
                        info = comm.last_call()
                        emit({"impl": "flextree", "algo": "nvls", "n": world, "bytes": nbytes, "div": div, "tile_max": tmax, "cap_ctas": ctas,
                              "reducers": red, "deep": deep, "ctas": info["ctas"], "tile_bytes": info["tile_bytes"], "us": round(us, 2),
                              "err": comm.error()}, win)
        comm.set_tuning("nvls_flags", 0)
    dist.barrier()
    ft.shutdown()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
