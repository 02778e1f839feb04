#!/usr/bin/env python
"""Wire formats for fp32 gradients on private buffers: native fp32, bf16 wire, block-scaled fp8 wire.

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/wire_scan.py --out gpurun_out/wire.jsonl

Per size: device time (CUDA events around back-to-back calls, max over ranks), clock record, and the error of the
result against an NCCL fp64 sum (max abs error / max |result| and the RMS error relative to the RMS of the result),
with the 1/N average fused in all three."""
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
    ap.add_argument("--out", default="gpurun_out/wire.jsonl")
    ap.add_argument("--sizes", default="16777216,67108864,268435456,1073741824")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = ft.init()
    sizes = [int(x) for x in args.sizes.split(",")]
    maxcount = max(sizes) // 4
    comm.reserve(max(sizes))
    gen = torch.Generator(device=dev).manual_seed(17 + rank)
    x = torch.randn(maxcount, device=dev, generator=gen) * 1e-2   # gradient-like magnitudes
    out = torch.empty_like(x)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    fout = open(args.out, "a") if rank == 0 else None

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
        xs, os_ = x[:count], out[:count]
        want = xs.double().clone()
        dist.all_reduce(want)
        want /= world
        us, win = timed(lambda: dist.all_reduce(os_.copy_(xs)), nbytes)   # NCCL incl. the copy it needs to stay out of place
        if rank == 0:
            rec = {"impl": "nccl (copy + in-place all_reduce, no fused average)", "n": world, "bytes": nbytes, "us": round(us, 2),
                   "clocks": sampler.window(win[0], win[1])}
            fout.write(json.dumps(rec) + "\n")
            print(json.dumps(rec), flush=True)
        for wire in (None, "bf16", "fp8"):
            comm.allreduce(xs, out=os_, scale=1.0 / world, wire_dtype=wire)
            torch.cuda.synchronize()
            d = (os_.double() - want)
            max_rel = (d.abs().max() / want.abs().max()).item()
            rms_rel = (d.pow(2).mean().sqrt() / want.pow(2).mean().sqrt()).item()
            us, win = timed(lambda: comm.allreduce(xs, out=os_, scale=1.0 / world, wire_dtype=wire), nbytes)
            info = comm.last_call()
            if rank == 0:
                rec = {"impl": "flextree", "wire": wire or "fp32 (native)", "algo": info["algo"], "widths": info["widths"], "n": world,
                       "bytes": nbytes, "us": round(us, 2), "max_err_over_max": max_rel, "rms_err_over_rms": rms_rel,
                       "device_error": comm.error(), "clocks": sampler.window(win[0], win[1])}
                fout.write(json.dumps(rec) + "\n")
                fout.flush()
                print(json.dumps(rec), flush=True)
    dist.barrier()
    ft.shutdown()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
