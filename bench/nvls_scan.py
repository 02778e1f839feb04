#!/usr/bin/env python
"""Where does the time of the NVLS allreduce go?  

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/nvls_scan.py \
      [--bytes 268435456] [--out gpurun_out/nvls_scan.jsonl] [--quick]

For one message size on registered (symmetric) buffers, fp32 SUM in place:
  * every inner-loop variant of kernels/nvls_kernel.cu  x  CTAs per rank  x  tile cap
  * the two measurement modes that take the switch path apart: only multimem.ld_reduce (no broadcast), only
    multimem.st (no reduce)  ->  is the in-switch reduce, the broadcast or their overlap the limiter?
  * exit barrier kept vs dropped
  * NCCL on the same buffers, and the multi-stage P2P trees for reference
Every row: device time (CUDA events around back-to-back calls, max over ranks), a clock record, and the NVLink
byte counters of rank 0's GPU (NVML, payload and raw) over the timed loop next to the algorithmic link bytes.
Correctness of every non-measurement row is checked against an NCCL fp64 sum.
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
from flextree.utils.clocks import ClockSampler  # noqa: E402
from flextree.utils.nvlink import NvlinkCounters  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bytes", type=int, default=256 << 20)
    ap.add_argument("--out", default="gpurun_out/nvls_scan.jsonl")
    ap.add_argument("--quick", action="store_true", help="variants at the default shape only")
    ap.add_argument("--dtype", default="fp32")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = ft.init()
    tdt = torch.float32 if args.dtype == "fp32" else torch.bfloat16
    es = 4 if args.dtype == "fp32" else 2
    count = args.bytes // es
    nbytes = count * es
    buf = comm.symm_empty(count, tdt)
    gen = torch.Generator(device=dev).manual_seed(99 + rank)
    pristine = torch.randn(count, device=dev, generator=gen).to(tdt)
    want = pristine.double().clone()
    dist.all_reduce(want)
    priv = torch.empty(count, device=dev, dtype=tdt)
    fout = open(args.out, "a") if rank == 0 else None
    sampler = ClockSampler(local)
    nvl = NvlinkCounters(local)
    if rank == 0:
        sampler.start()
        if nvl.h is None:
            print(f"nvls_scan: NVLink counters unavailable: {nvl.why}", file=sys.stderr)
    reps = 10 if nbytes >= (64 << 20) else 30

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0 = nvl.read() if rank == 0 else None
        t0 = time.perf_counter()
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        c1 = nvl.read() if rank == 0 else None
        t = torch.tensor([a.elapsed_time(b) / reps], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        link = NvlinkCounters.delta_bytes(c0, c1)
        if link:
            link = {k: v / reps for k, v in link.items()}
        return t.item() * 1e3, (t0, t1), link

    def emit(rec, win, link):
        if rank != 0:
            return
        us = rec["us"]
        rec["busbw_GBps"] = round(nbytes / us / 1e3 * 2 * (world - 1) / world, 2)
        rec["clocks"] = sampler.window(win[0], win[1])
        if link:
            rec["nvlink_per_call"] = {k: int(v) for k, v in link.items()}
            rec["nvlink_GBps"] = {k.replace("_bytes", ""): round(v / us / 1e3, 1) for k, v in link.items()}
        fout.write(json.dumps(rec) + "\n")
        fout.flush()
        print(json.dumps(rec), flush=True)

    def check(tag):
        buf.copy_(pristine)
        comm.allreduce(buf)
        torch.cuda.synchronize()
        err = (buf.double() - want).abs().max().item()
        tol = (1e-4 if tdt == torch.float32 else 0.25) * max(1.0, want.abs().max().item())
        bad = comm.error() != 0 or not err <= tol
        if bad and rank == 0:
            print(f"nvls_scan: {tag}: WRONG (err {err}, device error {comm.error():#x})", file=sys.stderr)
        return not bad

    alg_bytes = {"nvls": (1 + 1.0 / world) * nbytes, "p2p": 2.0 * (world - 1) / world * nbytes}

    # NCCL
    us, win, link = timed(lambda: dist.all_reduce(priv))
    emit({"impl": "nccl", "n": world, "bytes": nbytes, "dtype": args.dtype, "us": round(us, 2)}, win, link)

    if comm.has_multicast:
        comm.set_algo("nvls")
        shapes = [(0, 0)] if args.quick else [(0, 0), (8, 0), (16, 0), (32, 0), (64, 0), (32, 1 << 20), (64, 512 << 10), (148, 256 << 10)]
        names = {0: "stream 8", 1: "stream 16", 2: "stream 8 pipelined", 3: "stream 4 pipelined", 4: "round-1 loop (pipelined 8)",
                 5: "warp-specialised", 6: "plain 8 (per-vector addresses)"}
        for variant in (0, 1, 2, 3, 4, 5, 6):
            for ctas, tile in (shapes if variant != 6 else [(0, 0)]):
                comm.set_tuning("nvls_variant", variant)
                comm.set_tuning("nvls_flags", 0)
                comm.set_tuning("nvls_ctas", ctas)
                comm.set_tuning("nvls_max_tile_bytes", tile)
                ok = check(f"variant {variant} ctas {ctas} tile {tile}")
                us, win, link = timed(lambda: comm.allreduce(buf))
                info = comm.last_call()
                emit({"impl": "flextree", "algo": "nvls", "variant": variant, "what": names[variant], "flags": 0, "ctas": info["ctas"],
                      "threads": info["threads"], "tile_bytes": info["tile_bytes"], "n": world, "bytes": nbytes, "dtype": args.dtype,
                      "us": round(us, 2), "ok": ok, "alg_link_bytes_per_direction": int(alg_bytes["nvls"])}, win, link)
        # take the path apart (measurement kernels 7 / 8), and the exit barrier
        for ctas in ((0,) if args.quick else (0, 32, 64)):
            comm.set_tuning("nvls_ctas", ctas)
            comm.set_tuning("nvls_max_tile_bytes", 0)
            for variant, flags, what in ((7, 0, "only multimem.ld_reduce (no broadcast; wrong sums by design)"),
                                         (8, 0, "only multimem.st (no in-switch reduce; wrong sums by design)"),
                                         (0, 4, "stream 8, exit barrier kept")):
                comm.set_tuning("nvls_variant", variant)
                comm.set_tuning("nvls_flags", flags)
                us, win, link = timed(lambda: comm.allreduce(buf))
                info = comm.last_call()
                emit({"impl": "flextree", "algo": "nvls", "variant": variant, "flags": flags, "what": what, "ctas": info["ctas"],
                      "threads": info["threads"], "tile_bytes": info["tile_bytes"], "n": world, "bytes": nbytes, "dtype": args.dtype,
                      "us": round(us, 2)}, win, link)
                assert comm.error() == 0
        comm.set_tuning("nvls_ctas", 0)
        comm.set_tuning("nvls_flags", 0)
        comm.set_tuning("nvls_variant", -1)

    # the P2P trees for reference (TMA-staged)
    comm.set_algo("stages")
    for widths in ft.enumerate_widths(world):
        comm.set_widths(widths)
        ok = check(f"stages {widths}")
        us, win, link = timed(lambda: comm.allreduce(buf))
        info = comm.last_call()
        emit({"impl": "flextree", "algo": "stages", "widths": widths, "ctas": info["ctas"], "tile_bytes": info["tile_bytes"], "n": world,
              "bytes": nbytes, "dtype": args.dtype, "us": round(us, 2), "ok": ok, "alg_link_bytes_per_direction": int(alg_bytes["p2p"])}, win, link)
    comm.set_widths(None)
    comm.set_algo("auto")
    dist.barrier()
    ft.shutdown()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
