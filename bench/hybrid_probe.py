#!/usr/bin/env python
"""Probe: does P2P traffic fit next to the NVLS allreduce?

nvls_scan.py shows that on 8 GPUs the in-switch reduce alone and the multicast store alone each run at ~680 GB/s per
link direction, while together they settle at ~525 GB/s per direction whatever the kernel looks like; the two-phase
P2P tree sustains ~630 GB/s per direction in both directions at once. If the NVLS combination is limited by something
other than link bytes, part of the message can travel over the P2P tree AT THE SAME TIME for free. This script
measures exactly that with two communicators on two streams: fraction f of a 256 MiB message through the TMA stages
(communicator B), 1-f through NVLS (communicator A), launched back to back, timed from a common start event to the
later of the two ends (max over ranks).

  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench/hybrid_probe.py --out gpurun_out/hybrid.jsonl
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bytes", type=int, default=256 << 20)
    ap.add_argument("--out", default="gpurun_out/hybrid.jsonl")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ca = ft.Communicator.from_process_group(None, local)
    cb = ft.Communicator.from_process_group(None, local)
    count = args.bytes // 4
    a_buf = ca.symm_empty(count, torch.float32)
    b_buf = cb.symm_empty(count, torch.float32)
    a_buf.normal_()
    b_buf.normal_()
    ca.set_algo("nvls")
    cb.set_algo("stages")
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    fout = open(args.out, "a") if rank == 0 else None
    reps = 10

    def run(na, nb, widths, ctas_b):
        """na elements through NVLS, nb through the stages, concurrently; us per pair."""
        cb.set_widths(widths)
        cb.set_max_ctas(ctas_b)
        va, vb = a_buf[:na], b_buf[:nb]

        def pair():
            start = torch.cuda.Event()
            start.record()
            sa.wait_event(start)
            sb.wait_event(start)
            if na:
                ca.allreduce(va, stream=sa)
            if nb:
                cb.allreduce(vb, stream=sb)
            torch.cuda.current_stream().wait_stream(sa)
            torch.cuda.current_stream().wait_stream(sb)

        for _ in range(3):
            pair()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(reps):
            pair()
        e1.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        err = ca.error() | cb.error()
        return t.item() * 1e3, (t0, t1), err

    for f in (0.0, 0.05, 0.10, 0.15, 0.20, 0.25, 0.35, 1.0):
        nb = int(count * f) // (world * 1024) * (world * 1024)
        na = count - nb
        for widths in ([world],) if f in (0.0,) else ([world], [2, world // 2] if world >= 4 else [world]):
            for ctas_b in ((0,) if f == 0.0 else (16, 32, 64, 132)):
                us, win, err = run(na, nb, widths, ctas_b)
                if rank == 0:
                    rec = {"n": world, "bytes": count * 4, "p2p_fraction": f, "nvls_bytes": na * 4, "p2p_bytes": nb * 4, "p2p_widths": widths,
                           "p2p_max_ctas": ctas_b, "us": round(us, 2), "busbw_GBps": round(count * 4 / us / 1e3 * 2 * (world - 1) / world, 1),
                           "err": err, "clocks": sampler.window(win[0], win[1])}
                    fout.write(json.dumps(rec) + "\n")
                    fout.flush()
                    print(json.dumps(rec), flush=True)
    dist.barrier()
    ca.close()
    cb.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
