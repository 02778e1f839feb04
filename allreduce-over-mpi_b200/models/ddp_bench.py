#!/usr/bin/env python
"""DDP training-step driver: stock NCCL gradient allreduce vs the flextree comm hook (fused 1/N scale,
optional bf16 wire), same model, same synthetic data, same seeds.

  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 \
      allreduce-over-mpi_b200/models/ddp_bench.py --model resnet50 --batch 64 --steps 20 --comm both

Prints one JSON line per backend (step time = CUDA events, max over ranks; items/s is the whole job) and,
with --comm both, the relative difference of the final parameters (same data, same order, so the two
runs must agree to rounding). The reference has no model code; this is BASELINE.json config #4 / #5.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel as DDP

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import time  # noqa: E402

import flextree as ft  # noqa: E402
from flextree.models.zoo import build  # noqa: E402
from flextree.utils.clocks import ClockSampler  # noqa: E402


def run(args, backend, comm, dev, rank, world, local, sampler=None):
    torch.manual_seed(1234)
    pdtype = torch.bfloat16 if args.param_dtype == "bf16" else torch.float32
    model, make_batch, loss_fn, per_sample = build(args.model, dev, pdtype)
    state = None
    engine = None
    if backend == "flextree-dp":  # zero-copy engine: gradients live in symmetric buckets
        engine = ft.FlexTreeDataParallel(model, comm, bucket_mb=args.bucket_mb, max_ctas=args.max_ctas,
                                         wire_dtype=("fp8" if args.wire == "fp8" else None))
        ddp = engine
    else:
        ddp = DDP(model, device_ids=[local], bucket_cap_mb=args.bucket_mb, gradient_as_bucket_view=True)
    if backend == "flextree":
        wire = torch.bfloat16 if args.wire == "bf16" else ("fp8" if args.wire == "fp8" else None)
        state = ft.FlexTreeHookState(comm, wire_dtype=wire, max_ctas=args.max_ctas)
        ddp.register_comm_hook(state, ft.ddp_comm_hook)
    opt = torch.optim.SGD(ddp.parameters(), lr=args.lr, momentum=0.9)
    g = torch.Generator(device=dev).manual_seed(99 + rank)
    losses = []

    def step():
        x, y = make_batch(args.batch, g)
        if engine is not None:
            engine.zero_grad()
        else:
            opt.zero_grad(set_to_none=True)
        loss = loss_fn(ddp, x, y)
        loss.backward()
        if engine is not None:
            engine.synchronize()
        opt.step()
        return loss

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record()
    for _ in range(args.steps):
        losses.append(step().detach())
    ev1.record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    ms = torch.tensor([ev0.elapsed_time(ev1) / args.steps], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    nparam = sum(p.numel() for p in model.parameters())
    rec = {"model": args.model, "comm": backend, "n_gpus": world, "batch_per_gpu": args.batch, "steps": args.steps,
           "ms_per_step": round(ms.item(), 3), "items_per_s": round(args.batch * per_sample * world / (ms.item() / 1e3), 1),
           "params": nparam, "param_dtype": args.param_dtype, "wire": args.wire if backend == "flextree" else "native",
           "final_loss": round(torch.stack(losses)[-3:].float().mean().item(), 5)}
    if sampler is not None and rank == 0:
        rec["clocks"] = sampler.window(t0, t1)  # SM clock + throttle reasons sampled during THIS timed loop
    if engine is not None:
        rec["buckets"] = len(engine.buckets)
        rec["device_error"] = comm.error()
    if state is not None:
        rec["hook_calls"] = state.calls
        rec["hook_bytes_per_step"] = state.bytes // max(1, args.steps + args.warmup)
        rec["device_error"] = comm.error()
    flat = torch.cat([p.detach().float().flatten() for p in model.parameters()])
    del ddp, opt
    return rec, flat


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="mlp")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--comm", default="both", choices=["nccl", "flextree", "flextree-dp", "both", "all"])
    ap.add_argument("--wire", default="native", choices=["native", "bf16", "fp8"])
    ap.add_argument("--param-dtype", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--bucket-mb", type=int, default=25)
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--out", default="")
    ap.add_argument("--max-ctas", type=int, default=32)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    torch.backends.cudnn.benchmark = True
    comm = ft.init() if args.comm != "nccl" else None
    results, params = [], {}
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    order = {"both": ["nccl", "flextree"], "all": ["nccl", "flextree", "flextree-dp"]}.get(args.comm, [args.comm])
    for backend in order:
        rec, flat = run(args, backend, comm, dev, rank, world, local, sampler)
        results.append(rec)
        params[backend] = flat
        torch.cuda.empty_cache()
    if "nccl" in params and len(params) > 1:
        a = params["nccl"]
        for r in results:
            if r["comm"] != "nccl":
                b = params[r["comm"]]
                r["param_rel_diff_vs_nccl"] = ((a - b).norm() / a.norm().clamp_min(1e-12)).item()
    if rank == 0:
        for r in results:
            line = json.dumps(r)
            print(line, flush=True)
            if args.out:
                with open(args.out, "a") as f:
                    f.write(line + "\n")
    dist.barrier()
    if comm is not None:
        ft.shutdown()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
