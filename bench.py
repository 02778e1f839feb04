#!/usr/bin/env python
"""Headline benchmark: allreduce bus bandwidth of one large fp32 SUM message (BASELINE.json config
"fp32/bf16 allreduce sweep ... at 2/4/8 B200", metric "allreduce bus GB/s").

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
  python bench.py --impl reference --gpus N ...            (the unmodified reference on the box CPU)
  python bench.py --impl nccl --gpus N ...                 (dist.all_reduce in the same timing loop, N >= 2)

A "step" is ONE allreduce of --bytes (default 256 MiB, > L2) in place on a symmetric buffer through
the public API (flextree.Communicator.allreduce). The buffer is re-filled from a pristine copy between
steps, outside the timed events (the re-fill writes 256 MiB, which also flushes L2). Device time per
step comes from CUDA events on the launching stream; the job time is the MAX over ranks of the summed
step times.

Also in the JSON line, all measured in this process under the same clock record:
  vs_nccl   dist.all_reduce on the same buffers in the same timing loop (N >= 2)
  also      the same step on private (non-symmetric) buffers and in bf16
  e2e       Communicator.allreduce_host: the message starts and ends in pinned HOST memory (the reference's calling
            convention), chunk-pipelined H2D / allreduce / D2H; the WHOLE result is copied back every step
  roofline  fraction of the ceiling of the path that actually ran (NVLS: 2(N-1)/(N+1) x link rate; P2P: link rate)

N == 1: there is no peer, so the same kernels are run for R=8 logical ranks emulated on the one GPU
(one cooperative launch, every "peer" in local HBM) - this measures the kernel path against HBM instead
of NVLink and is labelled as such in `config`.

value = aggregate bus bandwidth over the job's GPUs = n_gpus * busbw,  busbw = (bytes / t) * 2 (R-1) / R
(NCCL definition, R = ranks). At N == 1 that is the one GPU's figure: the bus bandwidth each of its 8
emulated ranks sees (all traffic stays in HBM).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# NCCL (bootstrap only) prints its version banner on stdout when NCCL_DEBUG=VERSION: keep stdout to the one JSON line
if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
    os.environ["NCCL_DEBUG"] = "WARN"

METRIC = "allreduce_aggregate_busbw"
EMULATED_RANKS = 8


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="flextree", choices=["flextree", "reference", "nccl"])
    ap.add_argument("--bytes", type=int, default=256 << 20)
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--algo", default=os.environ.get("FT_BENCH_ALGO", "auto"))
    ap.add_argument("--widths", default=os.environ.get("FT_BENCH_WIDTHS", ""))
    ap.add_argument("--staging", action="store_true", help="private buffers (copy-in path) instead of registered")
    ap.add_argument("--sweep", action="store_true", help="also print a size sweep table to stderr")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-also", action="store_true", help="skip the extra rows (private buffers, bf16, NCCL)")
    return ap.parse_args()


from flextree.utils.clocks import ClockSampler  # noqa: E402


def busbw(nbytes, seconds, ranks):
    return nbytes / seconds * 2.0 * (ranks - 1) / ranks / 1e9 if ranks > 1 else nbytes / seconds / 1e9


# ---------------------------------------------------------------------------------------------------
def run_reference(args):
    """The unmodified reference (baseline/_ref, built by baseline/build_ref.sh) through its own CLI:
    ftmpirun -np R allreduce_over_mpi --size <count> --repeat K+W --comm-type flextree --to-file."""
    n = args.gpus
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0  # the reference is a CPU program: one launcher on rank 0 starts all its ranks
    ranks = n if n > 1 else EMULATED_RANKS
    binary = os.path.join(ROOT, "baseline", "_ref", "bin", "allreduce_over_mpi")
    runner = os.path.join(ROOT, "baseline", "_ref", "bin", "ftmpirun")
    if not (os.path.exists(binary) and os.path.exists(runner)):
        try:
            subprocess.run(["bash", os.path.join(ROOT, "baseline", "build_ref.sh")], check=True, capture_output=True)
        except Exception as e:
            print(json.dumps({"impl": "reference", "unavailable": f"baseline/_ref not built and build failed: {e}"}))
            return 0
    count = args.bytes // 4  # the reference benchmark is fp32 only (benchmark.cpp:119)
    reps = args.steps + args.warmup
    work = os.path.join(ROOT, "gpurun_out", "ref_run")
    os.makedirs(work, exist_ok=True)
    for f in os.listdir(work):
        if f.endswith(".txt"):
            os.remove(os.path.join(work, f))
    env = dict(os.environ)
    env["FT_TOPO"] = args.widths or ""
    env.setdefault("OMP_NUM_THREADS", "14")
    t0 = time.time()
    p = subprocess.run([runner, "-np", str(ranks), binary, "--size", str(count), "--repeat", str(reps),
                        "--comm-type", "flextree", "--to-file", "--tag", "bench"], cwd=work, env=env,
                       capture_output=True, text=True)
    wall = time.time() - t0
    files = [f for f in os.listdir(work) if f.startswith("bench.") and f.endswith(".txt")]
    if p.returncode != 0 or not files:
        print(json.dumps({"impl": "reference", "unavailable": f"reference run failed rc={p.returncode}: {p.stderr[-300:]}"}))
        return 0
    times = [float(x) for x in open(os.path.join(work, files[0])).read().split()]
    timed = times[args.warmup:]
    sec = sum(timed) / len(timed)
    bb = busbw(count * 4, sec, ranks)
    check = [ln for ln in p.stdout.splitlines() if ln.startswith("CHECK 0")]
    out = {
        "impl": "reference", "metric": METRIC, "value": bb * n, "unit": "GB/s", "n_gpus": n,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic (data[i]=i, reference benchmark.cpp:119-123)",
        "busbw_per_rank_GBps": bb, "algbw_GBps": count * 4 / sec / 1e9,
        "config": {"model": "allreduce fp32 SUM in place", "message_bytes": count * 4, "ranks": ranks,
                   "parallelism": f"{ranks} CPU ranks (reference has no GPU path; MPI subset shim over shared memory)",
                   "widths": env["FT_TOPO"] or f"{ranks}", "timing": "reference's own per-repeat MPI_Wtime on rank 0",
                   "wall_s": round(wall, 2)},
        "e2e": {"value": bb * n, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": "host-resident data: the reference's timed region is already end to end"},
        "gpu_launches": 0,
        "check": check[0] if check else None,
    }
    print(json.dumps(out))
    return 0


# ---------------------------------------------------------------------------------------------------
LINK_GBPS = 770.0   # measured one-directional peer copy rate of one B200 over NVLink 5 (B200_PROFILING.md)


def path_ceiling(algo, ranks):
    """busbw ceiling (NCCL definition, 2(N-1)/N x algbw) of the path that ran, at LINK_GBPS per direction:
    two-phase P2P (stages, ring) moves 2(N-1)/N S per direction -> LINK; NVLS moves (1+1/N) S -> 2(N-1)/(N+1) LINK;
    one-shot moves (N-1) S -> 2/N LINK; two-shot LL doubles the bytes of the two-phase exchange -> LINK/2."""
    if algo == "nvls":
        return 2.0 * (ranks - 1) / (ranks + 1) * LINK_GBPS, "NVLS: (1+1/N) S per direction"
    if algo == "oneshot":
        return 2.0 / ranks * LINK_GBPS, "one-shot: (N-1) S per direction"
    if algo == "twoshot":
        return LINK_GBPS / 2.0, "two-shot LL: 2 x 2(N-1)/N S per direction (16-byte lines carry 8)"
    return LINK_GBPS, "two-phase P2P: 2(N-1)/N S per direction"


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    # stdout must carry exactly one JSON line: native libraries (NCCL's version banner) also write to fd 1,
    # so everything except that line is sent to stderr
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(os.dup(2), "w")

    import torch
    import torch.distributed as dist
    import flextree as ft

    n = args.gpus
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if n > 1 and world != n:
        print(f"bench.py: --gpus {n} needs a torch.distributed.run launch with {n} ranks (WORLD_SIZE={world})", file=sys.stderr)
        return 2
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    emulated = n == 1
    ranks = EMULATED_RANKS if emulated else n
    nloc = ranks if emulated else 1
    nbytes = args.bytes // 16 * 16

    if not emulated:
        dist.init_process_group("nccl", device_id=dev)
        comm = ft.Communicator.from_process_group(None, local_rank)
    else:
        comm = ft.Communicator.emulated(ranks, local_rank)
    if args.algo != "auto":
        comm.set_algo(args.algo)
    if args.widths:
        comm.set_widths([int(x) for x in args.widths.replace("x", ",").split(",")])

    use_nccl = args.impl == "nccl"
    if use_nccl and emulated:
        os.write(real_stdout, (json.dumps({"impl": "nccl", "unavailable": "NCCL has nothing to do on one GPU"}) + "\n").encode())
        return 0

    def barrier():
        torch.cuda.synchronize()
        if not emulated:
            dist.barrier()
            torch.cuda.synchronize()

    class Arm:
        """One measured configuration: dtype x buffer kind x implementation, same timing loop for all."""

        def __init__(self, dtype, staging, nccl):
            self.tdt = torch.float32 if dtype == "fp32" else torch.bfloat16
            self.dtype, self.staging, self.nccl = dtype, staging, nccl
            esize = 4 if dtype == "fp32" else 2
            self.count = nbytes // esize
            gen = torch.Generator(device=dev).manual_seed(1234 + rank)
            self.pristine = [torch.randn(self.count, device=dev, dtype=torch.float32, generator=gen).to(self.tdt) for _ in range(nloc)]
            if staging or nccl:
                self.bufs = [torch.empty(self.count, device=dev, dtype=self.tdt) for _ in range(nloc)]
                if not nccl:
                    comm.reserve(nbytes)
            else:
                b = comm.symm_empty(self.count, self.tdt)
                self.bufs = b if emulated else [b]

        def refill(self):
            for b_, p_ in zip(self.bufs, self.pristine):
                b_.copy_(p_)

        def step(self):
            if self.nccl:
                dist.all_reduce(self.bufs[0])      # the library baseline: same buffers, same timing loop
            elif emulated:
                comm.allreduce_multi(self.bufs)
            else:
                comm.allreduce(self.bufs[0])

        def check(self):
            """Correctness gate on one step, against fp64 torch math (NCCL fp64 sum across ranks)."""
            self.refill()
            self.step()
            torch.cuda.synchronize()
            if not self.nccl and comm.error():
                raise RuntimeError(f"device error {comm.error():#x}")
            if emulated:
                want = torch.stack([p.double() for p in self.pristine]).sum(0)
            else:
                want = self.pristine[0].double().clone()
                dist.all_reduce(want)
            tol = 1e-4 if self.tdt == torch.float32 else 0.25
            err = (self.bufs[0].double() - want).abs().max().item()
            if not err <= tol * max(1.0, want.abs().max().item()):
                raise RuntimeError(f"WRONG RESULT max abs err {err}")

        def time(self, steps, warmup):
            """(ms per step = max over ranks of the summed CUDA-event times / steps, wall window of the loop)"""
            for _ in range(warmup):
                self.refill()
                self.step()
            barrier()
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            barrier()
            t0 = time.perf_counter()
            for i in range(steps):
                self.refill()          # untimed: restores the input and evicts the previous result from L2
                ev[i][0].record()
                self.step()
                ev[i][1].record()
            barrier()
            t1 = time.perf_counter()
            total = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], device=dev, dtype=torch.float64)
            if not emulated:
                dist.all_reduce(total, op=dist.ReduceOp.MAX)
            return total.item() / steps, (t0, t1)

        def free(self):
            self.bufs = self.pristine = None

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    windows = []

    main_arm = Arm(args.dtype, args.staging, use_nccl)
    try:
        main_arm.check()
    except RuntimeError as e:
        print(f"bench.py: {e}", file=sys.stderr)
        return 4
    ms_per_step, win = main_arm.time(args.steps, args.warmup)
    windows.append(win)
    t_wall = win[1] - win[0]
    info = {"algo": "nccl (torch.distributed.all_reduce)"} if use_nccl else comm.last_call()
    bb = busbw(nbytes, ms_per_step / 1e3, ranks)

    # ---- the same step in the other configurations, same process, same clock record -------------------------
    also, vs_nccl = [], None
    if not args.no_also and not use_nccl:
        extra = [(args.dtype, not args.staging, False), ("bf16" if args.dtype == "fp32" else "fp32", args.staging, False)]
        for dtype, staging, nccl in extra:
            try:
                arm = Arm(dtype, staging, nccl)
                arm.check()
                ms, w = arm.time(args.steps, args.warmup)
                windows.append(w)
                call = comm.last_call()
                ceil, what = path_ceiling(call["algo"], ranks)
                row = {"dtype": dtype, "buffers": "private (copy-in)" if staging else "symmetric (zero-copy)", "ms_per_step": ms,
                       "busbw_per_rank_GBps": busbw(nbytes, ms / 1e3, ranks), "algo": call["algo"], "widths": call["widths"],
                       "ctas": call["ctas"]}
                if not emulated:
                    row["roofline_fraction"] = round(row["busbw_per_rank_GBps"] / ceil, 4)
                also.append(row)
                arm.free()
            except Exception as e:  # an extra row must never cost the headline number
                also.append({"dtype": dtype, "buffers": "private" if staging else "symmetric", "error": str(e)[:200]})
        if not emulated:
            nccl_rows = []
            for dtype in (args.dtype, "bf16" if args.dtype == "fp32" else "fp32"):
                arm = Arm(dtype, True, True)
                arm.check()
                ms, w = arm.time(args.steps, args.warmup)
                windows.append(w)
                nccl_rows.append({"dtype": dtype, "ms_per_step": ms, "busbw_per_rank_GBps": busbw(nbytes, ms / 1e3, ranks)})
                arm.free()
            ours = {args.dtype: ms_per_step}
            for r_ in also:
                if "ms_per_step" in r_ and r_["buffers"].startswith("symmetric" if not args.staging else "private"):
                    ours.setdefault(r_["dtype"], r_["ms_per_step"])
            vs_nccl = {"impl": "torch.distributed.all_reduce (NCCL %s), same buffers size, same loop" % ".".join(map(str, torch.cuda.nccl.version())),
                       "ms_per_step": nccl_rows[0]["ms_per_step"], "busbw_per_rank_GBps": nccl_rows[0]["busbw_per_rank_GBps"],
                       "ratio": nccl_rows[0]["ms_per_step"] / ms_per_step, "rows": nccl_rows,
                       "ratio_private_buffers": next((nccl_rows[0]["ms_per_step"] / r_["ms_per_step"] for r_ in also
                                                      if r_.get("dtype") == args.dtype and str(r_.get("buffers", "")).startswith("private") and "ms_per_step" in r_), None),
                       "ratio_bf16": (nccl_rows[1]["ms_per_step"] / ours["bf16"]) if "bf16" in ours and nccl_rows[1]["dtype"] == "bf16" else None}

    clocks = None
    if rank == 0:
        lo = min(w[0] for w in windows)
        hi = max(w[1] for w in windows)
        clocks = sampler.stop(lo, hi)

    # ---- end to end through the public API: the message starts and ends in pinned HOST memory ----------------
    # Communicator.allreduce_host = ft_allreduce_host_multi: chunked H2D -> in-place allreduce -> D2H over three
    # streams; every step copies the whole input up and the WHOLE result back. Timed on the host clock around K
    # synchronous calls (the API returns when the result is in host memory), barrier + synchronize on both sides,
    # max over ranks.
    e2e = None
    if not use_nccl:
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        tdt = main_arm.tdt
        host_in = [torch.randn(main_arm.count, dtype=torch.float32).to(tdt).pin_memory() for _ in range(nloc)]
        host_out = [torch.empty(main_arm.count, dtype=tdt).pin_memory() for _ in range(nloc)]

        def e2e_step():
            if emulated:
                comm.allreduce_host(host_in, outs=host_out)
            else:
                comm.allreduce_host(host_in[0], outs=host_out[0])

        for _ in range(2):
            e2e_step()
        # gate: the host result equals the fp64 sum of the host inputs
        if emulated:
            want = torch.stack([h.double() for h in host_in]).sum(0)
        else:
            want = host_in[0].double().to(dev)
            dist.all_reduce(want)
            want = want.cpu()
        err = (host_out[0].double() - want).abs().max().item()
        tol = 1e-4 if tdt == torch.float32 else 0.25
        if not err <= tol * max(1.0, want.abs().max().item()):
            print(f"bench.py: WRONG e2e RESULT max abs err {err}", file=sys.stderr)
            return 5
        del want
        barrier()
        # device clock: each call returns only when its last D2H chunk has landed (it synchronises its output stream), so
        # two events recorded on the idle current stream before and after the K calls bracket all of their device work
        ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        ea.record()
        for _ in range(e2e_steps):
            e2e_step()
        eb.record()
        torch.cuda.synchronize()
        dt_wall = time.perf_counter() - t0
        barrier()
        e2e_ms = torch.tensor([ea.elapsed_time(eb) / e2e_steps], device=dev, dtype=torch.float64)
        if not emulated:
            dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
        e2e_bb = busbw(nbytes, e2e_ms.item() / 1e3, ranks)
        e2e = {"value": e2e_bb * n, "unit": "GB/s", "ms_per_step": e2e_ms.item(), "steps": e2e_steps,
               "h2d_bytes_per_step": nbytes * nloc, "d2h_bytes_per_step": nbytes * nloc,
               "pcie_GBps_per_direction": nbytes * nloc / (e2e_ms.item() / 1e3) / 1e9,
               "api": "Communicator.allreduce_host(pinned_in, pinned_out): 16 MiB chunks pipelined H2D / zero-copy allreduce / D2H "
                      "over three streams; whole result copied back; CUDA events around K synchronous calls, max over ranks",
               "wall_ms_per_step_rank0": round(dt_wall * 1e3 / e2e_steps, 3)}
        del host_in, host_out

    sweep = None
    if args.sweep:
        sweep = []
        esize = 4 if args.dtype == "fp32" else 2
        for sz in [1 << 10, 8 << 10, 64 << 10, 512 << 10, 4 << 20, 32 << 20, nbytes]:
            cnt = sz // esize
            views = [b_[:cnt] for b_ in main_arm.bufs]
            def one():
                if emulated:
                    comm.allreduce_multi(views)
                else:
                    comm.allreduce(views[0])
            for _ in range(5):
                one()
            barrier()
            a_, b2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 20
            a_.record()
            for _ in range(reps):
                one()
            b2.record()
            barrier()
            t = torch.tensor([a_.elapsed_time(b2) / reps], device=dev, dtype=torch.float64)
            if not emulated:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            sweep.append({"bytes": sz, "us": round(t.item() * 1e3, 2), "busbw_GBps": round(busbw(sz, t.item() / 1e3, ranks), 2),
                          "algo": comm.last_call()["algo"]})

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        if emulated:
            # One GPU: every "link" is HBM. The floor is the compulsory traffic - each logical rank's input is read once
            # and its output written once, 2 R S - at the measured copy rate. (The kernels issue (3R-1) S of loads and
            # stores for a flat plan on registered buffers; the all-gather re-reads largely hit in L2: ncu counts ~21 S
            # of DRAM traffic at 64 MiB, profiles/ncu_staged_tma_emulated.md.)
            roof = "HBM, single-GPU emulation: compulsory bytes 2*R*S at the measured copy rate (MEASURED_PEAKS.json hbm_gbs)"
            frac = (2 * ranks * nbytes / (ms_per_step / 1e3) / 1e9) / peaks.get("hbm_gbs", 6553.9)
            ceiling = None
        else:
            ceiling, what = path_ceiling(info.get("algo"), ranks)
            roof = f"NVLink {LINK_GBPS:.0f} GB/s per direction (measured peer copy); path ceiling {ceiling:.0f} GB/s busbw ({what})"
            frac = bb / ceiling
        out = {
            "metric": METRIC, "value": bb * n, "unit": "GB/s", "n_gpus": n, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic (randn, fixed seed)",
            "busbw_per_rank_GBps": bb, "algbw_GBps": nbytes / (ms_per_step / 1e3) / 1e9,
            "roofline": {"against": roof, "fraction": round(frac, 4), "ceiling_busbw_GBps": ceiling,
                         "link_bytes_per_direction": None if emulated else (
                             (1 + 1.0 / ranks) * nbytes if info.get("algo") == "nvls" else 2.0 * (ranks - 1) / ranks * nbytes),
                         "link_GBps_per_direction": None if emulated else round(
                             ((1 + 1.0 / ranks) if info.get("algo") == "nvls" else 2.0 * (ranks - 1) / ranks) * nbytes / (ms_per_step / 1e3) / 1e9, 1)},
            "config": {"model": f"allreduce {args.dtype} SUM in place", "message_bytes": nbytes, "ranks": ranks,
                       "global_batch": None, "seq_len": None,
                       "parallelism": (f"{ranks} logical ranks emulated on 1 GPU (one cooperative launch, HBM-bound, no NVLink)"
                                       if emulated else f"dp{n}: one process per GPU, NVLink P2P/NVLS kernels, no NCCL on the path"),
                       "buffers": "private (copy-in)" if args.staging else "symmetric (zero-copy)",
                       "l2": "message > L2 and re-filled (256 MiB written) between timed steps",
                       "call": info, "wall_s_timed_loop": round(t_wall, 3)},
            "clocks": clocks,
            "e2e": e2e,
            "gpu_launches": 0 if use_nccl else args.steps,
            "impl": args.impl,
        }
        if vs_nccl:
            out["vs_nccl"] = vs_nccl
        if also:
            out["also"] = also
        if sweep:
            out["sweep"] = sweep
        os.write(real_stdout, (json.dumps(out) + "\n").encode())
    comm.close()
    if not emulated:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
