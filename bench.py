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
    return ap.parse_args()


class ClockSampler:
    """SM clock + throttle reasons DURING the measured region (B200_PROFILING.md). NVML is polled in-process
    every ~2 ms from a background thread (an `nvidia-smi -lms` child cannot deliver several samples inside a
    timed loop that lasts a few tens of milliseconds); samples are time-stamped and the ones that fall inside
    the timed loop are the ones reported. Falls back to one-shot nvidia-smi queries without pynvml."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
               ("sw_power_cap", 0x4), ("hw_power_brake", 0x80))

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.samples = []   # (t, sm_mhz, max_mhz, power_w, reasons_bitmask)
        self._stop = threading.Event()
        self._t = None
        self._nvml = None
        self._h = None

    def _run(self):
        nv, h = self._nvml, self._h
        try:
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        except Exception:
            mx = None
        while not self._stop.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                except Exception:
                    pw = None
                self.samples.append((time.perf_counter(), float(sm), float(mx) if mx else None, pw, int(rs)))
            except Exception:
                pass
            self._stop.wait(0.002)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            # map the CUDA ordinal to NVML through the PCI bus id (CUDA_VISIBLE_DEVICES may reorder)
            import torch
            bus = torch.cuda.get_device_properties(self.idx).pci_bus_id if hasattr(torch.cuda.get_device_properties(self.idx), "pci_bus_id") else None
            self._h = None
            if bus is not None:
                for i in range(nv.nvmlDeviceGetCount()):
                    hh = nv.nvmlDeviceGetHandleByIndex(i)
                    if nv.nvmlDeviceGetPciInfo(hh).bus == bus:
                        self._h = hh
                        break
            if self._h is None:
                self._h = nv.nvmlDeviceGetHandleByIndex(self.idx)
            self._nvml = nv
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        except Exception:
            self._nvml = None

    def stop(self, t_begin=None, t_end=None):
        self._stop.set()
        if self._t:
            self._t.join(timeout=2)
        if self._nvml is None:   # fallback: one nvidia-smi query (not "during", flagged as such)
            try:
                out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm", "--format=csv,noheader,nounits",
                                      "-i", str(self.idx)], capture_output=True, text=True, timeout=5).stdout.split(",")
                return {"sm_mhz": float(out[0]), "sm_max_mhz": float(out[1]), "reasons": [], "samples": 1,
                        "samples_in_timed_loop": 0, "window": "after the run (pynvml unavailable)"}
            except Exception:
                return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        timed = [x for x in self.samples if t_begin is not None and t_begin <= x[0] <= t_end]
        use = timed if len(timed) >= 3 else self.samples
        sm = sorted(x[1] for x in use)
        mx = [x[2] for x in use if x[2]]
        pw = [x[3] for x in use if x[3] is not None]
        mask = 0
        for x in use:
            mask |= x[4]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": round(max(pw), 1) if pw else None,
                "reasons": [name for name, bit in self.REASONS if mask & bit], "samples": len(use),
                "samples_in_timed_loop": len(timed), "window": "timed loop" if len(timed) >= 3 else "warm-up + timed loop"}


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
    tdt = torch.float32 if args.dtype == "fp32" else torch.bfloat16
    esize = 4 if args.dtype == "fp32" else 2
    count = args.bytes // esize
    nbytes = count * esize
    emulated = n == 1
    ranks = EMULATED_RANKS if emulated else n

    if not emulated:
        dist.init_process_group("nccl", device_id=dev)
        comm = ft.Communicator.from_process_group(None, local_rank)
    else:
        comm = ft.Communicator.emulated(ranks, local_rank)
    if args.algo != "auto":
        comm.set_algo(args.algo)
    if args.widths:
        comm.set_widths([int(x) for x in args.widths.replace("x", ",").split(",")])

    nloc = ranks if emulated else 1
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    pristine = [torch.randn(count, device=dev, dtype=torch.float32, generator=gen).to(tdt) for _ in range(nloc)]
    if args.staging:
        bufs = [torch.empty(count, device=dev, dtype=tdt) for _ in range(nloc)]
    else:
        b = comm.symm_empty(count, tdt)
        bufs = b if emulated else [b]
    if args.staging:
        comm.reserve(nbytes)

    def refill():
        for b_, p_ in zip(bufs, pristine):
            b_.copy_(p_)

    use_nccl = args.impl == "nccl"
    if use_nccl and emulated:
        os.write(real_stdout, (json.dumps({"impl": "nccl", "unavailable": "NCCL has nothing to do on one GPU"}) + "\n").encode())
        return 0

    def step(stream=None):
        if use_nccl:
            dist.all_reduce(bufs[0])          # the library baseline: same buffers, same timing loop
        elif emulated:
            comm.allreduce_multi(bufs, stream=stream)
        else:
            comm.allreduce(bufs[0], stream=stream)

    def barrier():
        torch.cuda.synchronize()
        if not emulated:
            dist.barrier()
            torch.cuda.synchronize()

    # correctness gate on the first step (against fp64 torch math)
    refill()
    step()
    torch.cuda.synchronize()
    if comm.error():
        print(f"bench.py: device error {comm.error():#x}", file=sys.stderr)
        return 3
    if emulated:
        want = torch.stack([p.double() for p in pristine]).sum(0)
    else:
        want = pristine[0].double().clone()
        dist.all_reduce(want)
    tol = 1e-4 if tdt == torch.float32 else 0.25
    err = (bufs[0].double() - want).abs().max().item()
    if not err <= tol * max(1.0, want.abs().max().item()):
        print(f"bench.py: WRONG RESULT max abs err {err}", file=sys.stderr)
        return 4
    del want

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        refill()
        step()
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        refill()          # untimed: restores the input and evicts the previous result from L2
        ev[i][0].record()
        step()
        ev[i][1].record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop(t_wall0, t_wall0 + t_wall) if rank == 0 else None
    ms = [a.elapsed_time(b) for a, b in ev]
    total_ms = torch.tensor([sum(ms)], device=dev, dtype=torch.float64)
    if not emulated:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    ms_per_step = total_ms.item() / args.steps
    info = {"algo": "nccl (torch.distributed.all_reduce)"} if use_nccl else comm.last_call()
    bb = busbw(nbytes, ms_per_step / 1e3, ranks)

    # ---- end to end through the public API: pinned host -> device, allreduce, result back to host ---------
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    host_in = [torch.randn(count, dtype=torch.float32).to(tdt).pin_memory() for _ in range(nloc)]
    host_out = torch.empty(11, dtype=tdt).pin_memory()
    for _ in range(2):
        for b_, h_ in zip(bufs, host_in):
            b_.copy_(h_, non_blocking=True)
        step()
        host_out.copy_(bufs[0][9:20], non_blocking=True)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(e2e_steps):
        for b_, h_ in zip(bufs, host_in):
            b_.copy_(h_, non_blocking=True)                       # H2D of this step's input
        step()
        host_out.copy_(bufs[0][9:20], non_blocking=True)          # D2H of the result check (reference CHECK 9..19)
        torch.cuda.current_stream().synchronize()
    e1.record()
    barrier()
    e2e_ms = torch.tensor([e0.elapsed_time(e1) / e2e_steps], device=dev, dtype=torch.float64)
    if not emulated:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_bb = busbw(nbytes, e2e_ms.item() / 1e3, ranks)

    sweep = None
    if args.sweep:
        sweep = []
        for sz in [1 << 10, 8 << 10, 64 << 10, 512 << 10, 4 << 20, 32 << 20, nbytes]:
            cnt = sz // esize
            views = [b_[:cnt] for b_ in bufs]
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
            # HBM roofline of the emulation: copy-in none (registered) ; RS reads R*S/R.. total traffic below
            roof = "HBM (single GPU emulation)"
            # bytes moved per GPU-wide step: RS reads R*S, writes S ; AG reads S*(R... ) writes R*S  ~ (2R+1..2)S
            hbm_bytes = (2 * ranks + 2) * nbytes if not args.staging else (2 * ranks + 4) * nbytes
            frac = (hbm_bytes / (ms_per_step / 1e3) / 1e9) / peaks.get("hbm_gbs", 6650.0)
        else:
            roof = "NVLink 770 GB/s per direction (measured peer copy, B200_PROFILING.md)"
            frac = bb / 770.0
        out = {
            "metric": METRIC, "value": bb * n, "unit": "GB/s", "n_gpus": n, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic (randn, fixed seed)",
            "busbw_per_rank_GBps": bb, "algbw_GBps": nbytes / (ms_per_step / 1e3) / 1e9,
            "roofline": {"against": roof, "fraction": round(frac, 4)},
            "config": {"model": f"allreduce {args.dtype} SUM in place", "message_bytes": nbytes, "ranks": ranks,
                       "global_batch": None, "seq_len": None,
                       "parallelism": (f"{ranks} logical ranks emulated on 1 GPU (one cooperative launch, HBM-bound, no NVLink)"
                                       if emulated else f"dp{n}: one process per GPU, NVLink P2P/NVLS kernels, no NCCL on the path"),
                       "buffers": "private (copy-in)" if args.staging else "symmetric (zero-copy)",
                       "l2": "message > L2 and re-filled (256 MiB written) between timed steps",
                       "call": info, "wall_s_timed_loop": round(t_wall, 3)},
            "clocks": clocks,
            "e2e": {"value": e2e_bb * n, "unit": "GB/s", "ms_per_step": e2e_ms.item(), "steps": e2e_steps,
                    "h2d_bytes_per_step": nbytes * nloc, "d2h_bytes_per_step": 11 * esize,
                    "api": "Communicator.allreduce on a symmetric tensor; input copied from pinned host memory each step"},
            "gpu_launches": 0 if use_nccl else args.steps,
            "impl": args.impl,
        }
        if sweep:
            out["sweep"] = sweep
        os.write(real_stdout, (json.dumps(out) + "\n").encode())
    comm.close()
    if not emulated:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
