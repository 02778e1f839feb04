#!/usr/bin/env python
"""BASELINE.json config #1 on the CPU: "allreduce_over_mpi benchmark: mpirun -np 6, 1M fp32 sum, widths 2x3".

Four arms on the same machine, same world size, same widths, same message, run round-robin --rounds times:
(a) the UNMODIFIED reference through the in-repo MPI shim (its own benchmark binary, its own timing);
(b) "dropin": the same unmodified benchmark.cpp built against compat/mpi_mod.hpp (this library's CPU data plane);
(c) flextree.HostCommunicator on private numpy buffers; (d) the same on the registered (zero-copy) region.

  python bench/cpu_compare.py [--np 6] [--widths 2,3] [--size 1000000] [--repeat 20] [--rounds 3]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, time, json
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, {root!r})
import flextree as ft
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
comm = ft.HostCommunicator(window_bytes={win})
count, widths, repeat, ctas = {count}, {widths}, {repeat}, {ctas}
x = np.arange(count, dtype=np.float32)
buf = comm.registered(count, np.float32) if {registered} else x.copy()
times, times0 = [], []
tok = np.zeros(4, dtype=np.float32)
for i in range(repeat + 3):
    buf[:] = x
    comm.allreduce(tok, widths=widths, ctas=1)  # shared-memory barrier (the reference's MPI_Barrier is one too)
    t0 = time.perf_counter()
    comm.allreduce(buf, algo={algo!r}, widths=widths, ctas=ctas, tile_bytes={tile}, registered={registered})
    t1 = time.perf_counter()
    if i >= 3:
        times0.append(t1 - t0)
# max over ranks per repeat, taken after the loop (a gloo collective between repeats wakes its I/O threads and
# shows up in the next repeat)
t = torch.tensor(times0, dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
times = t.tolist()
assert buf[9] == 9.0 * world and buf[count - 1] == float(count - 1) * world, (buf[9], buf[count - 1])
if rank == 0:
    print("FT_RESULT " + json.dumps({{"avg_s": sum(times) / len(times), "min_s": min(times), "rank0_avg_s": sum(times0) / len(times0), "rank0_min_s": min(times0)}}))
comm.close()
dist.destroy_process_group()
'''


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--np", type=int, default=6)
    ap.add_argument("--widths", default="2,3")
    ap.add_argument("--size", type=int, default=1000000)
    ap.add_argument("--repeat", type=int, default=20)
    ap.add_argument("--ctas", type=int, default=1)
    ap.add_argument("--tile", type=int, default=262144)
    ap.add_argument("--algo", default="stages_bulk", help="stages_bulk (vectorised chunk tiles) or stages (the GPU text, one vector at a time)")
    ap.add_argument("--rounds", type=int, default=3, help="all arms are run this many times, interleaved; avg = median of the rounds")
    ap.add_argument("--port", type=int, default=29931)
    args = ap.parse_args()
    out = {"config": f"np={args.np}, {args.size} fp32 SUM in place, widths {args.widths}", "cpus": os.cpu_count()}

    ref_bin = os.path.join(ROOT, "baseline", "_ref", "bin")
    if not os.path.exists(os.path.join(ref_bin, "allreduce_over_mpi")):
        subprocess.run(["bash", os.path.join(ROOT, "baseline", "build_ref.sh")], capture_output=True)
    rounds = []
    for rnd in range(args.rounds):  # arms interleaved round by round: a load burst on a shared box hits all of them
        cur = {}
        # the reference's unmodified benchmark.cpp twice: with its own mpi_mod.hpp, and with compat/mpi_mod.hpp (this
        # library's CPU data plane behind the same signature) - same program, same clock, same launcher
        for key, exe, note in (("reference", "allreduce_over_mpi", f"{args.np} ranks x 14 OpenMP threads (hard-coded in mpi_mod.hpp:253)"),
                               ("dropin", "allreduce_over_mpi_dropin", f"{args.np} ranks x 1 thread; benchmark.cpp unmodified, compat/mpi_mod.hpp")):
            with tempfile.TemporaryDirectory() as tmp:
                env = dict(os.environ, FT_TOPO=args.widths)
                p = subprocess.run([os.path.join(ref_bin, "ftmpirun"), "-np", str(args.np), os.path.join(ref_bin, exe),
                                    "--size", str(args.size), "--repeat", str(args.repeat + 3), "--to-file", "--tag", "cmp"],
                                   cwd=tmp, env=env, capture_output=True, text=True)
                files = [f for f in os.listdir(tmp) if f.startswith("cmp.")]
                if p.returncode == 0 and files:
                    t = [float(v) for v in open(os.path.join(tmp, files[0])).read().split()][3:]
                    cur[key] = {"avg_s": sum(t) / len(t), "min_s": min(t), "timing": "rank-0 MPI_Wtime per repeat (benchmark.cpp's own method)",
                                "threads": note}
                else:
                    cur[key] = {"unavailable": p.stderr[-200:]}

        widths = [int(v) for v in args.widths.split(",")]
        for registered in (False, True):
            code = WORKER.format(root=ROOT, win=max(8 << 20, args.size * 8), count=args.size, widths=widths, repeat=args.repeat,
                                 ctas=args.ctas, tile=args.tile, algo=args.algo, registered=registered)
            with tempfile.NamedTemporaryFile("w", suffix=".py", delete=False) as f:
                f.write(code)
                script = f.name
            p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.np}",
                                "--master-addr", "127.0.0.1", "--master-port", str(args.port + 2 * rnd + int(registered)), script],
                               capture_output=True, text=True, env=dict(os.environ, OMP_NUM_THREADS="1", FT_ALLOW_HOST_ONLY="1"))
            os.unlink(script)
            key = "flextree_cpu_zero_copy" if registered else "flextree_cpu"
            line = [ln for ln in p.stdout.splitlines() if ln.startswith("FT_RESULT ")]
            if line:
                cur[key] = json.loads(line[0][len("FT_RESULT "):])
                cur[key]["timing"] = "wall clock around the call after a shared-memory barrier: max over ranks, and rank 0 alone (the reference's method)"
                cur[key]["threads"] = f"{args.np} ranks x {args.ctas} threads"
            else:
                cur[key] = {"unavailable": (p.stdout + p.stderr)[-300:]}
        rounds.append(cur)
    import statistics
    for key in rounds[0]:
        ok = [r[key] for r in rounds if "avg_s" in r.get(key, {})]
        if not ok:
            out[key] = rounds[0][key]
            continue
        out[key] = dict(ok[0])
        out[key]["avg_s"] = statistics.median(o["avg_s"] for o in ok)  # median over rounds of the per-round average
        out[key]["min_s"] = min(o["min_s"] for o in ok)
        out[key]["avg_s_rounds"] = [o["avg_s"] for o in ok]
        out[key].pop("rank0_avg_s", None)
        out[key].pop("rank0_min_s", None)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
