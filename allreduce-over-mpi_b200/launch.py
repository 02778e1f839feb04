"""``python -m flextree.launch -np N [--cpu] script.py [args...]`` - one process per rank on this host.

What ``mpirun -np N`` + the hostfile were to the reference (allreduce_over_mpi/Makefile:8-24, mpi_config_file): a thin
front end to ``torch.distributed.run`` with single-node defaults that work in containers (rendezvous on 127.0.0.1,
a free port, one OpenMP thread per rank). ``--cpu`` marks the job as CPU-only (gloo ranks for ``HostCommunicator`` /
``HostDataParallel``); otherwise N must not exceed the visible GPUs. ``FT_TOPO`` etc. pass through the environment.
"""
import argparse
import os
import socket
import subprocess
import sys


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def build_command(np_, script, script_args, port=None):
    return [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={np_}",
            "--master-addr", "127.0.0.1", "--master-port", str(port or free_port()), script, *script_args]


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m flextree.launch", description=__doc__.split("\n\n")[0])
    ap.add_argument("-np", "--np", type=int, required=True, dest="np_", help="ranks (processes) on this host")
    ap.add_argument("--cpu", action="store_true", help="CPU ranks: do not check the GPU count, allow the host-only library")
    ap.add_argument("--topo", help="sets FT_TOPO for the job, e.g. 2,4 or 2,3+1")
    ap.add_argument("--port", type=int)
    ap.add_argument("script")
    ap.add_argument("script_args", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    env = dict(os.environ)
    env.setdefault("OMP_NUM_THREADS", "1")
    if a.topo:
        env["FT_TOPO"] = a.topo
    if a.cpu:
        env["FT_ALLOW_HOST_ONLY"] = "1"
    else:
        import torch
        have = torch.cuda.device_count()
        if a.np_ > have:
            ap.error(f"-np {a.np_} but {have} GPU(s) visible (use --cpu for CPU ranks)")
    return subprocess.call(build_command(a.np_, a.script, a.script_args, a.port), env=env)


if __name__ == "__main__":
    sys.exit(main())
