#!/usr/bin/env python
"""Small single-GPU program for Nsight Compute: one launch of every kernel family that can run on one GPU (8 logical
ranks emulated, every "peer" in local HBM): TMA-staged stages (flat and 2x4), the cluster-pair variant, register-staged
stages, ring, one-shot LL, two-shot LL, the block-scaled fp8 wire. (The NVLS kernels need a real multicast object: their
evidence is profiles/nvls_limit.md.)  Run it plain first (must exit 0), then under ncu (see profiles/README.md)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("FT_TIMEOUT_S", "20")
import flextree as ft  # noqa: E402

n, count = 8, (64 << 20) // 4
comm = ft.Communicator.emulated(n, 0)
bufs = comm.symm_empty(count, torch.float32)
for b in bufs:
    b.normal_()
reps = int(os.environ.get("FT_NCU_REPS", "1"))


def run(label, tensors, **kw):
    for _ in range(reps):
        comm.allreduce_multi(tensors, **kw)
    torch.cuda.synchronize()
    assert comm.error() == 0, label
    print("ncu_target:", label, comm.last_call(), flush=True)


comm.set_algo("stages")
comm.set_tma(True)
for widths in ([8], [2, 4]):
    comm.set_widths(widths)
    run(f"tma stages {widths} 64 MiB", bufs)
comm.set_widths([8])
comm.set_tuning("cluster", 1)
run("cluster pair per tile [8] 64 MiB", bufs)
comm.set_tuning("cluster", 0)
comm.set_tma(False)
run("register-staged stages [8] 64 MiB", bufs)
comm.set_tma(None)
priv = [torch.randn(count // 4, device="cuda") for _ in range(n)]
comm.set_widths([8])
run("fp8 block-scaled wire, stages [8], 16 MiB fp32 private", priv, wire_dtype="fp8")
comm.set_widths(None)
comm.set_algo("ring")
run("ring 16 MiB", [b[: count // 4] for b in bufs])
comm.set_algo("oneshot")
run("one-shot LL 16 KiB", [b[:4096] for b in bufs])
comm.set_algo("twoshot")
run("two-shot LL 1 MiB", [b[: 1 << 18] for b in bufs])
print("ncu_target ok")
comm.close()
