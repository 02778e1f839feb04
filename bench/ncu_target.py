#!/usr/bin/env python
"""Small single-GPU program for Nsight Compute: the flagship kernel (TMA-staged multi-stage allreduce,
8 logical ranks emulated on one GPU) on a 64 MiB fp32 message, plus one one-shot call.
Run it plain first (must exit 0), then under ncu (see profiles/README.md)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("FT_TIMEOUT_S", "5")
import flextree as ft  # noqa: E402

n, count = 8, (64 << 20) // 4
comm = ft.Communicator.emulated(n, 0)
bufs = comm.symm_empty(count, torch.float32)
for b in bufs:
    b.normal_()
comm.set_algo("stages")
for widths in ([8], [2, 4]):
    comm.set_widths(widths)
    for _ in range(3):
        comm.allreduce_multi(bufs)
torch.cuda.synchronize()
comm.set_widths(None)
comm.set_algo("oneshot")
small = [b[:4096] for b in bufs]
for _ in range(3):
    comm.allreduce_multi(small)
torch.cuda.synchronize()
assert comm.error() == 0
print("ncu_target ok", comm.last_call())
comm.close()
