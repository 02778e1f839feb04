#!/usr/bin/env python
"""Tiny single-GPU workload for compute-sanitizer (one tool per run, see `make memcheck` / `make racecheck`):
every kernel family once on 4 emulated ranks with small messages, generous watchdog (the sanitizer slows the
kernels down by orders of magnitude)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["FT_TIMEOUT_S"] = "600"
import flextree as ft  # noqa: E402

n, count = 4, 24 * 1024
comm = ft.Communicator.emulated(n, 0, window_bytes=1 << 20)
comm.set_launch(ctas=2, threads=128, tile_bytes=8192)
for algo, widths, tma in (("stages", [2, 2], False), ("stages", [4], True), ("oneshot", None, None), ("ring", None, None)):
    comm.set_algo(algo)
    comm.set_widths(widths)
    if tma is not None:
        comm.set_tma(tma)
    xs = [torch.full((count if algo != "oneshot" else 2048,), float(r + 1), device="cuda") for r in range(n)]
    comm.allreduce_multi(xs)
    torch.cuda.synchronize()
    assert comm.error() == 0 and torch.all(xs[0] == 10.0), algo
print("sanitize_target ok")
comm.close()
