"""Fault injection across real GPUs: the last rank never joins a collective. The other ranks' kernels must
give up after the watchdog timeout (sticky error word) instead of hanging the GPU. Run as its own job: the
communicator is unusable afterwards (epochs diverge), exactly like after a peer crash."""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["FT_TIMEOUT_S"] = "2"
import flextree as ft  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = ft.init()
    x = torch.ones(1 << 16, device=dev)
    comm.set_algo("stages")
    comm.allreduce(x)                      # a healthy call first
    torch.cuda.synchronize()
    assert comm.error() == 0 and torch.all(x == world)
    dist.barrier()
    t0 = time.time()
    if rank != world - 1:                  # the last rank "dies"
        comm.allreduce(x)
        torch.cuda.synchronize()
        dt = time.time() - t0
        e = comm.error()
        assert e != 0 and (e & 0xFF) == 1, hex(e)            # ERR_TIMEOUT
        assert dt < 15, dt                                    # bounded, not a hang
        print(f"[rank {rank}] watchdog fired after {dt:.1f}s: error word {e:#x} "
              f"(phase {(e >> 8) & 0xff}, waiting for rank {(e >> 16) & 0xff})", flush=True)
    dist.barrier()
    if rank == 0:
        print("MP_FAULT_OK", flush=True)
    dist.destroy_process_group()
    os._exit(0)  # skip the collective teardown of a communicator with a missing peer


if __name__ == "__main__":
    main()
