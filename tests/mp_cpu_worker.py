"""CPU multi-process worker (torch.distributed gloo): the algorithm templates across real processes with
shared-memory windows, checked against gloo's all_reduce. Launched by tests/test_multiprocess_cpu.py."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flextree as ft  # noqa: E402


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    comm = ft.HostCommunicator()
    assert comm.rank == rank and comm.size == world
    g = torch.Generator().manual_seed(7 + rank)

    def ref_sum(t, op=dist.ReduceOp.SUM):
        r = t.clone()
        dist.all_reduce(r, op=op)
        return r

    # every factorization of the world size, a few sizes, in place / out of place
    for widths in ft.enumerate_widths(world):
        for count in (1, 7, 1000, 20001):
            x = torch.randn(count, generator=g)
            ref = ref_sum(x.double()).float()
            out = torch.zeros(count)
            comm.allreduce(x, out=out, widths=widths, ctas=3, tile_bytes=1024)
            assert torch.allclose(out, ref, rtol=1e-5, atol=1e-5), (widths, count)
            y = x.clone()
            comm.allreduce(y, widths=widths, ctas=2, tile_bytes=4096)
            assert torch.equal(y, out)
            chk = y.clone()
            dist.broadcast(chk, 0)
            assert torch.equal(chk, y), "ranks differ"
    # bulk tiles vs the per-vector text of the GPU kernels: same bits, private and registered buffers
    for widths in ft.enumerate_widths(world):
        x = torch.randn(40000, generator=g)
        a, b = x.clone(), x.clone()
        comm.allreduce(a, algo="stages", widths=widths, ctas=2, tile_bytes=4096)
        comm.allreduce(b, algo="stages_bulk", widths=widths, ctas=2, tile_bytes=4096)
        assert torch.equal(a, b), widths
        comm.allreduce(x, widths=widths)  # defaults: bulk, one thread, 256 KiB tiles
        assert torch.equal(a, x), widths
    # the per-call agreement on the code path lives in shared memory: one rank with a misaligned buffer moves
    # every rank to the scalar path (a disagreement would deadlock or corrupt, not just differ)
    raw = torch.zeros(1001 + 4)
    x = raw[1:1002] if rank == world - 1 else raw[4:1005] if raw.data_ptr() % 16 == 0 else raw[0:1001]
    x.copy_(torch.randn(1001, generator=g))
    ref = ref_sum(x.double()).float()
    comm.allreduce(x)
    assert torch.allclose(x, ref, rtol=1e-5, atol=1e-5)
    # other algorithms
    for algo in ("oneshot", "twoshot", "ring", "nvls"):
        x = torch.randn(3000, generator=g)
        ref = ref_sum(x.double()).float()
        comm.allreduce(x, algo=algo, ctas=2, tile_bytes=2048)
        assert torch.allclose(x, ref, rtol=1e-5, atol=1e-5), algo
    # dtypes / ops
    xi = torch.randint(-1000, 1000, (5000,), generator=g, dtype=torch.int32)
    ref = ref_sum(xi, dist.ReduceOp.MAX)
    comm.allreduce(xi, op="max")
    assert torch.equal(xi, ref)
    xb = torch.randint(0, 255, (777,), generator=g, dtype=torch.uint8)
    ref = ref_sum(xb, dist.ReduceOp.BAND)
    comm.allreduce(xb, op="band")
    assert torch.equal(xb, ref)
    # fused average, bf16 wire for fp32 data, fp8 block-scaled wire
    x = torch.randn(4096, generator=g)
    ref = ref_sum(x.double()).float() / world
    y = x.clone()
    comm.allreduce(y, scale=1.0 / world)
    assert torch.allclose(y, ref, rtol=1e-5, atol=1e-6)
    y = x.clone()
    comm.allreduce(y, scale=1.0 / world, wire=ft.BFLOAT16)
    assert torch.allclose(y, ref, rtol=3e-2, atol=3e-2)
    y = x.clone()
    comm.allreduce(y, scale=1.0 / world, wire=ft.FP8_BLOCK)
    assert (y - ref).abs().max() < 0.35
    # registered (zero-copy) region, repeated
    view = comm.registered(2048, np.float32)
    for it in range(3):
        src = torch.randn(2048, generator=g)
        ref = ref_sum(src.double()).float()
        view[:] = src.numpy()
        comm.allreduce(view, registered=True, ctas=2, tile_bytes=1024)
        assert np.allclose(view, ref.numpy(), rtol=1e-5, atol=1e-5)
    # pieces of the registered region: same offset on every rank works in place; different offsets (or a buffer
    # outside the region) are refused on EVERY rank, and the communicator stays usable
    piece = comm.registered(1000, np.float32, offset_bytes=8192)
    src = torch.randn(1000, generator=g)
    ref = ref_sum(src.double()).float()
    piece[:] = src.numpy()
    comm.allreduce(piece, registered=True)
    assert np.allclose(piece, ref.numpy(), rtol=1e-5, atol=1e-5)
    for bad in (comm.registered(256, np.float32, offset_bytes=16 * (1 + rank)), np.zeros(256, np.float32) if rank == 0 else piece[:256]):
        try:
            comm.allreduce(bad, registered=True)
            raise AssertionError("mismatched registered buffers must be refused")
        except RuntimeError as e:
            assert "-13" in str(e), e
    piece[:] = src.numpy()
    comm.allreduce(piece, registered=True)
    assert np.allclose(piece, ref.numpy(), rtol=1e-5, atol=1e-5)
    # randomized delays around back-to-back registered calls whose buffer is refilled right after each
    # return, with ranks entering at different times: without the entry waits a fast rank would reduce a peer's
    # stale buffer, without the exit barriers a slow peer would read the next iteration's input or a rank would
    # return before its peers' stores have landed. (Mutation-checked against staged_core.h.)
    import random
    import time
    lib = ft._lib.load()
    lib.ft_host_set_jitter(120)
    rnd = random.Random(1000 + rank)
    all_widths = ft.enumerate_widths(world)
    for it in range(36):
        algo = ("stages", "nvls", "ring", "stages_bulk")[it % 4]
        v = comm.registered(3000, np.float32)
        if rnd.random() < 0.5:
            time.sleep(rnd.random() * 2e-3)  # late to the call: my buffer still holds the poison below
        v[:] = float(it + 1) + rank
        comm.allreduce(v, algo=algo, registered=True, widths=all_widths[it % len(all_widths)], ctas=1 + it % 2,
                       tile_bytes=1024)
        want = world * (it + 1) + world * (world - 1) / 2
        assert np.all(v == want), (it, algo, v[:4], want)
        v[:] = -1e9  # poison: the next input arrives immediately
    lib.ft_host_set_jitter(0)
    # the per-call agreement numbers its calls modulo 2^32: run across the wrap
    import ctypes
    lib.ft_host_comm_debug_set_calls.argtypes = [ctypes.c_void_p, ctypes.c_ulonglong]
    dist.barrier()
    lib.ft_host_comm_debug_set_calls(comm._h, (1 << 32) - 3)
    for it in range(6):
        x = torch.randn(777, generator=g)
        ref = ref_sum(x.double()).float()
        comm.allreduce(x)
        assert torch.allclose(x, ref, rtol=1e-5, atol=1e-5), it
    # the halves
    per = 257
    x = torch.randn(per * world, generator=g)
    mine = torch.zeros(per)
    comm.reduce_scatter(x, mine)
    full = ref_sum(x.double()).float()
    assert torch.allclose(mine, full[rank * per:(rank + 1) * per], rtol=1e-5, atol=1e-5)
    gathered = torch.zeros(per * world)
    comm.all_gather(mine, gathered)
    assert torch.allclose(gathered, full, rtol=1e-5, atol=1e-5)
    # lonely ranks (tree smaller than the world)
    if world >= 3:
        tree = {3: [2], 4: [3], 5: [2, 2], 6: [5], 7: [2, 3], 8: [2, 3]}[world]
        lonely = world - int(np.prod(tree))
        x = torch.randn(999, generator=g)
        ref = ref_sum(x.double()).float()
        comm.allreduce(x, widths=tree, lonely=lonely)
        assert torch.allclose(x, ref, rtol=1e-5, atol=1e-5)
    # "-k" trees: more tree positions than ranks, the extra ones are phantom ranks hosted by the first processes
    ph = ft.HostCommunicator(window_bytes=1 << 20, max_phantoms=2)
    for k in (1, 2):
        for widths in ft.enumerate_widths(world + k):
            for algo in ("stages", "stages_bulk"):
                x = torch.randn(4099 if algo == "stages" else 4096, generator=g)
                ref = ref_sum(x.double()).float()
                y = x.clone()
                ph.allreduce(y, algo=algo, widths=widths, lonely=-k, ctas=2, tile_bytes=1024)
                assert torch.allclose(y, ref, rtol=1e-5, atol=1e-5), (k, widths, algo)
    xi = torch.randint(-1000, 1000, (3000,), generator=g, dtype=torch.int32)
    for op, rop in (("max", dist.ReduceOp.MAX), ("min", dist.ReduceOp.MIN), ("band", dist.ReduceOp.BAND)):
        y = xi.clone()
        ph.allreduce(y, op=op, widths=ft.enumerate_widths(world + 1)[-1], lonely=-1)
        assert torch.equal(y, ref_sum(xi, rop)), op
    xf = torch.randn(2048, generator=g)
    y = xf.clone()
    ph.allreduce(y, op="min", widths=[world + 1], lonely=-1, scale=None)
    assert torch.equal(y, ref_sum(xf, dist.ReduceOp.MIN))
    try:
        ph.allreduce(xf.clone(), widths=[world + 3], lonely=-3)  # more phantoms than the communicator has room for
        raise AssertionError("must be refused")
    except RuntimeError as e:
        assert "-14" in str(e)
    assert ph.error() == 0
    ph.close()
    # grow-only staging (mpi_mod.hpp:932-950): a communicator created small re-creates its windows collectively
    small = ft.HostCommunicator(window_bytes=1 << 16)
    for count in (1000, 200000, 5000):
        x = torch.randn(count, generator=g)
        ref = ref_sum(x.double()).float()
        small.allreduce(x)
        assert torch.allclose(x, ref, rtol=1e-5, atol=1e-5), count
    assert small.window_bytes >= 800000 and small.error() == 0
    small.close()
    assert comm.error() == 0
    dist.barrier()
    comm.close()
    if rank == 0:
        print("MP_CPU_WORKER_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
