"""Multi-process worker (one rank per GPU), launched by tests/test_multigpu.py through
torch.distributed.run. Every check compares the flextree kernels against torch.distributed (NCCL) or
exact fp64 math. Prints MP_WORKER_OK on rank 0 when everything passed."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("FT_TIMEOUT_S", "10")
import flextree as ft  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = ft.init()
    assert comm.rank == rank and comm.size == world
    if rank == 0:
        print(f"[worker] world {world}, multicast={comm.has_multicast}", flush=True)
    g = torch.Generator(device=dev).manual_seed(100 + rank)

    def check(t, ref, tol, what):
        torch.cuda.synchronize()
        e = comm.error()
        assert e == 0, f"{what}: device error {e:#x}"
        err = (t.double() - ref.double()).abs().max().item()
        scale = max(1.0, ref.double().abs().max().item())
        assert err <= tol * scale, f"{what}: max abs err {err} (scale {scale})"

    def exact_sum(x):
        y = x.double().clone()
        dist.all_reduce(y)
        return y

    # 1. every factorization, several sizes, in place and out of place (staging path)
    comm.set_algo("stages")
    for widths in ft.enumerate_widths(world):
        comm.set_widths(widths)
        for count in (1, 7, 1000, 65536 + 5, (1 << 22) + 3):
            x = torch.randn(count, device=dev, generator=g)
            ref = exact_sum(x)
            out = torch.zeros_like(x)
            comm.allreduce(x, out=out)
            check(out, ref, 1e-5, f"stages {widths} count {count} oop")
            y = x.clone()
            comm.allreduce(y)
            check(y, ref, 1e-5, f"stages {widths} count {count} in place")
            assert comm.last_call()["widths"] == widths
            # all ranks must hold bit-identical results
            chk = y.clone()
            dist.broadcast(chk, 0)
            assert torch.equal(chk, y), f"ranks differ for {widths}"
    comm.set_widths(None)

    # 1b. lonely ranks: a tree smaller than the world (the reference's "+k" trees)
    if world >= 3:
        tree = {3: [2], 4: [3], 5: [2, 2], 6: [5], 7: [2, 3], 8: [2, 3]}.get(world, [world - 1])
        comm.set_widths(tree)
        for count in (5, 100000):
            x = torch.randn(count, device=dev, generator=g)
            ref = exact_sum(x)
            out = torch.zeros_like(x)
            comm.allreduce(x, out=out)
            check(out, ref, 1e-5, f"lonely tree {tree} count {count}")
        comm.set_widths(None)

    # 1c. phantom ranks: a tree LARGER than the world (the reference's "-k" trees, e.g. 2*4-1 on 7 ranks): every ordered
    # factorization of world+1 (and of world+2), fp32 SUM in/out of place, int32 MAX (non-zero identity), bf16, odd counts
    comm.set_algo("stages")
    for k in (1, 2):
        if k > world:
            continue
        for widths in ft.enumerate_widths(world + k):
            if max(widths) > 32:
                continue
            comm.set_widths(widths)
            for count in (1, 1001, (1 << 20) + 3):
                x = torch.randn(count, device=dev, generator=g)
                ref = exact_sum(x)
                out = torch.zeros_like(x)
                comm.allreduce(x, out=out)
                check(out, ref, 1e-5, f"phantom tree {widths}-{k} count {count} oop")
                assert comm.last_call()["widths"] == widths
                y = x.clone()
                comm.allreduce(y, scale=0.5)
                check(y, ref * 0.5, 1e-5, f"phantom tree {widths}-{k} count {count} in place, scaled")
                chk = y.clone()
                dist.broadcast(chk, 0)
                assert torch.equal(chk, y), f"ranks differ for phantom tree {widths}"
            xi = torch.randint(-10**6, -10, (4097,), device=dev, dtype=torch.int32, generator=g)   # all negative: a zero identity would win MAX
            ri = xi.clone()
            dist.all_reduce(ri, op=dist.ReduceOp.MAX)
            comm.allreduce(xi, op="max")
            torch.cuda.synchronize()
            assert comm.error() == 0 and torch.equal(xi, ri), f"phantom tree {widths}: int32 MAX"
            xb = torch.randn(70000, device=dev, generator=g).bfloat16()
            rb = exact_sum(xb)
            comm.allreduce(xb)
            check(xb, rb, 3e-2, f"phantom tree {widths}: bf16")
    comm.set_widths(None)

    # 2. one-shot, ring, (NVLS), auto
    algos = ["oneshot", "twoshot", "ring", "auto"] + (["nvls"] if comm.has_multicast else [])
    for algo in algos:
        comm.set_algo(algo)
        for count in (3, 2048, 100000):
            if algo == "nvls" and count < 16:
                continue
            if algo == "oneshot" and count * 4 * 2 * world > (4 << 20):
                count = (4 << 20) // (8 * world) - 5  # largest message the default LL buffer takes
            x = torch.randn(count, device=dev, generator=g)
            ref = exact_sum(x)
            comm.allreduce(x)
            check(x, ref, 1e-5, f"{algo} count {count}")
    comm.set_algo("twoshot")   # several CTAs, ragged blocks, out of place + fused scale, bf16
    for count in (262144 + 3, 1 << 20):
        x = torch.randn(count, device=dev, generator=g)
        ref = exact_sum(x)
        out = torch.zeros_like(x)
        comm.allreduce(x, out=out, scale=0.5)
        check(out, ref * 0.5, 1e-5, f"twoshot count {count} out of place, scaled")
        xb = x.bfloat16()
        rb = exact_sum(xb)
        comm.allreduce(xb)
        check(xb, rb, 3e-2, f"twoshot count {count} bf16")
    comm.set_algo("auto")

    # 3. dtypes / ops
    comm.set_algo("stages")
    for dt, op, tor in ((torch.int32, "sum", dist.ReduceOp.SUM), (torch.int64, "max", dist.ReduceOp.MAX),
                        (torch.uint8, "band", None), (torch.float64, "min", dist.ReduceOp.MIN),
                        (torch.bfloat16, "sum", dist.ReduceOp.SUM), (torch.float16, "max", dist.ReduceOp.MAX)):
        if dt.is_floating_point:
            x = torch.randn(33333, device=dev, generator=g).to(dt)
        else:
            x = torch.randint(0, 100, (33333,), device=dev, generator=g).to(dt)
        if tor is None:  # NCCL has no bitwise AND: gather and fold locally
            parts = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(parts, x)
            ref = parts[0].clone()
            for q in parts[1:]:
                ref &= q
        else:
            ref = x.clone()
            dist.all_reduce(ref, op=tor)
        comm.allreduce(x, op=op)
        check(x, ref, 2e-2 if dt in (torch.bfloat16, torch.float16) else 0, f"{dt} {op}")

    # 4. fp32 gradients, bf16 wire, fused 1/N (DDP shape)
    x = torch.randn(250001, device=dev, generator=g)
    ref = exact_sum(x) / world
    comm.allreduce(x, scale=1.0 / world, wire_dtype=torch.bfloat16)
    check(x, ref, 2e-2, "bf16 wire + scale")

    # 5. registered (zero-copy) buffers, repeated, incl. NVLS if present
    for algo in ["stages"] + (["nvls"] if comm.has_multicast else []):
        comm.set_algo(algo)
        buf = comm.symm_empty(1 << 20)
        for it in range(4):
            src = torch.randn(1 << 20, device=dev, generator=g)
            ref = exact_sum(src)
            buf.copy_(src)
            comm.allreduce(buf)
            check(buf, ref, 1e-5, f"registered {algo} iter {it}")
            assert comm.last_call()["mode"] == "registered"
        buf.copy_(src)
        comm.allreduce(buf, scale=1.0 / world)  # fused average on a zero-copy buffer
        check(buf, ref / world, 1e-5, f"registered {algo} + scale")
        assert comm.last_call()["mode"] == "registered"
    comm.set_algo("auto")

    # 5b. the halves against torch.distributed
    comm.set_algo("stages")
    per = 50000
    x = torch.randn(per * world, device=dev, generator=g)
    mine = torch.empty(per, device=dev)
    comm.reduce_scatter(x, mine)
    ref = torch.empty(per, device=dev)
    dist.reduce_scatter_tensor(ref, x.clone())
    check(mine, ref, 1e-5, "reduce_scatter")
    full = torch.empty(per * world, device=dev)
    comm.all_gather(mine, full)
    ref_full = torch.empty(per * world, device=dev)
    dist.all_gather_into_tensor(ref_full, mine)
    check(full, ref_full, 0, "all_gather")
    comm.set_algo("auto")

    # 6. CUDA graph capture + replay
    comm.set_algo("stages")
    n = 1 << 18
    comm.reserve(n * 4)
    x = torch.full((n,), float(rank + 1), device=dev)
    y = torch.zeros_like(x)
    comm.allreduce(x, out=y)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        graph.capture_begin()
        comm.allreduce(x, out=y, stream=s)
        graph.capture_end()
    for it in range(10):
        x.fill_(float(rank + 1 + it))
        graph.replay()
        torch.cuda.synchronize()
        want = float(sum(r + 1 + it for r in range(world)))
        assert torch.all(y == want), f"graph replay {it}: {y[:4]} vs {want}"
    assert comm.error() == 0
    comm.set_algo("auto")

    # 6b. capture + 20 replays on real peers for the paths with their own replay hazards: registered NVLS (no exit
    # barrier, per-tile flags), registered TMA stages (exit barrier), one-shot / two-shot LL (parity halves by epoch)
    cases = [("oneshot", 2048, False), ("twoshot", 1 << 16, False), ("stages", 1 << 18, True)]
    if comm.has_multicast:
        cases += [("nvls", 1 << 20, True), ("nvls", 1 << 18, False)]
    for algo, cnt, registered in cases:
        comm.set_algo(algo)
        if algo == "nvls":
            comm.set_launch(4, 0, 64 << 10)   # several tiles per CTA
        buf = comm.symm_empty(cnt, torch.float32) if registered else torch.empty(cnt, device=dev)
        buf.fill_(1.0)
        comm.allreduce(buf)
        torch.cuda.synchronize()
        assert comm.last_call()["mode"] == ("registered" if registered else "staging"), (algo, comm.last_call())
        graph = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            graph.capture_begin()
            comm.allreduce(buf, stream=s)
            graph.capture_end()
        for it in range(20):
            buf.fill_(float(rank + 1 + it))
            graph.replay()
            torch.cuda.synchronize()
            want = float(sum(r + 1 + it for r in range(world)))
            assert torch.all(buf == want), f"graph replay {algo} registered={registered} it {it}: {buf[:4]} vs {want}"
        assert comm.error() == 0
        comm.set_launch(0, 0, 0)
    comm.set_algo("auto")

    # 7. DDP comm hook parity with stock NCCL DDP on a small MLP
    from torch.nn.parallel import DistributedDataParallel as DDP
    torch.manual_seed(0)

    def make():
        return torch.nn.Sequential(torch.nn.Linear(256, 512), torch.nn.GELU(), torch.nn.Linear(512, 512),
                                   torch.nn.GELU(), torch.nn.Linear(512, 10)).to(dev)
    m_ref, m_ft = make(), make()
    m_ft.load_state_dict(m_ref.state_dict())
    make_state = {k: v.clone() for k, v in m_ref.state_dict().items()}
    d_ref = DDP(m_ref, device_ids=[local], bucket_cap_mb=1)
    d_ft = DDP(m_ft, device_ids=[local], bucket_cap_mb=1)
    state = ft.FlexTreeHookState(comm)
    d_ft.register_comm_hook(state, ft.ddp_comm_hook)
    o_ref = torch.optim.SGD(d_ref.parameters(), lr=0.05)
    o_ft = torch.optim.SGD(d_ft.parameters(), lr=0.05)
    for step in range(5):
        xb = torch.randn(64, 256, device=dev, generator=g)
        yb = torch.randint(0, 10, (64,), device=dev, generator=g)
        for d, o in ((d_ref, o_ref), (d_ft, o_ft)):
            o.zero_grad()
            loss = torch.nn.functional.cross_entropy(d(xb), yb)
            loss.backward()
            o.step()
    torch.cuda.synchronize()
    assert state.calls > 0 and comm.error() == 0
    for p_ref, p_ft in zip(m_ref.parameters(), m_ft.parameters()):
        assert torch.allclose(p_ref, p_ft, rtol=1e-4, atol=1e-5), "DDP hook diverged from NCCL DDP"

    # 8. zero-copy data-parallel engine (gradients in symmetric buckets) vs the same NCCL DDP run
    m_dp = make()
    m_dp.load_state_dict(make_state)
    dp = ft.FlexTreeDataParallel(m_dp, comm, bucket_mb=1)
    o_dp = torch.optim.SGD(m_dp.parameters(), lr=0.05)
    g2 = torch.Generator(device=dev).manual_seed(4242 + rank)
    m_ref2 = make()
    m_ref2.load_state_dict(make_state)
    d_ref2 = DDP(m_ref2, device_ids=[local], bucket_cap_mb=1)
    o_ref2 = torch.optim.SGD(d_ref2.parameters(), lr=0.05)
    for step in range(5):
        xb = torch.randn(64, 256, device=dev, generator=g2)
        yb = torch.randint(0, 10, (64,), device=dev, generator=g2)
        o_ref2.zero_grad()
        torch.nn.functional.cross_entropy(d_ref2(xb), yb).backward()
        o_ref2.step()
        dp.zero_grad()
        torch.nn.functional.cross_entropy(dp(xb), yb).backward()
        dp.synchronize()
        o_dp.step()
    torch.cuda.synchronize()
    assert comm.error() == 0
    assert comm.last_call()["mode"] == "registered"
    for p_ref, p_dp in zip(m_ref2.parameters(), m_dp.parameters()):
        assert torch.allclose(p_ref, p_dp, rtol=1e-4, atol=1e-5), "FlexTreeDataParallel diverged from NCCL DDP"

    dist.barrier()
    ft.shutdown()
    if rank == 0:
        print("MP_WORKER_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
