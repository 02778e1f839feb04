"""CPU data-parallel worker (torch.distributed gloo): ft.HostDataParallel against gradients averaged with gloo's
all_reduce, several steps of SGD on a small MLP + conv net. Launched by tests/test_multiprocess_cpu.py."""
import copy
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flextree as ft  # noqa: E402


def make_model():
    torch.manual_seed(1234)  # same initial weights everywhere for the reference copy; dp broadcasts anyway
    return torch.nn.Sequential(
        torch.nn.Conv2d(3, 8, 3, padding=1), torch.nn.ReLU(), torch.nn.Flatten(),
        torch.nn.Linear(8 * 8 * 8, 37), torch.nn.Tanh(), torch.nn.Linear(37, 10))


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.set_num_threads(1)
    model = make_model()
    if rank != 0:  # perturb: the engine must broadcast rank 0's parameters
        with torch.no_grad():
            for p in model.parameters():
                p.add_(0.1 * rank)
    ref = make_model()
    comm = ft.HostCommunicator(window_bytes=ft.HostDataParallel.bytes_needed(model))
    dp = ft.HostDataParallel(model, comm, bucket_mb=0.05, widths=ft.enumerate_widths(world)[-1])
    assert len(dp.buckets) >= 2, len(dp.buckets)
    for a, b in zip(model.parameters(), ref.parameters()):
        assert torch.equal(a, b), "broadcast"
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    g = torch.Generator().manual_seed(100 + rank)
    for step in range(4):
        x = torch.randn(5, 3, 8, 8, generator=g)
        y = torch.randint(0, 10, (5,), generator=g)
        loss = torch.nn.functional.cross_entropy(dp(x), y)
        loss.backward()
        dp.synchronize()
        # reference: plain backward + gloo average
        ropt.zero_grad()
        torch.nn.functional.cross_entropy(ref(x), y).backward()
        for p in ref.parameters():
            dist.all_reduce(p.grad)
            p.grad /= world
        for (n, a), b in zip(model.named_parameters(), ref.parameters()):
            assert torch.allclose(a.grad, b.grad, rtol=1e-4, atol=1e-6), (step, n, (a.grad - b.grad).abs().max())
        opt.step()
        ropt.step()
        dp.zero_grad()
        for p in model.parameters():
            assert p.grad is not None and float(p.grad.abs().sum()) == 0.0
    for a, b in zip(model.parameters(), ref.parameters()):
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-6)
    # every rank holds the same parameters, bit for bit (every rank reads the same reduced blocks)
    for p in model.parameters():
        chk = p.detach().clone()
        dist.broadcast(chk, 0)
        assert torch.equal(chk, p.detach())
    # a message larger than the window cannot grow it while registered views are out
    try:
        comm.allreduce(torch.zeros(comm.window_bytes))
        raise AssertionError("growth with views out must be refused")
    except RuntimeError:
        pass
    assert comm.error() == 0
    dp.close()
    dist.barrier()
    comm.close()
    if rank == 0:
        print("MP_CPU_DP_WORKER_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
