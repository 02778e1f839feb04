"""DistributedDataParallel communication hook.

``model.register_comm_hook(FlexTreeHookState(comm), ddp_comm_hook)`` replaces DDP's NCCL allreduce of
each gradient bucket with the flextree kernel. The 1/world_size averaging (and, optionally, a bf16 wire
format for fp32 gradients) is fused into the last stage of the collective instead of running as
separate elementwise kernels. The collective runs on a side stream so that it overlaps the rest of
backward; the returned Future completes on that stream.

The reference has no training integration at all (its use case is the data-parallel gradient allreduce,
allreduce_over_mpi/mpi_mod.hpp:1167-1221); this is the torch-facing form of that call.
"""
import torch


class FlexTreeHookState:
    def __init__(self, comm, wire_dtype=None, average=True, side_stream=True, max_ctas=32):
        self.comm = comm
        if max_ctas:
            comm.set_max_ctas(max_ctas)  # the collective overlaps backward: do not take the whole GPU
        self.wire_dtype = wire_dtype
        self.average = average
        self.stream = torch.cuda.Stream(device=comm.device) if side_stream else None
        self.calls = 0
        self.bytes = 0


def ddp_comm_hook(state, bucket):
    buf = bucket.buffer()
    comm = state.comm
    # (comm.allreduce itself raises once an earlier bucket's collective has tripped the device watchdog: the
    # kernel mirrors its error word into mapped host memory and every call peeks it first)
    scale = 1.0 / comm.size if state.average else 1.0
    wire = None
    if state.wire_dtype is not None:
        if isinstance(state.wire_dtype, str) and state.wire_dtype.startswith("fp8"):
            wire = state.wire_dtype if buf.dtype in (torch.float32, torch.bfloat16) else None
        elif buf.dtype == torch.float32:
            wire = state.wire_dtype
    state.calls += 1
    state.bytes += buf.numel() * buf.element_size()
    fut = torch.futures.Future(devices=[comm.device])
    if state.stream is None:
        comm.allreduce(buf, scale=scale, wire_dtype=wire)
        fut.set_result(buf)
        return fut
    cur = torch.cuda.current_stream(comm.device)
    state.stream.wait_stream(cur)  # gradients of this bucket are ready on the current stream
    with torch.cuda.stream(state.stream):
        comm.allreduce(buf, scale=scale, wire_dtype=wire, stream=state.stream)
        buf.record_stream(state.stream)
        fut.set_result(buf)  # the Future records an event on the side stream
    return fut
