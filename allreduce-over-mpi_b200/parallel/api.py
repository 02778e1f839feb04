"""Module-level convenience API: ``flextree.init()`` then ``flextree.allreduce(tensor)``."""
import torch

from .comm import Communicator

_default = None


def init(group=None, device=None, **opts):
    """Create the default communicator from a torch.distributed process group (collective)."""
    global _default
    import torch.distributed as dist
    if not dist.is_initialized():
        raise RuntimeError("flextree.init: torch.distributed is not initialised")
    if _default is not None:
        _default.close()
    _default = Communicator.from_process_group(group, device, **opts)
    return _default


def get_communicator():
    if _default is None:
        raise RuntimeError("flextree: call flextree.init() first")
    return _default


def allreduce(tensor, op="sum", out=None, scale=1.0, wire_dtype=None, comm=None):
    """allreduce(tensor): in-place sum across the default communicator (see Communicator.allreduce)."""
    return (comm or get_communicator()).allreduce(tensor, op=op, out=out, scale=scale, wire_dtype=wire_dtype)


def shutdown():
    global _default
    if _default is not None:
        torch.cuda.synchronize()
        _default.close()
        _default = None
