"""Data-parallel integration: communicator bootstrap over torch.distributed and the DDP comm hook."""
from .comm import Communicator, FlexTreeError  # noqa: F401
from .api import init, allreduce, get_communicator, shutdown  # noqa: F401
from .ddp import ddp_comm_hook, FlexTreeHookState  # noqa: F401
from .data_parallel import FlexTreeDataParallel  # noqa: F401
from .host_comm import HostCommunicator  # noqa: F401
from .host_data_parallel import HostDataParallel  # noqa: F401
