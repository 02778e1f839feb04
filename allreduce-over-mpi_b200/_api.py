"""Public API (``import flextree``)."""
from .ops.planner import *  # noqa: F401,F403
from .ops.planner import __all__ as _planner_all
from .ops.constants import *  # noqa: F401,F403
from .ops.constants import __all__ as _const_all
from .utils.hostworld import HostWorld
from .utils.timer import Timer

from .parallel.comm import Communicator, FlexTreeError
from .parallel.api import init, allreduce, get_communicator, shutdown
from .parallel.ddp import ddp_comm_hook, FlexTreeHookState
from .parallel.data_parallel import FlexTreeDataParallel
from .parallel.host_comm import HostCommunicator
from .parallel.host_data_parallel import HostDataParallel

__all__ = list(_planner_all) + list(_const_all) + [
    "HostWorld", "Timer", "Communicator", "FlexTreeError", "init", "allreduce", "get_communicator", "shutdown",
    "ddp_comm_hook", "FlexTreeHookState", "FlexTreeDataParallel", "HostCommunicator", "HostDataParallel"]
