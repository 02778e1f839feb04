"""Public API (``import flextree``)."""
from .ops.planner import *  # noqa: F401,F403
from .ops.planner import __all__ as _planner_all
from .ops.constants import *  # noqa: F401,F403
from .ops.constants import __all__ as _const_all
from .utils.hostworld import HostWorld
from .utils.timer import Timer

__all__ = list(_planner_all) + list(_const_all) + ["HostWorld", "Timer"]
