"""Utilities: CPU world (functional model of the kernels), timers, logging, clock sampling."""
from .hostworld import HostWorld  # noqa: F401
from .timer import Timer  # noqa: F401
