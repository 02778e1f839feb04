"""Operators: the planner (factorizations, cost models, schedules) and the collectives."""
from .planner import *  # noqa: F401,F403
from .constants import *  # noqa: F401,F403
