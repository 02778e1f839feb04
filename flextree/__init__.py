"""Import shim: ``import flextree`` loads the package that physically lives in
``allreduce-over-mpi_b200/`` (a directory name Python cannot import directly)."""
import os as _os

_pkg_dir = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "allreduce-over-mpi_b200")
__path__.insert(0, _pkg_dir)

from ._api import *  # noqa: F401,F403,E402
from ._api import __all__ as _api_all  # noqa: E402
from . import _lib, ops, parallel, models, utils  # noqa: F401,E402

__all__ = list(_api_all)
