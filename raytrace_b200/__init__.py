"""Import alias: the package directory is `ray-trace_b200/` (not a valid Python identifier)."""
import os as _os

__path__.insert(0, _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "ray-trace_b200"))
from .api import *  # noqa: E402,F401,F403
from . import api, build, sweeps  # noqa: E402,F401
