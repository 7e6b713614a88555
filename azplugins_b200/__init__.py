"""azplugins_b200 — B200-native (sm_100a) force-compute engine behind the hoomd.azplugins pair/bond API."""

from . import bond, external, md, nlist, pair, params, state  # noqa: F401
from .state import Box, Snapshot, State  # noqa: F401

__version__ = "0.1.0"
