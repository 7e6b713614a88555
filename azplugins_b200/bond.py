"""Bond potentials — the ``hoomd.azplugins.bond`` API (/root/reference/src/bond.py:13-65 DoubleWell, :68-157 Quartic)."""

from . import params as _p
from .force import Bond


class DoubleWell(Bond):
    """Double-well bond; params ``r_0, r_1, U_1, U_tilt``."""

    _spec = _p.DOUBLE_WELL


class Quartic(Bond):
    """Quartic (breakable) bond with WCA core; params ``k, r_0, b_1, b_2, U_0, sigma, epsilon, delta`` (delta optional, 0)."""

    _spec = _p.QUARTIC
