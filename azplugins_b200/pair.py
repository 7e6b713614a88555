"""Pair potentials — the ``hoomd.azplugins.pair`` API (/root/reference/src/pair.py) on the B200 engine.

Class names, constructor arguments, parameter keys, accepted modes and ``_cpp_class_name`` follow the reference
(pair.py:14,106-118 Colloid; :121,213-239 DPDGeneralWeight; :242,287-297 ExpandedYukawa; :300,340-351 Hertz;
:354,413-426 PerturbedLennardJones; :429,507-525 TwoPatchMorse).
"""

import warnings

from . import params as _p
from .force import AnisotropicPair, DPDThermoPair, Pair


class Colloid(Pair):
    """Colloid (integrated Lennard-Jones) pair potential; params ``A, a_1, a_2, sigma``."""

    _spec = _p.COLLOID
    _accepted_modes = ("none", "shift", "xplor")

    def __init__(self, nlist, default_r_cut=None, default_r_on=0, mode="none"):
        super().__init__(nlist, default_r_cut, default_r_on, mode)


class DPDGeneralWeight(DPDThermoPair):
    """DPD thermostat with generalized weight ``w = (1 - r/r_cut)^s``; params ``A, gamma, s``; ``kT`` variant."""

    _spec = _p.DPD_GENERAL_WEIGHT
    _accepted_modes = ("none",)

    def __init__(self, nlist, kT, default_r_cut=None):
        super().__init__(nlist, kT, default_r_cut)

    def _attach(self, state):
        # pair.py:236-238: DPD uses RNGs, warn when the seed was never set
        if getattr(state, "seed_unset", False):
            warnings.warn("Simulation.seed is not set, using default seed=0", RuntimeWarning)
        super()._attach(state)


class DPDConservativeGeneralWeight(Pair):
    """Conservative part only (C++ class ``PotentialPairConservativeGeneralWeight``, no Python class upstream)."""

    _spec = _p.DPD_CONSERVATIVE
    _accepted_modes = ("none",)


class ExpandedYukawa(Pair):
    """Expanded Yukawa pair potential; params ``epsilon, kappa, delta``."""

    _spec = _p.EXPANDED_YUKAWA

    def __init__(self, nlist, default_r_cut=None, default_r_on=0, mode="none"):
        super().__init__(nlist, default_r_cut, default_r_on, mode)


class Hertz(Pair):
    """Hertz pair potential; params ``epsilon``."""

    _spec = _p.HERTZ
    _accepted_modes = ("none", "shift", "xplor")

    def __init__(self, nlist, default_r_cut=None, default_r_on=0, mode="none"):
        super().__init__(nlist, default_r_cut, default_r_on, mode)


class PerturbedLennardJones(Pair):
    """Perturbed Lennard-Jones; params ``epsilon, sigma, attraction_scale_factor``."""

    _spec = _p.PERTURBED_LENNARD_JONES
    _accepted_modes = ("none", "shift", "xplor")

    def __init__(self, nlist, default_r_cut=None, default_r_on=0, mode="none"):
        super().__init__(nlist, default_r_cut, default_r_on, mode)


class TwoPatchMorse(AnisotropicPair):
    """Two-patch Morse anisotropic potential; params ``M_d, M_r, r_eq, omega, alpha, repulsion``."""

    _spec = _p.TWO_PATCH_MORSE

    def __init__(self, nlist, default_r_cut=None, mode="none"):
        super().__init__(nlist, default_r_cut, mode)
