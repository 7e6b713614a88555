"""Shared builders for the parity tests: the reference's 2-particle configurations and seeded random systems."""

import numpy as np

from azplugins_b200 import params as P
from azplugins_b200.state import Box, build_bond_table

from oracle_api import HostSystem

SPEC_BY_API = {
    "Colloid": P.COLLOID,
    "DPDGeneralWeight": P.DPD_GENERAL_WEIGHT,
    "ExpandedYukawa": P.EXPANDED_YUKAWA,
    "Hertz": P.HERTZ,
    "PerturbedLennardJones": P.PERTURBED_LENNARD_JONES,
    "TwoPatchMorse": P.TWO_PATCH_MORSE,
    "DoubleWell": P.DOUBLE_WELL,
    "Quartic": P.QUARTIC,
}


def two_particle_box(r_cut, buffer=0.4):
    # test_pair.py:318-322: L = 2.1 * 2 * (r_cut + buffer)
    return Box.cube(2.1 * 2 * (r_cut + buffer))


def two_particle_pair_system(d, r_cut, bits, positions=None):
    box = two_particle_box(r_cut)
    pos = [[-d / 2, 0, 0], [d / 2, 0, 0]] if positions is None else positions
    hs = HostSystem(box, np.array(pos, dtype=np.float64), np.zeros(2, dtype=np.int64), 1, bits)
    hs.build_nlist(r_cut + 0.4)
    return hs


def two_particle_bond_system(d, bits):
    box = Box.cube(20.0)
    hs = HostSystem(box, np.array([[-d / 2, 0, 0], [d / 2, 0, 0]]), np.zeros(2, dtype=np.int64), 1, bits)
    table = build_bond_table(2, 2, np.array([[0, 1]]), np.array([0]))
    return hs, table


def random_fluid(n, rho, ntypes, seed, jitter=0.15):
    """Cubic-lattice fluid with uniform jitter; returns (box, positions, typeid)."""
    rng = np.random.default_rng(seed)
    m = int(np.ceil(n ** (1.0 / 3.0)))
    a = (1.0 / rho) ** (1.0 / 3.0)
    L = m * a
    g = np.stack(np.meshgrid(*(np.arange(m),) * 3, indexing="ij"), axis=-1).reshape(-1, 3)[:n]
    pos = (g + 0.5) * a - L / 2 + rng.uniform(-jitter, jitter, size=(n, 3)) * a
    pos = (pos + L / 2) % L - L / 2
    typeid = rng.integers(0, ntypes, size=n)
    return Box.cube(L), pos, typeid


def random_unit_quaternions(n, seed):
    rng = np.random.default_rng(seed)
    q = rng.normal(size=(n, 4))
    return q / np.linalg.norm(q, axis=1, keepdims=True)


def pair_tables(spec, param_by_pair, rcut_by_pair, ntypes, bits, ron_by_pair=None):
    """Symmetric (T*T) param / rcutsq / ronsq tables as the host class builds them."""
    ftype = np.float32 if bits == 32 else np.float64
    par = np.zeros(ntypes * ntypes, dtype=spec.dtype(bits))
    rc = np.zeros((ntypes, ntypes), dtype=ftype)
    ron = np.zeros((ntypes, ntypes), dtype=ftype)
    for i in range(ntypes):
        for j in range(ntypes):
            key = (min(i, j), max(i, j))
            par[i * ntypes + j] = spec.pack(param_by_pair[key], bits)
            rc[i, j] = rcut_by_pair[key]
            if ron_by_pair is not None:
                ron[i, j] = ron_by_pair[key]
    return par, (rc * rc).reshape(-1), (ron * ron).reshape(-1), rc.astype(np.float64)


def assert_golden(actual, desired, bits):
    """The reference's own tolerance (decimal=4) in double, HOOMD's default Scalar. A float build cannot hold
    1.5e-4 absolute on values of magnitude 400-1200 (float epsilon * 1200 = 1.4e-4), so float adds 2e-5 relative."""
    if bits == 64:
        np.testing.assert_array_almost_equal(actual, desired, decimal=4)
    else:
        np.testing.assert_allclose(actual, desired, rtol=2e-5, atol=1.5e-4)
