"""The oracle pinned against every known-answer row of the reference's own tests.

Both oracles are replayed: the plain-C restatement ("port") and, when present, the reference's evaluator headers
compiled unmodified ("ref"); both precisions; full-list and half-list (+ third law) host loops.
Tolerance is the reference's: numpy.testing.assert_array_almost_equal(decimal=4).
"""

import numpy as np
import pytest

import cases
import golden_cases as G
import oracle_api as O
from azplugins_b200 import _abi

WHICH = ["port"] + (["ref"] if O.have_ref() else [])


@pytest.mark.parametrize("which", WHICH)
@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("half", [False, True])
@pytest.mark.parametrize("case", G.PAIR_CASES, ids=lambda c: c[0])
def test_pair_golden(case, half, bits, which):
    api, params, r_cut, shift, d, energy, force = case
    spec = cases.SPEC_BY_API[api]
    hs = cases.two_particle_pair_system(d, r_cut, bits)
    if half:
        hs = hs.full_to_half()
    par, rcsq, ronsq, _ = cases.pair_tables(spec, {(0, 0): params}, {(0, 0): r_cut}, 1, bits)
    if api == "DPDGeneralWeight":
        hs.dpd(par, rcsq, seed=1, timestep=0, dt=0.005, kT=0.0, half_list=half, which=which)  # test_pair.py:331-332
    else:
        hs.pair(spec.potential_id, par, rcsq, ronsq, shift_mode=1 if shift else 0, half_list=half, which=which)
    cases.assert_golden(hs.force[:, 3], [0.5 * energy, 0.5 * energy], bits)
    cases.assert_golden(hs.force[:, :3], [[-force, 0, 0], [force, 0, 0]], bits)


@pytest.mark.parametrize("which", WHICH)
@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("half", [False, True])
@pytest.mark.parametrize("case", G.ANISO_CASES, ids=lambda c: f"rc{c[1]}")
def test_aniso_golden(case, half, bits, which):
    params, r_cut, shift, energy, force, torque = case
    spec = cases.SPEC_BY_API["TwoPatchMorse"]
    hs = cases.two_particle_pair_system(None, r_cut, bits, positions=G.ANISO_POSITIONS)
    if half:
        hs = hs.full_to_half()
    par, rcsq, _, _ = cases.pair_tables(spec, {(0, 0): params}, {(0, 0): r_cut}, 1, bits)
    hs.aniso(par, rcsq, shift_mode=1 if shift else 0, half_list=half, which=which)
    cases.assert_golden(hs.force[:, 3], [0.5 * energy, 0.5 * energy], bits)
    if force is not None:
        f = np.array(force)
        cases.assert_golden(hs.force[:, :3], [-f, f], bits)
    if torque is not None:
        cases.assert_golden(hs.torque[:, :3], [torque, torque], bits)


@pytest.mark.parametrize("which", WHICH)
@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("case", G.BOND_CASES, ids=lambda c: c[0])
def test_bond_golden(case, bits, which):
    api, params, d, energy, force = case
    spec = cases.SPEC_BY_API[api]
    hs, (n_bonds, bondlist, bond_pos, pitch, max_bonds) = cases.two_particle_bond_system(d, bits)
    par = np.array([spec.pack(params, bits)])
    hs.bond(spec.potential_id, par, n_bonds, bondlist, bond_pos, pitch, max_bonds, 1, which=which)
    assert hs.flags == 0
    cases.assert_golden(hs.force[:, 3], [0.5 * energy, 0.5 * energy], bits)
    cases.assert_golden(hs.force[:, :3], [[-force, 0, 0], [force, 0, 0]], bits)


@pytest.mark.parametrize("which", WHICH)
def test_bond_invalid_parameters_raise_flag(which):
    # BondEvaluatorDoubleWell.h:101-102 (r_diff == 0) and BondEvaluatorQuartic.h:134-135 (r_0 == 0) return false
    for api, params in (("DoubleWell", {"r_0": 1.0, "r_1": 1.0, "U_1": 1.0, "U_tilt": 0.0}),
                        ("Quartic", dict(k=1.0, r_0=0.0, b_1=0.0, b_2=0.0, U_0=1.0, sigma=1.0, epsilon=1.0, delta=0.0))):
        spec = cases.SPEC_BY_API[api]
        hs, (n_bonds, bondlist, bond_pos, pitch, max_bonds) = cases.two_particle_bond_system(1.0, 64)
        hs.bond(spec.potential_id, np.array([spec.pack(params, 64)]), n_bonds, bondlist, bond_pos, pitch, max_bonds, 1, which=which)
        assert hs.flags == 1
        assert np.all(hs.force == 0)


def test_philox_known_answers():
    import ctypes as C

    for ctr, key, out in G.PHILOX_KAT:
        c = (C.c_uint32 * 4)(*ctr)
        k = (C.c_uint32 * 2)(*key)
        o = (C.c_uint32 * 4)()
        O.port().azo_philox(c, k, o)
        assert tuple(o) == out


def test_dpd_alpha_range_and_symmetry():
    rng = np.random.default_rng(5)
    tags = rng.integers(0, 2**32, size=(200, 2), dtype=np.uint64)
    for bits in (32, 64):
        a = np.array([O.port().azo_dpd_alpha(bits, int(i), int(j), 12345, 42) for i, j in tags])
        b = np.array([O.port().azo_dpd_alpha(bits, int(j), int(i), 12345, 42) for i, j in tags])
        assert np.array_equal(a, b)  # ordered tags: same number for (i,j) and (j,i)
        assert np.all(a > -1.0) and np.all(a <= 1.0)
        assert abs(a.mean()) < 0.15
    # the float stream is the double stream's leading word
    a32 = O.port().azo_dpd_alpha(32, 7, 9, 3, 1)
    a64 = O.port().azo_dpd_alpha(64, 7, 9, 3, 1)
    assert abs(a32 - a64) < 2.0**-22
    if O.have_ref():
        for bits in (32, 64):
            fn = getattr(O.ref(), f"azref{bits}_dpd_alpha")
            for i, j in tags[:50]:
                assert fn(int(i), int(j), 99, 7) == O.port().azo_dpd_alpha(bits, int(i), int(j), 99, 7)


def test_oracle_cell_list_against_brute_force_scan():
    """The oracle's cell-list neighbour builder (the checker of the GPU builder) pinned by an O(N^2) numpy scan: 2 000 particles,
    two types with per-pair r_list (one pair switched off), periodic in y and z only, ghosts beyond N as candidates."""
    import azplugins_b200 as az

    rng = np.random.default_rng(17)
    n_local, n_ghost = 2000, 300
    L = np.array([9.0, 11.0, 10.0])
    pos = rng.uniform(-0.5, 0.5, size=(n_local + n_ghost, 3)) * L
    pos[n_local:, 0] = np.where(rng.random(n_ghost) < 0.5, rng.uniform(-L[0] / 2 - 1.5, -L[0] / 2, n_ghost), rng.uniform(L[0] / 2, L[0] / 2 + 1.5, n_ghost))
    typeid = rng.integers(0, 2, size=len(pos))
    r_list = np.array([[1.5, 1.1], [1.1, 0.0]])
    box = az.Box(*L, periodic=(False, True, True))
    for bits in (32, 64):
        hs = O.HostSystem(box, pos, typeid, 2, bits, n_local=n_local).build_nlist(r_list)
        p = hs.pos[:, :3].astype(np.float64)  # the coordinates the builder saw (rounded to Scalar)
        total = 0
        for i in range(n_local):
            d = p[i] - p
            d[:, 1:] -= L[1:] * np.rint(d[:, 1:] / L[1:])
            rsq = np.sum(d * d, axis=1)
            rl = r_list[typeid[i], typeid]
            near = (rl > 0) & (rsq < rl * rl)
            near[i] = False
            # pairs within float rounding of the list radius may go either way
            edge = (rl > 0) & (np.abs(rsq - rl * rl) < (1e-5 if bits == 32 else 1e-12) * rl * rl)
            row = hs.nlist[int(hs.head_list[i]) : int(hs.head_list[i]) + int(hs.n_neigh[i])]
            got = np.zeros(len(p), dtype=bool)
            got[row] = True
            assert len(np.unique(row)) == len(row)
            assert np.array_equal(got & ~edge, near & ~edge), (bits, i)
            total += len(row)
        assert total == hs.nlist.size and total > 10 * n_local
