"""The plain-C restatement against the reference's evaluator headers compiled unmodified (oracle/_ref), on seeded
random systems: every potential, both precisions, shift / xplor modes, virials, half and full lists."""

import numpy as np
import pytest

import cases
import oracle_api as O
from azplugins_b200 import _abi
from azplugins_b200.state import build_bond_table

pytestmark = pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref/libazref.so not built (needs /root/reference)")

PAIRS = {
    "PerturbedLennardJones": ({(0, 0): dict(epsilon=1.0, sigma=1.0, attraction_scale_factor=0.5), (0, 1): dict(epsilon=1.5, sigma=0.9, attraction_scale_factor=0.0),
                               (1, 1): dict(epsilon=0.0, sigma=1.0, attraction_scale_factor=1.0)}, {(0, 0): 3.0, (0, 1): 2.5, (1, 1): 2.0}),
    "Colloid": ({(0, 0): dict(A=144.0, a_1=0.0, a_2=0.0, sigma=1.0), (0, 1): dict(A=100.0, a_1=0.0, a_2=0.3, sigma=1.0),
                 (1, 1): dict(A=40.0, a_1=0.4, a_2=0.3, sigma=1.0)}, {(0, 0): 3.0, (0, 1): 3.0, (1, 1): 3.0}),
    "ExpandedYukawa": ({(0, 0): dict(epsilon=1.0, kappa=1.0, delta=0.2), (0, 1): dict(epsilon=2.0, kappa=0.5, delta=0.1),
                        (1, 1): dict(epsilon=0.0, kappa=1.0, delta=0.0)}, {(0, 0): 2.0, (0, 1): 2.5, (1, 1): 2.0}),
    "Hertz": ({(0, 0): dict(epsilon=100.0), (0, 1): dict(epsilon=50.0), (1, 1): dict(epsilon=0.0)}, {(0, 0): 1.2, (0, 1): 1.5, (1, 1): 1.0}),
}


def _close(a, b, bits):
    # same algorithm, same precision, different compilers' contraction choices: a few ulp on each term
    tol = 2e-5 if bits == 32 else 1e-11
    scale = max(np.abs(b).max(), 1e-30)
    np.testing.assert_allclose(a, b, rtol=tol, atol=tol * scale)


@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("api", list(PAIRS))
def test_pair_port_vs_ref(api, mode, bits):
    spec = cases.SPEC_BY_API[api]
    box, pos, typeid = cases.random_fluid(700, 0.7, 2, seed=11)
    par_by, rc_by = PAIRS[api]
    ron_by = {k: 0.8 * v for k, v in rc_by.items()}
    ron_by[(0, 1)] = rc_by[(0, 1)] + 0.5  # r_on > r_cut: xplor degenerates to shift for this pair
    par, rcsq, ronsq, rc = cases.pair_tables(spec, par_by, rc_by, 2, bits, ron_by)
    a = O.HostSystem(box, pos, typeid, 2, bits).build_nlist(rc + 0.3)
    b = O.HostSystem(box, pos, typeid, 2, bits)
    b.n_neigh, b.head_list, b.nlist = a.n_neigh, a.head_list, a.nlist
    a.pair(spec.potential_id, par, rcsq, ronsq, shift_mode=mode, which="port")
    b.pair(spec.potential_id, par, rcsq, ronsq, shift_mode=mode, which="ref")
    assert np.abs(b.force).max() > 0
    _close(a.force, b.force, bits)
    _close(a.virial, b.virial, bits)
    # half list + third law sums to the same thing
    h = a.full_to_half()
    h.pair(spec.potential_id, par, rcsq, ronsq, shift_mode=mode, half_list=True, which="ref")
    tol = 5e-4 if bits == 32 else 1e-9
    np.testing.assert_allclose(h.force, b.force, rtol=tol, atol=tol * np.abs(b.force).max())
    np.testing.assert_allclose(h.virial, b.virial, rtol=tol, atol=tol * np.abs(b.virial).max())


@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("s_exp", [2.0, 0.5])
def test_dpd_port_vs_ref(s_exp, bits):
    spec = cases.SPEC_BY_API["DPDGeneralWeight"]
    box, pos, typeid = cases.random_fluid(1500, 3.0, 2, seed=21, jitter=0.5)
    rng = np.random.default_rng(22)
    vel = rng.normal(size=pos.shape)
    tag = rng.permutation(len(pos)).astype(np.uint32)
    par_by = {(0, 0): dict(A=25.0, gamma=4.5, s=s_exp), (0, 1): dict(A=50.0, gamma=3.0, s=s_exp), (1, 1): dict(A=0.0, gamma=4.5, s=s_exp)}
    par, rcsq, _, rc = cases.pair_tables(spec, par_by, {(0, 0): 1.0, (0, 1): 1.0, (1, 1): 0.9}, 2, bits)
    for timestep in (0, 7, 2**32 - 1, 2**32 + 7):
        a = O.HostSystem(box, pos, typeid, 2, bits, velocity=vel, tag=tag).build_nlist(rc + 0.4)
        b = O.HostSystem(box, pos, typeid, 2, bits, velocity=vel, tag=tag)
        b.n_neigh, b.head_list, b.nlist = a.n_neigh, a.head_list, a.nlist
        a.dpd(par, rcsq, seed=42, timestep=timestep, dt=0.01, kT=1.0, which="port")
        b.dpd(par, rcsq, seed=42, timestep=timestep, dt=0.01, kT=1.0, which="ref")
        _close(a.force, b.force, bits)
        _close(a.virial, b.virial, bits)
        if timestep == 2**32 + 7:  # the evaluator truncates the timestep to 32 bits: same stream as timestep 7
            assert np.array_equal(b.force, prev7)
        if timestep == 7:
            prev7 = b.force.copy()


@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("mode", [0, 1])
def test_aniso_port_vs_ref(mode, bits):
    spec = cases.SPEC_BY_API["TwoPatchMorse"]
    box, pos, typeid = cases.random_fluid(800, 0.6, 2, seed=31)
    q = cases.random_unit_quaternions(len(pos), 32)
    par_by = {(0, 0): dict(M_d=1.8347, M_r=0.0302, r_eq=1.0043, omega=20.0, alpha=0.5, repulsion=True),
              (0, 1): dict(M_d=1.0, M_r=0.1, r_eq=1.0, omega=5.0, alpha=0.4, repulsion=False),
              (1, 1): dict(M_d=0.0, M_r=0.1, r_eq=1.0, omega=5.0, alpha=0.4, repulsion=False)}
    par, rcsq, _, rc = cases.pair_tables(spec, par_by, {(0, 0): 1.6, (0, 1): 2.0, (1, 1): 1.6}, 2, bits)
    a = O.HostSystem(box, pos, typeid, 2, bits, orientation=q).build_nlist(rc + 0.4)
    b = O.HostSystem(box, pos, typeid, 2, bits, orientation=q)
    b.n_neigh, b.head_list, b.nlist = a.n_neigh, a.head_list, a.nlist
    a.aniso(par, rcsq, shift_mode=mode, which="port")
    b.aniso(par, rcsq, shift_mode=mode, which="ref")
    assert np.abs(b.torque).max() > 0
    _close(a.force, b.force, bits)
    _close(a.torque, b.torque, bits)
    _close(a.virial, b.virial, bits)
    h = a.full_to_half()
    h.aniso(par, rcsq, shift_mode=mode, half_list=True, which="ref")
    tol = 5e-4 if bits == 32 else 1e-9
    np.testing.assert_allclose(h.force, b.force, rtol=tol, atol=tol * np.abs(b.force).max())
    np.testing.assert_allclose(h.torque, b.torque, rtol=tol, atol=tol * np.abs(b.torque).max())


@pytest.mark.parametrize("bits", [32, 64])
def test_bond_port_vs_ref(bits):
    rng = np.random.default_rng(41)
    n = 600
    box, pos, typeid = cases.random_fluid(n, 0.5, 1, seed=42)
    bonds = np.stack([np.arange(n - 1), np.arange(1, n)], axis=1)
    bonds = bonds[rng.random(len(bonds)) < 0.8]
    btype = rng.integers(0, 2, size=len(bonds))
    table = build_bond_table(n, n, bonds, btype)
    for api, plist in (("DoubleWell", [dict(r_0=1.0, r_1=2.0, U_1=1.0, U_tilt=0.5), dict(r_0=0.5, r_1=2.5, U_1=5.0, U_tilt=0.0)]),
                       ("Quartic", [dict(k=1434.3, r_0=1.5, b_1=-0.7589, b_2=0.0, U_0=67.2234, sigma=1.0, epsilon=1.0, delta=0.0),
                                    dict(k=100.0, r_0=1.2, b_1=-0.5, b_2=0.1, U_0=10.0, sigma=0.8, epsilon=1.0, delta=0.3)])):
        spec = cases.SPEC_BY_API[api]
        par = np.array([spec.pack(p, bits) for p in plist])
        a = O.HostSystem(box, pos, typeid, 1, bits)
        b = O.HostSystem(box, pos, typeid, 1, bits)
        a.bond(spec.potential_id, par, *table, 2, which="port")
        b.bond(spec.potential_id, par, *table, 2, which="ref")
        assert a.flags == 0 and b.flags == 0
        _close(a.force, b.force, bits)
        _close(a.virial, b.virial, bits)
