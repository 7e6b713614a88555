"""external.PlanarHarmonicBarrier / SphericalHarmonicBarrier.

Golden vectors: /root/reference/src/pytest/test_external.py:84-148 (spherical) and :151-223 (planar) — the same four
particles, per-type k / offset, and a barrier that sits at 5.0 for timesteps <= 1 and at 4.0 afterwards (the CustomVariant
of test_external.py:11-27); tolerance 1e-4 absolute as there. CPU tests pin the oracle (C port and the reference's own
evaluator headers in oracle/_ref) against them; the gpu tests run the CUDA path through the hoomd.azplugins.external API
and compare it with the oracle on random systems (triclinic boxes and particles that drifted out of the box included).
"""

import numpy as np
import pytest

import oracle_api as O
from azplugins_b200 import _abi

KA, DB = 50.0, 2.0
KB = KA * DB**2
PARAMS_1 = [(KA, 0.1), (KB, -0.1)]  # test_external.py:108-109
PARAMS_2 = [(KA, 0.1), (0.0, -0.1)]  # :132 — B switched off
TYPEID = [0, 1, 0, 0]
SPHERE_POS = [[0, 0, 4.6], [0, 0, -5.4], [0, 5.6, 0], [6.6, 0, 0]]  # :92-97
PLANE_POS = [[1, 4.6, 1], [-1, 5.4, 1], [1, 5.6, -1], [-1, 6.6, -1]]  # :159-164


def location(timestep):  # CustomVariant(z=5.0), test_external.py:16-21
    return 5.0 if timestep <= 1 else 4.0


# (geometry, positions, params, timestep) -> [fx, fy, fz, e] per particle
GOLDEN = [
    (_abi.BARRIER_SPHERICAL, SPHERE_POS, PARAMS_1, 1,
     [[0, 0, 0, 0], [0, 0, KB * 0.5, 0.5 * KB * 0.5**2], [0, -KA * 0.5, 0, 0.5 * KA * 0.5**2], [-KA * 1.5, 0, 0, 0.5 * KA * 1.5**2]]),  # :118-128
    (_abi.BARRIER_SPHERICAL, SPHERE_POS, PARAMS_2, 3,
     [[0, 0, -KA * 0.5, 0.5 * KA * 0.5**2], [0, 0, 0, 0], [0, -KA * 1.5, 0, 0.5 * KA * 1.5**2], [-KA * 2.5, 0, 0, 0.5 * KA * 2.5**2]]),  # :138-148
    (_abi.BARRIER_PLANAR, PLANE_POS, PARAMS_1, 1,
     [[0, 0, 0, 0], [0, -KB * 0.5, 0, 0.5 * KB * 0.5**2], [0, -KA * 0.5, 0, 0.5 * KA * 0.5**2], [0, -KA * 1.5, 0, 0.5 * KA * 1.5**2]]),  # :185-199
    (_abi.BARRIER_PLANAR, PLANE_POS, PARAMS_2, 3,
     [[0, -KA * 0.5, 0, 0.5 * KA * 0.5**2], [0, 0, 0, 0], [0, -KA * 1.5, 0, 0.5 * KA * 1.5**2], [0, -KA * 2.5, 0, 0.5 * KA * 2.5**2]]),  # :209-223
]


def _box20():
    import azplugins_b200 as az

    return az.Box(20, 20, 20)


def _random_system(seed, n, triclinic):
    import azplugins_b200 as az

    rng = np.random.default_rng(seed)
    box = az.Box(24.0, 20.0, 22.0, xy=0.2 if triclinic else 0.0, xz=-0.1 if triclinic else 0.0, yz=0.15 if triclinic else 0.0)
    # fractional coordinates in [-0.6, 0.6): one in six particles has drifted out of the box and must be wrapped back
    f = rng.uniform(-0.6, 0.6, size=(n, 3))
    pos = f * box.L
    pos[:, 0] += box.xy * pos[:, 1] + box.xz * pos[:, 2]
    pos[:, 1] += box.yz * pos[:, 2]
    typeid = rng.integers(0, 3, size=n)
    params = [(30.0, 0.25), (120.0, -0.5), (0.0, 0.0)]
    return box, pos, typeid, params


@pytest.mark.parametrize("which", ["port", "ref"])
@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("case", range(len(GOLDEN)))
def test_oracle_matches_reference_tests(which, bits, case):
    if which == "ref" and not O.have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    geometry, pos, params, timestep, want = GOLDEN[case]
    rc, force, virial = O.barrier(geometry, _box20(), pos, TYPEID, params, location(timestep), bits, which=which)
    assert rc == 0
    np.testing.assert_allclose(force, np.array(want), atol=1e-4)
    assert np.all(virial == 0)  # HarmonicBarrier.h:176


@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("triclinic", [False, True])
@pytest.mark.parametrize("geometry,loc", [(_abi.BARRIER_PLANAR, 3.0), (_abi.BARRIER_SPHERICAL, 7.5)])
def test_port_matches_reference_headers(bits, triclinic, geometry, loc):
    if not O.have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    box, pos, typeid, params = _random_system(11, 4000, triclinic)
    rc_p, f_p, _ = O.barrier(geometry, box, pos, typeid, params, loc, bits, which="port")
    rc_r, f_r, _ = O.barrier(geometry, box, pos, typeid, params, loc, bits, which="ref")
    assert rc_p == 0 and rc_r == 0
    assert np.count_nonzero(f_r[:, 3]) > 100
    # identical operation order; gcc may contract a*b+c differently in the two translation units, and an ulp on a wrapped
    # coordinate (|y| ~ 10) is k * 1e-6 ~ 1e-4 on the force
    np.testing.assert_allclose(f_p, f_r, rtol=1e-5 if bits == 32 else 1e-13, atol=5e-4 if bits == 32 else 1e-12)


@pytest.mark.parametrize("which", ["port", "ref"])
def test_invalid_location(which):
    if which == "ref" and not O.have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    box = _box20()
    pos, typeid, par = [[0, 0, 0]], [0], [(1.0, 0.0)]
    # plane outside [lo.y, hi.y) (PlanarBarrierEvaluator.h:51-56); sphere with 2R > L or R < 0 (SphericalBarrierEvaluator.h:54-60)
    for geometry, loc, ok in ((_abi.BARRIER_PLANAR, 10.0, False), (_abi.BARRIER_PLANAR, -10.0, True), (_abi.BARRIER_PLANAR, 9.99, True),
                              (_abi.BARRIER_SPHERICAL, 10.01, False), (_abi.BARRIER_SPHERICAL, 10.0, True), (_abi.BARRIER_SPHERICAL, -0.1, False)):
        rc, _, _ = O.barrier(geometry, box, pos, typeid, par, loc, 64, which=which)
        assert (rc == 0) == ok, (geometry, loc, rc)
        assert ok or rc == _abi.ERR_BARRIER_INVALID


# ---- the CUDA path ----------------------------------------------------------------------------------------------------
def _attach(cls, box, pos, typeid, types, bits, loc):
    import azplugins_b200 as az

    snap = az.Snapshot(box, pos, typeid=typeid, types=types)
    st = az.State(snap, device="cuda", scalar_bits=bits)
    f = cls(location=loc)
    return f, st


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("name", ["PlanarHarmonicBarrier", "SphericalHarmonicBarrier"])
def test_gpu_reference_test_cases(name, bits):
    """test_external.py:84-223 replayed through the external API: run(1) then B switched off and run(2)."""
    from azplugins_b200 import external

    cls = getattr(external, name)
    geometry = cls._geometry
    pos = SPHERE_POS if geometry == _abi.BARRIER_SPHERICAL else PLANE_POS
    f, st = _attach(cls, _box20(), pos, TYPEID, ["A", "B"], bits, location)
    f.params["A"] = dict(k=KA, offset=0.1)
    f.params["B"] = dict(k=KB, offset=-0.1)
    f._attach(st)
    golden = [g for g in GOLDEN if g[0] == geometry]
    for t in (0, 1):
        f.compute(t)
    np.testing.assert_allclose(np.c_[f.forces, f.energies], np.array(golden[0][4]), atol=1e-4)
    f.params["B"] = dict(k=0.0, offset=-0.1)
    for t in (2, 3):
        f.compute(t)
    np.testing.assert_allclose(np.c_[f.forces, f.energies], np.array(golden[1][4]), atol=1e-4)
    assert np.all(f.virials == 0)
    assert f.launches == 4


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["PlanarHarmonicBarrier", "SphericalHarmonicBarrier"])
def test_gpu_create(name):
    """test_external.py:55-81: constructor arguments and parameters survive attaching and running."""
    from azplugins_b200 import external

    cls = getattr(external, name)
    f, st = _attach(cls, _box20(), [[0, 0, 0], [1, 1, 1]], [0, 0], ["A"], 64, 3.0)
    f.params["A"].update(dict(k=10.0, offset=0.5)) if dict.__contains__(f.params, "A") else f.params.__setitem__("A", dict(k=10.0, offset=0.5))
    assert f.location(0) == 3.0
    assert f.params["A"] == dict(k=10.0, offset=0.5)
    f._attach(st)
    f.compute(0)
    assert f.location(0) == 3.0
    assert f.params["A"] == dict(k=10.0, offset=0.5)
    assert f._cpp_class_name == name + "GPU"
    with pytest.raises(TypeError):
        external.HarmonicBarrier(location=1.0)
    with pytest.raises(KeyError):
        f.params["A"] = dict(k=1.0)


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("triclinic", [False, True])
@pytest.mark.parametrize("name,loc", [("PlanarHarmonicBarrier", 3.0), ("SphericalHarmonicBarrier", 7.5)])
def test_gpu_matches_oracle(name, loc, triclinic, bits):
    from azplugins_b200 import external

    cls = getattr(external, name)
    box, pos, typeid, params = _random_system(23, 200_003, triclinic)
    f, st = _attach(cls, box, pos, typeid, ["A", "B", "C"], bits, loc)
    for t, (k, off) in zip("ABC", params):
        f.params[t] = dict(k=k, offset=off)
    f._attach(st)
    f._virial.fill_(1.0)  # must come back zeroed
    f.compute(5)
    rc, want, _ = O.barrier(cls._geometry, box, pos, typeid, params, loc, bits, which="ref" if O.have_ref() else "port")
    assert rc == 0
    got = np.c_[f.forces, f.energies]
    # same operations in the same order; only the contraction of a*b+c into an FMA may differ between gcc and nvcc
    # (an ulp on a wrapped coordinate |y| ~ 10 is k * 1e-6 ~ 1e-4 on the force)
    np.testing.assert_allclose(got, want, rtol=1e-5 if bits == 32 else 1e-13, atol=5e-4 if bits == 32 else 1e-12)
    assert np.mean((got[:, 3] == 0) != (want[:, 3] == 0)) < 1e-4  # the same particles feel the barrier (up to dy = +-ulp)
    assert np.count_nonzero(want[:, 3]) > 1000
    assert np.all(f.virials == 0)


@pytest.mark.gpu
def test_gpu_invalid_location_raises():
    from azplugins_b200 import external

    f, st = _attach(external.PlanarHarmonicBarrier, _box20(), [[0, 0, 0]], [0], ["A"], 32, 10.0)
    f.params["A"] = dict(k=1.0, offset=0.0)
    f._attach(st)
    with pytest.raises(RuntimeError, match="Barrier position is invalid"):
        f.compute(0)
    g, st = _attach(external.SphericalHarmonicBarrier, _box20(), [[0, 0, 0]], [0], ["A"], 32, 10.5)
    g.params["A"] = dict(k=1.0, offset=0.0)
    g._attach(st)
    with pytest.raises(RuntimeError, match="Barrier position is invalid"):
        g.compute(0)
