"""Host logic without a GPU: param struct packing against the reference's own pybind11 constructors (bit-exact),
asDict round trips (what test_pair.py:349 / test_bond.py:223 / test_pair_aniso.py:147-149 assert), the C-ABI
library's exported symbols, the bond table layout and the API surface."""

import ctypes as C
import re

import numpy as np
import pytest

import cases
import golden_cases as G
import oracle_api as O
from azplugins_b200 import _abi, bond, nlist, pair, params as P
from azplugins_b200.state import build_bond_table

ALL_PARAM_DICTS = ([(c[0], c[1]) for c in G.PAIR_CASES] + [("TwoPatchMorse", c[0]) for c in G.ANISO_CASES]
                   + [(c[0], c[1]) for c in G.BOND_CASES])


@pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("bits", [32, 64])
def test_pack_matches_reference_constructors(bits):
    for api, d in ALL_PARAM_DICTS:
        spec = cases.SPEC_BY_API[api]
        mine = spec.pack(d, bits)
        theirs = O.ref_make_params(spec, {k: (bool(v) if k == "repulsion" else float(v)) for k, v in d.items()}, bits)
        for name in mine.dtype.names:
            assert mine[name].tobytes() == theirs[name].tobytes(), (api, name, mine[name], theirs[name])
        back_ref = O.ref_params_as_dict(spec, theirs, bits)
        back = spec.unpack(mine)
        assert set(back) == set(back_ref)
        for k in back:
            assert back[k] == pytest.approx(back_ref[k], rel=1e-6, abs=1e-12)


@pytest.mark.parametrize("bits", [32, 64])
def test_param_round_trip(bits):
    for api, d in ALL_PARAM_DICTS:
        spec = cases.SPEC_BY_API[api]
        back = spec.unpack(spec.pack(d, bits))
        tol = 1e-5 if bits == 32 else 1e-12
        for k, v in d.items():
            assert back[k] == pytest.approx(float(v), rel=tol, abs=tol)


def test_param_layout_matches_library_and_reference():
    lib = _abi.load()
    for spec in P.ALL:
        for bits in (32, 64):
            assert spec.dtype(bits).itemsize == lib.azb_param_size(spec.kind, spec.potential_id, bits)
            if O.have_ref():
                assert spec.dtype(bits).itemsize == getattr(O.ref(), f"azref{bits}_param_size")(spec.kind, spec.potential_id)
        assert lib.azb_potential_name(spec.kind, spec.potential_id).decode() == spec.name
    assert lib.azb_param_size(0, 99, 32) == 0 and lib.azb_potential_name(3, 9) is None


def test_library_exports_every_declared_symbol():
    lib = _abi.load()
    header = open(cases.__file__.replace("tests/cases.py", "include/azb200.h")).read()
    declared = set(re.findall(r"AZB_API\s+[\w\s\*]+?\b(azb_\w+)\s*\(", header))
    assert declared == set(_abi.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name


def test_abi_rejects_bad_arguments_without_a_gpu():
    lib = _abi.load()
    a = _abi.PairArgs()
    assert lib.azb_pair_forces(0, 32, C.byref(a), None, None) == _abi.ERR_BAD_ARGUMENT  # ntypes == 0
    a.ntypes = 1
    assert lib.azb_pair_forces(0, 32, C.byref(a), None, None) == _abi.ERR_BAD_ARGUMENT  # no params
    dummy = C.c_double(0)
    assert lib.azb_pair_forces(77, 32, C.byref(a), C.addressof(dummy), None) == _abi.ERR_BAD_POTENTIAL
    assert lib.azb_pair_forces(0, 16, C.byref(a), C.addressof(dummy), None) == _abi.ERR_BAD_SCALAR
    assert lib.azb_pair_forces(0, 32, C.byref(a), C.addressof(dummy), None) == 0  # N == 0: nothing to launch
    b = _abi.BondArgs()
    assert lib.azb_bond_forces(5, 32, C.byref(b), C.addressof(dummy), None, None) == _abi.ERR_BAD_POTENTIAL


def test_bond_table_layout():
    n_bonds, bl, bp, pitch, mx = build_bond_table(4, 5, np.array([[0, 1], [1, 2], [3, 4], [1, 3]]), np.array([0, 1, 0, 1]))
    assert list(n_bonds) == [1, 3, 1, 2] and pitch == 4 and mx == 3
    bl = bl.reshape(mx, pitch, 2)
    assert sorted(map(tuple, bl[: n_bonds[1], 1])) == [(0, 0), (2, 1), (3, 1)]
    assert tuple(bl[0, 3]) in ((4, 0), (1, 1))  # partner 4 is a ghost (index >= N): still listed for local 3
    with pytest.raises(ValueError):
        build_bond_table(2, 2, np.array([[0, 5]]), np.array([0]))


def test_api_surface_matches_reference_python_module():
    # /root/reference/src/pair.py:106-108,213-215,287-288,340-342,413-415,507-508; bond.py:53-54,137-138
    nl = nlist.Cell(buffer=0.4)
    expect = {pair.Colloid: "PotentialPairColloid", pair.ExpandedYukawa: "PotentialPairExpandedYukawa", pair.Hertz: "PotentialPairHertz",
              pair.PerturbedLennardJones: "PotentialPairPerturbedLennardJones"}
    for cls, name in expect.items():
        f = cls(nlist=nl, default_r_cut=3.0, default_r_on=0, mode="none")
        assert f._cpp_class_name == name
    assert pair.Colloid._accepted_modes == ("none", "shift", "xplor")
    assert pair.DPDGeneralWeight._accepted_modes == ("none",)
    assert pair.DPDGeneralWeight(nlist=nl, kT=1.0, default_r_cut=1.0)._cpp_class_name == "PotentialPairDPDThermoGeneralWeight"
    assert pair.TwoPatchMorse(nlist=nl, default_r_cut=1.6, mode="shift")._cpp_class_name == "AnisoPotentialPairTwoPatchMorse"
    assert bond.DoubleWell()._cpp_class_name == "PotentialBondDoubleWell"
    assert bond.Quartic()._cpp_class_name == "PotentialBondQuartic"
    with pytest.raises(ValueError):
        pair.DPDGeneralWeight(nlist=nl, kT=1.0, default_r_cut=1.0).__class__.__mro__[1].__init__  # noqa: B018
        pair.Hertz(nlist=nl, default_r_cut=1.0, mode="bogus")
    plj = pair.PerturbedLennardJones(nlist=nl, default_r_cut=3.0)
    plj.params[("A", "B")] = dict(epsilon=1.0, sigma=1.0, attraction_scale_factor=0.5)
    assert plj.params[("B", "A")]["sigma"] == 1.0  # symmetric pair keys
    with pytest.raises(KeyError):
        plj.params[("A", "A")] = dict(epsilon=1.0, sigma=1.0)  # missing key
    q = bond.Quartic()
    q.params["A-A"] = dict(k=1.0, r_0=1.5, b_1=0.0, b_2=0.0, U_0=1.0, sigma=1.0, epsilon=1.0)  # delta defaults to 0 (bond.py:153)
    assert q.params["A-A"]["delta"] == 0.0


def test_force_refuses_cpu_state():
    import azplugins_b200 as az

    snap = az.Snapshot(az.Box.cube(10.0), np.array([[0.0, 0, 0], [1.0, 0, 0]]))
    st = az.State(snap, device="cpu", scalar_bits=32)
    f = pair.Hertz(nlist=nlist.Cell(), default_r_cut=1.5)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        f._attach(st)
