"""Committed reference outputs (tests/golden/ref_vectors.npz, written by tests/golden/make_golden.py from the reference's own
evaluator headers) against (i) the C port of the oracle — CPU, runs anywhere — and (ii) the CUDA path through the product API
(-m gpu). Double precision; per-particle error relative to max(|f_i|, rms |f|)."""

import os

import numpy as np
import pytest

import cases
import golden_systems as GS
import oracle_api as O

FIXTURE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vectors.npz")
CASES = GS.all_cases()
IDS = [c["name"] for c in CASES]


@pytest.fixture(scope="module")
def golden():
    return np.load(FIXTURE)


def _rel(got, want):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    scale = np.sqrt(np.mean(np.sum(want**2, axis=-1))) if want.ndim > 1 else np.sqrt(np.mean(want**2))
    den = np.maximum(np.linalg.norm(want, axis=-1) if want.ndim > 1 else np.abs(want), max(scale, 1e-300))
    num = np.linalg.norm(got - want, axis=-1) if want.ndim > 1 else np.abs(got - want)
    return float(np.max(num / den)) if num.size else 0.0


def _compare(name, golden, force, virial, torque, tol):
    want_f = golden[name + "/force"]
    assert _rel(np.asarray(force)[:, :3], want_f[:, :3]) <= tol, name
    assert abs(np.sum(np.asarray(force, dtype=np.float64)[:, 3]) - want_f[:, 3].sum()) <= tol * max(1.0, np.abs(want_f[:, 3]).sum()), name
    assert _rel(np.asarray(virial).T, golden[name + "/virial"].T) <= tol, name
    if name + "/torque" in golden.files:
        assert _rel(np.asarray(torque)[:, :3], golden[name + "/torque"][:, :3]) <= tol, name


@pytest.mark.parametrize("case", CASES, ids=IDS)
def test_port_reproduces_reference_outputs(case, golden):
    s = GS.system(case)
    assert np.array_equal(s["pos"], golden[case["name"] + "/pos"]), "input recipe changed: regenerate tests/golden/ref_vectors.npz"
    r = GS.run_oracle(case, "port")
    _compare(case["name"], golden, r["force"], r["virial"], r["torque"], 1e-11)
    assert np.abs(golden[case["name"] + "/force"]).max() > 0


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=IDS)
def test_cuda_reproduces_reference_outputs(case, golden):
    import azplugins_b200 as az
    import gpu_cases as GC
    from azplugins_b200 import bond, external, nlist, pair

    s = GS.system(case)
    kind = case["kind"]
    names = ["A", "B"]
    if kind == "barrier":
        cls = external.PlanarHarmonicBarrier if case["geometry"] == 0 else external.SphericalHarmonicBarrier
        st = az.State(az.Snapshot(s["box"], s["pos"], typeid=s["typeid"], types=names), device="cuda", scalar_bits=64)
        f = cls(location=case["location"])
        f.params["A"], f.params["B"] = dict(k=30.0, offset=0.25), dict(k=120.0, offset=-0.5)
        f._attach(st)
        f.compute(0)
        _compare(case["name"], golden, np.c_[f.forces, f.energies], f.virials.T, None, 1e-12)
        return
    r = GS.run_oracle(case, "port")  # for the neighbour list / bond selection the fixture was made with (deterministic)
    if kind == "bond":
        n = len(s["pos"])
        snap = az.Snapshot(s["box"], s["pos"], typeid=np.zeros(n, dtype=np.int64), types=["A"], bond_group=r["bonds"], bond_typeid=r["btype"],
                           bond_types=["b0", "b1"])
        st = az.State(snap, device="cuda", scalar_bits=64)
        f = getattr(bond, case["api"])()
        f.params["b0"], f.params["b1"] = GS.BOND_PARAMS[case["api"]]
        f._attach(st)
        f.compute(0).check_flags()
        _compare(case["name"], golden, np.c_[f.forces, f.energies], f.virials.T, None, 1e-9)
        return
    snap = az.Snapshot(s["box"], s["pos"], typeid=s["typeid"], types=names, velocity=s["vel"], tag=s["tag"], orientation=s["ori"])
    st = az.State(snap, device="cuda", scalar_bits=64, seed=42)
    nl = nlist.NeighborList(buffer=0.3)
    if kind == "pair":
        par_by, rc_by = GS.PAIR_PARAMS[case["api"]]
        f = getattr(pair, case["api"])(nlist=nl, default_r_cut=0.0, mode=case["mode"])
        for k in rc_by:
            f.r_on[(names[k[0]], names[k[1]])] = 0.8 * rc_by[k]
    elif kind == "dpd":
        par_by, rc_by = GS.DPD_PARAMS
        f = pair.DPDGeneralWeight(nlist=nl, kT=1.0, default_r_cut=0.0)
        f.dt = 0.01
    else:
        par_by, rc_by = GS.TPM_PARAMS
        f = pair.TwoPatchMorse(nlist=nl, default_r_cut=0.0, mode=case["mode"])
    for k, p in par_by.items():
        key = (names[k[0]], names[k[1]])
        f.params[key] = p
        f.r_cut[key] = rc_by[k]
    GC.attach_with_oracle_nlist(f, st, r["hs"])
    f.compute(case.get("timestep", 0))
    torque = f.torques if kind == "aniso" else None
    _compare(case["name"], golden, np.c_[f.forces, f.energies], f.virials.T, torque, 1e-9)
