"""Synthetic configurations (azplugins_b200/workloads.py): the properties the bench and the full-size parity tests rely on.
CPU only; the C4b mixtures are also evaluated once through the oracle at a small size (finite forces, third law)."""

import numpy as np

import cases
import oracle_api as O
from azplugins_b200 import workloads as W


def _tables(spec, pdict, rc, names, bits=64):
    T = len(names)
    key = lambda i, j: tuple(sorted((names[i], names[j])))  # noqa: E731
    par_by = {(min(i, j), max(i, j)): pdict[key(i, j)] for i in range(T) for j in range(T)}
    rc_by = {(min(i, j), max(i, j)): rc[key(i, j)] for i in range(T) for j in range(T)}
    return cases.pair_tables(spec, par_by, rc_by, T, bits)


def test_colloid_mixture_keeps_the_solvent_out_of_the_colloids():
    snap = W.colloid_mixture(m=16)
    L = snap.box.L
    pos, typ = snap.position, snap.typeid
    col, sol = pos[typ == 1], pos[typ == 0]
    assert len(col) > 8 and len(sol) > 10 * len(col)
    # nearest colloid of every solvent particle (minimum image): outside radius + 0.9, where the colloid-solvent potential
    # (divergent at r = radius) is steep but finite
    dmin = np.full(len(sol), np.inf)
    for c in col:
        d = sol - c
        d -= L * np.rint(d / L)
        dmin = np.minimum(dmin, np.sqrt(np.sum(d * d, axis=1)))
    assert dmin.min() > W.COLLOID_RADIUS + 0.9 - 1e-9
    # colloids sit on a cubic lattice whose spacing brings nearest neighbours inside the colloid-colloid cutoff
    d = col[1:] - col[0]
    d -= L * np.rint(d / L)
    nearest = np.sqrt(np.sum(d * d, axis=1)).min()
    assert 2 * W.COLLOID_RADIUS + 1.0 < nearest < W.COLLOID_RCUT[("C", "C")]
    assert np.all(np.abs(pos) <= L / 2 + 1e-9)


def test_hertz_mixture_species_and_cutoffs():
    snap = W.hertz_mixture(m=8)
    assert snap.N == 4 * 8**3 and sorted(snap.types) == ["A", "B", "C", "D"]
    counts = np.bincount(snap.typeid, minlength=4)
    assert counts.min() > 0.2 * snap.N
    par, rc = W.hertz_tables()
    assert len(par) == 10 and set(par) == set(rc)
    for (a, b), r in rc.items():
        assert abs(r - 0.5 * (W.HERTZ_DIAMETERS[a] + W.HERTZ_DIAMETERS[b])) < 1e-12


def _third_law_through_oracle(snap, spec, pdict, rc, buffer=0.4):
    names = snap.types
    par, rcsq, ronsq, rcm = _tables(spec, pdict, rc, names)
    hs = O.HostSystem(snap.box, snap.position, snap.typeid, len(names), 64)
    hs.build_nlist(np.where(rcm > 0, rcm + buffer, 0.0))
    hs.pair(spec.potential_id, par, rcsq, ronsq, compute_virial=True, which="port", nthreads=2)
    f = hs.force[:, :3]
    assert np.all(np.isfinite(hs.force)) and np.all(np.isfinite(hs.virial))
    assert np.abs(f.sum(axis=0)).max() <= 1e-9 * np.abs(f).sum()
    return hs


def test_colloid_mixture_through_the_oracle():
    snap = W.colloid_mixture(m=12)
    hs = _third_law_through_oracle(snap, cases.SPEC_BY_API["Colloid"], W.COLLOID_PARAMS, W.COLLOID_RCUT)
    # all three regimes are exercised: colloids feel forces (colloid-colloid and colloid-solvent pairs in range)
    assert np.abs(hs.force[snap.typeid == 1, :3]).max() > 0


def test_hertz_mixture_through_the_oracle():
    snap = W.hertz_mixture(m=6)
    par, rc = W.hertz_tables()
    hs = _third_law_through_oracle(snap, cases.SPEC_BY_API["Hertz"], par, rc)
    assert np.count_nonzero(hs.force[:, 3]) > 0.5 * snap.N
