"""Build the same seeded system twice — as device State + azplugins_b200 force object (the product path through
the C-ABI) and as an oracle HostSystem — and return both results for comparison."""

import numpy as np

import azplugins_b200 as az
import cases
import oracle_api as O
from azplugins_b200 import bond, nlist, pair

TYPE_NAMES = ["A", "B", "C", "D"]


def error_metrics(got, want):
    """Per-particle vector error normalised by max(|want_i|, rms(|want|)); returns (max, where)."""
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    if got.ndim == 1:
        got, want = got[:, None], want[:, None]
    mag = np.linalg.norm(want, axis=1)
    rms = np.sqrt(np.mean(mag**2)) if len(mag) else 0.0
    denom = np.maximum(mag, rms)
    denom[denom == 0] = 1.0
    err = np.linalg.norm(got - want, axis=1) / denom
    i = int(np.argmax(err)) if len(err) else 0
    return (float(err[i]) if len(err) else 0.0), i


def _with_periodic(box, periodic):
    """Same extent, other periodic flags (what a domain-decomposed rank hands to the kernels)."""
    return box if periodic is None else az.Box(box.Lx, box.Ly, box.Lz, periodic=tuple(periodic))


def make_state(box, pos, typeid, ntypes, bits, **kw):
    snap = az.Snapshot(box, pos, typeid=typeid, types=TYPE_NAMES[:ntypes], **kw)
    return az.State(snap, device="cuda", scalar_bits=bits)


def attach_with_oracle_nlist(force, state, hs):
    """Hand the oracle-built neighbour arrays (HOOMD layout) to the product force object."""
    force.nlist.set_arrays(hs.n_neigh[: state.N], hs.head_list[: state.N], hs.nlist, state.device)
    force.nlist._state = state
    force._attach(state)


def run_pair(api, par_by, rc_by, bits, n=4000, rho=0.7, ntypes=2, mode="none", ron_by=None, seed=101, tpp=0, virial=True, which=None,
             buffer=0.3, system=None, use_tiles=True, tile_size=128, periodic=None):
    spec = cases.SPEC_BY_API[api]
    box, pos, typeid = system if system is not None else cases.random_fluid(n, rho, ntypes, seed)
    box = _with_periodic(box, periodic)
    par, rcsq, ronsq, rc = cases.pair_tables(spec, par_by, rc_by, ntypes, bits, ron_by)
    r_list = np.where(rc > 0, rc + buffer, 0.0)
    # oracle in double on the same (bits-rounded) coordinates: convert through the Scalar type first
    hs_b = O.HostSystem(box, pos, typeid, ntypes, bits).build_nlist(r_list)
    pos_r = hs_b.pos[:, :3].astype(np.float64)
    par64, rcsq64, ronsq64, _ = cases.pair_tables(spec, par_by, rc_by, ntypes, 64, ron_by)
    hs = O.HostSystem(box, pos_r, typeid, ntypes, 64)
    hs.n_neigh, hs.head_list, hs.nlist = hs_b.n_neigh, hs_b.head_list, hs_b.nlist
    shift = {"none": 0, "shift": 1, "xplor": 2}[mode]
    hs.pair(spec.potential_id, par64, rcsq64, ronsq64, shift_mode=shift, compute_virial=True,
            which=which or ("ref" if O.have_ref() else "port"), nthreads=4)
    # same-precision oracle: what the reference computes when built with Scalar = float / double
    hs_b.pair(spec.potential_id, par, rcsq, ronsq, shift_mode=shift, compute_virial=True,
              which=which or ("ref" if O.have_ref() else "port"), nthreads=4)
    hs_b.hs64 = hs

    st = make_state(box, pos, typeid, ntypes, bits)
    cls = getattr(pair, api)
    f = cls(nlist=nlist.NeighborList(buffer=buffer), default_r_cut=0.0, mode=mode)
    for (i, j), p in par_by.items():
        key = (TYPE_NAMES[i], TYPE_NAMES[j])
        f.params[key] = p
        f.r_cut[key] = rc_by[(i, j)]
        if ron_by is not None:
            f.r_on[key] = ron_by[(i, j)]
    f.threads_per_particle = tpp
    f.compute_virial = virial
    f.use_tiles, f.tile_size = use_tiles, tile_size
    attach_with_oracle_nlist(f, st, hs_b)
    f.compute(0)
    assert (f._plan is not None) == bool(use_tiles)
    return f, hs_b


def run_dpd(par_by, rc_by, bits, n=6000, rho=3.0, ntypes=2, seed=201, timestep=0, sim_seed=42, dt=0.01, kT=1.0, tpp=0, buffer=0.4,
            use_tiles=True, tile_size=128, periodic=None):
    spec = cases.SPEC_BY_API["DPDGeneralWeight"]
    box, pos, typeid = cases.random_fluid(n, rho, ntypes, seed, jitter=0.5)
    box = _with_periodic(box, periodic)
    rng = np.random.default_rng(seed + 1)
    vel = rng.normal(size=pos.shape)
    tag = rng.permutation(len(pos)).astype(np.uint32)
    par, rcsq, _, rc = cases.pair_tables(spec, par_by, rc_by, ntypes, bits)
    hs_b = O.HostSystem(box, pos, typeid, ntypes, bits, velocity=vel, tag=tag).build_nlist(rc + buffer)
    par64, rcsq64, _, _ = cases.pair_tables(spec, par_by, rc_by, ntypes, 64)
    hs = O.HostSystem(box, hs_b.pos[:, :3].astype(np.float64), typeid, ntypes, 64, velocity=hs_b.vel[:, :3].astype(np.float64), tag=tag)
    hs.n_neigh, hs.head_list, hs.nlist = hs_b.n_neigh, hs_b.head_list, hs_b.nlist
    hs.dpd(par64, rcsq64, seed=sim_seed, timestep=timestep, dt=dt, kT=kT, which="ref" if O.have_ref() else "port", nthreads=4)
    # same-precision oracle (bit-level stream check: float alpha uses one 32-bit word)
    hs_b.dpd(par, rcsq, seed=sim_seed, timestep=timestep, dt=dt, kT=kT, which="ref" if O.have_ref() else "port", nthreads=4)

    snap = az.Snapshot(box, pos, typeid=typeid, types=TYPE_NAMES[:ntypes], velocity=vel, tag=tag)
    st = az.State(snap, device="cuda", scalar_bits=bits, seed=sim_seed)
    f = pair.DPDGeneralWeight(nlist=nlist.NeighborList(buffer=buffer), kT=kT, default_r_cut=0.0)
    f.dt = dt
    for (i, j), p in par_by.items():
        key = (TYPE_NAMES[i], TYPE_NAMES[j])
        f.params[key] = p
        f.r_cut[key] = rc_by[(i, j)]
    f.threads_per_particle = tpp
    f.use_tiles, f.tile_size = use_tiles, tile_size
    attach_with_oracle_nlist(f, st, hs_b)
    f.compute(timestep)
    assert (f._plan is not None) == bool(use_tiles)
    hs_b.hs64 = hs
    return f, hs_b, hs


def run_aniso(par_by, rc_by, bits, n=5000, rho=0.6, ntypes=2, mode="none", seed=301, tpp=0, buffer=0.4, use_tiles=True, tile_size=128):
    spec = cases.SPEC_BY_API["TwoPatchMorse"]
    box, pos, typeid = cases.random_fluid(n, rho, ntypes, seed)
    q = cases.random_unit_quaternions(len(pos), seed + 1)
    par, rcsq, _, rc = cases.pair_tables(spec, par_by, rc_by, ntypes, bits)
    hs_b = O.HostSystem(box, pos, typeid, ntypes, bits, orientation=q).build_nlist(rc + buffer)
    par64, rcsq64, _, _ = cases.pair_tables(spec, par_by, rc_by, ntypes, 64)
    hs = O.HostSystem(box, hs_b.pos[:, :3].astype(np.float64), typeid, ntypes, 64, orientation=hs_b.ori.astype(np.float64))
    hs.n_neigh, hs.head_list, hs.nlist = hs_b.n_neigh, hs_b.head_list, hs_b.nlist
    hs.aniso(par64, rcsq64, shift_mode=1 if mode == "shift" else 0, which="ref" if O.have_ref() else "port", nthreads=4)
    hs_b.aniso(par, rcsq, shift_mode=1 if mode == "shift" else 0, which="ref" if O.have_ref() else "port", nthreads=4)
    hs_b.hs64 = hs

    snap = az.Snapshot(box, pos, typeid=typeid, types=TYPE_NAMES[:ntypes], orientation=q)
    st = az.State(snap, device="cuda", scalar_bits=bits)
    f = pair.TwoPatchMorse(nlist=nlist.NeighborList(buffer=buffer), default_r_cut=0.0, mode=mode)
    for (i, j), p in par_by.items():
        key = (TYPE_NAMES[i], TYPE_NAMES[j])
        f.params[key] = p
        f.r_cut[key] = rc_by[(i, j)]
    f.threads_per_particle = tpp
    f.use_tiles, f.tile_size = use_tiles, tile_size
    attach_with_oracle_nlist(f, st, hs_b)
    f.compute(0)
    assert (f._plan is not None) == bool(use_tiles)
    return f, hs_b


def run_bond(api, plist, bits, n=20000, seed=401, bond_fraction=0.9):
    spec = cases.SPEC_BY_API[api]
    rng = np.random.default_rng(seed)
    box, pos, typeid = cases.random_fluid(n, 0.5, 1, seed + 1)
    bonds = np.stack([np.arange(n - 1), np.arange(1, n)], axis=1)
    bonds = bonds[rng.random(len(bonds)) < bond_fraction]
    btype = rng.integers(0, len(plist), size=len(bonds))
    names = [f"b{i}" for i in range(len(plist))]
    snap = az.Snapshot(box, pos, typeid=typeid, types=["A"], bond_group=bonds, bond_typeid=btype, bond_types=names)
    st = az.State(snap, device="cuda", scalar_bits=bits)
    f = getattr(bond, api)()
    for name, p in zip(names, plist):
        f.params[name] = p
    f._attach(st)
    f.compute(0)

    ftype = np.float32 if bits == 32 else np.float64
    pos_r = pos.astype(ftype).astype(np.float64)
    hs = O.HostSystem(box, pos_r, typeid, 1, 64)
    from azplugins_b200.state import build_bond_table

    table = build_bond_table(n, n, bonds, btype)
    par64 = np.array([spec.pack(p, 64) for p in plist])
    hs.bond(spec.potential_id, par64, *table, len(plist), which="ref" if O.have_ref() else "port", nthreads=4)
    hs_b = O.HostSystem(box, pos, typeid, 1, bits)
    hs_b.bond(spec.potential_id, np.array([spec.pack(p, bits) for p in plist]), *table, len(plist), which="ref" if O.have_ref() else "port", nthreads=4)
    hs_b.hs64 = hs
    return f, hs_b


def cutoff_ambiguous(hs, rcutsq, rel=2e-6):
    """Particles owning a listed pair whose r^2 lies within `rel` of r_cut^2: a float pipeline and a double
    pipeline may legitimately disagree on whether such a pair is inside the cutoff, so float-vs-double comparisons
    skip these particles (they are compared in the same-precision tests instead)."""
    N = hs.N
    i_idx = np.repeat(np.arange(N), hs.n_neigh[:N].astype(np.int64))
    starts = hs.head_list[:N].astype(np.int64)
    offs = np.arange(len(i_idx)) - np.repeat(np.cumsum(hs.n_neigh[:N].astype(np.int64)) - hs.n_neigh[:N].astype(np.int64), hs.n_neigh[:N].astype(np.int64))
    j_idx = hs.nlist[np.repeat(starts, hs.n_neigh[:N].astype(np.int64)) + offs].astype(np.int64)
    pos = hs.pos[:, :3].astype(np.float64)
    L = hs.box.L
    d = pos[i_idx] - pos[j_idx]
    d -= L * np.rint(d / L)
    rsq = np.sum(d * d, axis=1)
    itype = np.int64 if hs.bits == 64 else np.int32
    types = np.ascontiguousarray(hs.pos[:, 3]).view(itype).astype(np.int64)
    rc = np.asarray(rcutsq, dtype=np.float64).reshape(-1)[types[i_idx] * hs.ntypes + types[j_idx]]
    amb = np.abs(rsq - rc) <= rel * rc
    mask = np.zeros(N, dtype=bool)
    mask[i_idx[amb]] = True
    return mask
