"""Seeded small systems behind tests/golden/ref_vectors.npz: one case per potential family and mode, N = 400, double
precision. ``run_oracle(case, which)`` evaluates a case with the reference's own evaluator headers (which="ref",
oracle/_ref) or with the C port (which="port"); tests/golden/make_golden.py stores the "ref" results as the fixture."""

import numpy as np

import cases
import oracle_api as O
from azplugins_b200 import _abi
from azplugins_b200.state import build_bond_table

N = 400
BITS = 64

PAIR_PARAMS = {
    "PerturbedLennardJones": ({(0, 0): dict(epsilon=1.0, sigma=1.0, attraction_scale_factor=0.5), (0, 1): dict(epsilon=1.5, sigma=0.9, attraction_scale_factor=0.0),
                               (1, 1): dict(epsilon=0.7, sigma=1.1, attraction_scale_factor=1.0)}, {(0, 0): 3.0, (0, 1): 2.5, (1, 1): 2.0}),
    # radii small enough that no pair of the random fluid overlaps (the colloid-colloid energy is a log of (r^2 - 4a^2) terms)
    "Colloid": ({(0, 0): dict(A=144.0, a_1=0.0, a_2=0.0, sigma=1.0), (0, 1): dict(A=100.0, a_1=0.0, a_2=0.1, sigma=1.0),
                 (1, 1): dict(A=40.0, a_1=0.1, a_2=0.1, sigma=1.0)}, {(0, 0): 3.0, (0, 1): 3.0, (1, 1): 3.0}),
    "ExpandedYukawa": ({(0, 0): dict(epsilon=1.0, kappa=1.0, delta=0.2), (0, 1): dict(epsilon=2.0, kappa=0.5, delta=0.1),
                        (1, 1): dict(epsilon=0.5, kappa=1.5, delta=0.0)}, {(0, 0): 2.0, (0, 1): 2.5, (1, 1): 2.0}),
    "Hertz": ({(0, 0): dict(epsilon=100.0), (0, 1): dict(epsilon=50.0), (1, 1): dict(epsilon=25.0)}, {(0, 0): 1.2, (0, 1): 1.5, (1, 1): 1.0}),
}
DPD_PARAMS = ({(0, 0): dict(A=25.0, gamma=4.5, s=2.0), (0, 1): dict(A=50.0, gamma=4.5, s=1.0), (1, 1): dict(A=25.0, gamma=3.0, s=0.5)},
              {(0, 0): 1.0, (0, 1): 1.0, (1, 1): 1.0})
TPM_PARAMS = ({(0, 0): dict(M_d=1.8347, M_r=0.0302, r_eq=1.0043, omega=20.0, alpha=0.5, repulsion=True),
               (0, 1): dict(M_d=1.0, M_r=0.05, r_eq=1.1, omega=10.0, alpha=0.3, repulsion=False),
               (1, 1): dict(M_d=0.5, M_r=0.1, r_eq=0.9, omega=5.0, alpha=0.7, repulsion=True)}, {(0, 0): 1.6, (0, 1): 1.8, (1, 1): 1.5})
BOND_PARAMS = {
    "DoubleWell": [dict(r_0=1.0, r_1=2.0, U_1=1.0, U_tilt=0.5), dict(r_0=0.8, r_1=1.5, U_1=2.0, U_tilt=0.0)],
    "Quartic": [dict(k=1434.3, r_0=1.5, b_1=-0.7589, b_2=0.0, U_0=67.2234, sigma=1.0, epsilon=1.0, delta=0.0),
                dict(k=1434.3, r_0=1.5, b_1=-0.7589, b_2=0.0, U_0=67.2234, sigma=1.0, epsilon=1.0, delta=0.2)],
}


def all_cases():
    out = []
    for api in PAIR_PARAMS:
        for mode in ("none", "shift", "xplor"):
            out.append(dict(name=f"pair/{api}/{mode}", kind="pair", api=api, mode=mode))
    for timestep in (0, 2**32 + 5):
        out.append(dict(name=f"dpd/t{timestep}", kind="dpd", timestep=timestep))
    for mode in ("none", "shift"):
        out.append(dict(name=f"aniso/{mode}", kind="aniso", mode=mode))
    for api in BOND_PARAMS:
        out.append(dict(name=f"bond/{api}", kind="bond", api=api))
    for geometry, loc in ((_abi.BARRIER_PLANAR, 1.0), (_abi.BARRIER_SPHERICAL, 2.5)):
        out.append(dict(name=f"barrier/{geometry}", kind="barrier", geometry=geometry, location=loc))
    return out


def system(case):
    """Inputs of a case (deterministic in the seed)."""
    kind = case["kind"]
    rho = {"pair": 0.7, "dpd": 3.0, "aniso": 0.6, "bond": 0.5, "barrier": 0.5}[kind]
    box, pos, typeid = cases.random_fluid(N, rho, 2, seed=900 + len(case["name"]), jitter=0.5 if kind == "dpd" else 0.3)
    rng = np.random.default_rng(77)
    return dict(box=box, pos=pos, typeid=typeid, vel=rng.normal(size=pos.shape), tag=rng.permutation(len(pos)).astype(np.uint32),
                ori=cases.random_unit_quaternions(len(pos), 78))


def run_oracle(case, which, bits=BITS):
    """-> dict(force[N,4], virial[6,N], torque[N,4] | None)."""
    s = system(case)
    kind = case["kind"]
    if kind == "barrier":
        par = [(30.0, 0.25), (120.0, -0.5)]
        rc, force, virial = O.barrier(case["geometry"], s["box"], s["pos"], s["typeid"], par, case["location"], bits, which=which)
        assert rc == 0
        return dict(force=force, virial=virial, torque=None)
    if kind == "bond":
        spec = cases.SPEC_BY_API[case["api"]]
        n = len(s["pos"])
        bonds = np.stack([np.arange(n - 1), np.arange(1, n)], axis=1)
        btype = np.arange(len(bonds)) % 2
        # chain neighbours of a random fluid can be far apart: keep bonds shorter than the Quartic r_0
        d = s["pos"][bonds[:, 0]] - s["pos"][bonds[:, 1]]
        d -= s["box"].L * np.rint(d / s["box"].L)
        keep = np.linalg.norm(d, axis=1) < 1.2
        bonds, btype = bonds[keep], btype[keep]
        table = build_bond_table(n, n, bonds, btype)
        hs = O.HostSystem(s["box"], s["pos"], np.zeros(n, dtype=np.int64), 1, bits)
        hs.bond(spec.potential_id, np.array([spec.pack(p, bits) for p in BOND_PARAMS[case["api"]]]), *table, 2, which=which)
        assert hs.flags == 0
        return dict(force=hs.force, virial=hs.virial, torque=None, bonds=bonds, btype=btype)
    if kind == "pair":
        spec = cases.SPEC_BY_API[case["api"]]
        par_by, rc_by = PAIR_PARAMS[case["api"]]
        ron_by = {k: 0.8 * v for k, v in rc_by.items()}
        par, rcsq, ronsq, rc = cases.pair_tables(spec, par_by, rc_by, 2, bits, ron_by)
        hs = O.HostSystem(s["box"], s["pos"], s["typeid"], 2, bits).build_nlist(rc + 0.3)
        hs.pair(spec.potential_id, par, rcsq, ronsq, shift_mode={"none": 0, "shift": 1, "xplor": 2}[case["mode"]], which=which)
        return dict(force=hs.force, virial=hs.virial, torque=None, hs=hs)
    if kind == "dpd":
        spec = cases.SPEC_BY_API["DPDGeneralWeight"]
        par, rcsq, _, rc = cases.pair_tables(spec, DPD_PARAMS[0], DPD_PARAMS[1], 2, bits)
        hs = O.HostSystem(s["box"], s["pos"], s["typeid"], 2, bits, velocity=s["vel"], tag=s["tag"]).build_nlist(rc + 0.3)
        hs.dpd(par, rcsq, seed=42, timestep=case["timestep"], dt=0.01, kT=1.0, which=which)
        return dict(force=hs.force, virial=hs.virial, torque=None, hs=hs)
    if kind == "aniso":
        spec = cases.SPEC_BY_API["TwoPatchMorse"]
        par, rcsq, _, rc = cases.pair_tables(spec, TPM_PARAMS[0], TPM_PARAMS[1], 2, bits)
        hs = O.HostSystem(s["box"], s["pos"], s["typeid"], 2, bits, orientation=s["ori"]).build_nlist(rc + 0.3)
        hs.aniso(par, rcsq, shift_mode=1 if case["mode"] == "shift" else 0, which=which)
        return dict(force=hs.force, virial=hs.virial, torque=hs.torque, hs=hs)
    raise ValueError(kind)
