"""Synthetic configurations of BASELINE.json (SURVEY.md §8 d): seeded numpy generators returning Snapshots.

Particles are stored in cell order (coarse spatial sort), which is what HOOMD's SFC particle sorter maintains during
a run — neighbour gathers then hit nearby addresses.
"""

import numpy as np

from .state import Box, Snapshot


def spatial_sort(pos, box, cell=1.5):
    """Permutation that orders particles by a z-y-x raster of cubic cells of edge ~cell, serpentine-free."""
    L = box.L
    nc = np.maximum(1, np.floor(L / cell)).astype(np.int64)
    c = np.floor((pos + L / 2) / L * nc).astype(np.int64)
    c = np.clip(c, 0, nc - 1)
    # Morton-like blocking: sort by 4x4x4 super-cells first so that neighbouring cells in y/z stay close in memory
    sc = c // 4
    nsc = (nc + 3) // 4
    key = ((sc[:, 2] * nsc[1] + sc[:, 1]) * nsc[0] + sc[:, 0]) * 64 + ((c[:, 2] % 4) * 4 + (c[:, 1] % 4)) * 4 + (c[:, 0] % 4)
    return np.argsort(key, kind="stable")


def fcc_lattice(m, rho, jitter, rng):
    """FCC lattice of m^3 unit cells at number density rho with uniform jitter (fraction of the lattice constant)."""
    a = (4.0 / rho) ** (1.0 / 3.0)
    base = np.array([[0, 0, 0], [0.5, 0.5, 0], [0.5, 0, 0.5], [0, 0.5, 0.5]])
    g = np.stack(np.meshgrid(*(np.arange(m),) * 3, indexing="ij"), axis=-1).reshape(-1, 1, 3)
    pos = ((g + base[None]).reshape(-1, 3) + 0.25) * a
    L = m * a
    pos += rng.uniform(-jitter, jitter, size=pos.shape) * a
    pos = pos % L - L / 2
    return Box.cube(L), pos


def dpd_fluid(n=1_000_188, rho=3.0, seed=2001, kT=1.0):
    """C2: DPDGeneralWeight fluid, rho = 3, two types 50/50, uniform-random positions (DPD is soft), random tags."""
    rng = np.random.default_rng(seed)
    L = (n / rho) ** (1.0 / 3.0)
    box = Box.cube(L)
    pos = rng.uniform(-L / 2, L / 2, size=(n, 3))
    order = spatial_sort(pos, box, cell=1.4)
    pos = pos[order]
    typeid = rng.integers(0, 2, size=n)
    tag = np.random.default_rng(seed + 1).permutation(n).astype(np.uint32)
    vel = np.random.default_rng(seed + 2).normal(scale=np.sqrt(kT), size=(n, 3))
    return Snapshot(box, pos, typeid=typeid, types=["A", "B"], velocity=vel, tag=tag)


DPD_PARAMS = {("A", "A"): dict(A=25.0, gamma=4.5, s=2.0), ("A", "B"): dict(A=50.0, gamma=4.5, s=2.0), ("B", "B"): dict(A=25.0, gamma=4.5, s=2.0)}


def lj_fluid(m=20, rho=0.8, seed=1001, jitter=0.15):
    """C1: PerturbedLennardJones fluid on a jittered FCC lattice: N = 4 m^3 (m = 20 -> 32 000; m = 100 -> 4 000 000)."""
    rng = np.random.default_rng(seed)
    box, pos = fcc_lattice(m, rho, jitter, rng)
    order = spatial_sort(pos, box, cell=1.7)
    pos = pos[order]
    n = len(pos)
    tag = np.random.default_rng(seed + 1).permutation(n).astype(np.uint32)
    vel = np.random.default_rng(seed + 2).normal(size=(n, 3))
    return Snapshot(box, pos, types=["A"], velocity=vel, tag=tag)


PLJ_PARAMS = {("A", "A"): dict(epsilon=1.0, sigma=1.0, attraction_scale_factor=0.5)}


def yukawa_dpd_fluid(n=8_000_000, rho=1.0, seed=5001, jitter=0.12, kT=1.0):
    """C5: one system, two force objects on one neighbour list (ExpandedYukawa r_cut = 2.0 + DPDGeneralWeight r_cut = 1.0);
    jittered FCC lattice at rho = 1.0, N = 4 m^3 nearest to n. The jitter is 0.12 lattice constants (0.15 elsewhere): two nearest
    neighbours (1.122 apart) jittered towards each other then stay >= 0.58 apart (measured minimum of the 8 M system: 0.594), clear of
    the ExpandedYukawa pole at r = delta = 0.5; with 0.15, 14 212 particles of the 8 M system have a neighbour closer than 0.55 and
    some pairs sit beyond the pole (minimum 0.462). 93 % of the particles still have a neighbour inside the DPD cutoff."""
    rng = np.random.default_rng(seed)
    m = max(2, round((n / 4) ** (1.0 / 3.0)))
    box, pos = fcc_lattice(m, rho, jitter, rng)
    order = spatial_sort(pos, box, cell=1.2)
    pos = pos[order]
    n = len(pos)
    tag = np.random.default_rng(seed + 1).permutation(n).astype(np.uint32)
    vel = np.random.default_rng(seed + 2).normal(scale=np.sqrt(kT), size=(n, 3))
    return Snapshot(box, pos, types=["A"], velocity=vel, tag=tag)


EY_PARAMS = {("A", "A"): dict(epsilon=1.0, kappa=1.0, delta=0.5)}
C5_DPD_PARAMS = {("A", "A"): dict(A=25.0, gamma=4.5, s=2.0)}


def patchy_fluid(m=80, rho=0.6, seed=4001):
    """C4a: TwoPatchMorse colloids, N = 4 m^3 (m = 80 -> 2 048 000), uniform random unit quaternions."""
    rng = np.random.default_rng(seed)
    box, pos = fcc_lattice(m, rho, 0.15, rng)
    order = spatial_sort(pos, box, cell=2.0)
    pos = pos[order]
    q = rng.normal(size=(len(pos), 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return Snapshot(box, pos, types=["A"], orientation=q)


TPM_PARAMS = {("A", "A"): dict(M_d=1.8347, M_r=0.0302, r_eq=1.0043, omega=20.0, alpha=0.5, repulsion=True)}


def hertz_mixture(m=80, rho=1.2, seed=4101):
    """C4b: polydisperse soft spheres (pair.Hertz), N = 4 m^3 (m = 80 -> 2 048 000), four species of diameter
    0.85 / 0.95 / 1.05 / 1.15 in equal numbers; the cutoff of a type pair is the mean of the two diameters."""
    rng = np.random.default_rng(seed)
    box, pos = fcc_lattice(m, rho, 0.15, rng)
    order = spatial_sort(pos, box, cell=1.6)
    pos = pos[order]
    typeid = rng.integers(0, 4, size=len(pos))
    return Snapshot(box, pos, typeid=typeid, types=["A", "B", "C", "D"])


HERTZ_DIAMETERS = {"A": 0.85, "B": 0.95, "C": 1.05, "D": 1.15}
HERTZ_EPSILON = 100.0


def hertz_tables():
    """(params, r_cut) per type pair of the polydisperse Hertz mixture."""
    names = sorted(HERTZ_DIAMETERS)
    par = {(a, b): dict(epsilon=HERTZ_EPSILON) for i, a in enumerate(names) for b in names[i:]}
    rc = {(a, b): 0.5 * (HERTZ_DIAMETERS[a] + HERTZ_DIAMETERS[b]) for (a, b) in par}
    return par, rc


def colloid_mixture(m=75, rho=0.6, seed=4201, radius=1.5, spacing=5.0):
    """C4b: colloids (type C, radius 1.5) on a simple cubic lattice of the given spacing in a solvent (type S, radius 0) on a
    jittered FCC lattice at density rho; solvent particles closer than radius + 0.9 to a colloid centre are removed (the
    colloid-solvent potential diverges at r = radius). m = 75 -> ~1.0 M particles (1.69 M lattice sites, 41 % of them inside
    an exclusion zone)."""
    rng = np.random.default_rng(seed)
    box, pos = fcc_lattice(m, rho, 0.1, rng)
    L = box.L
    nc = np.maximum(1, np.floor(L / spacing)).astype(np.int64)
    g = np.stack(np.meshgrid(*[np.arange(k) for k in nc], indexing="ij"), axis=-1).reshape(-1, 3)
    cpos = (g + 0.5) * (L / nc) - L / 2
    # distance of every solvent particle to its nearest lattice site (the sites form a regular grid: round to it)
    rel = (pos + L / 2) / (L / nc) - 0.5
    d = (rel - np.rint(rel)) * (L / nc)
    keep = np.sum(d * d, axis=1) > (radius + 0.9) ** 2
    pos = np.concatenate([pos[keep], cpos])
    typeid = np.concatenate([np.zeros(int(keep.sum()), dtype=np.int64), np.ones(len(cpos), dtype=np.int64)])
    order = spatial_sort(pos, box, cell=1.7)
    return Snapshot(box, pos[order], typeid=typeid[order], types=["S", "C"])


COLLOID_RADIUS = 1.5
COLLOID_PARAMS = {("S", "S"): dict(A=144.0, a_1=0.0, a_2=0.0, sigma=1.0), ("C", "S"): dict(A=100.0, a_1=COLLOID_RADIUS, a_2=0.0, sigma=1.0),
                  ("C", "C"): dict(A=40.0, a_1=COLLOID_RADIUS, a_2=COLLOID_RADIUS, sigma=1.0)}
COLLOID_RCUT = {("S", "S"): 3.0, ("C", "S"): COLLOID_RADIUS + 3.0, ("C", "C"): 2 * COLLOID_RADIUS + 3.0}


def polymer_melt(n_chains=40_000, chain_len=100, rho=0.85, seed=3001, bond_len=0.97, double_well_fraction=0.1):
    """C3: bead-spring melt. Chains are laid along a boustrophedon (snake) path through a simple-cubic lattice of spacing
    ~rho^(-1/3) compressed to the bond length along the path direction, so consecutive beads sit one bond apart and no
    two beads overlap; a small jitter breaks the symmetry."""
    rng = np.random.default_rng(seed)
    n = n_chains * chain_len
    # lattice: spacing bond_len along x, and a_yz in y, z so that density is rho
    a_yz = np.sqrt(1.0 / (rho * bond_len))
    nx = int(np.ceil((n * bond_len * a_yz * a_yz) ** (1.0 / 3.0) / bond_len))
    nyz = int(np.ceil(np.sqrt(n / nx)))
    ny, nz = nyz, int(np.ceil(n / (nx * nyz)))
    idx = np.arange(n)
    iz = idx // (nx * ny)
    rem = idx % (nx * ny)
    iy = rem // nx
    ix = rem % nx
    iy = np.where(iz % 2 == 1, ny - 1 - iy, iy)  # snake in y
    ix = np.where((iz * ny + (idx % (nx * ny)) // nx) % 2 == 1, nx - 1 - ix, ix)  # snake in x
    Lx, Ly, Lz = nx * bond_len, ny * a_yz, nz * a_yz
    pos = np.stack([(ix + 0.5) * bond_len - Lx / 2, (iy + 0.5) * a_yz - Ly / 2, (iz + 0.5) * a_yz - Lz / 2], axis=1)
    pos += rng.uniform(-0.03, 0.03, size=pos.shape)
    box = Box(Lx, Ly, Lz)
    pos = (pos + box.L / 2) % box.L - box.L / 2
    first = np.arange(n).reshape(n_chains, chain_len)[:, :-1].reshape(-1)
    bonds = np.stack([first, first + 1], axis=1)
    btype = (rng.random(len(bonds)) < double_well_fraction).astype(np.int64)
    # the path order is already spatially coherent: keep it (bond partners are adjacent in memory)
    return Snapshot(box, pos, types=["A"], bond_group=bonds, bond_typeid=btype, bond_types=["quartic", "doublewell"])


QUARTIC_PARAMS = dict(k=1434.3, r_0=1.5, b_1=-0.7589, b_2=0.0, U_0=67.2234, sigma=1.0, epsilon=1.0, delta=0.0)
DOUBLE_WELL_PARAMS = dict(r_0=1.0, r_1=2.0, U_1=1.0, U_tilt=0.5)
# HOOMD evaluates every bond object on every bond: an object that does not own a bond type is given parameters for which the
# evaluator returns zero force and zero energy without tripping its validity check (r_0 != 0; r_0 != r_1)
QUARTIC_NEUTRAL = dict(k=0.0, r_0=1.5, b_1=0.0, b_2=0.0, U_0=0.0, sigma=0.0, epsilon=0.0, delta=0.0)
DOUBLE_WELL_NEUTRAL = dict(r_0=1.0, r_1=2.0, U_1=0.0, U_tilt=0.0)
