"""The step around the force path: NVE velocity Verlet + neighbour-list rebuild trigger on the GPU.

* C1 as BASELINE.json states it — PerturbedLennardJones fluid, NVE, dt = 0.005: the GPU trajectory (GPU integrator, GPU
  forces, GPU list rebuilds) against an oracle trajectory (numpy velocity Verlet in the same order of operations + oracle
  forces on a list rebuilt every step) in double, positions and forces compared at steps {1, 10, 100, 200} for N = 4000 and at
  {1, 10, 100, 1000} for the full 32 000 particles (an MD trajectory amplifies rounding differences exponentially, Lyapunov
  time ~ 1 tau = 200 steps: the bound widens with the step count).
* C1 at full size (N = 32 000, 1000 steps, float): size-independent invariants — total energy and linear momentum conserved,
  list rebuilt on the way.
* The reference's one statistical DPD test replayed (/root/reference/src/pytest/test_pair_dpd.py:12-49): N = 1000 lattice,
  A = 0 so only the thermostat acts, mean kinetic temperature over 100 samples within 10 % of kT = 1.5.
"""

import numpy as np
import pytest

import cases
import oracle_api as O

pytestmark = pytest.mark.gpu


def _oracle_forces(box, pos, vel_unused, params, rc, buffer, bits, shift_mode):
    spec = cases.SPEC_BY_API["PerturbedLennardJones"]
    par, rcsq, _, rcm = cases.pair_tables(spec, {(0, 0): params}, {(0, 0): rc}, 1, bits)
    hs = O.HostSystem(box, pos, np.zeros(len(pos), dtype=np.int64), 1, bits).build_nlist(rcm + buffer)
    hs.pair(spec.potential_id, par, rcsq, shift_mode=shift_mode, nthreads=8, which="ref" if O.have_ref() else "port")
    return hs.force.astype(np.float64)


def _wrap(pos, L):
    # BoxDim::wrap, orthorhombic: one image per direction, [-L/2, L/2)
    pos = np.where(pos >= L / 2, pos - L, pos)
    return np.where(pos < -L / 2, pos + L, pos)


@pytest.mark.parametrize("m,marks,slack", [(10, (1, 10, 100, 200), 1.0), (20, (1, 10, 100, 1000), 5.0)])
def test_nve_trajectory_matches_oracle(m, marks, slack):
    """m = 10: N = 4000, 200 steps. m = 20: C1 as BASELINE.json states it — 32 000 particles, 1000 NVE steps, parity at steps
    {1, 10, 100, 1000}; over five Lyapunov times rounding differences grow to ~1e-7 (the same growth the 1-GPU / 8-GPU
    trajectories show, profiles/r2_mgpu_md_check_8gpu.log), hence the wider bound."""
    import azplugins_b200 as az
    from azplugins_b200 import md, nlist, pair, workloads

    # jitter 0.05 a, not the 0.15 a of the force benchmarks: at 0.15 a some pairs start at r ~ 0.4 sigma (|F| ~ 1e5) and no
    # integrator survives that at dt = 0.005 — the oracle's trajectory blows up the same way
    snap = workloads.lj_fluid(m=m, jitter=0.05)  # N = 4000 (box 17.1) / 32 000
    snap.velocity -= snap.velocity.mean(axis=0)
    params = workloads.PLJ_PARAMS[("A", "A")]
    dt, rc, buffer, steps = 0.005, 3.0, 0.4, marks[-1]
    L = snap.box.Lx

    st = az.State(snap, device="cuda", scalar_bits=64)
    cell = nlist.Cell(buffer=buffer)
    plj = pair.PerturbedLennardJones(nlist=cell, default_r_cut=rc, mode="shift")
    plj.params[("A", "A")] = params
    ig = md.Integrator(dt=dt, methods=[md.ConstantVolume()], forces=[plj]).attach(st)

    # oracle trajectory: same order of operations as azb_nve_step_one / step_two
    x, v = snap.position.copy(), snap.velocity.copy()
    f = _oracle_forces(snap.box, x, None, params, rc, buffer, 64, 1)
    a = f[:, :3] / snap.mass[:, None]
    done = 0
    for mark in marks:
        for _ in range(mark - done):
            v = v + 0.5 * dt * a
            x = _wrap(x + dt * v, L)
            f = _oracle_forces(snap.box, x, None, params, rc, buffer, 64, 1)
            a = f[:, :3] / snap.mass[:, None]
            v = v + 0.5 * dt * a
        ig.run(mark - done)
        done = mark
        gx = st.pos[:, :3].cpu().numpy()
        gv = st.vel[:, :3].cpu().numpy()
        d = gx - x
        d -= L * np.rint(d / L)  # a particle on the boundary may sit on either side
        print(f"NVE parity N={snap.N} step {mark}: max|dx|={np.abs(d).max():.3e} max|dv|={np.abs(gv - v).max():.3e} "
              f"max|df|={np.abs(ig._net.cpu().numpy()[:, :3] - f[:, :3]).max():.3e} (max|f|={np.abs(f[:, :3]).max():.3e})")
        assert np.abs(d).max() < slack * 1e-9 * (1 + mark), (mark, np.abs(d).max())
        assert np.abs(gv - v).max() < slack * 1e-8 * (1 + mark), (mark, np.abs(gv - v).max())
        # forces on the oracle's snapshot of this step
        np.testing.assert_allclose(ig._net.cpu().numpy()[:, :3], f[:, :3], rtol=0, atol=slack * 1e-6 * (1 + mark) * np.abs(f[:, :3]).max() / 100)
        assert abs(ig.potential_energy - f[:, 3].sum()) <= 1e-8 * abs(f[:, 3].sum()) * (1 + mark)
    assert ig.timestep == steps
    assert ig.nlist_rebuilds >= (1 if steps <= 200 else 20)  # buffer / 2 = 0.2 is crossed within 200 steps at kT ~ 1
    assert np.abs(st.image.cpu().numpy()).max() <= 1


@pytest.mark.parametrize("bits", [32, 64])
def test_nve_c1_conserves_energy_and_momentum(bits):
    import azplugins_b200 as az
    from azplugins_b200 import md, nlist, pair, workloads

    snap = workloads.lj_fluid(m=20, jitter=0.05)  # C1: N = 32 000 (jitter: see above)
    snap.velocity -= snap.velocity.mean(axis=0)
    st = az.State(snap, device="cuda", scalar_bits=bits)
    cell = nlist.Cell(buffer=0.4)
    plj = pair.PerturbedLennardJones(nlist=cell, default_r_cut=3.0, mode="shift")
    plj.params[("A", "A")] = workloads.PLJ_PARAMS[("A", "A")]
    ig = md.Integrator(dt=0.005, methods=[md.ConstantVolume()], forces=[plj]).attach(st)
    ig.run(0)
    e0 = ig.kinetic_energy + ig.potential_energy
    ke0 = ig.kinetic_energy
    ig.run(1000)
    e1 = ig.kinetic_energy + ig.potential_energy
    p1 = ig.linear_momentum
    print(f"C1 NVE bits={bits}: E0={e0:.6f} E1={e1:.6f} drift/KE={abs(e1 - e0) / ke0:.3e} |P|={np.abs(p1).max():.3e} "
          f"rebuilds={ig.nlist_rebuilds} launches={ig.launches}")
    # velocity Verlet at dt = 0.005 on a jittered-lattice start (which melts and converts potential into kinetic energy)
    # holds the total to a few 1e-4 of the kinetic energy; a wrong half step, a stale list or a lost pair shows up as percents
    assert abs(e1 - e0) < 2e-3 * ke0
    assert np.abs(p1).max() < (1e-2 if bits == 32 else 1e-8) * np.sqrt(snap.N)
    assert ig.nlist_rebuilds >= 5
    assert ig.timestep == 1000


def test_dpd_temperature():
    """test_pair_dpd.py:12-49."""
    import azplugins_b200 as az
    from azplugins_b200 import md, nlist, pair

    n, a, kT = 10, 0.6, 1.5
    g = np.stack(np.meshgrid(*(np.arange(n),) * 3, indexing="ij"), axis=-1).reshape(-1, 3)
    L = n * a
    pos = (g + 0.5) * a - L / 2
    rng = np.random.default_rng(7)
    vel = rng.normal(scale=np.sqrt(kT), size=pos.shape)  # thermalize_particle_momenta(kT = 1.5), unit masses
    vel -= vel.mean(axis=0)
    snap = az.Snapshot(az.Box.cube(L), pos, types=["A"], velocity=vel)
    st = az.State(snap, device="cuda", scalar_bits=64, seed=11)
    cell = nlist.Cell(buffer=0.4)
    dpd = pair.DPDGeneralWeight(nlist=cell, kT=kT, default_r_cut=1.0)
    dpd.params[("A", "A")] = dict(A=0.0, gamma=4.5, s=0.5)
    ig = md.Integrator(dt=0.01, methods=[md.ConstantVolume()], forces=[dpd]).attach(st)
    ig.run(10)
    samples = np.zeros(100)
    dof = 3 * snap.N - 3
    for k in range(100):
        samples[k] = 2.0 * ig.kinetic_energy / dof
        ig.run(1)
    assert samples.mean() == pytest.approx(1.5, 0.1)
    assert dpd.dt == 0.01
