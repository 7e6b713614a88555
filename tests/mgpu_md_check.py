"""Multi-GPU molecular dynamics with migration: run under torchrun with N = 2, 4 or 8 ranks (one GPU each).

(1) C1 — 32 000-particle PerturbedLennardJones fluid, Scalar = double, 1000 NVE steps through md.Integrator on the decomposed
    system: every neighbour-list rebuild migrates particles between the ranks, re-selects the ghost layer and rewrites the
    peer-memory routes. Per-tag unwrapped positions must equal the 1-GPU trajectory of the same system (which every rank
    integrates for itself on its own GPU) to 1e-9 x steps; summation order is the only difference between the two.
(2) C3 at test size — bead-spring melt (PerturbedLJ with bonded pairs excluded + bond.Quartic + bond.DoubleWell), bonds whose
    partner is a ghost, beads that migrate with their bonds: 300 steps, same comparison.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 tests/mgpu_md_check.py
"""

import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def make_forces(kind, nl):
    from azplugins_b200 import bond, pair, workloads as W

    f = pair.PerturbedLennardJones(nlist=nl, default_r_cut=3.0, mode="shift")
    for k, v in W.PLJ_PARAMS.items():
        f.params[k] = v
    if kind == "lj":
        return [f]
    q = bond.Quartic()
    q.params["quartic"], q.params["doublewell"] = W.QUARTIC_PARAMS, W.QUARTIC_NEUTRAL
    dw = bond.DoubleWell()
    dw.params["quartic"], dw.params["doublewell"] = W.DOUBLE_WELL_NEUTRAL, W.DOUBLE_WELL_PARAMS
    return [f, q, dw]


def unwrapped_by_tag(st, n_global, L):
    """(n_global, 3) float64: pos + image * L of the particles this rank owns, zero elsewhere."""
    out = torch.zeros((n_global, 3), dtype=torch.float64, device=st.device)
    tags = st.tag[: st.N].long() & 0xFFFFFFFF
    out[tags] = st.pos[: st.N, :3].double() + st.image[: st.N].double() * torch.tensor(L, dtype=torch.float64, device=st.device)
    return out


def run_case(kind, steps, dt, dev, rank, world):
    import azplugins_b200 as az
    from azplugins_b200 import md, nlist, workloads as W
    from azplugins_b200.domain import Decomposition, HaloExchange

    snap = W.lj_fluid(m=20, seed=1001, jitter=0.05) if kind == "lj" else W.polymer_melt(n_chains=320, seed=3001)
    if kind == "melt":
        snap.velocity = np.random.default_rng(7).normal(scale=0.5, size=(snap.N, 3))
        snap.tag = np.random.default_rng(8).permutation(snap.N).astype(np.uint32)
    n, L = snap.N, snap.box.L
    # ---- the undecomposed system on this rank's GPU ---------------------------------------------------------------------------
    st1 = az.State(snap, device=dev, scalar_bits=64, seed=42)
    ig1 = md.Integrator(dt=dt, methods=[md.ConstantVolume()], forces=make_forces(kind, nlist.Cell(buffer=0.4))).attach(st1)
    ig1.run(steps)
    want = unwrapped_by_tag(st1, n, L)
    e1 = ig1.kinetic_energy + ig1.potential_energy
    # ---- the same system decomposed over the ranks -----------------------------------------------------------------------------
    dec = Decomposition(world, rank, snap.box)
    own = np.nonzero(dec.owns(snap.position))[0]
    sub = az.Snapshot(snap.box, snap.position[own], typeid=snap.typeid[own], types=snap.types, velocity=snap.velocity[own], tag=snap.tag[own])
    st = az.State(sub, device=dev, scalar_bits=64, seed=42)
    st.bond_types = list(snap.bond_types)
    if kind == "melt" and rank == 0:
        st.bond_tags, st.bond_typeid = snap.tag[snap.bond_group].astype(np.uint32), snap.bond_typeid.copy()
    halo = HaloExchange(dec, st, 3.4, fields=("pos",))
    ig = md.Integrator(dt=dt, methods=[md.ConstantVolume()], forces=make_forces(kind, nlist.Cell(buffer=0.4))).attach(st, domain=halo)
    ig.run(steps)
    halo.check()
    got = unwrapped_by_tag(st, n, L)
    dist.all_reduce(got)
    cnt = torch.tensor([st.N], dtype=torch.int64, device=dev)
    dist.all_reduce(cnt)
    assert int(cnt.item()) == n, (kind, int(cnt.item()), n)
    err = float((got - want).abs().max().item())
    en = torch.tensor([ig.kinetic_energy + ig.potential_energy], dtype=torch.float64, device=dev)
    dist.all_reduce(en)
    mig = torch.tensor([halo.migrated], dtype=torch.int64, device=dev)
    dist.all_reduce(mig)
    tol = 1e-9 * steps
    assert err <= tol, (kind, rank, err, tol)
    assert abs(float(en.item()) - e1) <= 1e-9 * abs(e1) + 1e-6, (kind, float(en.item()), e1)
    assert ig.nlist_rebuilds == ig1.nlist_rebuilds or abs(ig.nlist_rebuilds - ig1.nlist_rebuilds) <= 2
    assert int(mig.item()) > 0, (kind, "no particle migrated: the run did not exercise the path")
    return err, int(mig.item()), ig.nlist_rebuilds, halo.transport


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    out = {}
    out["lj"] = run_case("lj", 1000, 0.005, dev, rank, world)
    out["melt"] = run_case("melt", 300, 0.002, dev, rank, world)
    if rank == 0:
        for k, (err, mig, reb, transport) in out.items():
            print(f"mgpu md check ok: {k} world={world} max |dx| vs 1 GPU = {err:.3g}, {mig} migrations over {reb} rebuilds, halo={transport}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
