"""Multi-GPU parity: run under torchrun with N = 2, 4 or 8 ranks (one GPU each). Every rank decomposes the same seeded global
DPD system, exchanges ghosts over NCCL, builds its neighbour list on the GPU and evaluates DPDGeneralWeight through the
C-ABI; forces are compared per tag with the CPU oracle evaluated on the undecomposed periodic system.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tests/mgpu_force_check.py
"""

import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import azplugins_b200 as az
    import cases
    import domain_worker
    import oracle_api as O
    from azplugins_b200 import nlist, pair
    from azplugins_b200.domain import Decomposition, HaloExchange

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    box, pos, vel, tag, typeid = domain_worker.global_system(n=20000)
    params = {(0, 0): dict(A=25.0, gamma=4.5, s=2.0), (0, 1): dict(A=50.0, gamma=4.5, s=2.0), (1, 1): dict(A=25.0, gamma=4.5, s=0.5)}
    worst = 0.0
    transports = set()
    # float: a ghost that crossed the global boundary is stored pre-shifted (x_j +/- L rounded to float), the undecomposed
    # system wraps dx instead; the two roundings differ by ~ulp(L) on dx, i.e. ~1e-5 on a force — inherent to float
    # coordinates under any decomposition (HOOMD's MPI ghosts included), so float is held to 1e-4 here
    for bits, tol in ((64, 1e-9), (32, 1e-4)):
        # oracle on the whole periodic system, same Scalar
        spec = cases.SPEC_BY_API["DPDGeneralWeight"]
        par, rcsq, _, rc = cases.pair_tables(spec, params, {k: 1.0 for k in params}, 2, bits)
        hs = O.HostSystem(box, pos, typeid, 2, bits, velocity=vel, tag=tag).build_nlist(rc + 0.4)
        hs.dpd(par, rcsq, seed=42, timestep=7, dt=0.01, kT=1.0, nthreads=2, which="ref" if O.have_ref() else "port")
        f_by_tag = np.zeros((len(pos), 4))
        f_by_tag[tag] = hs.force
        amb = np.zeros(len(pos), dtype=bool)
        amb[tag] = cases_amb(hs, rcsq, bits)

        dec = Decomposition(world, rank, box)
        own = np.nonzero(dec.owns(pos))[0]
        if bits == 32:
            # spatially sorted locals (what HOOMD's SFC sorter maintains): most tiles then read no ghost
            from azplugins_b200.workloads import spatial_sort

            own = own[spatial_sort(pos[own], box, cell=1.4)]
        snap = az.Snapshot(box, pos[own], typeid=typeid[own], types=["A", "B"], velocity=vel[own], tag=tag[own])
        st = az.State(snap, device=dev, scalar_bits=bits, seed=42)
        cell = nlist.Cell(buffer=0.4)
        f = pair.DPDGeneralWeight(nlist=cell, kT=1.0, default_r_cut=1.0)
        f.dt = 0.01
        names = ["A", "B"]
        for (i, j), p in params.items():
            f.params[(names[i], names[j])] = p
        halo = HaloExchange(dec, st, 1.4, fields=("pos", "vel")) if world > 1 else None
        f._attach(st)
        f.compute(7)
        got = np.concatenate([f.forces, f.energies[:, None]], axis=1).astype(np.float64)
        if halo is not None:
            # the overlapped step (ghost-free tiles while the halo travels, boundary tiles after) must give the same forces
            plain = f._force.clone()
            f._force.zero_()
            st.pos[st.N:].zero_()  # the step has to bring the ghosts itself
            halo.step(f, 7)
            torch.cuda.synchronize()
            assert torch.equal(f._force, plain), (rank, bits, "overlapped step differs from exchange-then-evaluate")
            order, n_interior = halo.tile_order(f._plan)
            assert sorted(order.tolist()) == list(range(f._plan.n_tiles))
            assert 0 <= n_interior <= f._plan.n_tiles  # (in a 10 000-particle domain every tile touches a face)
            # several steps with moving particles: the peer-memory push (flags, epochs, acknowledgements) must deliver exactly
            # the rows the grouped NCCL send/recv delivers
            transports.add(halo.transport)
            if halo.transport == "peer":
                gen = torch.Generator(device="cpu").manual_seed(100 + rank)
                for it in range(4):
                    st.pos[: st.N, :3] += (torch.rand((st.N, 3), generator=gen, dtype=torch.float64) * 2e-3 - 1e-3).to(st.pos)
                    st.vel[: st.N, :3] += (torch.rand((st.N, 3), generator=gen, dtype=torch.float64) * 2e-3 - 1e-3).to(st.vel)
                    halo.step(f, 8 + it)
                    torch.cuda.synchronize()
                    by_peer = (f._force.clone(), st.pos[st.N:].clone(), st.vel[st.N:].clone())
                    halo.transport = "nccl"
                    st.pos[st.N:].zero_()
                    st.vel[st.N:].zero_()
                    halo.step(f, 8 + it)
                    torch.cuda.synchronize()
                    halo.transport = "peer"
                    assert torch.equal(st.pos[st.N:], by_peer[1]) and torch.equal(st.vel[st.N:], by_peer[2]), (rank, bits, it, "ghost rows differ")
                    assert torch.equal(f._force, by_peer[0]), (rank, bits, it, "forces differ between transports")
                halo.update()  # the non-overlapped refresh uses the same flags
                torch.cuda.synchronize()
                assert torch.equal(st.pos[st.N:], by_peer[1])
                halo.check()
        want = f_by_tag[tag[own]]
        keep = ~amb[tag[own]]
        scale = np.sqrt(np.mean(np.sum(want[:, :3] ** 2, axis=1)))
        err = np.max(np.linalg.norm(got[keep, :3] - want[keep, :3], axis=1) / np.maximum(np.linalg.norm(want[keep, :3], axis=1), scale))
        worst = max(worst, err / tol)
        assert err <= tol, (rank, bits, err)
        assert st.n_ghost > 0 or world == 1
    t = torch.tensor([worst], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"mgpu force check ok: world={world} worst err/tol={t.item():.3g} halo transports={sorted(transports)}")
    dist.destroy_process_group()


def cases_amb(hs, rcsq, bits):
    import gpu_cases as GC

    if bits == 64:
        return np.zeros(hs.N, dtype=bool)
    return GC.cutoff_ambiguous(hs, rcsq, rel=2e-4)  # s = 0.5 on the BB pair: see test_gpu_parity.test_dpd_random_system


if __name__ == "__main__":
    main()
