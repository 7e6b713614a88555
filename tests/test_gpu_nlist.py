"""GPU cell-list builder against the oracle's builder: bit-exact neighbour indexing as per-particle SET equality
(row order differs between builders by design), head_list consistency, bond exclusions,
per-type-pair r_cut, non-periodic directions with ghosts outside the box."""

import numpy as np
import pytest
import torch

import azplugins_b200 as az
import cases
import oracle_api as O
from azplugins_b200 import bond, nlist, pair

pytestmark = pytest.mark.gpu


def _rows(n_neigh, head, nl, N):
    return [np.sort(nl[int(head[i]): int(head[i]) + int(n_neigh[i])]) for i in range(N)]


def _ordered_rows(cell, N):
    nn = cell.n_neigh.cpu().numpy().view(np.uint32)
    head = cell.head_list.cpu().numpy().view(np.uint64)
    nl = cell.nlist.cpu().numpy().view(np.uint32)
    return [nl[int(head[i]): int(head[i]) + int(nn[i])].copy() for i in range(N)]


def _gpu_rows(cell, N):
    nn = cell.n_neigh.cpu().numpy().view(np.uint32)
    head = cell.head_list.cpu().numpy().view(np.uint64)
    nl = cell.nlist.cpu().numpy().view(np.uint32)
    assert np.all(np.diff(head[:N].astype(np.int64)) >= nn[: N - 1].astype(np.int64))
    assert int(head[N - 1]) + int(nn[N - 1]) <= cell.size_nlist
    return _rows(nn, head, nl, N), nn


@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("n,rho,rcut", [(5000, 0.8, 3.0), (20000, 3.0, 1.0), (300, 0.8, 3.0)])
def test_cell_nlist_matches_oracle(n, rho, rcut, bits):
    box, pos, typeid = cases.random_fluid(n, rho, 2, seed=7, jitter=0.5)
    st = az.State(az.Snapshot(box, pos, typeid=typeid, types=["A", "B"]), device="cuda", scalar_bits=bits)
    cell = nlist.Cell(buffer=0.4)
    f = pair.PerturbedLennardJones(nlist=cell, default_r_cut=rcut)
    f.r_cut[("A", "B")] = 0.8 * rcut
    f.r_cut[("B", "B")] = 0.0  # no interaction: pair never listed
    cell.compute(st)
    r_list = cell.r_list(st.types)
    assert r_list[1, 1] == 0 and r_list[0, 1] == pytest.approx(0.8 * rcut + 0.4)
    hs = O.HostSystem(box, pos, typeid, 2, bits).build_nlist(r_list)
    want = _rows(hs.n_neigh, hs.head_list, hs.nlist, n)
    got, nn = _gpu_rows(cell, n)
    assert np.array_equal(nn[:n], hs.n_neigh[:n])
    for i in range(n):
        assert np.array_equal(got[i], want[i]), i
    # rebuilding is deterministic (same row order, not just the same sets). The first build walked twice (count, fill); the
    # second knows how long rows get and walks ONCE, writing the rows in place at a fixed stride
    # (azb_nlist_args.row_stride_hint + rows_in_place): head_list[i] = i * stride, same rows in the same order.
    first = (cell.nlist.clone(), cell.n_neigh.clone(), cell.head_list.clone())
    first_rows = _ordered_rows(cell, n)
    assert cell._row_max == int(hs.n_neigh[:n].max()) and cell.n_entries == int(hs.n_neigh[:n].sum()) == cell.size_nlist
    stride = int(cell._row_max * 1.2) + 8
    cell.invalidate()
    cell.compute(st)
    assert cell.size_nlist == n * stride and cell.n_entries == int(hs.n_neigh[:n].sum())
    assert bool((cell.head_list.cpu() == torch.arange(n, dtype=torch.int64) * stride).all()) and bool((first[1] == cell.n_neigh).all())
    rows = _ordered_rows(cell, n)
    assert all(np.array_equal(a, b) for a, b in zip(first_rows, rows))
    assert cell.mean_neighbors(n) == pytest.approx(hs.n_neigh[:n].mean())

    def same_arrays():
        return bool((first[0] == cell.nlist[: first[0].numel()]).all()) and bool((first[1] == cell.n_neigh).all()) and bool((first[2] == cell.head_list).all())

    # a bound that is too small (rows outgrew it): the packed head_list comes back and the fill pass walks again, same arrays
    cell._row_max = 3
    cell.invalidate()
    cell.compute(st)
    assert same_arrays() and cell.size_nlist == first[0].numel()
    # one walk into temporary rows + compaction (tightly packed list) on request
    cell.rows_in_place = False
    cell.invalidate()
    cell.compute(st)
    assert same_arrays() and cell.size_nlist == first[0].numel()
    # and the classic two-walk build
    cell.single_walk = False
    cell.invalidate()
    cell.compute(st)
    assert same_arrays()


def test_cell_nlist_bond_exclusions_and_force_through_cell():
    n = 4000
    box, pos, typeid = cases.random_fluid(n, 0.85, 1, seed=9)
    bonds = np.stack([np.arange(n - 1), np.arange(1, n)], axis=1)
    snap = az.Snapshot(box, pos, typeid=typeid, bond_group=bonds, bond_typeid=np.zeros(len(bonds), dtype=int), bond_types=["A-A"])
    st = az.State(snap, device="cuda", scalar_bits=32)
    cell = nlist.Cell(buffer=0.4)  # exclusions=('bond',) by default, like hoomd.md.nlist.Cell
    f = pair.PerturbedLennardJones(nlist=cell, default_r_cut=2.5)
    f.params[("A", "A")] = dict(epsilon=1.0, sigma=1.0, attraction_scale_factor=0.5)
    f._attach(st)
    f.compute(0)
    head, lst = st.exclusions_csr()
    hs = O.HostSystem(box, pos, typeid, 1, 32).build_nlist(2.9, excl_head=head, excl_list=lst)
    got, nn = _gpu_rows(cell, n)
    want = _rows(hs.n_neigh, hs.head_list, hs.nlist, n)
    for i in range(n):
        assert np.array_equal(got[i], want[i]), i
        assert (i + 1) not in got[i] and (i - 1) not in got[i]
    spec = cases.SPEC_BY_API["PerturbedLennardJones"]
    par, rcsq, ronsq, _ = cases.pair_tables(spec, {(0, 0): dict(epsilon=1.0, sigma=1.0, attraction_scale_factor=0.5)}, {(0, 0): 2.5}, 1, 32)
    hs.pair(spec.potential_id, par, rcsq, ronsq, which="ref" if O.have_ref() else "port")
    scale = np.sqrt(np.mean(np.sum(hs.force[:, :3].astype(np.float64) ** 2, axis=1)))
    assert np.max(np.abs(f.forces - hs.force[:, :3])) <= 2e-5 * scale


def test_cell_nlist_nonperiodic_with_ghosts():
    """Domain-decomposition shape: non-periodic x, ghosts (indices >= N) outside the local box are candidates only."""
    rng = np.random.default_rng(3)
    n_local, n_ghost = 3000, 900
    L = 12.0
    local = rng.uniform(-L / 2, L / 2, size=(n_local, 3))
    ghost = rng.uniform(-L / 2, L / 2, size=(n_ghost, 3))
    ghost[:, 0] = np.where(rng.random(n_ghost) < 0.5, rng.uniform(-L / 2 - 1.4, -L / 2, n_ghost), rng.uniform(L / 2, L / 2 + 1.4, n_ghost))
    pos = np.concatenate([local, ghost])
    box = az.Box(L, L, L, periodic=(False, True, True))
    typeid = np.zeros(len(pos), dtype=np.int64)
    st = az.State(az.Snapshot(box, pos, typeid=typeid), device="cuda", scalar_bits=32)
    st.N, st.n_ghost = n_local, n_ghost
    cell = nlist.Cell(buffer=0.4)
    pair.DPDConservativeGeneralWeight(nlist=cell, default_r_cut=1.0)
    cell.compute(st)
    hs = O.HostSystem(box, pos, typeid, 1, 32, n_local=n_local).build_nlist(1.4)
    got, nn = _gpu_rows(cell, n_local)
    want = _rows(hs.n_neigh, hs.head_list, hs.nlist, n_local)
    assert sum(int((r >= n_local).sum()) for r in want) > 0
    for i in range(n_local):
        assert np.array_equal(got[i], want[i]), i
