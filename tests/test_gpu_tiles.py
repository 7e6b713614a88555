"""Tile plan (azb_tiles_build): the derived view of a neighbour list the staged kernels consume, against numpy."""

import numpy as np
import pytest

import azplugins_b200 as az
import cases
import oracle_api as O
from azplugins_b200 import nlist

pytestmark = pytest.mark.gpu


def _list_on_device(n, rho, r_list, seed, shuffle=False):
    box, pos, typeid = cases.random_fluid(n, rho, 1, seed, jitter=0.4)
    if shuffle:
        pos = pos[np.random.default_rng(seed).permutation(len(pos))]
    hs = O.HostSystem(box, pos, typeid, 1, 32).build_nlist(np.array([[r_list]]))
    nl = nlist.NeighborList().set_arrays(hs.n_neigh[:n], hs.head_list[:n], hs.nlist, "cuda")
    return hs, nl


@pytest.mark.parametrize("rotate", [False, True, "sorted"])
@pytest.mark.parametrize("lane_width", [1, 2, 4, 8, 16])
@pytest.mark.parametrize("tile_size", [32, 64, 128, 256])
def test_tile_plan_matches_numpy(tile_size, lane_width, rotate):
    n = 5000
    hs, nl = _list_on_device(n, 3.0, 1.4, 11)
    sort_rows = rotate == "sorted"  # slot rows in ascending slot order (azb_tile_build_args.sort_rows)
    rotate = rotate is True
    plan = nl.tile_plan(n, tile_size, 2048, lane_width, group_width=lane_width if rotate else 0, sort_rows=sort_rows)
    assert plan.sorted_rows == sort_rows
    W, M = lane_width, 4 * lane_width
    A = max(M, 8)  # AZB_SLOT_ROW_ALIGN
    assert plan is not None and plan.tile_size == tile_size and plan.n_tiles == (n + tile_size - 1) // tile_size
    count = plan.uniq_count.cpu().numpy().view(np.uint32)
    uniq = plan.uniq.cpu().numpy().view(np.uint32).reshape(plan.n_tiles, plan.uniq_capacity)
    slots = plan.slots.cpu().numpy().view(np.uint16)
    nn, head, lst = hs.n_neigh[:n].astype(np.int64), hs.head_list[:n].astype(np.int64), hs.nlist
    worst = 0
    for t in range(plan.n_tiles):
        rows = range(t * tile_size, min((t + 1) * tile_size, n))
        entries = np.concatenate([lst[head[i] : head[i] + nn[i]] for i in rows]) if rows else np.zeros(0, np.uint32)
        want = np.unique(entries)
        assert count[t] == len(want)
        assert np.array_equal(uniq[t, : len(want)], want)  # ascending, no duplicates
        for i in rows:
            sh = (head[i] + A * i) & ~(A - 1)  # AZB_SLOT_HEAD
            e = np.arange((nn[i] + M - 1) // M * M)
            r = e % M
            stored = slots[sh + (e - r) + 4 * (r % W) + r // W]  # entry e of the row in the lane-transposed layout
            assert np.all(stored[nn[i] :] == 0)  # padding
            if sort_rows:
                assert np.all(np.diff(stored[: nn[i]].astype(np.int64)) > 0)  # ascending slots, each neighbour exactly once
                assert np.array_equal(want[stored[: nn[i]]], np.sort(lst[head[i] : head[i] + nn[i]]))
                continue
            if not rotate:
                assert np.array_equal(want[stored[: nn[i]]], lst[head[i] : head[i] + nn[i]])  # slot -> the listed neighbour, in order
                continue
            # group_width > 0: the row's neighbours, each exactly once, in bank-rotated order
            assert np.array_equal(np.sort(want[stored[: nn[i]]]), np.sort(lst[head[i] : head[i] + nn[i]]))
            # the deal: position e carries bank group (start + e) mod 8 until the first bucket runs dry
            if nn[i]:
                groups = stored[: nn[i]] & 7
                dry = np.bincount(groups, minlength=8).min() * 8
                start = (i * W) & 7
                assert np.array_equal(groups[:dry], (start + np.arange(dry)) & 7)
        worst = max(worst, len(want))
    assert plan.max_uniq == worst
    spans = []
    for t in range(plan.n_tiles):
        i0, i1 = t * tile_size, min((t + 1) * tile_size, n) - 1
        spans.append(((head[i1] + A * i1) & ~(A - 1)) + (nn[i1] + M - 1) // M * M - ((head[i0] + A * i0) & ~(A - 1)))
    assert plan.max_slot_span == max(spans)


def test_tile_plan_overflow_falls_back_to_smaller_tiles():
    n = 4000
    # shuffled particle order: a tile's neighbours are scattered over the whole system
    hs, nl = _list_on_device(n, 3.0, 1.4, 12, shuffle=True)
    from azplugins_b200 import _nlist_gpu

    assert _nlist_gpu.build_tile_plan(nl, n, 256, 1024) is None  # 256 rows x ~34 entries, nearly all distinct
    plan = nl.tile_plan(n, 256, 1024)  # halves the tile until the distinct set fits
    assert plan is not None and plan.tile_size < 256 and plan.max_uniq <= 1024


def test_tile_plan_is_rebuilt_with_the_list():
    n = 2000
    hs, nl = _list_on_device(n, 3.0, 1.4, 13)
    p0 = nl.tile_plan(n, 128)
    assert nl.tile_plan(n, 128) is p0  # cached while the arrays are unchanged
    nl.set_arrays(hs.n_neigh[:n], hs.head_list[:n], hs.nlist, "cuda")
    assert nl.tile_plan(n, 128) is not p0


@pytest.mark.parametrize("sort_rows", [False, True])
@pytest.mark.parametrize("bits", [32, 64])
def test_pruned_plan_gives_bit_identical_forces(bits, sort_rows):
    """Two forces on one list (C5): the DPD force (r_cut 1.0) walks a view of the list (built for 2.0 + 0.4) without the entries
    that were beyond r_cut + buffer when the list was built. Same pair set inside the cutoff, same order, so the forces must
    equal the unpruned plan's bit for bit — also after every particle has moved by up to buffer / 2."""
    import torch

    from azplugins_b200 import pair, workloads as W

    snap = W.yukawa_dpd_fluid(n=20000, seed=21, jitter=0.15)
    st = az.State(snap, device="cuda", scalar_bits=bits, seed=42)
    cell = nlist.Cell(buffer=0.4)
    cell.sort_rows = sort_rows  # (slot rows in ascending slot order: the pruned view keeps the relative order of its base plan)
    forces = []
    for prune in (True, False):
        d = pair.DPDGeneralWeight(nlist=cell, kT=1.0, default_r_cut=1.0)
        d.dt = 0.01
        d.params[("A", "A")] = W.C5_DPD_PARAMS[("A", "A")]
        d.prune_lists = prune
        d.threads_per_particle = 4  # same lanes per particle = same summation order (left to itself the pruned view, with its
        forces.append(d)            # shorter rows, would pick narrower groups: equal to rounding only)
    y = pair.ExpandedYukawa(nlist=cell, default_r_cut=2.0)
    y.params[("A", "A")] = W.EY_PARAMS[("A", "A")]
    for f in forces + [y]:
        f._attach(st)
        f.compute(3)
    pruned, full = forces
    assert pruned._plan.row_count is not None and full._plan.row_count is None and y._plan.row_count is None
    assert pruned._plan is not full._plan
    # the view holds exactly the entries within r_cut + buffer = 1.4 at build time
    nn = cell.n_neigh.cpu().numpy().view(np.uint32).astype(np.int64)
    head = cell.head_list.cpu().numpy().view(np.uint64).astype(np.int64)
    lst = cell.nlist.cpu().numpy().view(np.uint32).astype(np.int64)
    pos = st.pos[:, :3].double().cpu().numpy()
    L = snap.box.L
    i_idx = np.repeat(np.arange(st.N), nn)
    j_idx = lst[np.repeat(head, nn) + (np.arange(len(i_idx)) - np.repeat(np.cumsum(nn) - nn, nn))]
    d = pos[i_idx] - pos[j_idx]
    d -= L * np.rint(d / L)
    rsq = np.sum(d * d, axis=1)
    sure_in, maybe = rsq <= 1.4**2 * (1 - 1e-4), rsq <= 1.4**2 * (1 + 1e-4)
    rows = pruned._plan.row_count.cpu().numpy().astype(np.int64)
    assert np.all(rows >= np.bincount(i_idx[sure_in], minlength=st.N)) and np.all(rows <= np.bincount(i_idx[maybe], minlength=st.N))
    assert pruned._plan.entries == rows.sum() and pruned._plan.entries < 0.35 * cell.size_nlist
    assert torch.equal(pruned._force, full._force) and torch.equal(pruned._virial, full._virial)
    # move everything by up to buffer / 2 (the most a particle may move before HOOMD rebuilds the list): still identical
    gen = torch.Generator(device="cpu").manual_seed(5)
    step = torch.rand((st.N, 3), generator=gen, dtype=torch.float64) * 2 - 1
    step = 0.199 * step / step.norm(dim=1, keepdim=True).clamp(min=1.0)
    st.pos[: st.N, :3] += step.to(st.pos)
    for f in forces:
        f.compute(4)
    assert pruned._plan.row_count is not None  # same plans: the list was not rebuilt
    assert torch.equal(pruned._force, full._force) and torch.equal(pruned._virial, full._virial)
    assert float(pruned._force[:, :3].abs().sum()) > 0


@pytest.mark.parametrize("bits", [32, 64])
@pytest.mark.parametrize("periodic", [(True, True, True), (False, True, True)])
def test_plans_emitted_by_the_list_build(bits, periodic):
    """From the second build on the list is built in one walk and emits the tile plans its consumers ask for (cell-box staging
    order instead of a hash of the entries): every slot must name the listed neighbour, pruned views must hold the rows the
    hash builder's pruned plan holds, and the forces must equal the hash-built plans' bit for bit."""
    import torch

    from azplugins_b200 import pair, workloads as W

    snap = W.yukawa_dpd_fluid(n=30000, seed=31, jitter=0.15)
    if not all(periodic):
        snap.box = az.Box(snap.box.Lx, snap.box.Ly, snap.box.Lz, periodic=periodic)
    st = az.State(snap, device="cuda", scalar_bits=bits, seed=42)
    cell = nlist.Cell(buffer=0.4)
    cell.fused_plans = True  # opt-in (off by default: see nlist.NeighborList.plan_requests)
    cell.sort_rows = False  # the plans the list build emits are in list order: compare with hash plans in list order
    d = pair.DPDGeneralWeight(nlist=cell, kT=1.0, default_r_cut=1.0)
    d.dt = 0.01
    d.params[("A", "A")] = W.C5_DPD_PARAMS[("A", "A")]
    d.threads_per_particle = 4
    y = pair.ExpandedYukawa(nlist=cell, default_r_cut=2.0)
    y.params[("A", "A")] = W.EY_PARAMS[("A", "A")]
    for f in (d, y):
        f._attach(st)
        f.compute(2)  # first build: two walks, plans from azb_tiles_build
    assert not d._plan.from_list_build and not y._plan.from_list_build and d._plan.row_count is not None
    want = [(f._force.clone(), f._virial.clone(), None if f._plan.row_count is None else f._plan.row_count.clone()) for f in (d, y)]
    cell.invalidate()
    for f in (d, y):
        f.compute(2)  # second build: one walk that also emits both views
    n = st.N
    nn = cell.n_neigh.cpu().numpy().view(np.uint32).astype(np.int64)
    head = cell.head_list.cpu().numpy().view(np.uint64).astype(np.int64)
    lst = cell.nlist.cpu().numpy().view(np.uint32).astype(np.int64)
    for f, (force, virial, rows) in zip((d, y), want):
        plan = f._plan
        assert plan.from_list_build and plan.uniq is d._plan.uniq
        W_, M = plan.lane_width, 4 * plan.lane_width
        A = max(M, 8)
        count = plan.uniq_count.cpu().numpy().view(np.uint32)
        uniq = plan.uniq.cpu().numpy().view(np.uint32).reshape(plan.n_tiles, plan.uniq_capacity).astype(np.int64)
        slots = plan.slots.cpu().numpy().view(np.uint16).astype(np.int64)
        rc = nn if plan.row_count is None else plan.row_count.cpu().numpy().astype(np.int64)
        if rows is not None:
            assert np.array_equal(rc, rows.cpu().numpy())  # the pruned view holds the rows the hash builder's pruned plan holds
        assert plan.entries == rc.sum() and plan.max_uniq == count.max() <= plan.uniq_capacity
        tmax = plan.tile_max().cpu().numpy()
        for t in range(0, plan.n_tiles, 7):
            assert tmax[t] == uniq[t, : count[t]].max()
            for i in range(t * plan.tile_size, min((t + 1) * plan.tile_size, n)):
                sh = i * plan.slot_row_stride if plan.slot_row_stride else (head[i] + A * i) & ~(A - 1)
                e = np.arange((rc[i] + M - 1) // M * M)
                r = e % M
                stored = slots[sh + (e - r) + 4 * (r % W_) + r // W_]
                assert np.all(stored[rc[i]:] == 0) and np.all(stored[: rc[i]] < count[t])
                named = uniq[t, stored[: rc[i]]]
                row = lst[head[i] : head[i] + nn[i]]
                if plan.row_count is None:
                    assert np.array_equal(named, row)  # slot -> the listed neighbour, in list order
                else:
                    assert np.all(np.isin(named, row)) and len(np.unique(named)) == len(named)
        assert torch.equal(f._force, force) and torch.equal(f._virial, virial)
    # the A/B switch: same list, plans from the hash builder
    cell.fused_plans = False
    cell.invalidate()
    d.compute(2)
    assert not d._plan.from_list_build and torch.equal(d._force, want[0][0])
