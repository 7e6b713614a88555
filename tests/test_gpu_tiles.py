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


@pytest.mark.parametrize("rotate", [False, True])
@pytest.mark.parametrize("lane_width", [1, 2, 4, 8, 16])
@pytest.mark.parametrize("tile_size", [32, 64, 128, 256])
def test_tile_plan_matches_numpy(tile_size, lane_width, rotate):
    n = 5000
    hs, nl = _list_on_device(n, 3.0, 1.4, 11)
    plan = nl.tile_plan(n, tile_size, 2048, lane_width, group_width=lane_width if rotate else 0)
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
