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


@pytest.mark.parametrize("tile_size", [32, 64, 128, 256])
def test_tile_plan_matches_numpy(tile_size):
    n = 5000
    hs, nl = _list_on_device(n, 3.0, 1.4, 11)
    plan = nl.tile_plan(n, tile_size, 2048)
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
            sh = (head[i] + 4 * i) & ~3  # AZB_SLOT_HEAD
            s = slots[sh : sh + nn[i]]
            assert np.array_equal(want[s], lst[head[i] : head[i] + nn[i]])  # slot -> the listed neighbour, entry for entry
            assert np.all(slots[sh + nn[i] : sh + ((nn[i] + 3) & ~3)] == 0)  # padding
        worst = max(worst, len(want))
    assert plan.max_uniq == worst


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
