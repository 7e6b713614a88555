"""Host logic of the pipelined host step (azplugins_b200/hostio.py::plan_schedule) — runs without a GPU.

The schedule decides which tiles a launch holds and when a block of output rows may travel back; a mistake here would not
crash, it would read inputs that have not arrived or copy forces that are not written yet. The invariants are checked on
synthetic "highest index a tile reads" tables shaped like the real ones (monotone with a look-ahead, periodic-boundary
tiles reaching the far end of the index range), on random ones, and on the degenerate sizes."""

import numpy as np
import pytest

from azplugins_b200.hostio import plan_schedule


def _check(hi, in_edges, tile_size, N):
    hi = np.asarray(hi, dtype=np.int64)
    n_tiles = len(hi)
    order, launches, outs = plan_schedule(hi, in_edges, tile_size, N)
    # every tile is launched exactly once
    assert sorted(order.tolist()) == list(range(n_tiles))
    assert sum(n for _, _, n in launches) == n_tiles
    # launches are contiguous position ranges in launch order, waiting for non-decreasing input chunks
    at = 0
    for wait, begin, count in launches:
        assert begin == at and count > 0
        at += count
    waits = [w for w, _, _ in launches]
    assert waits == sorted(waits) and len(set(waits)) == len(waits)
    # a tile is launched only after the chunk holding the highest row it reads has been copied
    launch_of_tile = np.empty(n_tiles, dtype=np.int64)
    for k, (wait, begin, count) in enumerate(launches):
        tiles = order[begin : begin + count]
        launch_of_tile[tiles] = k
        assert np.all(hi[tiles] < in_edges[wait + 1])
        assert np.all(np.diff(tiles) > 0)  # ascending tile index inside a launch
    # output ranges: sorted by launch, disjoint, covering [0, N); none travels before its last writer has run
    assert [o[0] for o in outs] == sorted(o[0] for o in outs)
    covered = np.zeros(N, dtype=np.int32)
    for after, lo, up in outs:
        assert 0 <= lo < up <= N and lo % tile_size == 0
        covered[lo:up] += 1
        t0, t1 = lo // tile_size, (up + tile_size - 1) // tile_size
        assert launch_of_tile[t0:t1].max() <= after < len(launches)
    assert np.all(covered == 1)
    return order, launches, outs


@pytest.mark.parametrize("K", [1, 3, 8, 16])
def test_slab_ordered_system(K):
    """Index order = spatial slabs: a tile reads ~9 % ahead of itself; the first 8 % of the tiles also read the last rows
    (periodic boundary) and must land in the last launch."""
    tile_size, N = 128, 100_000
    n_tiles = (N + tile_size - 1) // tile_size
    own_hi = np.minimum((np.arange(n_tiles) + 1) * tile_size, N) - 1
    hi = np.minimum(own_hi + int(0.09 * N), N - 1)
    hi[: int(0.08 * n_tiles)] = N - 1
    in_edges = np.linspace(0, N, K + 1).astype(np.int64)
    order, launches, outs = _check(hi, in_edges, tile_size, N)
    last_wait, last_begin, last_count = launches[-1]
    assert last_wait == K - 1
    assert set(range(int(0.08 * n_tiles))) <= set(order[last_begin:].tolist())
    if K >= 8:
        # the point of the pipeline: most tiles run before the last input chunk has arrived
        assert last_count < 0.45 * n_tiles


@pytest.mark.parametrize("seed", range(5))
def test_random_tables(seed):
    rng = np.random.default_rng(seed)
    tile_size = int(rng.choice([32, 64, 128, 256]))
    N = int(rng.integers(tile_size, 50_000))
    n_tiles = (N + tile_size - 1) // tile_size
    own_hi = np.minimum((np.arange(n_tiles) + 1) * tile_size, N) - 1
    hi = np.maximum(own_hi, rng.integers(0, N, size=n_tiles))
    K = int(rng.integers(1, 12))
    # ghosts behind the local particles: the input arrays have more rows than N
    n_rows = N + int(rng.integers(0, 2000))
    hi = np.where(rng.random(n_tiles) < 0.1, n_rows - 1, hi)
    in_edges = np.linspace(0, n_rows, K + 1).astype(np.int64)
    _check(hi, in_edges, tile_size, N)


def test_degenerate_sizes():
    _check([5], np.array([0, 6]), 32, 6)  # one tile, one chunk
    _check([31, 40, 40], np.array([0, 10, 20, 41]), 16, 41)  # ragged last tile, empty readiness classes
    order, launches, outs = _check(np.arange(8) * 4 + 3, np.linspace(0, 32, 9).astype(np.int64), 4, 32)
    assert len(launches) == 8 and order.tolist() == list(range(8))  # no look-ahead: tile c runs right after chunk c


def test_tapered_chunk_edges():
    """chunk_edges: ascending, covering [0, n_rows], the last chunks holding the tapered shares; the schedule's invariants hold on
    them, and the last launch of a periodic raster is smaller than with equal chunks."""
    from azplugins_b200.hostio import chunk_edges

    for n_rows, K in ((1000, 8), (8_001_504, 8), (37, 8), (5, 8), (1000, 1), (1000, 2), (1000, 3), (1000, 4)):
        e = chunk_edges(n_rows, K)
        assert len(e) == K + 1 and e[0] == 0 and e[-1] == n_rows and np.all(np.diff(e) >= 0)
    e = chunk_edges(1_000_000, 8)
    assert abs((e[-1] - e[-2]) / 1e6 - 0.02) < 1e-3 and abs((e[-2] - e[-3]) / 1e6 - 0.04) < 1e-3 and abs((e[-3] - e[-4]) / 1e6 - 0.08) < 1e-3
    assert np.array_equal(chunk_edges(1000, 4, taper=()), np.linspace(0, 1000, 5).astype(np.int64))
    # a raster of 29 layers of 100 tiles: a tile reads up to one layer above it, the first layer wraps around to the last one
    tile_size, layers, per = 128, 29, 100
    n_tiles = layers * per
    N = n_tiles * tile_size
    t = np.arange(n_tiles)
    hi = np.minimum((t // per + 2) * per * tile_size, N) - 1
    hi[:per] = N - 1
    last = {}
    for name, edges in (("equal", chunk_edges(N, 8, taper=())), ("tapered", chunk_edges(N, 8))):
        order, launches, outs = _check(hi, edges, tile_size, N)
        last[name] = launches[-1][2]
    assert last["tapered"] <= 0.5 * last["equal"], last
