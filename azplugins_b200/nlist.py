"""Neighbour lists in HOOMD-blue's array layout (``n_neigh`` u32, ``head_list`` size_t, ``nlist`` u32; FULL mode).

``Cell`` mirrors ``hoomd.md.nlist.Cell(buffer=..., exclusions=('bond',))`` as used by every reference test
(/root/reference/src/pytest/test_pair.py:337). The list itself is what the force kernels consume through
``pair_args_t`` (SURVEY.md §8 b); how it is produced is outside the reference (HOOMD's NeighborListGPU*).
"""

import os

import numpy as np
import torch


class NeighborList:
    """Device arrays of one neighbour list plus the per-type-pair r_cut bookkeeping of its consumers."""

    def __init__(self, buffer=0.4, exclusions=("bond",)):
        self.buffer = float(buffer)
        self.exclusions = tuple(exclusions)
        self._consumers = []
        self.n_neigh = None
        self.head_list = None
        self.nlist = None
        self.size_nlist = 0  # extent of the nlist array (entries); rows may have gaps between them
        self.n_entries = 0  # sum of the row lengths
        self._state = None
        self._tiles = {}
        # slot rows of the tile plans in ascending slot order instead of list order (azb_tile_build_args.sort_rows)
        self.sort_rows = os.environ.get("AZB_SORT_ROWS", "0") == "1"

    # consumers (pair forces) register so the list covers max r_cut per type pair, like hoomd's nlist subscribers
    def _subscribe(self, force):
        if force not in self._consumers:
            self._consumers.append(force)
            self.invalidate()

    def invalidate(self):
        self.n_neigh = None
        self._tiles = {}
        self._last_pos = None

    def plan_requests(self, state):
        """Tile plans the consumers attached to ``state`` will ask for, by cache key — known BEFORE the list is built (from the
        previous build's mean row length), so that the build can emit them from its own walk (``_nlist_gpu.build_cell_nlist``)."""
        # OFF by default (opt in: ``cell.fused_plans = True`` or AZB_FUSED_PLANS=1). Measured on B200 (profiles/r2_rebuild_paths.txt):
        # the list build + plans get cheaper (C5 8 M: 24.0 ms against 13.6 + 14.6 ms; LJ 4 M: 16.9 against 15.0 + 7.0 ms) but the
        # cell-box staged sets are supersets of the hash builder's distinct sets, the force kernels stage more and lose
        # occupancy, and a whole MD step gets SLOWER (C5: 11.6 against 10.7 ms per step; DPD 1 M: 1.45 against 1.24 ms).
        on = getattr(self, "fused_plans", None)
        if on is None:
            on = os.environ.get("AZB_FUSED_PLANS", "0") == "1"
        if not self.size_nlist or os.environ.get("AZB_NO_FUSED_PLANS", "0") == "1" or not on:
            return {}
        nbar = self.mean_neighbors(state.N)
        out = {}
        for f in self._consumers:
            if getattr(f, "_state", None) is not state or not getattr(f, "use_tiles", False) or not getattr(f, "_tiled_kernels", False):
                continue
            req = f._plan_request(nbar, require_built=False)
            if req["prune"] is not None and os.environ.get("AZB_FUSED_PRUNED", "1") != "1":
                continue  # A/B switch: pruned views from the hash builder (tighter staged sets) instead of from the list build
            key = (int(req["tile_size"]), 2048, int(req["lane_width"]), int(req["group_width"]), req["prune_key"], bool(req.get("sort_rows", False)))
            out[key] = req
        return out

    def tile_plan(self, N, tile_size=128, uniq_capacity=2048, lane_width=1, group_width=0, prune=None, prune_key=None, sort_rows=False):
        """Tile plan of the current arrays (``azb_tiles_build``), built once per list rebuild and cached; ``None`` when a
        tile holds more than ``uniq_capacity`` distinct neighbours even at the smallest tile size. ``prune`` /
        ``prune_key``: see ``_nlist_gpu.build_tile_plan`` (a pruned plan is specific to the cutoffs of one force). ``lane_width`` is the
        slot-row layout: 1 for the DPD kernels, the lanes per particle for the pair kernels; ``group_width`` the lanes per
        particle of the kernel that walks the rows (bank rotation of the row order; 0 = ``lane_width``)."""
        sort_rows = bool(sort_rows) and int(group_width) == 0
        key = (int(tile_size), int(uniq_capacity), int(lane_width), int(group_width), prune_key, sort_rows)
        if key not in self._tiles:
            from . import _nlist_gpu

            plan = None
            ts = int(tile_size)
            if prune is not None and int(group_width) == 0 and os.environ.get("AZB_NO_DERIVED_PLANS", "0") != "1":
                # a pruned view is derived from the unpruned plan of the same list when some consumer needs that one anyway
                # (azb_tile_build_args.base_plan: no second hash build)
                base = self._unpruned_plan(N, ts, int(uniq_capacity), sort_rows)
                if base is not None:
                    plan = _nlist_gpu.build_tile_plan(self, N, ts, int(uniq_capacity), int(lane_width), 0, prune, base_plan=base)
            while plan is None and ts >= 32:
                plan = _nlist_gpu.build_tile_plan(self, N, ts, int(uniq_capacity), int(lane_width), int(group_width), prune, sort_rows=sort_rows)
                ts //= 2
            self._tiles[key] = plan
        return self._tiles[key]

    def _unpruned_plan(self, N, tile_size, uniq_capacity, sort_rows=False):
        """An unpruned, list-ordered plan of the current arrays with this tile size: one that is cached already, or the one
        another consumer of the list is going to ask for (built now). ``None`` when nobody needs one."""
        for key, plan in self._tiles.items():
            if plan is not None and key[0] == tile_size and key[1] == uniq_capacity and key[3] == 0 and key[4] is None \
                    and plan.tile_size == tile_size and not plan.from_list_build and plan.sorted_rows == sort_rows:
                return plan
        nbar = self.mean_neighbors(N)
        for f in self._consumers:
            if getattr(f, "_state", None) is None or not getattr(f, "use_tiles", False) or not getattr(f, "_tiled_kernels", False):
                continue
            req = f._plan_request(nbar)
            if req["prune"] is None and req["group_width"] == 0 and int(req["tile_size"]) == tile_size and bool(req.get("sort_rows", False)) == sort_rows:
                plan = self.tile_plan(N, **req)
                if plan is not None and plan.tile_size == tile_size:
                    return plan
        return None

    def r_list(self, types):
        """(T,T) matrix of r_cut_max + buffer over all consumers; 0 where no consumer interacts."""
        T = len(types)
        r = np.zeros((T, T))
        for f in self._consumers:
            r = np.maximum(r, f._r_cut_matrix(types))
        return np.where(r > 0, r + self.buffer, 0.0)

    def set_arrays(self, n_neigh, head_list, nlist, device):
        """Adopt externally built arrays (numpy, HOOMD layout)."""
        n_neigh = np.ascontiguousarray(n_neigh, dtype=np.uint32)
        head_list = np.ascontiguousarray(head_list, dtype=np.uint64)
        nlist = np.ascontiguousarray(nlist, dtype=np.uint32)
        self.n_neigh = torch.from_numpy(n_neigh.view(np.int32)).to(device)
        self.head_list = torch.from_numpy(head_list.view(np.int64)).to(device)
        self.nlist = torch.from_numpy(nlist.view(np.int32) if nlist.size else np.zeros(1, np.int32)).to(device)
        self.size_nlist = int(nlist.size)
        self.n_entries = int(n_neigh.sum(dtype=np.int64))
        self._tiles = {}
        self._last_pos = None  # positions at build time unknown: no displacement check, no pruned plans
        return self

    def compute(self, state):
        if self.n_neigh is None or self._state is not state:
            self.build(state)
            self._state = state
        return self

    def build(self, state):
        raise RuntimeError("this NeighborList has no builder: use nlist.Cell or set_arrays()")

    def mean_neighbors(self, N):
        return (self.n_entries or self.size_nlist) / max(N, 1)


class Cell(NeighborList):
    """Cell-list neighbour list built on the GPU (``hoomd.md.nlist.Cell``)."""

    def build(self, state):
        from . import _nlist_gpu

        _nlist_gpu.build_cell_nlist(self, state)
