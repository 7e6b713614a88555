"""Host side of the GPU cell-list builder (azb_nlist_count / azb_nlist_fill in csrc/azb_nlist.cu)."""

import ctypes as C
import os

import numpy as np
import torch

from . import _abi
from .state import scalar_dtype


def build_cell_nlist(nl, state, row_align=1):
    if state.device.type != "cuda":
        raise RuntimeError("nlist.Cell builds on a CUDA device only (no CPU fallback)")
    lib = _abi.load()
    ftype, _, tdtype = scalar_dtype(state.scalar_bits)
    r_list = nl.r_list(state.types)
    if not np.any(r_list > 0):
        raise RuntimeError("neighbour list has no consumer with a positive r_cut")
    dev = state.device
    a = _abi.NlistArgs()
    a.d_pos = state.pos.data_ptr()
    a.N, a.n_max = state.N, state.n_max
    a.box = state.box.as_abi()
    if not all(state.box.periodic):
        p = state.pos[:, :3]
        lo, hi = p.min(dim=0).values.double().cpu().numpy(), p.max(dim=0).values.double().cpu().numpy()
        for d in range(3):
            a.ext_lo[d], a.ext_hi[d] = float(lo[d]) - 1e-6, float(hi[d]) + 1e-6 + 1e-6 * abs(float(hi[d]))
    a.r_list_max = float(r_list.max())
    rl = r_list.astype(ftype)
    rlsq = torch.from_numpy(np.where(r_list > 0, rl * rl, ftype(-1)).reshape(-1).astype(ftype)).to(dev)
    a.d_r_listsq = rlsq.data_ptr()
    a.ntypes = len(state.types)
    a.row_align = row_align
    # one walk instead of two once the previous build has told how long rows get (azb_nlist_args.row_stride_hint)
    prev_max = getattr(nl, "_row_max", 0)
    a.row_stride_hint = (int(prev_max * 1.2) + 8) if (prev_max and getattr(nl, "single_walk", True)) else 0
    # tile plans the consumers will ask for, emitted by the same walk (views of azb_nlist_args)
    requests = nl.plan_requests(state) if a.row_stride_hint else {}
    usable = {k: r for k, r in requests.items() if k[3] == 0 and not k[5]}
    if len(usable) > 4 or len({k[0] for k in usable}) > 1:
        usable = {}
    view_keys = list(usable)
    uniq_count = uniq = uniq_max = None
    if view_keys:
        a.n_views, a.tile_size, a.uniq_capacity = len(view_keys), view_keys[0][0], 2048
        n_tiles = (state.N + a.tile_size - 1) // a.tile_size
        uniq_count = torch.zeros(n_tiles, dtype=torch.int32, device=dev)
        uniq = torch.empty(n_tiles * 2048, dtype=torch.int32, device=dev)
        uniq_max = torch.zeros(n_tiles, dtype=torch.int32, device=dev)
        a.d_uniq_count, a.d_uniq, a.d_uniq_max = uniq_count.data_ptr(), uniq.data_ptr(), uniq_max.data_ptr()
        view_rows, view_slots, view_strides = [], [], []
        for v, key in enumerate(view_keys):
            prune = usable[key]["prune"]
            a.views[v].lane_width = key[2]
            align = max(4 * key[2], 8)  # AZB_SLOT_ROW_ALIGN
            stride = (a.row_stride_hint + align - 1) // align * align  # AZB_VIEW_SLOT_STRIDE
            sl = torch.empty(max(state.N, 1) * stride, dtype=torch.int16, device=dev)
            a.views[v].d_slots = sl.data_ptr()
            view_slots.append(sl)
            view_strides.append(stride)
            rc_t = torch.zeros(max(state.N, 1), dtype=torch.int32, device=dev) if prune is not None else None
            a.views[v].d_keep_rsq = prune[1].data_ptr() if prune is not None else None
            a.views[v].d_row_count = rc_t.data_ptr() if rc_t is not None else None
            view_rows.append(rc_t)
    keep = [rlsq]
    if "bond" in nl.exclusions and state.has_bonds:
        d_head, d_lst = state.exclusions_device()
        a.d_excl_head, a.d_excl_list = d_head.data_ptr(), d_lst.data_ptr()
        keep += [d_head, d_lst]
    n_neigh = torch.zeros(max(state.N, 1), dtype=torch.int32, device=dev)
    head_list = torch.zeros(max(state.N, 1), dtype=torch.int64, device=dev)
    a.d_n_neigh, a.d_head_list = n_neigh.data_ptr(), head_list.data_ptr()
    ws_bytes = C.c_size_t(0)
    _abi.check(lib.azb_nlist_workspace_size(C.byref(a), C.byref(ws_bytes)), "azb_nlist_workspace_size")
    ws = torch.empty(ws_bytes.value, dtype=torch.uint8, device=dev)
    a.d_workspace, a.workspace_bytes = ws.data_ptr(), ws_bytes.value
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    total = C.c_size_t(0)
    # rows in place (azb_nlist_args.rows_in_place): from the second build on, the one walk writes the rows straight into a list
    # array of N * row_stride_hint entries (head_list[i] = i * stride: disjoint, ordered rows with gaps) — no temporary rows, no
    # scan, no compaction. The array is kept between rebuilds. AZB_NLIST_PACKED=1 keeps the tightly packed list (A/B).
    in_place = bool(a.row_stride_hint) and not view_keys and os.environ.get("AZB_NLIST_PACKED", "0") != "1" and getattr(nl, "rows_in_place", True)
    nlist = None
    if in_place:
        want = max(state.N, 1) * a.row_stride_hint
        nlist = getattr(nl, "_nlist_buf", None)
        if nlist is None or nlist.numel() < want or nlist.device != dev:
            nlist = torch.empty(want, dtype=torch.int32, device=dev)
        a.rows_in_place = 1
        a.d_nlist, a.capacity = nlist.data_ptr(), nlist.numel()
    _abi.check(lib.azb_nlist_count(state.scalar_bits, C.byref(a), C.byref(total), stream), "azb_nlist_count")
    if nlist is None or nlist.numel() < max(total.value, 1):
        nlist = torch.empty(max(total.value, 1), dtype=torch.int32, device=dev)
    a.d_nlist, a.capacity = nlist.data_ptr(), nlist.numel()
    _abi.check(lib.azb_nlist_fill(state.scalar_bits, C.byref(a), stream), "azb_nlist_fill")
    nl._nlist_buf = nlist if in_place else None
    plans = {}
    nl_entries = int(n_neigh.sum(dtype=torch.int64).item()) if view_keys else 0
    if view_keys:
        status = _abi.NlistPlanStatus()
        _abi.check(lib.azb_nlist_plan_result(state.scalar_bits, C.byref(a), C.byref(status), stream), "azb_nlist_plan_result")
        n_tiles = (state.N + a.tile_size - 1) // a.tile_size
        good = bool(status.ok) and status.n_bad_tiles <= max(8, n_tiles // 5)
        max_uniq = status.max_uniq
        if good and status.n_bad_tiles:
            # tiles whose particles sit in distant clumps (cell box too large): their rows come from the hash builder, with the
            # distinct set of ALL their list entries so that every view can share it
            bad = torch.nonzero(uniq_count == -1).flatten().to(torch.int32).contiguous()
            scratch = torch.zeros(4, dtype=torch.int32, device=dev)
            for v, key in enumerate(view_keys):
                b = _abi.TileBuildArgs()
                b.d_n_neigh, b.d_nlist, b.d_head_list = n_neigh.data_ptr(), nlist.data_ptr(), head_list.data_ptr()
                b.size_nlist, b.N, b.tile_size, b.uniq_capacity, b.lane_width = int(total.value), state.N, key[0], 2048, key[2]
                b.d_uniq_count, b.d_uniq, b.d_slots, b.d_status = uniq_count.data_ptr(), uniq.data_ptr(), view_slots[v].data_ptr(), scratch.data_ptr()
                b.d_tile_list, b.n_tile_list = bad.data_ptr(), int(bad.numel())
                b.slot_row_stride = view_strides[v]
                prune = usable[key]["prune"]
                if prune is not None:
                    b.d_prune_pos, b.d_prune_rsq = state.pos.data_ptr(), prune[1].data_ptr()
                    b.prune_scalar_bits, b.prune_ntypes, b.prune_box = state.scalar_bits, len(state.types), state.box.as_abi()
                    b.d_row_count, b.prune_full_set = view_rows[v].data_ptr(), 1
                mu, ms, ov = C.c_uint(0), C.c_uint(0), C.c_uint(0)
                _abi.check(lib.azb_tiles_build(C.byref(b), C.byref(mu), C.byref(ms), C.byref(ov), stream), "azb_tiles_build")
                good = good and not ov.value
                max_uniq = max(max_uniq, mu.value)
            if good:
                cnt = uniq_count[bad.long()].long()
                uniq_max[bad.long()] = uniq[bad.long() * 2048 + torch.clamp(cnt - 1, min=0)]  # ascending there: the last entry
        if good:
            for v, key in enumerate(view_keys):
                entries = int(view_rows[v].sum().item()) if view_rows[v] is not None else nl_entries
                plans[key] = TilePlanHandle(key[0], n_tiles, 2048, max_uniq, key[2], status.max_slot_span[v], uniq_count, uniq, view_slots[v], 0.0,
                                            view_rows[v], entries, uniq_max=uniq_max, slot_row_stride=view_strides[v])
    torch.cuda.current_stream(dev).synchronize()
    del keep
    nl.n_neigh, nl.head_list, nl.nlist, nl.size_nlist = n_neigh, head_list, nlist, int(total.value)
    if state.N:
        mx_sum = torch.stack((n_neigh.max().long(), n_neigh.sum(dtype=torch.int64))).cpu()
        nl._row_max, nl.n_entries = int(mx_sum[0]), int(mx_sum[1])
    else:
        nl._row_max, nl.n_entries = 0, 0
    nl._tiles = plans  # plans the build emitted itself; anything else is built on demand (azb_tiles_build)
    # positions the list was built from (locals + ghosts): reference point of the displacement check (md.Integrator) and of
    # the pruning test of force-specific tile plans
    nl._last_pos = state.pos.clone()
    nl.launches = getattr(nl, "launches", 0) + 1
    return nl


class TilePlanHandle:
    """Device arrays of one ``azb_tile_plan`` plus the C struct that points at them (kept alive together)."""

    def __init__(self, tile_size, n_tiles, uniq_capacity, max_uniq, lane_width, max_slot_span, uniq_count, uniq, slots, build_ms, row_count=None,
                 entries=None, uniq_max=None, slot_row_stride=0):
        self.uniq_count, self.uniq, self.slots, self.row_count = uniq_count, uniq, slots, row_count
        self.uniq_max = uniq_max  # plans emitted by the list build: the staged sets are not ascending, the largest index is kept apart
        self.from_list_build = uniq_max is not None
        self.sorted_rows = False  # slot rows in ascending slot order instead of list order (azb_tile_build_args.sort_rows)
        self.build_ms = build_ms
        self.entries = entries  # list entries the plan's rows hold (fewer than the list's when pruned)
        self.c = _abi.TilePlan()
        self.c.tile_size, self.c.n_tiles, self.c.uniq_capacity, self.c.max_uniq = tile_size, n_tiles, uniq_capacity, max_uniq
        self.c.lane_width, self.c.max_slot_span = lane_width, max_slot_span
        self.c.d_uniq_count, self.c.d_uniq, self.c.d_slots = uniq_count.data_ptr(), uniq.data_ptr(), slots.data_ptr()
        self.c.d_row_count = row_count.data_ptr() if row_count is not None else None
        self.c.n_entries = int(entries) if entries is not None else 0  # (a list with gaps between rows: size_nlist overstates it)
        self.c.slot_row_stride = int(slot_row_stride)

    tile_size = property(lambda self: self.c.tile_size)
    n_tiles = property(lambda self: self.c.n_tiles)
    uniq_capacity = property(lambda self: self.c.uniq_capacity)
    max_uniq = property(lambda self: self.c.max_uniq)
    lane_width = property(lambda self: self.c.lane_width)
    max_slot_span = property(lambda self: self.c.max_slot_span)
    slot_row_stride = property(lambda self: self.c.slot_row_stride)

    def pointer(self):
        return C.pointer(self.c)

    def tile_max(self):
        """int64 tensor [n_tiles]: the largest particle index each tile stages (it reads a ghost iff that is >= N)."""
        if self.uniq_max is not None:
            return self.uniq_max.long()
        count = self.uniq_count.long()
        t = torch.arange(self.n_tiles, device=count.device, dtype=torch.int64)
        last = self.uniq[t * self.uniq_capacity + torch.clamp(count - 1, min=0)].long()  # ascending list: the last entry
        return torch.where(count > 0, last, torch.zeros_like(last))


def build_tile_plan(nl, N, tile_size, uniq_capacity, lane_width=1, group_width=0, prune=None, base_plan=None, sort_rows=False):
    """``azb_tiles_build`` over the list's current arrays; returns ``None`` on overflow (caller retries smaller tiles).
    ``prune = (state, keep_rsq)``: drop the entries that cannot come inside the force's cutoff before the next rebuild
    (``keep_rsq``: device Scalar tensor [T*T] of (r_cut + buffer)^2; the state's positions must be the ones the list was built
    from)."""
    dev = nl.n_neigh.device
    if dev.type != "cuda":
        raise RuntimeError("tile plans are built on a CUDA device only (no CPU fallback)")
    if N == 0:
        return None
    lib = _abi.load()
    n_tiles = (N + tile_size - 1) // tile_size
    uniq_count = torch.zeros(n_tiles, dtype=torch.int32, device=dev)
    uniq = torch.empty(n_tiles * uniq_capacity, dtype=torch.int32, device=dev)
    # AZB_SLOTS_SIZE; not cleared: the builder writes every slot row including its padding, nothing reads the gaps between rows
    slots = torch.empty(nl.size_nlist + max(4 * lane_width, 8) * (N + 1), dtype=torch.int16, device=dev)
    status = torch.zeros(4, dtype=torch.int32, device=dev)
    a = _abi.TileBuildArgs()
    a.d_n_neigh, a.d_nlist, a.d_head_list = nl.n_neigh.data_ptr(), nl.nlist.data_ptr(), nl.head_list.data_ptr()
    a.size_nlist, a.N, a.tile_size, a.uniq_capacity, a.lane_width = nl.size_nlist, N, tile_size, uniq_capacity, lane_width
    a.group_width = group_width
    a.sort_rows = 1 if (sort_rows and group_width == 0) else 0
    a.d_uniq_count, a.d_uniq, a.d_slots, a.d_status = uniq_count.data_ptr(), uniq.data_ptr(), slots.data_ptr(), status.data_ptr()
    row_count = None
    if prune is not None:
        state, keep_rsq = prune
        row_count = torch.zeros(max(N, 1), dtype=torch.int32, device=dev)
        a.d_prune_pos, a.d_prune_rsq = nl._last_pos.data_ptr(), keep_rsq.data_ptr()
        a.prune_scalar_bits, a.prune_ntypes = state.scalar_bits, len(state.types)
        a.prune_box = state.box.as_abi()
        a.d_row_count = row_count.data_ptr()
        if base_plan is not None:
            a.base_plan = C.cast(base_plan.pointer(), C.c_void_p)
    max_uniq, max_span, overflow = C.c_uint(0), C.c_uint(0), C.c_uint(0)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    _abi.check(lib.azb_tiles_build(C.byref(a), C.byref(max_uniq), C.byref(max_span), C.byref(overflow), stream), "azb_tiles_build")
    e1.record()
    e1.synchronize()
    if overflow.value:
        return None
    entries = int(row_count.sum().item()) if row_count is not None else (nl.n_entries or nl.size_nlist)
    plan = TilePlanHandle(tile_size, n_tiles, uniq_capacity, max_uniq.value, lane_width, max_span.value, uniq_count, uniq, slots,
                          e0.elapsed_time(e1), row_count, entries)
    plan.sorted_rows = bool(a.sort_rows) or (base_plan is not None and getattr(base_plan, "sorted_rows", False))
    return plan
