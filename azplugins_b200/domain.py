"""Spatial domain decomposition with a ghost-particle halo (SURVEY.md §8 e) — the multi-GPU form of the force path.

One process per GPU on a Cartesian grid (1) / (2,1,1) / (2,2,1) / (2,2,2) of the periodic global box. Every rank owns the
particles inside its sub-box; copies of neighbours' particles within ``r_ghost`` of a face / edge / corner are stored behind
the locals at indices ``[N, N + n_ghost)`` — HOOMD's convention: the force kernels read ``pos[j]`` for ``j >= N`` and never
write there. With the FULL neighbour list there is no reverse (force) communication. The local box is non-periodic along
decomposed directions; ghosts arriving across the global boundary are shifted by -/+L so they sit next to the sub-box.

Per step ``HaloExchange.update()`` refreshes the dynamic ghost fields (positions; velocities for DPD; orientations for
anisotropic pairs) with one grouped point-to-point exchange (``torch.distributed`` P2P: NCCL over NVLink on GPUs, gloo in the CPU
tests). Tags and types are static and travel once at setup. DPD random numbers depend only on (seed, timestep, tag_i, tag_j),
so forces are independent of the decomposition up to summation order.
"""

import ctypes as C
import itertools
import os

import numpy as np
import torch
import torch.distributed as dist

from . import _abi
from .state import Box, scalar_dtype

GRIDS = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}
DIRECTIONS = [d for d in itertools.product((-1, 0, 1), repeat=3) if d != (0, 0, 0)]


class Decomposition:
    """Cartesian split of a periodic, origin-centred, orthorhombic global box."""

    def __init__(self, world, rank, global_box, grid=None):
        if grid is None:
            if world not in GRIDS:
                raise ValueError(f"no default domain grid for {world} ranks (use 1, 2, 4 or 8)")
            grid = GRIDS[world]
        if int(np.prod(grid)) != world:
            raise ValueError("grid does not match the number of ranks")
        if global_box.xy or global_box.xz or global_box.yz:
            raise ValueError("domain decomposition supports orthorhombic boxes only")
        self.world, self.rank, self.grid, self.global_box = world, rank, tuple(grid), global_box
        self.coords = self.coords_of(rank)
        L = global_box.L
        w = L / np.array(self.grid)
        self.lo = -L / 2 + w * np.array(self.coords)
        self.hi = self.lo + w
        self.width = w

    def coords_of(self, rank):
        gx, gy, gz = self.grid
        return (rank % gx, (rank // gx) % gy, rank // (gx * gy))

    def rank_of(self, coords):
        gx, gy, gz = self.grid
        return (coords[0] % gx) + gx * ((coords[1] % gy) + gy * (coords[2] % gz))

    def neighbor(self, d):
        """(rank, shift) of the domain in direction d; shift moves MY particles into the neighbour's frame."""
        c = [self.coords[k] + d[k] for k in range(3)]
        shift = np.zeros(3)
        for k in range(3):
            if c[k] < 0:
                shift[k] = +self.global_box.L[k]  # leaving through the low face: reappear above the neighbour's box
            elif c[k] >= self.grid[k]:
                shift[k] = -self.global_box.L[k]
        return self.rank_of(c), shift

    def active_directions(self):
        return [d for d in DIRECTIONS if all(d[k] == 0 or self.grid[k] > 1 for k in range(3))]

    def local_box(self):
        """Box handed to the kernels: global extent + periodic where not decomposed, non-periodic where decomposed."""
        gb = self.global_box
        return Box(gb.Lx, gb.Ly, gb.Lz, periodic=tuple(self.grid[k] == 1 for k in range(3)))

    def owns(self, pos):
        pos = np.asarray(pos)
        m = np.ones(len(pos), dtype=bool)
        for k in range(3):
            if self.grid[k] > 1:
                m &= (pos[:, k] >= self.lo[k]) & (pos[:, k] < self.hi[k])
        return m

    def send_mask(self, pos, d, r_ghost):
        """Local particles that the neighbour in direction d needs as ghosts."""
        m = np.ones(len(pos), dtype=bool)
        for k in range(3):
            if d[k] > 0:
                m &= pos[:, k] >= self.hi[k] - r_ghost
            elif d[k] < 0:
                m &= pos[:, k] < self.lo[k] + r_ghost
        return m


def _gather_rows(out, src, idx, shift):
    """out[e] = src[idx[e]] (+ shift[e] on xyz): CUDA kernel through the C-ABI on device tensors; torch indexing on CPU tensors
    (the CPU form exists for the gloo tests of the exchange logic, the force path itself has no CPU form)."""
    n = idx.numel()
    if n == 0:
        return
    if src.is_cuda:
        lib = _abi.load()
        bits = 32 if src.dtype == torch.float32 else 64
        code = lib.azb_gather_scalar4(bits, src.data_ptr(), idx.data_ptr(), shift.data_ptr() if shift is not None else None, out.data_ptr(), n,
                                      C.c_void_p(torch.cuda.current_stream(src.device).cuda_stream))
        _abi.check(code, "azb_gather_scalar4")
    else:
        torch.index_select(src, 0, idx.long(), out=out)
        if shift is not None:
            out[:, :3] += shift[:, :3]


class HaloExchange:
    """Ghost bookkeeping of one rank. At construction and on every neighbour-list rebuild (``rebuild()``): particles that left
    the sub-box migrate to the neighbouring ranks, the ghost layer is re-selected, bonds are re-indexed (a partner may be a
    ghost), and the route table of the peer-memory push is rewritten — all on device tensors with variable-size messages whose
    sizes travel first. Between rebuilds ``update()`` / ``step()`` refresh the dynamic ghost fields every step.

    The particle arrays of the State become views ``[: N + n_ghost]`` of fixed-capacity stores, so that their addresses (and
    the CUDA IPC mappings the neighbours hold on them) survive migration; a store grows only when a rank outgrows it."""

    STORE_FIELDS = ("pos", "vel", "orientation")

    def __init__(self, decomp, state, r_ghost, fields=("pos",), group=None, transport=None, slack=1.25):
        self.decomp, self.state, self.group = decomp, state, group
        self.fields = tuple(fields)
        self.transport = "nccl"  # "peer" once the neighbours' arrays are mapped (below)
        self.timeout_ms = 30000
        self.r_ghost = float(r_ghost)
        self.slack = float(slack)
        self.launches = 0  # pack / push / flag kernels issued
        self.rebuilds = 0
        self.migrated = 0  # particles this rank has sent away so far
        self._steps = 0
        if np.any(decomp.width[np.array(decomp.grid) > 1] < 2 * r_ghost):
            raise ValueError("sub-domain narrower than 2 r_ghost")
        dev = state.device
        self.dirs = decomp.active_directions()
        nd = len(self.dirs)
        self.peers_send = [decomp.neighbor(d)[0] for d in self.dirs]
        self.peers_recv = [decomp.neighbor(tuple(-x for x in d))[0] for d in self.dirs]
        ftype = scalar_dtype(state.scalar_bits)[0]
        self._shifts = torch.tensor(np.array([decomp.neighbor(d)[1] for d in self.dirs]).reshape(nd, 3), dtype=state.pos.dtype, device=dev)
        # direction lookup: (dx + 1) * 9 + (dy + 1) * 3 + (dz + 1) -> index into self.dirs (-1: not an active direction)
        lut = np.full(27, -1, dtype=np.int64)
        for k, d in enumerate(self.dirs):
            lut[(d[0] + 1) * 9 + (d[1] + 1) * 3 + (d[2] + 1)] = k
        self._dir_lut = torch.from_numpy(lut).to(dev)
        self._lo = torch.tensor(decomp.lo, dtype=torch.float64, device=dev)
        self._hi = torch.tensor(decomp.hi, dtype=torch.float64, device=dev)
        self._split = torch.tensor([g > 1 for g in decomp.grid], device=dev)
        # every rank's table of peers: who fills rank r's slot k (needed to derive everybody's ghost offsets from one gather)
        self._recv_peer_of = np.zeros((decomp.world, nd), dtype=np.int64)
        for r in range(decomp.world):
            other = Decomposition(decomp.world, r, decomp.global_box, decomp.grid)
            for k, d in enumerate(self.dirs):
                self._recv_peer_of[r, k] = other.neighbor(tuple(-x for x in d))[0]
        state.global_box = decomp.global_box  # what external.HarmonicBarrier wraps positions into
        state.box = decomp.local_box()
        if getattr(state, "image", None) is None:
            state.image = torch.zeros((state.N, 3), dtype=torch.int32, device=dev)
        self._store = {}
        self._cap = 0
        self._bufs = {}
        self._peer = None  # opened base pointers per (rank, field) once the peer transport is up
        self._gather_global_bonds()
        self._n_global = self._count_global()
        self.send_idx, self.recv_counts = [], []
        self.recv_offsets = np.zeros(nd + 1, dtype=np.int64)
        self.reselect()
        want = transport or os.environ.get("AZB_HALO_TRANSPORT", "auto")
        if want not in ("auto", "peer", "nccl"):
            raise ValueError("transport must be 'auto', 'peer' or 'nccl'")
        self._want_peer = want != "nccl" and decomp.world > 1 and dev.type == "cuda"
        self._peer_required = want == "peer"
        if self._want_peer:
            self._setup_peer(required=self._peer_required)

    # ---- storage ------------------------------------------------------------------------------------------------------------
    def _views(self, n_local, n_ghost):
        st = self.state
        n = n_local + n_ghost
        for name in self.STORE_FIELDS:
            setattr(st, name, self._store[name][:n])
        st.tag = self._store["tag"][:n]
        st.image = self._store["image"][:n_local]
        st.N, st.n_ghost = n_local, n_ghost

    def _ensure_capacity(self, rows):
        """Stores of at least `rows` rows (locals + ghosts). Growing reallocates, which invalidates the neighbours' IPC mappings:
        returns True when it grew."""
        if rows <= self._cap:
            return False
        self._grew = True
        st, dev = self.state, self.state.device
        cap = int(rows * self.slack) + 1024
        n_keep = st.N
        new = {name: torch.zeros((cap, 4), dtype=getattr(st, name).dtype, device=dev) for name in self.STORE_FIELDS}
        new["tag"] = torch.zeros(cap, dtype=st.tag.dtype, device=dev)
        new["image"] = torch.zeros((cap, 3), dtype=torch.int32, device=dev)
        for name in self.STORE_FIELDS + ("tag", "image"):
            new[name][:n_keep] = getattr(st, name)[:n_keep]
        self._store, self._cap = new, cap
        self._views(st.N, 0)
        return True

    def _count_global(self):
        n = torch.tensor([self.state.N], dtype=torch.int64, device=self.state.device)
        if self.decomp.world > 1:
            dist.all_reduce(n, group=self.group)
        return int(n.item())

    def _gather_counts(self, row):
        """One collective: every rank contributes a row of int64, everyone gets the (world, len) table."""
        mine = torch.tensor(row, dtype=torch.int64, device=self.state.device)
        if self.decomp.world == 1:
            return mine.reshape(1, -1).cpu().numpy()
        out = [torch.empty_like(mine) for _ in range(self.decomp.world)]
        dist.all_gather(out, mine, group=self.group)
        return torch.stack(out).cpu().numpy()

    # ---- bonds: global identity by tag, local indices per rank ---------------------------------------------------------------
    def _gather_global_bonds(self):
        """The bond topology (tag pairs + type) is replicated on every rank — 12 B per bond — and re-indexed locally after
        every migration; HOOMD distributes the bond table with the particles instead, same result."""
        st, dev = self.state, self.state.device
        mine = torch.from_numpy(np.concatenate([st.bond_tags.astype(np.int64), st.bond_typeid.astype(np.int64)[:, None]], axis=1)
                                if len(st.bond_tags) else np.zeros((0, 3), dtype=np.int64)).to(dev)
        if self.decomp.world > 1:
            sizes = self._gather_counts([mine.shape[0]])[:, 0]
            if sizes.sum() == 0:
                self._bonds = None
                return
            pad = torch.zeros((int(sizes.max()), 3), dtype=torch.int64, device=dev)
            pad[: mine.shape[0]] = mine
            parts = [torch.empty_like(pad) for _ in range(self.decomp.world)]
            dist.all_gather(parts, pad, group=self.group)
            allb = torch.cat([p[: int(n)] for p, n in zip(parts, sizes)])
            # a bond known to both owners of its members appears twice: keep one copy
            key = torch.minimum(allb[:, 0], allb[:, 1]) * (1 << 32) + torch.maximum(allb[:, 0], allb[:, 1])
            _, first = np.unique(key.cpu().numpy(), return_index=True)
            allb = allb[torch.from_numpy(np.sort(first)).to(dev)]
        else:
            allb = mine
        self._bonds = allb if allb.shape[0] else None

    def _index_bonds(self):
        """Local indices of every bond with a local member (tag -> index through a reverse map; locals win over ghosts)."""
        st = self.state
        if self._bonds is None:
            return
        dev = st.device
        n_max = st.N + st.n_ghost
        tags = st.tag[:n_max].long() & 0xFFFFFFFF
        size = max(self._n_global, int(tags.max().item()) + 1 if n_max else 1, int(self._bonds[:, :2].max().item()) + 1)
        rtag = torch.full((size,), -1, dtype=torch.int64, device=dev)
        idx = torch.arange(n_max, device=dev)
        rtag[tags[st.N:]] = idx[st.N:]
        rtag[tags[: st.N]] = idx[: st.N]
        a, b = rtag[self._bonds[:, 0]], rtag[self._bonds[:, 1]]
        mine = ((a >= 0) & (a < st.N)) | ((b >= 0) & (b < st.N))
        a, b, t = a[mine], b[mine], self._bonds[mine, 2]
        if bool(((a < 0) | (b < 0)).any()):
            raise RuntimeError("domain decomposition: a bond's partner is neither local nor in the ghost layer "
                               "(bond longer than r_cut + buffer, or the bond list does not match the particle tags)")
        st.set_local_bonds(torch.stack([a, b], dim=1), t)

    # ---- migration --------------------------------------------------------------------------------------------------------------
    def _axis_codes(self, p, lo, hi):
        """(n, 3) int64 in {-1, 0, +1}: below lo / inside / at or above hi, along decomposed directions only."""
        p = p.double()
        code = (p >= hi).long() - (p < lo).long()
        return code * self._split.long()

    def migrate(self):
        """Particles outside the sub-box move to the rank that owns their position (at most one domain away per rebuild:
        a rebuild happens before anything has moved half a buffer). Leavers' rows are replaced in place by arrivals and by
        particles from the tail, so the arrays stay dense. Returns (sent, received)."""
        st, dev, nd = self.state, self.state.device, len(self.dirs)
        N = st.N
        code = self._axis_codes(st.pos[:N, :3], self._lo, self._hi)
        cand = torch.nonzero((code != 0).any(dim=1)).flatten()
        k_of = self._dir_lut[(code[cand, 0] + 1) * 9 + (code[cand, 1] + 1) * 3 + (code[cand, 2] + 1)]
        send_idx = [cand[k_of == k] for k in range(nd)]
        table = self._gather_counts([int(i.numel()) for i in send_idx])
        recv_counts = [int(table[self.peers_recv[k], k]) for k in range(nd)]
        n_leave, n_arrive = int(sum(int(i.numel()) for i in send_idx)), int(sum(recv_counts))
        if int(table.sum()) == 0:
            return 0, 0
        sends, recvs, arrivals = [], [], {}
        names = self.STORE_FIELDS + ("tag", "image")
        for name in names:
            src = self._store[name]
            arrivals[name] = torch.empty((n_arrive,) + tuple(src.shape[1:]), dtype=src.dtype, device=dev)
            off = 0
            for k in range(nd):
                buf = src[send_idx[k]].clone()
                if name == "pos" and buf.numel():
                    buf[:, :3] += self._shifts[k]
                if name == "image" and buf.numel():
                    buf -= torch.sign(self._shifts[k]).to(torch.int32)  # crossing the global face: position shifted by +-L, image by -+1
                sends.append((buf, self.peers_send[k]))
                recvs.append((arrivals[name][off : off + recv_counts[k]], self.peers_recv[k]))
                off += recv_counts[k]
        self._p2p(sends, recvs)
        # new dense layout: holes (leavers) <- arrivals, then tail particles; surplus arrivals are appended
        holes = torch.sort(torch.cat(send_idx))[0] if n_leave else torch.zeros(0, dtype=torch.int64, device=dev)
        n_new = N - n_leave + n_arrive
        self._ensure_capacity(n_new)  # (growth is remembered: reselect() re-maps the peers)
        n_fill = min(n_leave, n_arrive)
        plan = []
        if n_fill:
            plan.append((holes[:n_fill], ("arrivals", 0, n_fill)))
        if n_arrive > n_leave:
            plan.append((torch.arange(N, N + n_arrive - n_leave, device=dev), ("arrivals", n_fill, n_arrive)))
        elif n_leave > n_arrive:
            rest = holes[n_arrive:]
            is_hole = torch.zeros(N, dtype=torch.bool, device=dev)
            is_hole[rest] = True
            tail = torch.arange(n_new, N, device=dev)
            movers = tail[~is_hole[n_new:N]]
            low = rest[rest < n_new]
            plan.append((low, ("locals", movers)))
        for name in names:
            store = self._store[name]
            for dst, src in plan:
                store[dst] = arrivals[name][src[1] : src[2]] if src[0] == "arrivals" else store[src[1]]
        self._views(n_new, 0)
        self.migrated += n_leave
        return n_leave, n_arrive

    # ---- ghost selection ---------------------------------------------------------------------------------------------------------
    def reselect(self):
        """Choose the ghost layer from the current local positions: per direction the local particles within r_ghost of the
        face / edge / corner; sizes are exchanged first, then tags and every particle field travel once (the per-step
        refresh moves only the dynamic fields). Re-indexes the bonds and rewrites the peer route table."""
        st, dev, nd = self.state, self.state.device, len(self.dirs)
        N = st.N
        rg = self.r_ghost
        p = st.pos[:N, :3]
        near = self._axis_codes(p, self._lo + rg, self._hi - rg)  # -1: within r_ghost of the low face, +1: of the high face
        lo_flag, hi_flag = (near < 0), (near > 0)
        cand = torch.nonzero((lo_flag | hi_flag).any(dim=1)).flatten()
        lo_c, hi_c = lo_flag[cand], hi_flag[cand]
        self.send_idx = []
        for d in self.dirs:
            m = torch.ones(cand.numel(), dtype=torch.bool, device=dev)
            for k in range(3):
                if d[k] > 0:
                    m &= hi_c[:, k]
                elif d[k] < 0:
                    m &= lo_c[:, k]
            self.send_idx.append(cand[m].to(torch.int32).contiguous())
        table = self._gather_counts([N] + [int(i.numel()) for i in self.send_idx] + [self._cap])
        self._table = table
        counts = table[:, 1 : 1 + nd]
        self.recv_counts = [int(counts[self.peers_recv[k], k]) for k in range(nd)]
        self.recv_offsets = np.concatenate([[0], np.cumsum(self.recv_counts)]).astype(np.int64)
        n_ghost = int(self.recv_offsets[-1])
        self._ensure_capacity(N + n_ghost)
        grew, self._grew = getattr(self, "_grew", False), False
        self._views(N, n_ghost)
        self._bufs = {}
        self._order_cache = None
        # tags (static) + all particle fields once
        sends, recvs = [], []
        for k in range(nd):
            sends.append((st.tag[self.send_idx[k].long()].contiguous(), self.peers_send[k]))
            o0, o1 = N + int(self.recv_offsets[k]), N + int(self.recv_offsets[k + 1])
            recvs.append((st.tag[o0:o1], self.peers_recv[k]))
        self._p2p(sends, recvs)
        self._exchange_fields(self.STORE_FIELDS)
        self._index_bonds()
        # peer transport: everyone learns whether anybody's stores moved; then either re-map or just rewrite the routes
        if self._peer is not None:
            flag = torch.tensor([1 if grew else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=self.group)
            if int(flag.item()):
                self._setup_peer(required=self._peer_required)
            else:
                self._write_routes()
        self.rebuilds += 1

    def rebuild(self):
        """Neighbour-list rebuild of a decomposed system: migration, then ghost re-selection (collective: every rank calls it
        at the same step)."""
        self.migrate()
        self.reselect()

    # ---- ghost refresh over peer memory: one push kernel per step instead of pack kernels + grouped send/recv ---------
    def _setup_peer(self, required):
        """Map the neighbours' particle stores and flag words into this process (CUDA IPC over NVLink). Every rank decides
        together: if any mapping fails, all stay on the NCCL exchange. The stores have a fixed address and capacity, so this
        runs once (and again only if some rank had to grow its stores); the per-rebuild part is ``_write_routes``."""
        st, lib, nd = self.state, _abi.load(), len(self.dirs)
        dev = st.device
        # flag words: [0, nd) data epochs written by the ranks that send to me; [nd, 2 nd) acknowledgements written by the ranks
        # I send to; then the push kernel's arrival counter and the timeout indicator
        self._sync = torch.zeros(2 * nd + 2, dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)

        def export(t):
            h = C.create_string_buffer(_abi.IPC_HANDLE_BYTES)
            off = C.c_size_t(0)
            code = lib.azb_ipc_export(C.c_void_p(t.data_ptr()), h, C.byref(off))
            return (code, h.raw, int(off.value))

        mine = {"sync": export(self._sync), "fields": {name: export(self._store[name]) for name in self.STORE_FIELDS}}
        everyone = [None] * self.decomp.world
        dist.all_gather_object(everyone, mine, group=self.group)
        ok, opened = True, {}

        def open_(rank, what):
            nonlocal ok
            code, handle, off = what
            if rank == self.decomp.rank:
                return None  # caller uses the local pointer
            if code != 0:
                ok = False
                return 0
            key = (rank, handle, off)
            if key not in opened:
                ptr = C.c_void_p(0)
                if lib.azb_ipc_open(handle, off, C.byref(ptr)) != 0:
                    ok = False
                    return 0
                opened[key] = int(ptr.value)
            return opened[key]

        peer = {"sync": {}, "fields": {}}
        for r in set(self.peers_send) | set(self.peers_recv):
            base = open_(r, everyone[r]["sync"])
            peer["sync"][r] = self._sync.data_ptr() if base is None else base
            for name in self.fields:
                base = open_(r, everyone[r]["fields"][name])
                peer["fields"][(r, name)] = self._store[name].data_ptr() if base is None else base
        flag = torch.tensor([0 if ok else 1], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, group=self.group)
        if int(flag.item()) != 0:
            if required:
                raise RuntimeError("halo transport 'peer': the neighbours' memory cannot be mapped (no P2P access between the GPUs?)")
            self._peer, self.transport = None, "nccl"
            return
        self._peer = peer
        data_flags, ack_flags = [], []
        for k in range(nd):
            # the rank in direction d stores what I send for direction k in ITS slot k; its data flag k is mine to raise. The rank
            # behind me filled my slot k: its acknowledgement word k is mine to raise once I have read the ghosts
            data_flags.append(peer["sync"][self.peers_send[k]] + 4 * k)
            ack_flags.append(peer["sync"][self.peers_recv[k]] + 4 * (nd + k))
        self._data_ptrs = torch.tensor(data_flags, dtype=torch.int64, device=dev)
        self._ack_ptrs = torch.tensor(ack_flags, dtype=torch.int64, device=dev)
        self._epoch = 0
        self._d_error = self._sync.data_ptr() + 4 * (2 * nd + 1)
        self._write_routes()
        dist.barrier(group=self.group)  # nobody raises a flag before everybody's words exist
        self.transport = "peer"

    def _write_routes(self):
        """Route table of ``azb_halo_push`` for the current selection: one route per (direction, field) = (my rows to gather,
        where they land behind the receiver's locals). Needs only the count table every rank already holds."""
        st, nd = self.state, len(self.dirs)
        table = self._table
        counts = table[:, 1 : 1 + nd]
        row_bytes = 4 * (st.scalar_bits // 8)
        routes = []
        for k, d in enumerate(self.dirs):
            to = self.peers_send[k]
            # the receiver's ghost offset of slot k: everything its slots 0..k-1 receive
            before = sum(int(counts[self._recv_peer_of[to, j], j]) for j in range(k))
            first = int(table[to, 0]) + before
            for name in self.fields:
                shift = self.decomp.neighbor(d)[1] if name == "pos" else np.zeros(3)
                routes.append((self._store[name].data_ptr(), self._peer["fields"][(to, name)] + first * row_bytes, self.send_idx[k].data_ptr(),
                               int(self.send_idx[k].numel()), shift))
        tab = (_abi.HaloRoute * max(len(routes), 1))()
        begin = 0
        for r, (src, dst, idx, n, shift) in zip(tab, routes):
            r.d_src, r.d_dst, r.d_idx, r.n, r.row_begin = src, dst, idx, n, begin
            r.shift[0], r.shift[1], r.shift[2] = (float(x) for x in shift)
            begin += n
        self._routes = torch.frombuffer(bytearray(bytes(tab)), dtype=torch.uint8).to(st.device)
        a = _abi.HaloPushArgs()
        a.d_routes, a.n_routes, a.total_rows = self._routes.data_ptr(), len(routes), begin
        a.d_peer_flags, a.n_flags = self._data_ptrs.data_ptr(), nd
        a.d_counter = self._sync.data_ptr() + 4 * (2 * nd)
        self._push_args = a

    def _peer_begin(self, stream):
        """Start of a refresh on ``stream``: whatever was enqueued before has finished reading the previous ghosts."""
        lib, nd = _abi.load(), len(self.dirs)
        # a free-running 32-bit counter: the device compares epochs by signed difference, so the wrap needs no special case
        self._epoch = (self._epoch + 1) & 0xFFFFFFFF
        _abi.check(lib.azb_halo_signal(self._ack_ptrs.data_ptr(), nd, (self._epoch - 1) & 0xFFFFFFFF, stream), "azb_halo_signal")
        self.launches += 1

    def _peer_push(self, stream):
        """Wait until the receivers have released their ghost rows, then store this step's rows into them."""
        lib, nd = _abi.load(), len(self.dirs)
        _abi.check(lib.azb_halo_wait(self._sync.data_ptr() + 4 * nd, nd, (self._epoch - 1) & 0xFFFFFFFF, self.timeout_ms, self._d_error, stream),
                   "azb_halo_wait")
        self._push_args.epoch = self._epoch
        _abi.check(lib.azb_halo_push(self.state.scalar_bits, C.byref(self._push_args), stream), "azb_halo_push")
        self.launches += 2

    def _peer_arrived(self, stream):
        """Stream-ordered wait for this step's ghosts from every neighbour."""
        lib, nd = _abi.load(), len(self.dirs)
        _abi.check(lib.azb_halo_wait(self._sync.data_ptr(), nd, self._epoch, self.timeout_ms, self._d_error, stream), "azb_halo_wait")
        self.launches += 1

    def check(self):
        """Raise if a flag wait ever timed out (a neighbour stopped stepping); synchronises."""
        if self.transport == "peer" and int(self._sync[2 * len(self.dirs) + 1].item()) != 0:
            raise RuntimeError("halo exchange: timed out waiting for a neighbouring rank")

    def _p2p(self, sends, recvs):
        """One grouped exchange; per peer, sends and receives are posted in direction order on both sides so they match."""
        ops = []
        for buf, peer in sends:
            if buf.numel():
                ops.append(dist.P2POp(dist.isend, buf, peer, group=self.group))
        for buf, peer in recvs:
            if buf.numel():
                ops.append(dist.P2POp(dist.irecv, buf, peer, group=self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    def _exchange_fields(self, fields):
        """Grouped point-to-point refresh of the given ghost fields (NCCL over NVLink on GPUs, gloo in the CPU tests): one pack
        kernel per direction and field, received in place behind the locals (no unpack)."""
        st = self.state
        N = st.N
        sends, recvs = [], []
        for name in fields:
            arr = getattr(st, name)
            for k in range(len(self.dirs)):
                n = self.send_idx[k].numel()
                key = (name, k)
                if key not in self._bufs:
                    self._bufs[key] = torch.empty((n, 4), dtype=arr.dtype, device=arr.device)
                    if name == "pos":
                        sh = torch.zeros((n, 4), dtype=arr.dtype, device=arr.device)
                        sh[:, :3] = self._shifts[k]
                        self._bufs[("shift", k)] = sh
                buf = self._bufs[key]
                _gather_rows(buf, arr, self.send_idx[k], self._bufs[("shift", k)] if name == "pos" else None)
                self.launches += 1 if (n and arr.is_cuda) else 0
                sends.append((buf, self.peers_send[k]))
                o0, o1 = N + int(self.recv_offsets[k]), N + int(self.recv_offsets[k + 1])
                recvs.append((arr[o0:o1], self.peers_recv[k]))  # contiguous slice: received in place, no unpack kernel
        self._p2p(sends, recvs)

    def _poll(self):
        """A flag wait that timed out leaves stale ghosts behind: look at the error word every 64 refreshes (one 4-byte read)
        and stop the run instead of integrating garbage."""
        self._steps += 1
        if self.transport == "peer" and self._steps % 64 == 0:
            self.check()

    def update(self, all_fields=False):
        """Refresh ghost copies of the dynamic fields (every step, before the force kernels)."""
        st = self.state
        if self.transport == "peer" and not all_fields:
            stream = C.c_void_p(torch.cuda.current_stream(st.device).cuda_stream)
            self._peer_begin(stream)
            self._peer_push(stream)
            self._peer_arrived(stream)
            self._poll()
            return
        self._exchange_fields(self.STORE_FIELDS if all_fields else self.fields)

    # ---- ghost exchange overlapped with the ghost-free part of the force evaluation ---------------------------------
    def tile_order(self, plan):
        """(order, n_interior): launch order of the plan's tiles with the ghost-free ones first. A tile reads a ghost iff
        the last (largest) entry of its ascending distinct-neighbour list is >= N. Cached per plan."""
        cached = getattr(self, "_order_cache", None)
        if cached is not None and cached[0] is plan:
            return cached[1], cached[2]
        N = self.state.N
        boundary = plan.tile_max() >= N
        order = torch.cat([torch.nonzero(~boundary).flatten(), torch.nonzero(boundary).flatten()]).to(torch.int32).contiguous()
        n_interior = int((~boundary).sum().item())
        self._order_cache = (plan, order, n_interior)
        return order, n_interior

    def _comm(self):
        """High-priority side stream: the pack / push / flag kernels (and the boundary tiles behind them) must get SM slots
        ahead of the queued CTAs of the ghost-free launch."""
        if getattr(self, "_comm_stream", None) is None:
            self._comm_stream = torch.cuda.Stream(self.state.device, priority=-1)
        return self._comm_stream

    def step(self, force, timestep):
        """One force evaluation of a decomposed system: the ghost refresh runs on a side stream while the tiles that read
        no ghost are evaluated; the tiles on the domain boundary follow once the halo has arrived. Without a tile plan:
        exchange, then evaluate."""
        force._prepare()
        plan = force._current_plan()
        if plan is None or not self.state.device.type == "cuda":
            self.update()
            force.compute(timestep)
            return
        order, n_interior = self.tile_order(plan)
        cur = torch.cuda.current_stream(self.state.device)
        comm = self._comm()
        if self.transport == "peer":
            s_cur, s_comm = C.c_void_p(cur.cuda_stream), C.c_void_p(comm.cuda_stream)
            self._peer_begin(s_cur)  # acknowledge: the previous step's kernels are done with the old ghosts
            comm.wait_stream(cur)    # ... and with the local rows the push reads
            force.compute(timestep, tile_range=(0, n_interior), tile_order=order)
            # the boundary tiles go on the (high-priority) side stream right behind the arrival wait: they start as soon as
            # the neighbours' rows have landed and share the GPU with the tail of the ghost-free launch
            with torch.cuda.stream(comm):
                self._peer_push(s_comm)
                self._peer_arrived(s_comm)
                force.compute(timestep, tile_range=(n_interior, plan.n_tiles - n_interior), tile_order=order)
            cur.wait_stream(comm)
            self._poll()
            return
        comm.wait_stream(cur)  # the previous step's boundary tiles still read the old ghosts
        # ghost-free tiles first: whatever the exchange does on the host side (torch's P2P wait may block), the GPU is busy
        force.compute(timestep, tile_range=(0, n_interior), tile_order=order)
        with torch.cuda.stream(comm):
            self.update()
        cur.wait_stream(comm)
        force.compute(timestep, tile_range=(n_interior, plan.n_tiles - n_interior), tile_order=order)

    def bytes_per_update(self):
        st = self.state
        s4 = 4 * (st.scalar_bits // 8)
        return sum(int(i.numel()) for i in self.send_idx) * s4 * len(self.fields)
