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
    """Ghost bookkeeping of one rank: builds the send lists, appends ghost storage to the State, refreshes it per step."""

    def __init__(self, decomp, state, r_ghost, fields=("pos",), group=None, transport=None):
        self.decomp, self.state, self.group = decomp, state, group
        self.fields = tuple(fields)
        self.transport = "nccl"  # "peer" once the neighbours' arrays are mapped (below)
        self.timeout_ms = 30000
        self.r_ghost = float(r_ghost)
        self.launches = 0  # pack kernels issued (azb_gather_scalar4)
        if np.any(decomp.width[np.array(decomp.grid) > 1] < 2 * r_ghost):
            raise ValueError("sub-domain narrower than 2 r_ghost")
        dev = state.device
        ftype = scalar_dtype(state.scalar_bits)[0]
        pos_h = state.pos[: state.N, :3].double().cpu().numpy()
        self.dirs = decomp.active_directions()
        self.send_idx, self.send_shift, self.peers_send, self.peers_recv = [], [], [], []
        for d in self.dirs:
            peer, shift = decomp.neighbor(d)
            idx = np.nonzero(decomp.send_mask(pos_h, d, self.r_ghost))[0].astype(np.int32)
            sh = np.zeros((len(idx), 4), dtype=ftype)
            sh[:, :3] = shift
            self.send_idx.append(torch.from_numpy(idx).to(dev))
            self.send_shift.append(torch.from_numpy(sh).to(dev))
            self.peers_send.append(peer)
            self.peers_recv.append(decomp.neighbor(tuple(-x for x in d))[0])
        # counts: what I receive for direction k is what the rank behind me (direction -d_k) selected for ITS direction k
        mine = [len(i) for i in self.send_idx]
        if decomp.world > 1:
            everyone = [None] * decomp.world
            dist.all_gather_object(everyone, mine, group=group)
        else:
            everyone = [mine]
        recv_counts = [everyone[self.peers_recv[k]][k] for k in range(len(self.dirs))]
        self.recv_counts = [int(c) for c in recv_counts]
        self.recv_offsets = np.concatenate([[0], np.cumsum(self.recv_counts)]).astype(np.int64)
        n_ghost = int(self.recv_offsets[-1])
        # grow the state arrays: locals first, ghosts behind them
        N = state.N
        for name in ("pos", "vel", "orientation"):
            t = getattr(state, name)
            grown = torch.zeros((N + n_ghost, 4), dtype=t.dtype, device=dev)
            grown[:N] = t[:N]
            setattr(state, name, grown)
        tag = torch.zeros(N + n_ghost, dtype=state.tag.dtype, device=dev)
        tag[:N] = state.tag[:N]
        state.tag = tag
        state.n_ghost = n_ghost
        state.global_box = decomp.global_box  # what external.HarmonicBarrier wraps positions into
        state.box = decomp.local_box()
        self._bufs = {}
        # static data once: tags (as a 1-column exchange) — types travel inside pos.w with every position update
        sends, recvs = [], []
        for k in range(len(self.dirs)):
            sends.append((state.tag[self.send_idx[k].long()].contiguous(), self.peers_send[k]))
            o0, o1 = N + int(self.recv_offsets[k]), N + int(self.recv_offsets[k + 1])
            recvs.append((state.tag[o0:o1], self.peers_recv[k]))
        self._p2p(sends, recvs)
        self.update(all_fields=True)
        want = transport or os.environ.get("AZB_HALO_TRANSPORT", "auto")
        if want not in ("auto", "peer", "nccl"):
            raise ValueError("transport must be 'auto', 'peer' or 'nccl'")
        if want != "nccl" and decomp.world > 1 and dev.type == "cuda":
            self._setup_peer(required=(want == "peer"))

    # ---- ghost refresh over peer memory: one push kernel per step instead of pack kernels + grouped send/recv ---------
    def _setup_peer(self, required):
        """Map the neighbours' particle arrays and flag words into this process (CUDA IPC over NVLink) and build the route
        table of ``azb_halo_push``. Every rank decides together: if any mapping fails, all stay on the NCCL exchange."""
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

        mine = {"N": st.N, "recv_offsets": [int(x) for x in self.recv_offsets], "sync": export(self._sync),
                "fields": {name: export(getattr(st, name)) for name in ("pos", "vel", "orientation")}}
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

        row_bytes = 4 * (st.scalar_bits // 8)
        routes, data_flags, ack_flags = [], [], []
        for k, d in enumerate(self.dirs):
            to, frm = self.peers_send[k], self.peers_recv[k]
            # the rank in direction d stores what I send for direction k in ITS slot k; its data flag k is mine to raise
            base = open_(to, everyone[to]["sync"])
            data_flags.append((self._sync.data_ptr() if base is None else base) + 4 * k)
            # the rank behind me filled my slot k: its acknowledgement word k is mine to raise once I have read the ghosts
            base = open_(frm, everyone[frm]["sync"])
            ack_flags.append((self._sync.data_ptr() if base is None else base) + 4 * (nd + k))
            first = everyone[to]["N"] + everyone[to]["recv_offsets"][k]
            for name in self.fields:
                base = open_(to, everyone[to]["fields"][name])
                if base is None:
                    base = getattr(st, name).data_ptr()
                shift = self.decomp.neighbor(d)[1] if name == "pos" else np.zeros(3)
                routes.append((getattr(st, name).data_ptr(), base + first * row_bytes, self.send_idx[k].data_ptr(),
                               int(self.send_idx[k].numel()), shift))
        flag = torch.tensor([0 if ok else 1], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, group=self.group)
        if int(flag.item()) != 0:
            if required:
                raise RuntimeError("halo transport 'peer': the neighbours' memory cannot be mapped (no P2P access between the GPUs?)")
            return
        table = (_abi.HaloRoute * max(len(routes), 1))()
        begin = 0
        for r, (src, dst, idx, n, shift) in zip(table, routes):
            r.d_src, r.d_dst, r.d_idx, r.n, r.row_begin = src, dst, idx, n, begin
            r.shift[0], r.shift[1], r.shift[2] = (float(x) for x in shift)
            begin += n
        self._routes = torch.frombuffer(bytearray(bytes(table)), dtype=torch.uint8).to(dev)
        self._data_ptrs = torch.tensor(data_flags, dtype=torch.int64, device=dev)
        self._ack_ptrs = torch.tensor(ack_flags, dtype=torch.int64, device=dev)
        a = _abi.HaloPushArgs()
        a.d_routes, a.n_routes, a.total_rows = self._routes.data_ptr(), len(routes), begin
        a.d_peer_flags, a.n_flags = self._data_ptrs.data_ptr(), nd
        a.d_counter = self._sync.data_ptr() + 4 * (2 * nd)
        self._push_args, self._epoch = a, 0
        self._d_error = self._sync.data_ptr() + 4 * (2 * nd + 1)
        dist.barrier(group=self.group)  # nobody raises a flag before everybody's words exist
        self.transport = "peer"

    def _peer_begin(self, stream):
        """Start of a refresh on ``stream``: whatever was enqueued before has finished reading the previous ghosts."""
        lib, nd = _abi.load(), len(self.dirs)
        self._epoch = (self._epoch + 1) & 0x7FFFFFFF
        _abi.check(lib.azb_halo_signal(self._ack_ptrs.data_ptr(), nd, self._epoch - 1, stream), "azb_halo_signal")
        self.launches += 1

    def _peer_push(self, stream):
        """Wait until the receivers have released their ghost rows, then store this step's rows into them."""
        lib, nd = _abi.load(), len(self.dirs)
        _abi.check(lib.azb_halo_wait(self._sync.data_ptr() + 4 * nd, nd, self._epoch - 1, self.timeout_ms, self._d_error, stream), "azb_halo_wait")
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

    def update(self, all_fields=False):
        """Refresh ghost copies of the dynamic fields (every step, before the force kernels)."""
        st = self.state
        N = st.N
        if self.transport == "peer" and not all_fields:
            stream = C.c_void_p(torch.cuda.current_stream(st.device).cuda_stream)
            self._peer_begin(stream)
            self._peer_push(stream)
            self._peer_arrived(stream)
            return
        fields = ("pos", "vel", "orientation") if all_fields else self.fields
        sends, recvs = [], []
        for name in fields:
            arr = getattr(st, name)
            for k in range(len(self.dirs)):
                n = self.send_idx[k].numel()
                key = (name, k)
                if key not in self._bufs:
                    self._bufs[key] = torch.empty((n, 4), dtype=arr.dtype, device=arr.device)
                buf = self._bufs[key]
                _gather_rows(buf, arr, self.send_idx[k], self.send_shift[k] if name == "pos" else None)
                self.launches += 1 if (n and arr.is_cuda) else 0
                sends.append((buf, self.peers_send[k]))
                o0, o1 = N + int(self.recv_offsets[k]), N + int(self.recv_offsets[k + 1])
                recvs.append((arr[o0:o1], self.peers_recv[k]))  # contiguous slice: received in place, no unpack kernel
        self._p2p(sends, recvs)

    # ---- ghost exchange overlapped with the ghost-free part of the force evaluation ---------------------------------
    def tile_order(self, plan):
        """(order, n_interior): launch order of the plan's tiles with the ghost-free ones first. A tile reads a ghost iff
        the last (largest) entry of its ascending distinct-neighbour list is >= N. Cached per plan."""
        cached = getattr(self, "_order_cache", None)
        if cached is not None and cached[0] is plan:
            return cached[1], cached[2]
        N = self.state.N
        count = plan.uniq_count.long()
        t = torch.arange(plan.n_tiles, device=count.device, dtype=torch.int64)
        last = plan.uniq[t * plan.uniq_capacity + torch.clamp(count - 1, min=0)].long()
        boundary = (count > 0) & (last >= N)
        order = torch.cat([torch.nonzero(~boundary).flatten(), torch.nonzero(boundary).flatten()]).to(torch.int32).contiguous()
        n_interior = int((~boundary).sum().item())
        self._order_cache = (plan, order, n_interior)
        return order, n_interior

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
        if getattr(self, "_comm_stream", None) is None:
            # high priority: the pack and NCCL kernels must get SM slots ahead of the queued CTAs of the force kernel
            self._comm_stream = torch.cuda.Stream(self.state.device, priority=-1)
        comm = self._comm_stream
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
