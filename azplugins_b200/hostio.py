"""One force evaluation driven from HOST buffers, pipelined inside the step.

The reference-facing call a host code makes per timestep is "here are this step's positions (+ velocities /
orientations), give me the forces". With the arrays in pinned host memory the step is PCIe-bound, so the three legs are
overlapped instead of run back to back:

  H2D   the inputs are copied in ascending index chunks on a copy stream;
  force one launch per input chunk, holding exactly the tiles whose inputs are complete once that chunk has arrived (the
        tile plan knows the highest particle index a tile reads: the last entry of its ascending distinct-neighbour
        list; ``azb_pair_args.d_tile_order`` + ``tile_begin / tile_count`` select them) — the tiles on the periodic
        boundary, whose neighbours sit at the far end of the index range, fall into the last launch;
  D2H   blocks of output rows are copied back on a third stream as soon as the last launch writing into them is done,
        while later launches compute.

Neighbour list, tile plan, tags and parameter tables stay resident between list rebuilds, as they do in HOOMD.
"""

import ctypes as C

import numpy as np
import torch

from . import _abi


def chunk_edges(n_rows, chunks, taper=(0.08, 0.04, 0.02)):
    """Row boundaries of the input chunks. What a step pays after its last input byte is the last launch and the copy of that
    launch's output rows, and the last launch holds every tile whose highest input row sits in the LAST chunk: that chunk's own
    tiles, the layer of tiles below it and the tiles on the periodic boundary. So the chunks shrink towards the end — the last
    ``len(taper)`` chunks hold these fractions of the rows, the others share the rest equally — and the tail shrinks with them
    (8 equal chunks: 22 % of the tiles in the last launch; tapered: ~9 %). ``taper=()`` gives equal chunks."""
    K = max(1, int(chunks))
    taper = tuple(taper)[-(K - 1):] if K > 1 else ()
    if len(taper) and K - len(taper) >= 1:
        head = np.linspace(0.0, 1.0 - float(sum(taper)), K - len(taper) + 1)
        frac = np.concatenate([head, head[-1] + np.cumsum(taper)])
        frac[-1] = 1.0
    else:
        frac = np.linspace(0.0, 1.0, K + 1)
    edges = np.round(frac * n_rows).astype(np.int64)
    edges[0], edges[-1] = 0, n_rows
    return np.maximum.accumulate(edges)


def plan_schedule(hi, in_edges, tile_size, N, out_blocks_per_chunk=8, late=None):
    """Pure host logic of the pipelined step (no device access; tests/test_hostio_schedule.py).

    ``hi[t]`` is the highest particle index tile t reads, ``in_edges`` the ascending row boundaries of the K input chunks.
    A tile is ready when the input chunk holding ``hi[t]`` has arrived: launch e holds exactly the tiles that become ready
    with chunk e (ascending tile index inside) through the launch-position -> tile map ``order``; tiles on the periodic
    boundary, whose neighbours sit at the far end of the index range, fall into the last launch. Outputs go back in blocks
    of consecutive tiles, each after the last launch that writes into it; neighbouring blocks finished by the same launch
    travel as one copy. Returns ``(order, launches, outs)`` with launches = [(wait_chunk, first_position, n_tiles)] and
    outs = [(launch_index, row_lo, row_hi)] sorted by launch index.

    ``late`` (bool per tile): tiles that cannot run from the host inputs alone — in a decomposed run, the tiles that read ghost
    particles. They go to the END of ``order`` (positions ``[n_tiles - n_late, n_tiles)``), belong to no launch, and the output
    blocks they touch are returned separately: the result is then ``(order, launches, outs, late_outs)`` with late_outs =
    [(row_lo, row_hi)]."""
    hi = np.asarray(hi, dtype=np.int64)
    n_tiles = len(hi)
    K = len(in_edges) - 1
    ready = np.clip(np.searchsorted(in_edges, hi, side="right") - 1, 0, K - 1)
    if late is not None:
        ready = np.where(np.asarray(late, dtype=bool), K, ready)
    order = np.argsort(ready, kind="stable")
    per_class = np.bincount(ready, minlength=K + 1)
    sched, launch_of_class, at = [], {}, 0
    for e in range(K):
        if per_class[e]:
            launch_of_class[e] = len(sched)
            sched.append((e, at, int(per_class[e])))
            at += int(per_class[e])
    B = max(1, min(out_blocks_per_chunk * K, n_tiles))
    tile_edges = np.linspace(0, n_tiles, B + 1).astype(np.int64)
    outs, late_outs = [], []
    for b in range(B):
        tb, te = int(tile_edges[b]), int(tile_edges[b + 1])
        if te > tb:
            cls = int(ready[tb:te].max())
            if cls == K:
                late_outs.append((tb * tile_size, min(te * tile_size, N)))
            else:
                outs.append((launch_of_class[cls], tb * tile_size, min(te * tile_size, N)))
    outs.sort(key=lambda x: (x[0], x[1]))
    merged = []
    for o in outs:
        if merged and merged[-1][0] == o[0] and merged[-1][2] == o[1]:
            merged[-1] = (o[0], merged[-1][1], o[2])
        else:
            merged.append(o)
    if late is None:
        return order, sched, merged
    late_merged = []
    for lo, up in late_outs:
        if late_merged and late_merged[-1][1] == lo:
            late_merged[-1] = (late_merged[-1][0], up)
        else:
            late_merged.append((lo, up))
    return order, sched, merged, late_merged


class HostStep:
    """``HostStep(force_or_forces, chunks)``: several tile-staged force objects over the same particles and the same neighbour
    list (C5: DPD + ExpandedYukawa) are evaluated by the same launch ranges of one ``azb_host_step`` call."""

    def __init__(self, force, chunks=8, halo=None, taper=(0.08, 0.04, 0.02)):
        """``halo``: the ``domain.HaloExchange`` of a decomposed state. The step then pipelines across the ghost exchange too:
        the tiles that read no ghost run as their host inputs arrive, the halo waits only for the last input chunk, and the
        tiles on the domain boundary (and the output rows they own) follow behind it on a side stream."""
        forces = list(force) if isinstance(force, (list, tuple)) else [force]
        self.halo = halo
        if not forces or any(f._state is None for f in forces):
            raise RuntimeError("force is not attached to a state")
        if any(f._state is not forces[0]._state for f in forces):
            raise ValueError("HostStep: every force object must be attached to the same state")
        self.forces = forces
        self.force = forces[0]
        self.chunks = int(chunks)
        self.taper = tuple(taper)  # chunk_edges: the last chunks' shares of the rows; () = equal chunks
        st = self.force._state
        dev = st.device
        self.s_in = torch.cuda.Stream(dev)
        self.s_out = torch.cuda.Stream(dev)
        self._input_names = ["pos"] + (["vel"] if any(f._kind == _abi.KIND_DPD for f in forces) else []) + (
            ["orientation"] if any(f._kind == _abi.KIND_ANISO for f in forces) else [])
        self.outputs = []
        for f in forces:
            self.outputs.append(f._force)
            if f._torque is not None:
                self.outputs.append(f._torque)
        if len(self.outputs) > 8:
            raise ValueError("HostStep: at most 8 output arrays")
        self._plans = None
        self._schedule = None

    @property
    def inputs(self):
        """The local rows of the device arrays the host writes every step (ghost rows are the halo's)."""
        st = self.force._state
        return [getattr(st, name)[: st.N] for name in self._input_names]

    # bytes moved per step (the bench reports them)
    @property
    def h2d_bytes(self):
        return sum(t.numel() * t.element_size() for t in self.inputs)

    @property
    def d2h_bytes(self):
        return sum(t.numel() * t.element_size() for t in self.outputs)

    def pinned_inputs(self):
        return [torch.empty(t.shape, dtype=t.dtype, pin_memory=True).copy_(t.cpu()) for t in self.inputs]

    def pinned_outputs(self):
        return [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in self.outputs]

    def _current_plans(self):
        plans = [f._current_plan() for f in self.forces]
        if any(p is None for p in plans):
            return None
        if any(p.tile_size != plans[0].tile_size or p.n_tiles != plans[0].n_tiles for p in plans):
            return None  # different tilings: no common launch ranges
        return plans

    @staticmethod
    def _tile_hi(plan, N):
        """Highest particle index a tile reads: its own last particle, or the largest index it stages."""
        hi = plan.tile_max().cpu().numpy().astype(np.int64)
        own_hi = np.minimum((np.arange(plan.n_tiles) + 1) * plan.tile_size, N) - 1
        return np.maximum(hi, own_hi)

    def _make_schedule(self, plans):
        """Which tiles each launch holds (by the input chunk that completes their inputs), and which launch completes
        each block of output rows."""
        f, st = self.force, self.force._state
        self._plans = plans
        if plans is None:
            self._schedule = None
            return
        plan = plans[0]
        K = max(1, min(self.chunks, plan.n_tiles))
        in_edges = chunk_edges(st.N, K, self.taper)
        hi = self._tile_hi(plan, st.N)
        for p in plans[1:]:
            if p.uniq is not plan.uniq:
                hi = np.maximum(hi, self._tile_hi(p, st.N))
        late_outs, n_late, d_order_holder = [], 0, [None]
        if st.n_ghost:
            if self.halo is None:
                raise RuntimeError("HostStep: a decomposed state (ghost particles) needs its HaloExchange passed as `halo`")
            late = hi >= st.N  # the tile reads a ghost row
            n_late = int(late.sum())
            order, sched, merged, late_outs = plan_schedule(hi, in_edges, plan.tile_size, st.N, late=late)
        else:
            order, sched, merged = plan_schedule(hi, in_edges, plan.tile_size, st.N)
        self._late = (n_late, late_outs, d_order_holder)
        # the native call's tables, built once per plan: a step only patches the timestep and the host pointers
        d_order = torch.from_numpy(order.astype(np.int32)).to(st.device)
        d_order_holder[0] = d_order
        wholes = [g._build_args(0, tile_order=d_order) for g in self.forces]
        hs = _abi.HostStepArgs()
        hs.kind = f._kind
        hs.potential = f._spec.potential_id
        hs.scalar_bits = st.scalar_bits
        hs.force_args = C.addressof(wholes[0])
        hs.d_params = f._d_params.data_ptr()
        more = (_abi.HostForce * max(len(self.forces) - 1, 1))()
        for k, g in enumerate(self.forces[1:]):
            more[k].kind, more[k].potential = g._kind, g._spec.potential_id
            more[k].force_args, more[k].d_params = C.addressof(wholes[1 + k]), g._d_params.data_ptr()
        hs.n_more, hs.more = len(self.forces) - 1, more
        hs.n_inputs, hs.n_outputs = len(self.inputs), len(self.outputs)
        for k, t in enumerate(self.inputs):
            hs.inputs[k].d_ptr, hs.inputs[k].row_bytes = t.data_ptr(), t.shape[1] * t.element_size()
        for k, t in enumerate(self.outputs):
            hs.outputs[k].d_ptr, hs.outputs[k].row_bytes = t.data_ptr(), t.shape[1] * t.element_size()
        edges = (C.c_size_t * (K + 1))(*[int(e) for e in in_edges])
        launches = (_abi.HostLaunch * len(sched))()
        for k, (wait_chunk, tb, tn) in enumerate(sched):
            launches[k].wait_chunk, launches[k].tile_begin, launches[k].tile_count = wait_chunk, tb, tn
        out_ranges = (_abi.HostOutRange * len(merged))()
        for k, (after, lo, up) in enumerate(merged):
            out_ranges[k].after_launch, out_ranges[k].row_lo, out_ranges[k].row_hi = after, lo, up
        hs.flags = _abi.HOST_STEP_NO_SYNC if n_late else 0
        hs.n_in_chunks, hs.in_edges = K, edges
        hs.n_launches, hs.launches = len(sched), launches
        hs.n_outs, hs.outs = len(merged), out_ranges
        hs.stream_in, hs.stream_out = self.s_in.cuda_stream, self.s_out.cuda_stream
        # everything the native call dereferences stays referenced from the schedule: the ctypes tables, the launch order, and
        # the plans + list arrays the argument structs point into (a rebuilt list makes new plan objects -> new schedule)
        nl = f.nlist
        keep = (edges, launches, out_ranges, d_order, more, plans, nl.n_neigh, nl.head_list, nl.nlist)
        self._schedule = (hs, wholes, keep, merged, sched)

    def run(self, timestep, host_in, host_out):
        """host_in / host_out: lists of pinned host tensors matching ``inputs`` / ``outputs``. Returns after the forces are
        in ``host_out`` (the host consumes them every step). With a tile plan this is ONE call into the C-ABI
        (``azb_host_step``, csrc/azb_hostio.cu), which issues the chunked copies and the tile-range launches itself."""
        f = self.force
        dev = f._state.device
        cur = torch.cuda.current_stream(dev)
        for g in self.forces:
            g._prepare()
        plans = self._current_plans()
        stale = self._plans is None or plans is None or len(plans) != len(self._plans) or any(a is not b for a, b in zip(plans, self._plans))
        if self._schedule is None or stale:
            self._make_schedule(plans)
        if self._schedule is None:
            # no tile plan (e.g. a kernel family without a tiled flavour): plain copy - (exchange) - compute - copy
            for dst, src in zip(self.inputs, host_in):
                dst.copy_(src, non_blocking=True)
            if self.halo is not None:
                self.halo.update()
            for g in self.forces:
                g.compute(timestep)
            for dst, src in zip(host_out, self.outputs):
                dst.copy_(src, non_blocking=True)
            cur.synchronize()
            return
        hs, wholes, _, _, sched = self._schedule
        for g, whole in zip(self.forces, wholes):
            g._set_timestep(whole, timestep)
        for k, t in enumerate(host_in):
            hs.inputs[k].h_ptr = t.data_ptr()
        for k, t in enumerate(host_out):
            hs.outputs[k].h_ptr = t.data_ptr()
        hs.stream = cur.cuda_stream
        n_late, late_outs, order = self._late[0], self._late[1], self._late[2][0]
        halo = self.halo
        if n_late and halo.transport == "peer":
            halo._peer_begin(C.c_void_p(cur.cuda_stream))  # acknowledge: the previous step's kernels are done with the old ghosts
        _abi.check(f._lib.azb_host_step(C.byref(hs)), "azb_host_step")
        for g in self.forces:
            g.launches += len(sched)
        if not n_late:
            return
        # ---- behind the (unsynchronised) native call: ghost exchange, the tiles that read ghosts, their output rows -----------
        plan = self._plans[0]
        comm = halo._comm()
        inputs_in = torch.cuda.Event()
        inputs_in.record(self.s_in)  # every input chunk of this step has been queued on s_in before this point
        comm.wait_event(inputs_in)
        with torch.cuda.stream(comm):
            if halo.transport == "peer":
                s_comm = C.c_void_p(comm.cuda_stream)
                halo._peer_push(s_comm)
                halo._peer_arrived(s_comm)
            else:
                halo.update()
            for g in self.forces:
                g.compute(timestep, tile_range=(plan.n_tiles - n_late, n_late), tile_order=order)
            boundary_done = torch.cuda.Event()
            boundary_done.record(comm)
        self.s_out.wait_event(boundary_done)
        with torch.cuda.stream(self.s_out):
            for lo, up in late_outs:
                for dst, src in zip(host_out, self.outputs):
                    dst[lo:up].copy_(src[lo:up], non_blocking=True)
        cur.wait_stream(comm)
        self.s_out.synchronize()
        cur.synchronize()
        halo._poll()

    def release(self):
        """Drop the native call's per-thread event pool (``azb_host_step_release``) and the cached schedule."""
        self._schedule = self._plans = None
        self.force._lib.azb_host_step_release()
