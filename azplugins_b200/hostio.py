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


def plan_schedule(hi, in_edges, tile_size, N, out_blocks_per_chunk=8):
    """Pure host logic of the pipelined step (no device access; tests/test_hostio_schedule.py).

    ``hi[t]`` is the highest particle index tile t reads, ``in_edges`` the ascending row boundaries of the K input chunks.
    A tile is ready when the input chunk holding ``hi[t]`` has arrived: launch e holds exactly the tiles that become ready
    with chunk e (ascending tile index inside) through the launch-position -> tile map ``order``; tiles on the periodic
    boundary, whose neighbours sit at the far end of the index range, fall into the last launch. Outputs go back in blocks
    of consecutive tiles, each after the last launch that writes into it; neighbouring blocks finished by the same launch
    travel as one copy. Returns ``(order, launches, outs)`` with launches = [(wait_chunk, first_position, n_tiles)] and
    outs = [(launch_index, row_lo, row_hi)] sorted by launch index."""
    hi = np.asarray(hi, dtype=np.int64)
    n_tiles = len(hi)
    K = len(in_edges) - 1
    ready = np.clip(np.searchsorted(in_edges, hi, side="right") - 1, 0, K - 1)
    order = np.argsort(ready, kind="stable")
    per_class = np.bincount(ready, minlength=K)
    sched, launch_of_class, at = [], {}, 0
    for e in range(K):
        if per_class[e]:
            launch_of_class[e] = len(sched)
            sched.append((e, at, int(per_class[e])))
            at += int(per_class[e])
    B = max(1, min(out_blocks_per_chunk * K, n_tiles))
    tile_edges = np.linspace(0, n_tiles, B + 1).astype(np.int64)
    outs = []
    for b in range(B):
        tb, te = int(tile_edges[b]), int(tile_edges[b + 1])
        if te > tb:
            outs.append((launch_of_class[int(ready[tb:te].max())], tb * tile_size, min(te * tile_size, N)))
    outs.sort(key=lambda x: (x[0], x[1]))
    merged = []
    for o in outs:
        if merged and merged[-1][0] == o[0] and merged[-1][2] == o[1]:
            merged[-1] = (o[0], merged[-1][1], o[2])
        else:
            merged.append(o)
    return order, sched, merged


class HostStep:
    def __init__(self, force, chunks=8):
        if force._state is None:
            raise RuntimeError("force is not attached to a state")
        self.force = force
        self.chunks = int(chunks)
        st = force._state
        dev = st.device
        self.s_in = torch.cuda.Stream(dev)
        self.s_out = torch.cuda.Stream(dev)
        self.inputs = [st.pos]
        if force._kind == _abi.KIND_DPD:
            self.inputs.append(st.vel)
        if force._kind == _abi.KIND_ANISO:
            self.inputs.append(st.orientation)
        self.outputs = [force._force] + ([force._torque] if force._torque is not None else [])
        self._plan_id = None
        self._schedule = None

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

    def _make_schedule(self):
        """Which tiles each launch holds (by the input chunk that completes their inputs), and which launch completes
        each block of output rows."""
        f, st = self.force, self.force._state
        f.nlist.compute(st)
        plan = f._current_plan()
        if plan is None:
            self._schedule = None
            return
        K = max(1, min(self.chunks, plan.n_tiles))
        n_rows = st.pos.shape[0]
        in_edges = np.linspace(0, n_rows, K + 1).astype(np.int64)
        count = plan.uniq_count.cpu().numpy().astype(np.int64)
        # highest index a tile reads: its own last particle, or the last (largest) entry of its ascending uniq list
        last = torch.arange(plan.n_tiles, device=plan.uniq.device, dtype=torch.int64) * plan.uniq_capacity + torch.from_numpy(
            np.maximum(count - 1, 0)).to(plan.uniq.device)
        hi = plan.uniq[last].cpu().numpy().astype(np.int64)
        hi = np.where(count > 0, hi, 0)
        own_hi = np.minimum((np.arange(plan.n_tiles) + 1) * plan.tile_size, st.N) - 1
        hi = np.maximum(hi, own_hi)
        order, sched, merged = plan_schedule(hi, in_edges, plan.tile_size, st.N)
        # the native call's tables, built once per plan: a step only patches the timestep and the host pointers
        d_order = torch.from_numpy(order.astype(np.int32)).to(st.device)
        whole = f._build_args(0, tile_order=d_order)
        hs = _abi.HostStepArgs()
        hs.kind = f._kind
        hs.potential = f._spec.potential_id
        hs.scalar_bits = st.scalar_bits
        hs.force_args = C.addressof(whole)
        hs.d_params = f._d_params.data_ptr()
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
        hs.n_in_chunks, hs.in_edges = K, edges
        hs.n_launches, hs.launches = len(sched), launches
        hs.n_outs, hs.outs = len(merged), out_ranges
        hs.stream_in, hs.stream_out = self.s_in.cuda_stream, self.s_out.cuda_stream
        self._schedule = (hs, whole, (edges, launches, out_ranges, d_order), merged, sched)  # the ctypes objects must outlive the calls
        self._plan_id = id(plan)

    def run(self, timestep, host_in, host_out):
        """host_in / host_out: lists of pinned host tensors matching ``inputs`` / ``outputs``. Returns after the forces are
        in ``host_out`` (the host consumes them every step). With a tile plan this is ONE call into the C-ABI
        (``azb_host_step``, csrc/azb_hostio.cu), which issues the chunked copies and the tile-range launches itself."""
        f = self.force
        dev = f._state.device
        cur = torch.cuda.current_stream(dev)
        f._prepare()
        plan = f._current_plan()
        if self._schedule is None or plan is None or id(plan) != self._plan_id:
            self._make_schedule()
        if self._schedule is None:
            # no tile plan (e.g. a kernel family without a tiled flavour): plain copy - compute - copy
            for dst, src in zip(self.inputs, host_in):
                dst.copy_(src, non_blocking=True)
            f.compute(timestep)
            for dst, src in zip(host_out, self.outputs):
                dst.copy_(src, non_blocking=True)
            cur.synchronize()
            return
        hs, whole, _, _, sched = self._schedule
        f._set_timestep(whole, timestep)
        for k, t in enumerate(host_in):
            hs.inputs[k].h_ptr = t.data_ptr()
        for k, t in enumerate(host_out):
            hs.outputs[k].h_ptr = t.data_ptr()
        hs.stream = cur.cuda_stream
        _abi.check(f._lib.azb_host_step(C.byref(hs)), "azb_host_step")
        f.launches += len(sched)
