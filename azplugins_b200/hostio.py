"""One force evaluation driven from HOST buffers, pipelined inside the step.

The reference-facing call a host code makes per timestep is "here are this step's positions (+ velocities /
orientations), give me the forces". With the arrays in pinned host memory the step is PCIe-bound, so the three legs are
overlapped instead of run back to back:

  H2D   the inputs are copied in ascending index chunks on a copy stream;
  force the tile kernel is launched chunk of tiles by chunk of tiles (``azb_pair_args.tile_begin / tile_count``), each
        launch waiting only for the input chunk that holds the highest particle index its tiles read (the tile plan knows
        it: the last entry of each tile's ascending distinct-neighbour list). Chunks are launched in the order their
        inputs complete — the tiles on the periodic boundary, whose neighbours sit at the far end of the index range, go
        last;
  D2H   each chunk's forces are copied back on a third stream while later chunks compute.

Neighbour list, tile plan, tags and parameter tables stay resident between list rebuilds, as they do in HOOMD.
"""

import numpy as np
import torch

from . import _abi


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
        """Tile chunks, the input chunk each one has to wait for, and the launch order (by input arrival)."""
        f, st = self.force, self.force._state
        f.nlist.compute(st)
        plan = f._current_plan()
        if plan is None:
            self._schedule = None
            return
        K = max(1, min(self.chunks, plan.n_tiles))
        n_rows = st.pos.shape[0]
        in_edges = np.linspace(0, n_rows, K + 1).astype(np.int64)
        tile_edges = np.linspace(0, plan.n_tiles, K + 1).astype(np.int64)
        count = plan.uniq_count.cpu().numpy().astype(np.int64)
        # highest index a tile reads: its own last particle, or the last (largest) entry of its ascending uniq list
        last = torch.arange(plan.n_tiles, device=plan.uniq.device, dtype=torch.int64) * plan.uniq_capacity + torch.from_numpy(
            np.maximum(count - 1, 0)).to(plan.uniq.device)
        hi = plan.uniq[last].cpu().numpy().astype(np.int64)
        hi = np.where(count > 0, hi, 0)
        own_hi = np.minimum((np.arange(plan.n_tiles) + 1) * plan.tile_size, st.N) - 1
        hi = np.maximum(hi, own_hi)
        sched = []
        for c in range(K):
            tb, te = int(tile_edges[c]), int(tile_edges[c + 1])
            if te <= tb:
                continue
            need = int(hi[tb:te].max())
            wait_chunk = int(np.searchsorted(in_edges, need, side="right") - 1)
            lo, up = tb * plan.tile_size, min(te * plan.tile_size, st.N)
            sched.append((wait_chunk, tb, te - tb, lo, up))
        sched.sort(key=lambda x: (x[0], x[1]))
        # one argument struct per chunk, built once per plan: a step only patches the timestep into them
        args = [f._build_args(0, tile_range=(tb, tn)) for _, tb, tn, _, _ in sched]
        self._schedule = (in_edges, sched, args)
        self._plan_id = id(plan)

    def run(self, timestep, host_in, host_out):
        """host_in / host_out: lists of pinned host tensors matching ``inputs`` / ``outputs``. Returns after the forces are
        in ``host_out`` (the host consumes them every step)."""
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
        in_edges, sched, chunk_args = self._schedule
        self.s_in.wait_stream(cur)
        arrived = []
        with torch.cuda.stream(self.s_in):
            for c in range(len(in_edges) - 1):
                lo, up = int(in_edges[c]), int(in_edges[c + 1])
                for dst, src in zip(self.inputs, host_in):
                    dst[lo:up].copy_(src[lo:up], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.s_in)
                arrived.append(ev)
        for (wait_chunk, tb, tn, lo, up), a in zip(sched, chunk_args):
            cur.wait_event(arrived[wait_chunk])
            f._set_timestep(a, timestep)
            f._launch(a)
            done = torch.cuda.Event()
            done.record(cur)
            self.s_out.wait_event(done)
            with torch.cuda.stream(self.s_out):
                for dst, src in zip(host_out, self.outputs):
                    dst[lo:up].copy_(src[lo:up], non_blocking=True)
        cur.wait_stream(self.s_in)
        self.s_out.synchronize()
        cur.synchronize()
