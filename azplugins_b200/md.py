"""The step around the force path: ``hoomd.md.Integrator`` with ``hoomd.md.methods.ConstantVolume`` (no
thermostat = NVE velocity Verlet), the integrator every reference test drives its forces with (test_pair.py:325-327,
test_pair_dpd.py:31-32, test_external.py:36-40), plus the neighbour-list rebuild trigger (``buffer / 2`` displacement).

Per step: ``azb_nve_step_one`` (half kick, drift, wrap), displacement check (+ list and tile-plan rebuild when a particle
has moved half the buffer), one launch per attached force at ``timestep + 1``, ``azb_nve_step_two`` (net-force sum over the
forces, acceleration, half kick) — all on the device; the host reads one flag word per step, as HOOMD does.
HOOMD-blue is not under /root/reference: the order of operations follows TwoStepConstantVolume / Integrator (recalled).
"""

import ctypes as C

import numpy as np
import torch

from . import _abi


class ConstantVolume:
    """``hoomd.md.methods.ConstantVolume(filter, thermostat=None)``: all particles, no thermostat."""

    def __init__(self, filter=None, thermostat=None):
        if thermostat is not None:
            raise NotImplementedError("only ConstantVolume without a thermostat (NVE) is implemented")
        if filter is not None and filter != "all":
            raise NotImplementedError("only the All filter is implemented")


class Integrator:
    """``hoomd.md.Integrator(dt, methods=[...], forces=[...])`` for one State."""

    def __init__(self, dt, methods=None, forces=None):
        self.dt = float(dt)
        self.methods = list(methods) if methods is not None else []
        self.forces = list(forces) if forces is not None else []
        self.timestep = 0
        self.launches = 0
        self.nlist_rebuilds = 0
        self._state = None

    def attach(self, state, timestep=0, domain=None):
        """``domain``: the ``domain.HaloExchange`` of a decomposed state (one rank of a multi-GPU run). Every list rebuild
        then migrates particles between the ranks and re-selects the ghost layer (HOOMD: Communicator::migrateParticles /
        exchangeGhosts before the neighbour list is rebuilt); between rebuilds the ghosts are refreshed every step."""
        if state.device.type != "cuda":
            raise RuntimeError("azplugins_b200 integrates on a CUDA device only (no CPU fallback)")
        if state.n_ghost and domain is None:
            raise ValueError("Integrator: a decomposed state (ghost particles) needs its HaloExchange passed as `domain`")
        self._lib = _abi.load()
        self._state = state
        self._domain = domain
        self.timestep = int(timestep)
        dev = state.device
        if getattr(state, "image", None) is None:
            state.image = torch.zeros((state.N, 3), dtype=torch.int32, device=dev)
        self._flag = torch.zeros(1, dtype=torch.int32, device=dev)
        for f in self.forces:
            if f._state is not state:
                f._attach(state)
            if hasattr(f, "dt"):
                f.dt = self.dt  # DPD reads the integrator's step size (HOOMD: m_deltaT)
        self._size_arrays()
        self._ready = False
        return self

    def _size_arrays(self):
        """Per-particle work arrays for the current number of local particles (it changes when particles migrate)."""
        st = self._state
        state_accel = getattr(st, "accel", None)
        if state_accel is None or state_accel.shape[0] != st.N:
            st.accel = torch.zeros((st.N, 4), dtype=st.pos.dtype, device=st.device)
        if getattr(self, "_net", None) is None or self._net.shape[0] != st.N:
            self._net = torch.zeros((st.N, 4), dtype=st.pos.dtype, device=st.device)
        for f in self.forces:
            f._sync_size()

    # ---- helpers ---------------------------------------------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self._state.device).cuda_stream)

    def _args(self, with_forces):
        st = self._state
        a = _abi.NveArgs()
        a.d_pos, a.d_vel, a.d_accel = st.pos.data_ptr(), st.vel.data_ptr(), st.accel.data_ptr()
        a.d_image = st.image.data_ptr()
        a.N = st.N
        a.box = st.box.as_abi()
        a.deltaT = self.dt
        if with_forces:
            self._ptrs = (C.c_void_p * max(len(self.forces), 1))(*[f._force.data_ptr() for f in self.forces])
            a.d_forces = C.cast(self._ptrs, C.c_void_p)
            a.n_forces = len(self.forces)
            a.d_net_force = self._net.data_ptr()
        return a

    def _nlists(self):
        out = []
        for f in self.forces:
            nl = getattr(f, "nlist", None)
            if nl is not None and nl not in out:
                out.append(nl)
        return out

    def _update_nlists(self):
        """NeighborList::distanceCheck: rebuild when any particle has moved buffer / 2 since the last build. In a decomposed run
        the decision is collective (one MAX all-reduce of the flag word): every rank migrates and re-selects ghosts together.
        Returns True when the lists were invalidated."""
        st, dom = self._state, self._domain
        rebuild = False
        lists = self._nlists()
        self._flag.zero_()
        for nl in lists:
            last = getattr(nl, "_last_pos", None)
            if nl.n_neigh is None or last is None or nl._state is not st or last.shape[0] != st.pos.shape[0]:
                continue  # the force builds it on first use
            box = st.box.as_abi()
            half = 0.5 * nl.buffer
            _abi.check(self._lib.azb_displacement_check(st.scalar_bits, st.pos.data_ptr(), last.data_ptr(), st.N, C.byref(box), half * half,
                                                        self._flag.data_ptr(), self._stream()), "azb_displacement_check")
            self.launches += 1
        if dom is not None and dom.decomp.world > 1:
            import torch.distributed as dist

            dist.all_reduce(self._flag, op=dist.ReduceOp.MAX, group=dom.group)
        if int(self._flag.item()) != 0:
            rebuild = True
            if dom is not None:
                dom.rebuild()
                self._size_arrays()
            for nl in lists:
                nl.invalidate()
            self.nlist_rebuilds += 1
        elif dom is not None:
            dom.update()
        return rebuild

    # ---- stepping --------------------------------------------------------------------------------------------------
    def run(self, steps):
        """``Simulation.run(steps)``: forces at the current timestep on the first call (Integrator::prepRun), then
        ``steps`` velocity-Verlet steps."""
        if self._state is None:
            raise RuntimeError("integrator is not attached to a state")
        if not any(isinstance(m, ConstantVolume) for m in self.methods):
            raise RuntimeError("Integrator needs a ConstantVolume method")
        st, lib, bits = self._state, self._lib, self._state.scalar_bits
        if not self._ready:
            if self._domain is not None:
                self._domain.update()
            for f in self.forces:
                f.compute(self.timestep)
            a = self._args(True)
            a.deltaT = 0.0  # a = F / m only
            _abi.check(lib.azb_nve_step_two(bits, C.byref(a), self._stream()), "azb_nve_step_two")
            self.launches += 1
            self._ready = True
        one, two = self._args(False), self._args(True)
        for _ in range(int(steps)):
            _abi.check(lib.azb_nve_step_one(bits, C.byref(one), self._stream()), "azb_nve_step_one")
            if self._update_nlists() and self._domain is not None:
                one, two = self._args(False), self._args(True)  # N and the work arrays changed with the migration
            for f in self.forces:
                f.compute(self.timestep + 1)
            _abi.check(lib.azb_nve_step_two(bits, C.byref(two), self._stream()), "azb_nve_step_two")
            self.launches += 2
            self.timestep += 1
        return self

    # ---- thermodynamic quantities (hoomd.md.compute.ThermodynamicQuantities) ------------------------------------------
    @property
    def kinetic_energy(self):
        v = self._state.vel[: self._state.N].double()
        return float((0.5 * v[:, 3] * (v[:, :3] ** 2).sum(dim=1)).sum().item())

    @property
    def potential_energy(self):
        return float(self._net[:, 3].double().sum().item())

    @property
    def linear_momentum(self):
        v = self._state.vel[: self._state.N].double()
        return (v[:, 3:4] * v[:, :3]).sum(dim=0).cpu().numpy()
