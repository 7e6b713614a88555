"""Host classes that call the C-ABI — the B200 stand-ins for HOOMD's ``PotentialPairGPU<E>``,
``PotentialPairDPDThermoGPU<E>``, ``AnisoPotentialPairGPU<E>`` and ``PotentialBondGPU<E, BondData>`` (the templates
the reference instantiates in /root/reference/src/export_*GPU.cc.inc) plus the slice of ``hoomd.md.pair.Pair`` /
``hoomd.md.bond.Bond`` their Python classes inherit: ``params`` / ``r_cut`` / ``r_on`` type parameters, ``mode``,
``forces`` / ``energies`` / ``torques`` / ``virials`` / ``energy``.

Per ``compute(timestep)`` exactly one kernel launch is issued through ``azb_*_forces``; outputs are overwritten.
There is no CPU path: a state that does not live on a CUDA device raises.
"""

import ctypes as C
import itertools
import os

import numpy as np
import torch

from . import _abi
from .state import scalar_dtype


class TypeParameter(dict):
    """``params[(a, b)] = dict(...)`` store with hoomd's symmetric pair keys and schema validation."""

    def __init__(self, owner, validate, pairwise):
        super().__init__()
        self._owner, self._validate, self._pairwise = owner, validate, pairwise

    def _keys(self, key):
        if not self._pairwise:
            return [key]
        a, b = key
        a_list = [a] if isinstance(a, str) else list(a)
        b_list = [b] if isinstance(b, str) else list(b)
        return [tuple(sorted(k)) for k in itertools.product(a_list, b_list)]

    def __setitem__(self, key, value):
        for k in self._keys(key):
            super().__setitem__(k, self._validate(value))
        self._owner._dirty = True

    def __getitem__(self, key):
        return super().__getitem__(self._keys(key)[0])


class Force:
    """Common output handling (``hoomd.md.force.Force`` loggables)."""

    _kind = None

    def __init__(self):
        self._dirty = True
        self._state = None
        self._force = None
        self._virial = None
        self._torque = None
        self.compute_virial = True  # HOOMD sets this from the pressure_tensor flag each step
        self.block_size = 0
        self.threads_per_particle = 0
        self.launches = 0
        # neighbour data staged per tile in shared memory (azb_tile_plan) or gathered straight from global memory
        self.use_tiles = True
        self.tile_size = 0  # particles per tile of the tile plan; 0 = chosen from the mean row length (_auto_tile_size)
        self._plan = None
        self._tile_range = None
        self._tile_order = None

    def _attach(self, state):
        if state.device.type != "cuda":
            raise RuntimeError("azplugins_b200 forces run on a CUDA device only (no CPU fallback)")
        self._lib = _abi.load()
        self._state = state
        _, _, td = scalar_dtype(state.scalar_bits)
        self._force = torch.zeros((state.N, 4), dtype=td, device=state.device)
        self._virial = torch.zeros((6, max(state.N, 1)), dtype=td, device=state.device)
        self._dirty = True

    def _sync_size(self):
        """Output arrays for the state's current number of local particles (domain decomposition: N changes when particles
        migrate on a list rebuild)."""
        st = self._state
        if st is not None and self._force is not None and self._force.shape[0] != st.N:
            td = self._force.dtype
            self._force = torch.zeros((st.N, 4), dtype=td, device=st.device)
            self._virial = torch.zeros((6, max(st.N, 1)), dtype=td, device=st.device)
            if self._torque is not None:
                self._torque = torch.zeros((st.N, 4), dtype=td, device=st.device)

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self._state.device).cuda_stream)

    # ---- loggables -----------------------------------------------------------------------------------------
    @property
    def forces(self):
        return self._force[:, :3].cpu().numpy()

    @property
    def energies(self):
        return self._force[:, 3].cpu().numpy()

    @property
    def energy(self):
        return float(self._force[:, 3].double().sum().item())

    @property
    def virials(self):
        return self._virial[:, : self._state.N].T.cpu().numpy()

    @property
    def torques(self):
        if self._torque is None:
            return np.zeros((self._state.N, 3))
        return self._torque[:, :3].cpu().numpy()


class Pair(Force):
    """Isotropic pair force over a neighbour list (``hoomd.md.pair.Pair`` + ``PotentialPairGPU<E>``)."""

    _spec = None
    _accepted_modes = ("none", "shift", "xplor")
    _kind = _abi.KIND_PAIR
    _tiled_kernels = True  # which kernel families have a tile-staged flavour

    def __init__(self, nlist, default_r_cut=None, default_r_on=0.0, mode="none"):
        super().__init__()
        if mode not in self._accepted_modes:
            raise ValueError(f"mode must be one of {self._accepted_modes}")
        self.mode = mode
        self.nlist = nlist
        self.params = TypeParameter(self, self._spec.validate, True)
        self.r_cut = TypeParameter(self, float, True)
        self.r_on = TypeParameter(self, float, True)
        self._default_r_cut = default_r_cut
        self._default_r_on = default_r_on
        nlist._subscribe(self)

    @property
    def _cpp_class_name(self):
        return self._spec.cpp_class

    def _slot_lane_width(self, nbar):
        """Slot-row layout of the tile plan = lanes per particle of the tile kernel: narrow groups (positions come from
        shared memory, so nothing is lost on coalescing) unless rows are long."""
        if self.threads_per_particle:
            return self.threads_per_particle
        return 2 if nbar <= 200 else 4

    def _plan_request(self, nbar, require_built=True):
        """What this force asks of a tile plan of its list: ``nl.tile_plan`` keyword arguments. The list build itself can emit
        the plan when it knows the request beforehand (``nlist.NeighborList.build`` collects them from its consumers)."""
        prune, prune_key = self._prune_spec(require_built)
        gw = self._group_width(nbar) if prune is None else 0
        return dict(tile_size=self.tile_size or self._auto_tile_size(nbar), lane_width=self._slot_lane_width(nbar),
                    group_width=gw, prune=prune, prune_key=prune_key, sort_rows=bool(getattr(self.nlist, "sort_rows", False)) and gw == 0)

    def _current_plan(self):
        """Tile plan of the current neighbour list for this force's kernel (None: gather from global memory)."""
        if not (self.use_tiles and self._tiled_kernels) or self._state is None:
            return None
        st, nl = self._state, self.nlist
        return nl.tile_plan(st.N, **self._plan_request(nl.mean_neighbors(st.N)))

    # A force whose cutoffs are smaller than the list's (several forces sharing one list) walks a PRUNED view of it: entries
    # that were farther apart than r_cut + buffer when the list was built cannot be inside r_cut before the next rebuild
    # (HOOMD rebuilds before anything has moved buffer / 2). Only worth a plan of its own when it drops a fair share.
    prune_lists = True
    _PRUNE_MIN_GAIN = 0.85

    def _prune_spec(self, require_built=True):
        st, nl = self._state, self.nlist
        if not self.prune_lists or os.environ.get("AZB_NO_PRUNE", "0") == "1" or st is None:
            return None, None
        if require_built and getattr(nl, "_last_pos", None) is None:
            return None, None  # positions at build time unknown (externally built list)
        rc = self._r_cut_matrix(st.types)
        rl = nl.r_list(st.types)  # r_cut_max + buffer over all consumers of the list
        keep = np.where(rc > 0, rc + nl.buffer, 0.0)
        active = rl > 0
        if not np.any(active) or np.all(keep[active] ** 3 >= self._PRUNE_MIN_GAIN * rl[active] ** 3):
            return None, None
        cached = getattr(self, "_prune_cache", None)
        key = (id(st), st.scalar_bits, tuple(np.round(keep.reshape(-1), 12)))
        if cached is None or cached[0] != key:
            ftype = scalar_dtype(st.scalar_bits)[0]
            k = keep.astype(ftype)
            self._prune_cache = (key, torch.from_numpy((k * k).reshape(-1)).to(st.device))
        return (st, self._prune_cache[1]), key[1:]

    def _auto_tile_size(self, nbar):
        """128 particles per tile for every kernel family. (With 256-thread CTAs, 256-particle tiles were 2 % faster for
        long rows; with the 128-thread CTAs the pair kernels now use, 128 wins: 4 M LJ 0.931 ms against 0.949 ms.)"""
        return 128

    def _group_width(self, nbar):
        """Bank rotation of the slot rows (``azb_tile_build_args.group_width``): off. It halves the shared-memory bank conflicts
        of the neighbour reads but the kernels are not bound by them, and it doubles the plan build. ``AZB_BANK_ROTATE=1``
        turns it on for A/B measurements."""
        if os.environ.get("AZB_BANK_ROTATE", "0") == "1":
            return self._slot_lane_width(nbar)
        return 0

    def _pair_value(self, store, a, b, default):
        key = tuple(sorted((a, b)))
        if dict.__contains__(store, key):
            return dict.__getitem__(store, key)
        if default is None:
            raise KeyError(f"{type(self).__name__}: no value set for type pair {key}")
        return default

    def _r_cut_matrix(self, types):
        T = len(types)
        r = np.zeros((T, T))
        for i, a in enumerate(types):
            for j, b in enumerate(types):
                r[i, j] = self._pair_value(self.r_cut, a, b, self._default_r_cut)
        return r

    def _tables(self):
        """Device tables indexed typei*T+typej: param structs, rcutsq, ronsq (symmetric fill)."""
        st = self._state
        types = st.types
        T = len(types)
        ftype = scalar_dtype(st.scalar_bits)[0]
        dt = self._spec.dtype(st.scalar_bits)
        par = np.zeros(T * T, dtype=dt)
        rc = self._r_cut_matrix(types)
        ron = np.zeros((T, T))
        for i, a in enumerate(types):
            for j, b in enumerate(types):
                par[i * T + j] = self._spec.pack(self._pair_value(self.params, a, b, None), st.scalar_bits)
                ron[i, j] = self._pair_value(self.r_on, a, b, self._default_r_on)
        rcs = (rc.astype(ftype) * rc.astype(ftype)).reshape(-1)
        rons = (ron.astype(ftype) * ron.astype(ftype)).reshape(-1)
        self._d_params = torch.from_numpy(par.view(np.uint8).copy()).to(st.device)
        self._d_rcutsq = torch.from_numpy(rcs).to(st.device)
        self._d_ronsq = torch.from_numpy(rons).to(st.device)
        self._dirty = False

    def _fill_pair_args(self, a):
        st, nl = self._state, self.nlist
        a.d_force = self._force.data_ptr()
        a.d_virial = self._virial.data_ptr()
        a.virial_pitch = self._virial.shape[1]
        a.N = st.N
        a.n_max = st.n_max
        a.d_pos = st.pos.data_ptr()
        a.d_charge = None
        a.box = st.box.as_abi()
        a.d_n_neigh = nl.n_neigh.data_ptr()
        a.d_nlist = nl.nlist.data_ptr()
        a.d_head_list = nl.head_list.data_ptr()
        a.d_rcutsq = self._d_rcutsq.data_ptr()
        a.d_ronsq = self._d_ronsq.data_ptr()
        a.size_nlist = nl.size_nlist
        a.ntypes = len(st.types)
        a.block_size = self.block_size
        a.shift_mode = {"none": _abi.SHIFT_NONE, "shift": _abi.SHIFT_SHIFT, "xplor": _abi.SHIFT_XPLOR}[self.mode]
        a.compute_virial = 1 if self.compute_virial else 0
        a.threads_per_particle = self.threads_per_particle
        self._plan = self._current_plan()
        a.tiles = self._plan.pointer() if self._plan is not None else None
        a.tile_begin, a.tile_count = self._tile_range if (self._plan is not None and self._tile_range) else (0, 0)
        a.d_tile_order = self._tile_order.data_ptr() if (self._plan is not None and self._tile_order is not None) else None

    def _prepare(self):
        if self._state is None:
            raise RuntimeError("force is not attached to a state")
        self._sync_size()
        self.nlist.compute(self._state)
        if self._dirty:
            self._tables()

    def _build_args(self, timestep, tile_range=None, tile_order=None):
        """The C-ABI argument struct of one launch (kept alive by the caller until the launch is issued)."""
        self._tile_range, self._tile_order = tile_range, tile_order
        args = _abi.PairArgs()
        self._fill_pair_args(args)
        return args

    def _set_timestep(self, args, timestep):
        pass

    def _launch(self, args):
        code = self._lib.azb_pair_forces(self._spec.potential_id, self._state.scalar_bits, C.byref(args),
                                         self._d_params.data_ptr(), self._stream())
        _abi.check(code, f"azb_pair_forces[{self._spec.name}]")
        self.launches += 1

    def compute(self, timestep=0, tile_range=None, tile_order=None):
        """One evaluation (one kernel launch). ``tile_range=(first, count)`` restricts a tile-staged launch to that range
        of launch positions and leaves the outputs of all other particles untouched; ``tile_order`` (device int32 tensor)
        maps launch positions to tiles. Together they let a step overlap with copies (``hostio.HostStep``) or with the
        ghost exchange (``domain.HaloExchange.step``)."""
        self._prepare()
        if tile_range is not None and tile_range[1] == 0:
            return  # an empty range (the ABI reads tile_count == 0 as "every tile")
        self._launch(self._build_args(timestep, tile_range, tile_order))


class DPDThermoPair(Pair):
    """``PotentialPairDPDThermoGPU<E>``: pair force + velocity/tag gathers + Philox random force."""

    _accepted_modes = ("none",)
    _kind = _abi.KIND_DPD
    _tiled_kernels = True

    def __init__(self, nlist, kT, default_r_cut=None):
        super().__init__(nlist, default_r_cut, 0.0, "none")
        self.kT = kT
        self.dt = 0.005  # set by the integrator (hoomd passes m_deltaT to the force)

    def _slot_lane_width(self, nbar):
        return 1  # the DPD queue kernels read slots one at a time in row order

    def _group_width(self, nbar):
        """Bank rotation for the DPD walk: rows keep the plain layout (lane_width 1) but are dealt over the bank groups for
        the kernel's lanes per particle (same table as csrc/azb_dpd.cu)."""
        if os.environ.get("AZB_BANK_ROTATE", "0") != "1":
            return 0
        if self.threads_per_particle:
            return self.threads_per_particle
        return 2 if nbar <= 16 else (4 if nbar <= 96 else (8 if nbar <= 256 else 16))

    def _kT_at(self, timestep):
        return float(self.kT(timestep)) if callable(self.kT) else float(self.kT)

    def _build_args(self, timestep, tile_range=None, tile_order=None):
        st = self._state
        self._tile_range, self._tile_order = tile_range, tile_order
        args = _abi.DPDArgs()
        self._fill_pair_args(args.pair)
        args.d_vel = st.vel.data_ptr()
        args.d_tag = st.tag.data_ptr()
        args.seed = st.seed
        args.deltaT = float(self.dt)
        self._set_timestep(args, timestep)
        return args

    def _set_timestep(self, args, timestep):
        args.timestep = int(timestep)
        args.T = self._kT_at(timestep)

    def _launch(self, args):
        code = self._lib.azb_dpd_forces(self._spec.potential_id, self._state.scalar_bits, C.byref(args), self._d_params.data_ptr(), self._stream())
        _abi.check(code, f"azb_dpd_forces[{self._spec.name}]")
        self.launches += 1


class AnisotropicPair(Pair):
    """``AnisoPotentialPairGPU<E>``: force + torque (``hoomd.md.pair.aniso.AnisotropicPair``)."""

    _accepted_modes = ("none", "shift")
    _kind = _abi.KIND_ANISO

    def __init__(self, nlist, default_r_cut=None, mode="none"):
        super().__init__(nlist, default_r_cut, 0.0, mode)
        # in-range compaction (csrc/azb_ring_walk.cuh): the evaluator runs on queued in-range pairs only. Needs plain-order slot
        # rows (plan lane_width 1) and 2, 4 or 8 lanes per particle. OFF by default — measured on B200 (C4a, 2 048 000 particles,
        # 20 list entries per particle): 0.350 ms at 4 lanes per particle, 0.304 ms at 2, against 0.288 ms for the branchy tile
        # kernel: rows this short spend more on the walk's votes than the compaction saves. AZB_ANISO_RING=1 turns it on (A/B).
        self.ring_walk = os.environ.get("AZB_ANISO_RING", "0") == "1"

    def _slot_lane_width(self, nbar):
        if self.ring_walk and self.threads_per_particle in (0, 2, 4, 8):
            return 1
        return super()._slot_lane_width(nbar)

    def _attach(self, state):
        super()._attach(state)
        self._torque = torch.zeros((state.N, 4), dtype=self._force.dtype, device=state.device)

    def _build_args(self, timestep, tile_range=None, tile_order=None):
        st = self._state
        self._tile_range, self._tile_order = tile_range, tile_order
        args = _abi.AnisoArgs()
        self._fill_pair_args(args.pair)
        args.d_orientation = st.orientation.data_ptr()
        args.d_torque = self._torque.data_ptr()
        args.d_shape_params = None
        return args

    def _launch(self, args):
        code = self._lib.azb_aniso_forces(self._spec.potential_id, self._state.scalar_bits, C.byref(args), self._d_params.data_ptr(), self._stream())
        _abi.check(code, f"azb_aniso_forces[{self._spec.name}]")
        self.launches += 1


class Bond(Force):
    """``PotentialBondGPU<E, BondData>`` (``hoomd.md.bond.Bond``)."""

    _spec = None
    _kind = _abi.KIND_BOND

    def __init__(self):
        super().__init__()
        self.params = TypeParameter(self, self._spec.validate, False)

    @property
    def _cpp_class_name(self):
        return self._spec.cpp_class

    def _attach(self, state):
        super()._attach(state)
        self._flags = torch.zeros(1, dtype=torch.int32, device=state.device)

    def _tables(self):
        st = self._state
        dt = self._spec.dtype(st.scalar_bits)
        par = np.zeros(max(len(st.bond_types), 1), dtype=dt)
        for i, name in enumerate(st.bond_types):
            if not dict.__contains__(self.params, name):
                raise KeyError(f"{type(self).__name__}: no parameters set for bond type {name!r}")
            par[i] = self._spec.pack(dict.__getitem__(self.params, name), st.scalar_bits)
        self._d_params = torch.from_numpy(par.view(np.uint8).copy()).to(st.device)
        self._dirty = False

    def compute(self, timestep=0):
        if self._state is None:
            raise RuntimeError("force is not attached to a state")
        st = self._state
        self._sync_size()
        if self._dirty:
            self._tables()
        tb = st.bond_table()
        a = _abi.BondArgs()
        a.d_force = self._force.data_ptr()
        a.d_virial = self._virial.data_ptr()
        a.virial_pitch = self._virial.shape[1]
        a.N = st.N
        a.n_max = st.n_max
        a.d_pos = st.pos.data_ptr()
        a.d_charge = None
        a.box = st.box.as_abi()
        a.d_gpu_bondlist = tb["bondlist"].data_ptr()
        a.d_gpu_bond_pos = tb["bond_pos"].data_ptr()
        a.d_gpu_n_bonds = tb["n_bonds"].data_ptr()
        a.pitch = tb["pitch"]
        a.max_bonds = tb["max_bonds"]
        a.n_bond_types = max(len(st.bond_types), 1)
        a.block_size = self.block_size
        a.compute_virial = 1 if self.compute_virial else 0
        self._flags.zero_()
        code = self._lib.azb_bond_forces(self._spec.potential_id, st.scalar_bits, C.byref(a), self._d_params.data_ptr(),
                                         self._flags.data_ptr(), self._stream())
        _abi.check(code, f"azb_bond_forces[{self._spec.name}]")
        self.launches += 1
        return self

    def check_flags(self):
        """Host-side error check HOOMD's PotentialBondGPU performs after the kernel (device flag -> exception)."""
        if int(self._flags.item()) != 0:
            raise RuntimeError(f"bond.{self._spec.name}: bond out of bounds / invalid parameters")
