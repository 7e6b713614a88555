"""Test-side access to the CPU oracles (tests/ is one of the few places allowed to load anything under oracle/).

``port``: oracle/_build/libazoracle.so — the plain-C restatement (always buildable: gcc only).
``ref`` : oracle/_ref/libazref{32,64}.so — the reference's evaluator headers compiled unmodified (built in the container
          where /root/reference exists; the prebuilt .so travels to the GPU box).

Both take the structs of include/azb200.h with HOST pointers; this module fills them from numpy arrays.
"""

import ctypes as C
import os
import subprocess
import types

import numpy as np

from azplugins_b200 import _abi
from azplugins_b200.state import pack_scalar4, scalar_dtype

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PORT_PATH = os.path.join(ROOT, "oracle", "_build", "libazoracle.so")
REF_PATHS = {b: os.path.join(ROOT, "oracle", "_ref", f"libazref{b}.so") for b in (32, 64)}

_port = None
_ref = None


def build_port():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "port"], check=True)


def port():
    global _port
    if _port is None:
        if not os.path.exists(PORT_PATH):
            build_port()
        lib = C.CDLL(PORT_PATH)
        lib.azo_dpd_alpha.restype = C.c_double
        lib.azo_dpd_alpha.argtypes = [C.c_int, C.c_uint, C.c_uint, C.c_uint, C.c_uint]
        for name in ("azo_pair_forces", "azo_dpd_forces", "azo_aniso_forces"):
            fn = getattr(lib, name)
            fn.restype = C.c_int
            fn.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        lib.azo_bond_forces.restype = C.c_int
        lib.azo_bond_forces.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        lib.azo_barrier_forces.restype = C.c_int
        lib.azo_barrier_forces.argtypes = [C.c_int, C.c_int, C.c_void_p]
        lib.azo_nlist_build.restype = C.c_int
        lib.azo_nlist_build.argtypes = [C.c_int, C.c_void_p, C.c_uint, C.c_uint, C.POINTER(_abi.Box), C.c_void_p, C.c_uint,
                                        C.c_void_p, C.c_void_p, C.c_uint, C.c_void_p, C.c_void_p,
                                        C.POINTER(C.POINTER(C.c_uint)), C.POINTER(C.c_size_t)]
        lib.azo_free.argtypes = [C.c_void_p]
        lib.azo_philox.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _port = lib
    return _port


def have_ref():
    return all(os.path.exists(p) for p in REF_PATHS.values())


def ref():
    global _ref
    if _ref is None:
        lib = types.SimpleNamespace()
        for bits in (32, 64):
            # PyDLL keeps the GIL: make_params / params_as_dict touch Python objects. One library per precision.
            dll = C.PyDLL(REF_PATHS[bits])
            p = f"azref{bits}_"
            for name in ("pair_forces", "dpd_forces", "aniso_forces", "bond_forces", "barrier_forces", "dpd_alpha", "param_size", "make_params",
                         "params_as_dict"):
                setattr(lib, p + name, getattr(dll, p + name))
            for name in ("pair_forces", "dpd_forces", "aniso_forces"):
                fn = getattr(lib, p + name)
                fn.restype = C.c_int
                fn.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
            fn = getattr(lib, p + "bond_forces")
            fn.restype = C.c_int
            fn.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
            fn = getattr(lib, p + "barrier_forces")
            fn.restype = C.c_int
            fn.argtypes = [C.c_int, C.c_void_p]
            fn = getattr(lib, p + "dpd_alpha")
            fn.restype = C.c_double
            fn.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_uint]
            fn = getattr(lib, p + "param_size")
            fn.restype = C.c_size_t
            fn.argtypes = [C.c_int, C.c_int]
            fn = getattr(lib, p + "make_params")
            fn.restype = C.c_int
            fn.argtypes = [C.c_int, C.c_int, C.py_object, C.c_void_p]
            fn = getattr(lib, p + "params_as_dict")
            fn.restype = C.py_object
            fn.argtypes = [C.c_int, C.c_int, C.c_void_p]
        _ref = lib
    return _ref


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class HostSystem:
    """numpy arrays in HOOMD layout + the filled C structs, for driving an oracle."""

    def __init__(self, box, position, typeid, ntypes, bits, velocity=None, mass=None, orientation=None, tag=None, n_local=None):
        self.bits = bits
        self.ftype = scalar_dtype(bits)[0]
        self.box = box
        self.n_max = len(position)
        self.N = self.n_max if n_local is None else n_local
        self.ntypes = ntypes
        self.pos = pack_scalar4(np.asarray(position, dtype=np.float64), np.asarray(typeid), bits, True)
        v = np.zeros((self.n_max, 3)) if velocity is None else np.asarray(velocity, dtype=np.float64)
        m = np.ones(self.n_max) if mass is None else np.asarray(mass, dtype=np.float64)
        self.vel = pack_scalar4(v, m, bits, False)
        q = np.tile([1.0, 0, 0, 0], (self.n_max, 1)) if orientation is None else np.asarray(orientation, dtype=np.float64)
        self.ori = np.ascontiguousarray(q, dtype=self.ftype)
        self.tag = np.arange(self.n_max, dtype=np.uint32) if tag is None else np.ascontiguousarray(tag, dtype=np.uint32)
        self.force = np.zeros((self.N, 4), dtype=self.ftype)
        self.torque = np.zeros((self.N, 4), dtype=self.ftype)
        self.virial = np.zeros((6, max(self.N, 1)), dtype=self.ftype)
        self.n_neigh = self.head_list = self.nlist = None

    # ---- neighbour list through the oracle's builder -----------------------------------------------------
    def build_nlist(self, r_list, excl_head=None, excl_list=None, row_align=1):
        r_list = np.ascontiguousarray(np.broadcast_to(np.asarray(r_list, dtype=np.float64), (self.ntypes, self.ntypes)))
        self.n_neigh = np.zeros(max(self.N, 1), dtype=np.uint32)
        self.head_list = np.zeros(max(self.N, 1), dtype=np.uint64)
        out = C.POINTER(C.c_uint)()
        size = C.c_size_t(0)
        b = self.box.as_abi()
        eh = None if excl_head is None else np.ascontiguousarray(excl_head, dtype=np.uint64)
        el = None if excl_list is None else np.ascontiguousarray(excl_list, dtype=np.uint32)
        rc = port().azo_nlist_build(self.bits, _ptr(self.pos), self.N, self.n_max, C.byref(b), _ptr(r_list), self.ntypes,
                                    _ptr(eh), _ptr(el), row_align, _ptr(self.n_neigh), _ptr(self.head_list),
                                    C.byref(out), C.byref(size))
        assert rc == 0
        n = size.value
        self.nlist = np.ctypeslib.as_array(out, shape=(max(n, 1),)).copy()[:n] if n else np.zeros(0, dtype=np.uint32)
        port().azo_free(out)
        return self

    def full_to_half(self):
        """Half-storage copy of the list (pairs i<j kept once, plus all ghost neighbours), HOOMD CPU mode."""
        nn = np.zeros_like(self.n_neigh)
        head = np.zeros_like(self.head_list)
        rows = []
        tot = 0
        for i in range(self.N):
            r = self.nlist[int(self.head_list[i]): int(self.head_list[i]) + int(self.n_neigh[i])]
            r = r[(r > i) | (r >= self.N)]
            head[i] = tot
            nn[i] = len(r)
            tot += len(r)
            rows.append(r)
        half = HostSystem.__new__(HostSystem)
        half.__dict__.update(self.__dict__)
        half.n_neigh, half.head_list = nn, head
        half.nlist = np.concatenate(rows).astype(np.uint32) if rows and tot else np.zeros(0, dtype=np.uint32)
        half.force = np.zeros_like(self.force)
        half.torque = np.zeros_like(self.torque)
        half.virial = np.zeros_like(self.virial)
        return half

    # ---- struct filling -------------------------------------------------------------------------------------
    def pair_args(self, a, rcutsq, ronsq, shift_mode, compute_virial):
        self._rcutsq = np.ascontiguousarray(rcutsq, dtype=self.ftype).reshape(-1)
        self._ronsq = np.ascontiguousarray(ronsq if ronsq is not None else np.zeros_like(self._rcutsq), dtype=self.ftype).reshape(-1)
        self._nl = self.nlist if self.nlist.size else np.zeros(1, dtype=np.uint32)
        a.d_force = self.force.ctypes.data
        a.d_virial = self.virial.ctypes.data
        a.virial_pitch = self.virial.shape[1]
        a.N, a.n_max = self.N, self.n_max
        a.d_pos = self.pos.ctypes.data
        a.box = self.box.as_abi()
        a.d_n_neigh = self.n_neigh.ctypes.data
        a.d_nlist = self._nl.ctypes.data
        a.d_head_list = self.head_list.ctypes.data
        a.d_rcutsq = self._rcutsq.ctypes.data
        a.d_ronsq = self._ronsq.ctypes.data
        a.size_nlist = self.nlist.size
        a.ntypes = self.ntypes
        a.shift_mode = shift_mode
        a.compute_virial = 1 if compute_virial else 0
        return a

    def _call(self, which, kind, potential, args, params, half_list, nthreads):
        params = np.ascontiguousarray(params)
        if which == "port":
            fn = {_abi.KIND_PAIR: port().azo_pair_forces, _abi.KIND_DPD: port().azo_dpd_forces, _abi.KIND_ANISO: port().azo_aniso_forces}[kind]
            rc = fn(potential, self.bits, C.addressof(args), params.ctypes.data, int(half_list), nthreads)
        else:
            name = {_abi.KIND_PAIR: "pair_forces", _abi.KIND_DPD: "dpd_forces", _abi.KIND_ANISO: "aniso_forces"}[kind]
            rc = getattr(ref(), f"azref{self.bits}_{name}")(potential, C.addressof(args), params.ctypes.data, int(half_list), nthreads)
        assert rc == 0, rc

    def pair(self, potential, params, rcutsq, ronsq=None, shift_mode=0, compute_virial=True, half_list=False, nthreads=1, which="port"):
        a = self.pair_args(_abi.PairArgs(), rcutsq, ronsq, shift_mode, compute_virial)
        self._call(which, _abi.KIND_PAIR, potential, a, params, half_list, nthreads)
        return self

    def dpd(self, params, rcutsq, seed, timestep, dt, kT, compute_virial=True, half_list=False, nthreads=1, which="port"):
        a = _abi.DPDArgs()
        self.pair_args(a.pair, rcutsq, None, 0, compute_virial)
        a.d_vel = self.vel.ctypes.data
        a.d_tag = self.tag.ctypes.data
        a.seed, a.timestep, a.deltaT, a.T = seed, timestep, dt, kT
        self._call(which, _abi.KIND_DPD, 0, a, params, half_list, nthreads)
        return self

    def aniso(self, params, rcutsq, shift_mode=0, compute_virial=True, half_list=False, nthreads=1, which="port"):
        a = _abi.AnisoArgs()
        self.pair_args(a.pair, rcutsq, None, shift_mode, compute_virial)
        a.d_orientation = self.ori.ctypes.data
        a.d_torque = self.torque.ctypes.data
        self._call(which, _abi.KIND_ANISO, 0, a, params, half_list, nthreads)
        return self

    def bond(self, potential, params, n_bonds, bondlist, bond_pos, pitch, max_bonds, n_bond_types, nthreads=1, which="port"):
        a = _abi.BondArgs()
        params = np.ascontiguousarray(params)
        self._bt = (np.ascontiguousarray(n_bonds, dtype=np.uint32), np.ascontiguousarray(bondlist, dtype=np.uint32),
                    np.ascontiguousarray(bond_pos, dtype=np.uint32))
        a.d_force = self.force.ctypes.data
        a.d_virial = self.virial.ctypes.data
        a.virial_pitch = self.virial.shape[1]
        a.N, a.n_max = self.N, self.n_max
        a.d_pos = self.pos.ctypes.data
        a.box = self.box.as_abi()
        a.d_gpu_n_bonds = self._bt[0].ctypes.data
        a.d_gpu_bondlist = self._bt[1].ctypes.data
        a.d_gpu_bond_pos = self._bt[2].ctypes.data
        a.pitch, a.max_bonds, a.n_bond_types = pitch, max_bonds, n_bond_types
        a.compute_virial = 1
        flags = C.c_uint(0)
        if which == "port":
            rc = port().azo_bond_forces(potential, self.bits, C.addressof(a), params.ctypes.data, C.addressof(flags), nthreads)
        else:
            rc = getattr(ref(), f"azref{self.bits}_bond_forces")(potential, C.addressof(a), params.ctypes.data, C.addressof(flags), nthreads)
        assert rc == 0, rc
        self.flags = flags.value
        return self


def barrier(geometry, box, position, typeid, params, location, bits, which="port"):
    """HarmonicBarrier oracle: returns (rc, force[N,4], virial[6,N]); params = [(k, offset)] per type."""
    ftype = scalar_dtype(bits)[0]
    pos = pack_scalar4(np.asarray(position, dtype=np.float64), np.asarray(typeid), bits, True)
    n = len(pos)
    par = np.ascontiguousarray(params, dtype=ftype).reshape(-1, 2)
    force = np.full((n, 4), np.nan, dtype=ftype)
    virial = np.full((6, max(n, 1)), np.nan, dtype=ftype)
    a = _abi.BarrierArgs()
    a.d_force, a.d_virial, a.virial_pitch = force.ctypes.data, virial.ctypes.data, virial.shape[1]
    a.N, a.ntypes = n, len(par)
    a.d_pos, a.d_params = pos.ctypes.data, par.ctypes.data
    a.global_box = box.as_abi()
    a.location = float(location)
    if which == "port":
        rc = port().azo_barrier_forces(geometry, bits, C.addressof(a))
    else:
        rc = getattr(ref(), f"azref{bits}_barrier_forces")(geometry, C.addressof(a))
    return rc, force, virial


def ref_make_params(spec, d, bits):
    """dict -> struct bytes through the reference's own pybind11 constructor."""
    out = np.zeros((), dtype=spec.dtype(bits))
    rc = getattr(ref(), f"azref{bits}_make_params")(spec.kind, spec.potential_id, dict(d), out.ctypes.data)
    assert rc == 0, rc
    return out


def ref_params_as_dict(spec, rec, bits):
    rec = np.ascontiguousarray(rec)
    return getattr(ref(), f"azref{bits}_params_as_dict")(spec.kind, spec.potential_id, rec.ctypes.data)
