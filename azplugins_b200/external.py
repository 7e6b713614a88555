"""External potentials — the ``hoomd.azplugins.external`` API (/root/reference/src/external.py:16-151) over
``azb_barrier_forces`` (the B200 stand-in for ``HarmonicBarrierGPU<BarrierEvaluatorT>``, HarmonicBarrierGPU.h:47-72).

``location`` is variant-like as in HOOMD: a number (constant) or any callable ``location(timestep) -> float``
(``hoomd.variant.Variant`` instances are callables). ``params[type] = dict(k=..., offset=...)`` per particle type
(external.py:63-68). The virial contribution is not computed and reads as zero (HarmonicBarrier.h:176).
"""

import ctypes as C

import numpy as np
import torch

from . import _abi
from .force import Force, TypeParameter
from .state import scalar_dtype


class _Constant:
    """``hoomd.variant.Constant``: what a plain number becomes (external.py:61, variant_preprocessing)."""

    def __init__(self, value):
        self.value = float(value)

    def __call__(self, timestep):
        return self.value


def _validate(value):
    value = dict(value)
    unknown = set(value) - {"k", "offset"}
    if unknown:
        raise KeyError(f"unknown HarmonicBarrier parameter(s) {sorted(unknown)}")
    missing = {"k", "offset"} - set(value)
    if missing:
        raise KeyError(f"HarmonicBarrier parameter(s) {sorted(missing)} are required")
    return {"k": float(value["k"]), "offset": float(value["offset"])}


class HarmonicBarrier(Force):
    """Repulsive barrier implemented as a harmonic potential (external.py:16-91). Use a derived type."""

    _geometry = None

    def __init__(self, location):
        if self._geometry is None:
            raise TypeError("HarmonicBarrier should not be instantiated directly; use a derived type")
        super().__init__()
        self.location = location
        self.params = TypeParameter(self, _validate, False)

    @property
    def location(self):
        return self._location

    @location.setter
    def location(self, value):
        if callable(value):
            self._location = value
        elif isinstance(value, (int, float, np.integer, np.floating)):
            self._location = _Constant(value)
        else:
            raise TypeError("location must be a number or a callable location(timestep)")

    @property
    def _cpp_class_name(self):
        return type(self).__name__ + "GPU"  # external.py:87-91

    def _tables(self):
        st = self._state
        ftype = scalar_dtype(st.scalar_bits)[0]
        par = np.zeros((max(len(st.types), 1), 2), dtype=ftype)
        for i, name in enumerate(st.types):
            if not dict.__contains__(self.params, name):
                raise KeyError(f"{type(self).__name__}: no parameters set for particle type {name!r}")
            p = dict.__getitem__(self.params, name)
            par[i] = (p["k"], p["offset"])
        self._d_params = torch.from_numpy(par).to(st.device)
        self._dirty = False

    def compute(self, timestep=0):
        if self._state is None:
            raise RuntimeError("force is not attached to a state")
        st = self._state
        if self._dirty:
            self._tables()
        a = _abi.BarrierArgs()
        a.d_force = self._force.data_ptr()
        a.d_virial = self._virial.data_ptr()
        a.virial_pitch = self._virial.shape[1]
        a.N = st.N
        a.ntypes = max(len(st.types), 1)
        a.d_pos = st.pos.data_ptr()
        a.d_params = self._d_params.data_ptr()
        a.global_box = getattr(st, "global_box", st.box).as_abi()
        a.location = float(self._location(int(timestep)))
        a.block_size = self.block_size
        code = self._lib.azb_barrier_forces(self._geometry, st.scalar_bits, C.byref(a), self._stream())
        if code == _abi.ERR_BARRIER_INVALID:
            raise RuntimeError("Barrier position is invalid")  # HarmonicBarrier.h:131
        _abi.check(code, f"azb_barrier_forces[{type(self).__name__}]")
        self.launches += 1
        return self


class PlanarHarmonicBarrier(HarmonicBarrier):
    """Planar harmonic barrier normal to *y* (external.py:94-124): U = k/2 (y - H)^2 for y > H = location + offset."""

    _geometry = _abi.BARRIER_PLANAR


class SphericalHarmonicBarrier(HarmonicBarrier):
    """Spherical harmonic barrier (external.py:127-151): U = k/2 (r - R)^2 for r > R = location + offset."""

    _geometry = _abi.BARRIER_SPHERICAL
