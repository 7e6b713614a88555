"""ctypes mirror of ``include/azb200.h`` and the loader of the CUDA library.

The structures here are field-for-field the C structs of the drop-in boundary; the product library
(``libazb200.so``, hand-written sm_100a kernels) is loaded from the package directory and nothing else is ever
substituted for it: if it is missing, importing a force class fails loudly.
"""

import ctypes as C
import os

ABI_VERSION = 3

ERR_BAD_POTENTIAL = 10001
ERR_BAD_SCALAR = 10002
ERR_BAD_ARGUMENT = 10003
ERR_STATIC_SCRATCH_TOO_SMALL = 10005
STATIC_WORKSPACE_BYTES = 16 * 1024 * 1024

KIND_PAIR, KIND_DPD, KIND_ANISO, KIND_BOND = 0, 1, 2, 3
PAIR_PERTURBED_LENNARD_JONES, PAIR_COLLOID, PAIR_EXPANDED_YUKAWA, PAIR_HERTZ, PAIR_DPD_CONSERVATIVE = range(5)
DPD_GENERAL_WEIGHT = 0
ANISO_TWO_PATCH_MORSE = 0
BOND_DOUBLE_WELL, BOND_QUARTIC = 0, 1
SHIFT_NONE, SHIFT_SHIFT, SHIFT_XPLOR = 0, 1, 2


class Box(C.Structure):
    _fields_ = [
        ("Lx", C.c_double),
        ("Ly", C.c_double),
        ("Lz", C.c_double),
        ("xy", C.c_double),
        ("xz", C.c_double),
        ("yz", C.c_double),
        ("periodic", C.c_ubyte * 3),
        ("pad_", C.c_ubyte * 5),
    ]


class TilePlan(C.Structure):
    _fields_ = [
        ("tile_size", C.c_uint),
        ("n_tiles", C.c_uint),
        ("uniq_capacity", C.c_uint),
        ("max_uniq", C.c_uint),
        ("lane_width", C.c_uint),
        ("max_slot_span", C.c_uint),
        ("d_uniq_count", C.c_void_p),
        ("d_uniq", C.c_void_p),
        ("d_slots", C.c_void_p),
        ("d_row_count", C.c_void_p),
        ("n_entries", C.c_size_t),
        ("slot_row_stride", C.c_uint),
    ]


class TileBuildArgs(C.Structure):
    _fields_ = [
        ("d_n_neigh", C.c_void_p),
        ("d_nlist", C.c_void_p),
        ("d_head_list", C.c_void_p),
        ("size_nlist", C.c_size_t),
        ("N", C.c_uint),
        ("tile_size", C.c_uint),
        ("uniq_capacity", C.c_uint),
        ("lane_width", C.c_uint),
        ("group_width", C.c_uint),
        ("d_uniq_count", C.c_void_p),
        ("d_uniq", C.c_void_p),
        ("d_slots", C.c_void_p),
        ("d_status", C.c_void_p),
        ("d_prune_pos", C.c_void_p),
        ("d_prune_rsq", C.c_void_p),
        ("prune_scalar_bits", C.c_int),
        ("prune_ntypes", C.c_uint),
        ("prune_box", Box),
        ("d_row_count", C.c_void_p),
        ("prune_full_set", C.c_uint),
        ("d_tile_list", C.c_void_p),
        ("n_tile_list", C.c_uint),
        ("slot_row_stride", C.c_uint),
        ("base_plan", C.c_void_p),
        ("sort_rows", C.c_uint),
    ]


class PairArgs(C.Structure):
    _fields_ = [
        ("d_force", C.c_void_p),
        ("d_virial", C.c_void_p),
        ("virial_pitch", C.c_size_t),
        ("N", C.c_uint),
        ("n_max", C.c_uint),
        ("d_pos", C.c_void_p),
        ("d_charge", C.c_void_p),
        ("box", Box),
        ("d_n_neigh", C.c_void_p),
        ("d_nlist", C.c_void_p),
        ("d_head_list", C.c_void_p),
        ("d_rcutsq", C.c_void_p),
        ("d_ronsq", C.c_void_p),
        ("size_nlist", C.c_size_t),
        ("ntypes", C.c_uint),
        ("block_size", C.c_uint),
        ("shift_mode", C.c_uint),
        ("compute_virial", C.c_uint),
        ("threads_per_particle", C.c_uint),
        ("tiles", C.POINTER(TilePlan)),
        ("tile_begin", C.c_uint),
        ("tile_count", C.c_uint),
        ("d_tile_order", C.c_void_p),
    ]


class DPDArgs(C.Structure):
    _fields_ = [
        ("pair", PairArgs),
        ("d_vel", C.c_void_p),
        ("d_tag", C.c_void_p),
        ("seed", C.c_uint16),
        ("timestep", C.c_uint64),
        ("deltaT", C.c_double),
        ("T", C.c_double),
    ]


class AnisoArgs(C.Structure):
    _fields_ = [
        ("pair", PairArgs),
        ("d_orientation", C.c_void_p),
        ("d_torque", C.c_void_p),
        ("d_shape_params", C.c_void_p),
    ]


class BondArgs(C.Structure):
    _fields_ = [
        ("d_force", C.c_void_p),
        ("d_virial", C.c_void_p),
        ("virial_pitch", C.c_size_t),
        ("N", C.c_uint),
        ("n_max", C.c_uint),
        ("d_pos", C.c_void_p),
        ("d_charge", C.c_void_p),
        ("box", Box),
        ("d_gpu_bondlist", C.c_void_p),
        ("d_gpu_bond_pos", C.c_void_p),
        ("d_gpu_n_bonds", C.c_void_p),
        ("pitch", C.c_uint),
        ("max_bonds", C.c_uint),
        ("n_bond_types", C.c_uint),
        ("block_size", C.c_uint),
        ("compute_virial", C.c_uint),
    ]


class NlistView(C.Structure):
    _fields_ = [("lane_width", C.c_uint), ("d_keep_rsq", C.c_void_p), ("d_slots", C.c_void_p), ("d_row_count", C.c_void_p)]


class NlistPlanStatus(C.Structure):
    _fields_ = [("ok", C.c_uint), ("max_uniq", C.c_uint), ("max_slot_span", C.c_uint * 4), ("n_bad_tiles", C.c_uint)]


class NlistArgs(C.Structure):
    _fields_ = [
        ("d_pos", C.c_void_p),
        ("N", C.c_uint),
        ("n_max", C.c_uint),
        ("box", Box),
        ("ext_lo", C.c_double * 3),
        ("ext_hi", C.c_double * 3),
        ("r_list_max", C.c_double),
        ("d_r_listsq", C.c_void_p),
        ("ntypes", C.c_uint),
        ("row_align", C.c_uint),
        ("d_excl_head", C.c_void_p),
        ("d_excl_list", C.c_void_p),
        ("d_n_neigh", C.c_void_p),
        ("d_head_list", C.c_void_p),
        ("d_nlist", C.c_void_p),
        ("capacity", C.c_size_t),
        ("d_workspace", C.c_void_p),
        ("workspace_bytes", C.c_size_t),
        ("row_stride_hint", C.c_uint),
        ("n_views", C.c_uint),
        ("views", NlistView * 4),
        ("tile_size", C.c_uint),
        ("uniq_capacity", C.c_uint),
        ("d_uniq_count", C.c_void_p),
        ("d_uniq", C.c_void_p),
        ("d_uniq_max", C.c_void_p),
        ("rows_in_place", C.c_uint),
    ]


# every symbol include/azb200.h declares (tests check the library exports exactly these)
EXPORTS = (
    "azb_host_step",
    "azb_host_step_release",
    "azb_tile_cache_get",
    "azb_tile_cache_invalidate",
    "azb_tile_cache_clear",
    "azb_abi_version",
    "azb_param_size",
    "azb_param_align",
    "azb_potential_name",
    "azb_pair_forces",
    "azb_dpd_forces",
    "azb_aniso_forces",
    "azb_bond_forces",
    "azb_nlist_workspace_size",
    "azb_nlist_count",
    "azb_nlist_fill",
    "azb_nlist_plan_result",
    "azb_nlist_build_static",
    "azb_gather_scalar4",
    "azb_tiles_build",
    "azb_barrier_forces",
    "azb_nve_step_one",
    "azb_nve_step_two",
    "azb_displacement_check",
    "azb_ipc_export",
    "azb_ipc_open",
    "azb_halo_push",
    "azb_halo_wait",
    "azb_halo_signal",
)

IPC_HANDLE_BYTES = 64
ERR_BARRIER_INVALID = 10004
BARRIER_PLANAR, BARRIER_SPHERICAL = 0, 1


class NveArgs(C.Structure):
    """``azb_nve_args``."""

    _fields_ = [
        ("d_pos", C.c_void_p),
        ("d_vel", C.c_void_p),
        ("d_accel", C.c_void_p),
        ("d_image", C.c_void_p),
        ("N", C.c_uint),
        ("box", Box),
        ("deltaT", C.c_double),
        ("d_forces", C.c_void_p),
        ("n_forces", C.c_uint),
        ("d_net_force", C.c_void_p),
    ]


class BarrierArgs(C.Structure):
    """``azb_barrier_args``."""

    _fields_ = [
        ("d_force", C.c_void_p),
        ("d_virial", C.c_void_p),
        ("virial_pitch", C.c_size_t),
        ("N", C.c_uint),
        ("ntypes", C.c_uint),
        ("d_pos", C.c_void_p),
        ("d_params", C.c_void_p),
        ("global_box", Box),
        ("location", C.c_double),
        ("block_size", C.c_uint),
    ]



class HaloRoute(C.Structure):
    """``azb_halo_route``: rows ``d_src[d_idx[0..n)]`` (+ shift) -> ``d_dst[0..n)``."""

    _fields_ = [
        ("d_src", C.c_void_p),
        ("d_dst", C.c_void_p),
        ("d_idx", C.c_void_p),
        ("n", C.c_uint),
        ("row_begin", C.c_uint),
        ("shift", C.c_double * 3),
    ]


class HaloPushArgs(C.Structure):
    _fields_ = [
        ("d_routes", C.c_void_p),
        ("n_routes", C.c_uint),
        ("total_rows", C.c_uint),
        ("d_peer_flags", C.c_void_p),
        ("n_flags", C.c_uint),
        ("d_counter", C.c_void_p),
        ("epoch", C.c_uint),
    ]


LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libazb200.so")

class HostBuffer(C.Structure):
    _fields_ = [("d_ptr", C.c_void_p), ("h_ptr", C.c_void_p), ("row_bytes", C.c_size_t)]


class HostLaunch(C.Structure):
    _fields_ = [("wait_chunk", C.c_uint), ("tile_begin", C.c_uint), ("tile_count", C.c_uint)]


class HostOutRange(C.Structure):
    _fields_ = [("after_launch", C.c_uint), ("row_lo", C.c_size_t), ("row_hi", C.c_size_t)]


class HostForce(C.Structure):
    _fields_ = [("kind", C.c_int), ("potential", C.c_int), ("force_args", C.c_void_p), ("d_params", C.c_void_p)]


class HostStepArgs(C.Structure):
    _fields_ = [
        ("kind", C.c_int),
        ("potential", C.c_int),
        ("scalar_bits", C.c_int),
        ("force_args", C.c_void_p),
        ("d_params", C.c_void_p),
        ("n_inputs", C.c_uint),
        ("n_outputs", C.c_uint),
        ("inputs", HostBuffer * 8),
        ("outputs", HostBuffer * 8),
        ("n_in_chunks", C.c_uint),
        ("in_edges", C.POINTER(C.c_size_t)),
        ("n_launches", C.c_uint),
        ("launches", C.POINTER(HostLaunch)),
        ("n_outs", C.c_uint),
        ("outs", C.POINTER(HostOutRange)),
        ("stream", C.c_void_p),
        ("stream_in", C.c_void_p),
        ("stream_out", C.c_void_p),
        ("n_more", C.c_uint),
        ("more", C.POINTER(HostForce)),
        ("flags", C.c_uint),
    ]


HOST_STEP_NO_SYNC = 1


_lib = None


def load():
    """Load ``libazb200.so`` (built by ``__graft_entry__.build()`` / ``make -C azplugins_b200/csrc``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA extension has not been built "
            "(run `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    lib.azb_abi_version.restype = C.c_int
    lib.azb_param_size.restype = C.c_size_t
    lib.azb_param_size.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.azb_param_align.restype = C.c_size_t
    lib.azb_param_align.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.azb_potential_name.restype = C.c_char_p
    lib.azb_potential_name.argtypes = [C.c_int, C.c_int]
    lib.azb_pair_forces.restype = C.c_int
    lib.azb_pair_forces.argtypes = [C.c_int, C.c_int, C.POINTER(PairArgs), C.c_void_p, C.c_void_p]
    lib.azb_dpd_forces.restype = C.c_int
    lib.azb_dpd_forces.argtypes = [C.c_int, C.c_int, C.POINTER(DPDArgs), C.c_void_p, C.c_void_p]
    lib.azb_aniso_forces.restype = C.c_int
    lib.azb_aniso_forces.argtypes = [C.c_int, C.c_int, C.POINTER(AnisoArgs), C.c_void_p, C.c_void_p]
    lib.azb_bond_forces.restype = C.c_int
    lib.azb_bond_forces.argtypes = [C.c_int, C.c_int, C.POINTER(BondArgs), C.c_void_p, C.c_void_p, C.c_void_p]
    lib.azb_nlist_workspace_size.restype = C.c_int
    lib.azb_nlist_workspace_size.argtypes = [C.POINTER(NlistArgs), C.POINTER(C.c_size_t)]
    lib.azb_nlist_count.restype = C.c_int
    lib.azb_nlist_count.argtypes = [C.c_int, C.POINTER(NlistArgs), C.POINTER(C.c_size_t), C.c_void_p]
    lib.azb_nlist_fill.restype = C.c_int
    lib.azb_nlist_fill.argtypes = [C.c_int, C.POINTER(NlistArgs), C.c_void_p]
    lib.azb_nlist_build_static.restype = C.c_int
    lib.azb_nlist_build_static.argtypes = [C.c_int, C.POINTER(NlistArgs), C.POINTER(C.c_size_t), C.c_void_p]
    lib.azb_nlist_plan_result.restype = C.c_int
    lib.azb_nlist_plan_result.argtypes = [C.c_int, C.POINTER(NlistArgs), C.POINTER(NlistPlanStatus), C.c_void_p]
    lib.azb_gather_scalar4.restype = C.c_int
    lib.azb_gather_scalar4.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.azb_tiles_build.restype = C.c_int
    lib.azb_tiles_build.argtypes = [C.POINTER(TileBuildArgs), C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.c_void_p]
    lib.azb_barrier_forces.restype = C.c_int
    lib.azb_barrier_forces.argtypes = [C.c_int, C.c_int, C.POINTER(BarrierArgs), C.c_void_p]
    for fn in (lib.azb_nve_step_one, lib.azb_nve_step_two):
        fn.restype = C.c_int
        fn.argtypes = [C.c_int, C.POINTER(NveArgs), C.c_void_p]
    lib.azb_displacement_check.restype = C.c_int
    lib.azb_displacement_check.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint, C.POINTER(Box), C.c_double, C.c_void_p, C.c_void_p]
    lib.azb_ipc_export.restype = C.c_int
    lib.azb_ipc_export.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_size_t)]
    lib.azb_ipc_open.restype = C.c_int
    lib.azb_ipc_open.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_void_p)]
    lib.azb_halo_push.restype = C.c_int
    lib.azb_halo_push.argtypes = [C.c_int, C.POINTER(HaloPushArgs), C.c_void_p]
    lib.azb_halo_wait.restype = C.c_int
    lib.azb_halo_wait.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_void_p, C.c_void_p]
    lib.azb_halo_signal.restype = C.c_int
    lib.azb_halo_signal.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_void_p]
    lib.azb_host_step.restype = C.c_int
    lib.azb_host_step.argtypes = [C.POINTER(HostStepArgs)]
    lib.azb_tile_cache_get.restype = C.c_int
    lib.azb_tile_cache_get.argtypes = [C.POINTER(PairArgs), C.c_uint, C.c_uint, C.c_void_p, C.POINTER(C.POINTER(TilePlan))]
    lib.azb_tile_cache_invalidate.restype = None
    lib.azb_tile_cache_invalidate.argtypes = [C.c_void_p]
    lib.azb_tile_cache_clear.restype = None
    lib.azb_tile_cache_clear.argtypes = []
    lib.azb_host_step_release.restype = None
    lib.azb_host_step_release.argtypes = []
    if lib.azb_abi_version() != ABI_VERSION:
        raise ImportError("libazb200.so ABI version mismatch")
    _lib = lib
    return lib


def check(code, what):
    """Turn a C-ABI return code into an exception (the ABI itself never throws)."""
    if code == 0:
        return
    names = {ERR_BAD_POTENTIAL: "unknown potential id", ERR_BAD_SCALAR: "scalar_bits must be 32 or 64", ERR_BAD_ARGUMENT: "bad argument",
             ERR_STATIC_SCRATCH_TOO_SMALL: "system too large for the library's built-in scratch array / output capacity"}
    raise RuntimeError(f"{what} failed: {names.get(code, 'cudaError_t ' + str(code))}")
