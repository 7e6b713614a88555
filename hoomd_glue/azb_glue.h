// azb_glue.h — shared by the four kernel-driver stubs of the B200 build of hoomd.azplugins (hoomd_glue/*GPUKernel.cu.inc):
// conversions from HOOMD's argument structs to the C-ABI structs of include/azb200.h, and the error mapping.
// The stubs replace the explicit instantiations of /root/reference/src/PotentialPairGPUKernel.cu.inc:25-28,
// PotentialPairDPDThermoGPUKernel.cu.inc:21-24, AnisoPotentialPairGPUKernel.cu.inc:21-25 and PotentialBondGPUKernel.cu.inc:25-29
// by full specialisations that forward to libazb200.so. Nothing above that line of HOOMD / azplugins changes.
//
// Field names of pair_args_t / dpd_pair_args_t / a_pair_args_t / bond_args_t<2> follow HOOMD-blue v5 as recalled (HOOMD is not
// under /root/reference); hoomd_glue/mock/ declares exactly the members used here, and tests/test_hoomd_glue.py compiles the
// stubs against it and links them with libazb200.so. The build against a real HOOMD install is unverified on this box.
#pragma once
#include "azb200.h"
#include "hoomd/BoxDim.h"
#include "hoomd/HOOMDMath.h"

namespace hoomd
    {
namespace azplugins
    {
namespace glue
    {
inline azb_box to_azb_box(const BoxDim& box)
    {
    azb_box b;
    const Scalar3 L = box.getL();
    b.Lx = L.x;
    b.Ly = L.y;
    b.Lz = L.z;
    b.xy = box.getTiltFactorXY();
    b.xz = box.getTiltFactorXZ();
    b.yz = box.getTiltFactorYZ();
    const uchar3 p = box.getPeriodic();
    b.periodic[0] = p.x;
    b.periodic[1] = p.y;
    b.periodic[2] = p.z;
    return b;
    }

// 0 -> hipSuccess; CUDA error codes pass through; AZB_ERR_* (>= 10000) -> hipErrorInvalidValue
inline hipError_t to_hip_error(int rc)
    {
    return rc == 0 ? hipSuccess : (rc < 10000 ? (hipError_t)rc : hipErrorInvalidValue);
    }

// the members shared by pair_args_t, dpd_pair_args_t and a_pair_args_t
template<class HoomdArgs> inline void fill_common(azb_pair_args& a, const HoomdArgs& h)
    {
    a = azb_pair_args();
    a.d_force = h.d_force;
    a.d_virial = h.d_virial;
    a.virial_pitch = h.virial_pitch;
    a.N = h.N;
    a.n_max = h.n_max;
    a.d_pos = h.d_pos;
    a.d_charge = nullptr; // needsCharge() is false for every azplugins evaluator (PairEvaluator.h:67-75)
    a.box = to_azb_box(h.box);
    a.d_n_neigh = h.d_n_neigh;
    a.d_nlist = h.d_nlist;
    a.d_head_list = h.d_head_list;
    a.d_rcutsq = h.d_rcutsq;
    a.d_ronsq = nullptr;
    a.size_nlist = 0; // filled where HOOMD's struct carries it; 0 lets the tile cache read it from the device
    a.ntypes = h.ntypes;
    a.block_size = h.block_size;
    a.shift_mode = h.shift_mode;
    a.compute_virial = h.compute_virial;
    a.threads_per_particle = h.threads_per_particle;
    }

// Tile plan of the list in `a` from the library's cache (built on first use, rebuilt after azb_tile_cache_invalidate): the
// fast, shared-memory-staged kernels. AZB_GLUE_NO_TILES keeps the direct kernels.
inline int attach_tiles(azb_pair_args& a, unsigned int lane_width)
    {
#ifdef AZB_GLUE_NO_TILES
    (void)lane_width;
    a.tiles = nullptr;
    return 0;
#else
    const azb_tile_plan* plan = nullptr;
    const int rc = azb_tile_cache_get(&a, lane_width, /*tile_size*/ 0, /*stream*/ nullptr, &plan);
    a.tiles = plan; // NULL: no plan possible for this list -> direct kernels
    return rc;
#endif
    }

// lanes per particle of the pair / aniso tile kernels = slot layout of their plan: HOOMD's autotuner value is replaced by the
// engine's own choice (2 for rows up to 200 entries, else 4) so that one plan serves every launch between two list rebuilds
inline unsigned int tile_lanes(const azb_pair_args& a)
    {
    const size_t nbar = a.N ? a.size_nlist / a.N : 0;
    return nbar <= 200 ? 2u : 4u;
    }
    } // namespace glue
    } // namespace azplugins
    } // namespace hoomd
