// azb_launch.cuh — host-side glue shared by the per-family translation units: converts the C-ABI argument
// structs (include/azb200.h) into the typed device argument structs and picks launch shapes.
#pragma once
#include <atomic>
#include <cuda_runtime.h>

#include "azb200.h"
#include "azb_math.cuh"
#include "azb_pair_kernel.cuh"

namespace azb
    {
template<class S> inline Box<S> make_box(const azb_box& b)
    {
    Box<S> o;
    o.Lx = (S)b.Lx;
    o.Ly = (S)b.Ly;
    o.Lz = (S)b.Lz;
    o.Lxinv = S(1.0) / o.Lx;
    o.Lyinv = S(1.0) / o.Ly;
    o.Lzinv = S(1.0) / o.Lz;
    o.xy = (S)b.xy;
    o.xz = (S)b.xz;
    o.yz = (S)b.yz;
    o.flags = (b.periodic[0] ? 1u : 0u) | (b.periodic[1] ? 2u : 0u) | (b.periodic[2] ? 4u : 0u);
    if (b.xy != 0.0 || b.xz != 0.0 || b.yz != 0.0)
        o.flags |= 8u;
    return o;
    }

template<class S> inline PairArgsDev<S> make_pair_args(const azb_pair_args& a)
    {
    PairArgsDev<S> d;
    d.force = (typename vec4<S>::type*)a.d_force;
    d.virial = (S*)a.d_virial;
    d.virial_pitch = a.virial_pitch;
    d.N = a.N;
    d.pos = (const typename vec4<S>::type*)a.d_pos;
    d.box = make_box<S>(a.box);
    // the tile kernels take a row's length from n_neigh; a pruned plan carries its own (shorter) row lengths
    d.n_neigh = (a.tiles && a.tiles->d_row_count) ? a.tiles->d_row_count : a.d_n_neigh;
    d.nlist = a.d_nlist;
    d.head_list = a.d_head_list;
    d.rcutsq = (const S*)a.d_rcutsq;
    d.ronsq = (const S*)a.d_ronsq;
    d.ntypes = a.ntypes;
    return d;
    }

// threads cooperating on one particle: the caller's choice, else from the mean neighbour-row length
inline unsigned int pick_tpp(const azb_pair_args& a)
    {
    unsigned int tpp = a.threads_per_particle;
    if (tpp == 0)
        {
        const double nbar = a.N ? (double)a.size_nlist / (double)a.N : 0.0;
        tpp = nbar <= 8 ? 1 : (nbar <= 24 ? 4 : (nbar <= 72 ? 8 : (nbar <= 200 ? 16 : 32)));
        }
    return tpp;
    }
inline bool valid_tpp(unsigned int tpp)
    {
    return tpp == 1 || tpp == 2 || tpp == 4 || tpp == 8 || tpp == 16 || tpp == 32;
    }
inline unsigned int pick_block(unsigned int requested)
    {
    unsigned int b = requested ? requested : 128;
    if (b < 32)
        b = 32;
    if (b > 256)
        b = 256;
    return (b / 32) * 32;
    }

// opt-in dynamic shared memory above 48 KB. The attribute belongs to (device, kernel): the granted size is tracked per device
// of this process (one atomic per device and kernel instantiation — a concurrent duplicate call is harmless), and a refusal
// comes back to the caller as the CUDA error code.
#define AZB_MAX_SMEM (227 * 1024)
#define AZB_MAX_DEVICES 64
template<auto KERNEL> inline int ensure_smem(size_t bytes)
    {
    static std::atomic<size_t> granted[AZB_MAX_DEVICES];
    if (bytes <= 48 * 1024)
        return 0;
    int dev = 0;
    cudaError_t err = cudaGetDevice(&dev);
    if (err != cudaSuccess)
        return (int)err;
    const bool tracked = dev >= 0 && dev < AZB_MAX_DEVICES;
    if (tracked && bytes <= granted[dev].load(std::memory_order_acquire))
        return 0;
    err = cudaFuncSetAttribute(KERNEL, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (err != cudaSuccess)
        return (int)err;
    if (tracked)
        {
        size_t seen = granted[dev].load(std::memory_order_relaxed);
        while (seen < bytes && !granted[dev].compare_exchange_weak(seen, bytes, std::memory_order_release))
            {
            }
        }
    return 0;
    }

// Validates a.tiles against the launch and picks the CTA size: blockDim / tpp groups, each walking
// tile_size / groups particles (both powers of two, so every warp runs the same number of passes).
inline int make_tile_plan(const azb_pair_args& a, unsigned int tpp, unsigned int slot_lane_width, TilePlanDev& plan, unsigned int& block,
                          unsigned int& grid, unsigned int default_block = 256u)
    {
    const azb_tile_plan& t = *a.tiles;
    const unsigned int ts = t.tile_size;
    if (!(ts == 32 || ts == 64 || ts == 128 || ts == 256) || t.n_tiles != (a.N + ts - 1) / ts || t.max_uniq > t.uniq_capacity
        || !t.d_uniq_count || !t.d_uniq || !t.d_slots || t.lane_width != slot_lane_width)
        return AZB_ERR_BAD_ARGUMENT;
    plan.tile_size = ts;
    plan.uniq_capacity = t.uniq_capacity;
    plan.smem_uniq = ((t.max_uniq > 0 ? t.max_uniq : 1u) + 7u) & ~7u;
    plan.smem_span = (t.max_slot_span + 15u) & ~15u;
    plan.tile_begin = a.tile_count ? a.tile_begin : 0u;
    plan.tile_order = a.d_tile_order;
    grid = a.tile_count ? a.tile_count : t.n_tiles;
    if (plan.tile_begin >= t.n_tiles || grid > t.n_tiles - plan.tile_begin)
        return AZB_ERR_BAD_ARGUMENT;
    plan.uniq_count = t.d_uniq_count;
    plan.uniq = t.d_uniq;
    plan.slots = t.d_slots;
    plan.slot_stride = t.slot_row_stride;
    block = a.block_size ? pick_block(a.block_size) : default_block;
    if (block > ts * tpp)
        block = ts * tpp;
    if (block < 32 || (block & (block - 1u)))
        return AZB_ERR_BAD_ARGUMENT;
    return 0;
    }

// Lazy minimum image (azb_pair_kernel.cuh, separation()): usable in every orthorhombic box; the threshold is half the
// shortest periodic edge, squared (a value no r^2 exceeds when no direction is periodic).
template<class S> inline bool lazy_wrap(const Box<S>& box, S& wrap_sq)
    {
    S lmin = S(1e18);
    if (box.flags & 1u)
        lmin = box.Lx < lmin ? box.Lx : lmin;
    if (box.flags & 2u)
        lmin = box.Ly < lmin ? box.Ly : lmin;
    if (box.flags & 4u)
        lmin = box.Lz < lmin ? box.Lz : lmin;
    wrap_sq = S(0.25) * lmin * lmin;
    return (box.flags & 8u) == 0u;
    }

inline int check_pair_args(const azb_pair_args* a)
    {
    if (!a || a->ntypes == 0 || a->shift_mode > 2)
        return AZB_ERR_BAD_ARGUMENT;
    if (a->N && (!a->d_force || !a->d_pos || !a->d_n_neigh || !a->d_nlist || !a->d_head_list || !a->d_rcutsq))
        return AZB_ERR_BAD_ARGUMENT;
    if (a->compute_virial && a->N && (!a->d_virial || a->virial_pitch < a->N))
        return AZB_ERR_BAD_ARGUMENT;
    if (a->shift_mode == AZB_SHIFT_XPLOR && !a->d_ronsq)
        return AZB_ERR_BAD_ARGUMENT;
    return 0;
    }

// Expands `CALL(TPPVALUE)` for the run-time tpp
#define AZB_DISPATCH_TPP(tpp, CALL) \
    switch (tpp)                    \
        {                           \
    case 1:                         \
        CALL(1);                    \
        break;                      \
    case 2:                         \
        CALL(2);                    \
        break;                      \
    case 4:                         \
        CALL(4);                    \
        break;                      \
    case 8:                         \
        CALL(8);                    \
        break;                      \
    case 16:                        \
        CALL(16);                   \
        break;                      \
    default:                        \
        CALL(32);                   \
        break;                      \
        }

// Generic isotropic launcher: one instantiation set per (potential, scalar); defined in azb_pair_launch.cuh
template<class Pot, class S> int launch_pair(const azb_pair_args& a, const void* d_params, cudaStream_t stream);
    } // namespace azb
