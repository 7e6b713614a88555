// azb_pair_kernel.cuh — neighbour-list driven isotropic pair force kernel (replaces HOOMD's
// gpu_compute_pair_forces<E> as instantiated by /root/reference/src/PotentialPairGPUKernel.cu.inc:25-28).
//
// Work decomposition: TPP lanes of a warp cooperate on one particle. Lane l walks neighbour-row entries
// l, l+TPP, ...: the row read is a coalesced 4*TPP-byte segment, each entry is followed by one 16 B (float) or
// 32 B (double) read-only gather of pos[j]. Two row entries are in flight per lane and iteration (both index loads,
// then both position gathers, then the arithmetic) so that gather latency overlaps. Per-type-pair coefficients
// (hoisted evaluator-constructor math and the energy shift) live in shared memory — or in registers when the system
// has a single particle type (SINGLE). FASTBOX selects straight-line minimum-image code for the fully periodic
// orthorhombic box. Force, energy and the six virial rows are reduced across the TPP lanes with shuffles; lane 0
// writes Scalar4 force (energy in .w) and the virial rows.
//
// pair_tile_kernel is the same evaluation with the neighbour positions taken from a shared-memory copy of the tile's
// distinct neighbours (azb_tile_plan): one CTA per tile stages them once with coalesced loads, then addresses them
// through 16-bit slots read four at a time.
#pragma once
#include "azb200.h"
#include "azb_math.cuh"
#include "azb_pair_potentials.cuh"

namespace azb
    {
template<class S> struct PairArgsDev
    {
    typename vec4<S>::type* force;
    S* virial;
    size_t virial_pitch;
    unsigned int N;
    const typename vec4<S>::type* pos;
    Box<S> box;
    const unsigned int* n_neigh;
    const unsigned int* nlist;
    const size_t* head_list;
    const S* rcutsq;
    const S* ronsq;
    unsigned int ntypes;
    };

// device view of azb_tile_plan
struct TilePlanDev
    {
    unsigned int tile_size, uniq_capacity, smem_uniq; // smem_uniq: capacity of the staging arrays (>= max_uniq)
    unsigned int smem_span;                           // capacity of the staged slot rows, entries (>= max_slot_span)
    unsigned int tile_begin;                          // first tile of this launch (blockIdx.x = 0)
    const unsigned int* tile_order;                   // optional permutation: tile = tile_order[tile_begin + blockIdx.x]
    const unsigned int* uniq_count;
    const unsigned int* uniq;
    const uint16_t* slots;
    unsigned int slot_stride; // > 0: the slot row of particle i starts at i * slot_stride instead of AZB_SLOT_HEAD(head_list[i], i, W)
    };

// first entry of a particle's slot row
__device__ __forceinline__ size_t slot_row_begin(const TilePlanDev& plan, const size_t* __restrict__ head_list, unsigned int idx, unsigned int W)
    {
    if (plan.slot_stride)
        return (size_t)idx * plan.slot_stride;
    return AZB_SLOT_HEAD(__ldg(head_list + idx), idx, W);
    }

template<int TPP, class S> __device__ __forceinline__ S group_sum(S v)
    {
#pragma unroll
    for (int o = TPP / 2; o > 0; o >>= 1)
        v += __shfl_down_sync(0xffffffffu, v, o, TPP);
    return v;
    }

// per-lane accumulators of one particle
template<class S, bool VIRIAL> struct PairAccum
    {
    S fx = S(0), fy = S(0), fz = S(0), pe = S(0);
    S vxx = S(0), vxy = S(0), vxz = S(0), vyy = S(0), vyz = S(0), vzz = S(0);
    __device__ __forceinline__ void add(S dx, S dy, S dz, S fdivr, S fdivr_virial, S e)
        {
        fx += dx * fdivr;
        fy += dy * fdivr;
        fz += dz * fdivr;
        pe += e;
        if (VIRIAL) // sums f/r dx_a dx_b; the virial's factor 1/2 is applied once, in store()
            {
            const S hx = fdivr_virial * dx, hy = fdivr_virial * dy, hz = fdivr_virial * dz;
            vxx += hx * dx;
            vxy += hx * dy;
            vxz += hx * dz;
            vyy += hy * dy;
            vyz += hy * dz;
            vzz += hz * dz;
            }
        }
    template<int TPP> __device__ __forceinline__ void reduce()
        {
        if (TPP > 1)
            {
            fx = group_sum<TPP>(fx);
            fy = group_sum<TPP>(fy);
            fz = group_sum<TPP>(fz);
            pe = group_sum<TPP>(pe);
            if (VIRIAL)
                {
                vxx = group_sum<TPP>(vxx);
                vxy = group_sum<TPP>(vxy);
                vxz = group_sum<TPP>(vxz);
                vyy = group_sum<TPP>(vyy);
                vyz = group_sum<TPP>(vyz);
                vzz = group_sum<TPP>(vzz);
                }
            }
        }
    __device__ __forceinline__ void store(const PairArgsDev<S>& a, unsigned int idx) const
        {
        st4(a.force + idx, fx, fy, fz, S(0.5) * pe);
        if (VIRIAL)
            {
            a.virial[0 * a.virial_pitch + idx] = S(0.5) * vxx;
            a.virial[1 * a.virial_pitch + idx] = S(0.5) * vxy;
            a.virial[2 * a.virial_pitch + idx] = S(0.5) * vxz;
            a.virial[3 * a.virial_pitch + idx] = S(0.5) * vyy;
            a.virial[4 * a.virial_pitch + idx] = S(0.5) * vyz;
            a.virial[5 * a.virial_pitch + idx] = S(0.5) * vzz;
            }
        }
    };

// XPLOR smoothing between r_on and r_cut (HOOMD PotentialPair)
template<class S> __device__ __forceinline__ void xplor(S rsq, S rcutsq, S ronsq, S& fdivr, S& e)
    {
    if (rsq >= ronsq && rsq < rcutsq)
        {
        const S d = rcutsq - ronsq;
        const S denom_inv = S(1.0) / (d * d * d);
        const S q = rsq - rcutsq;
        const S s = q * q * (rcutsq + S(2.0) * rsq - S(3.0) * ronsq) * denom_inv;
        const S ds = S(12.0) * (rsq - ronsq) * q * denom_inv;
        const S old_e = e;
        e = old_e * s;
        fdivr = s * fdivr - ds * old_e;
        }
    }

// SHIFT: 0 none, 1 shift, 2 xplor (azb_shift_mode)
template<class Pot, class S, int SHIFT, bool VIRIAL, int TPP, bool FASTBOX, bool SINGLE>
__global__ void __launch_bounds__(256) pair_force_kernel(const PairArgsDev<S> a, const typename Pot::template Params<S>* __restrict__ params)
    {
    typedef typename Pot::template Coeff<S> coeff_t;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int T2 = a.ntypes * a.ntypes;
    coeff_t* s_coeff = reinterpret_cast<coeff_t*>(smem_raw);
    S* s_rcutsq = reinterpret_cast<S*>(s_coeff + T2);
    S* s_ronsq = s_rcutsq + T2;
    for (unsigned int t = threadIdx.x; t < T2; t += blockDim.x)
        {
        const S rc = a.rcutsq[t];
        S ron = S(0);
        bool energy_shift = (SHIFT == 1);
        if (SHIFT == 2)
            {
            ron = a.ronsq[t];
            energy_shift = ron > rc;
            }
        s_coeff[t] = Pot::make(params[t], rc, energy_shift);
        s_rcutsq[t] = Pot::active(s_coeff[t]) ? rc : S(-1); // inactive type pair: no distance passes the cutoff test
        s_ronsq[t] = ron;
        }
    __syncthreads();

    const unsigned int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int idx = gtid / TPP;
    const unsigned int lane = threadIdx.x % TPP;
    const bool active = idx < a.N;

    PairAccum<S, VIRIAL> acc;
    if (active)
        {
        const unsigned int nn = __ldg(a.n_neigh + idx);
        const unsigned int* row = a.nlist + __ldg(a.head_list + idx);
        const auto pi = ld4(a.pos + idx);
        const unsigned int trow = SINGLE ? 0u : type_of(pi.w) * a.ntypes;
        // single-type systems: the one coefficient set lives in registers for the whole loop
        const coeff_t c0 = s_coeff[0];
        const S rc0 = s_rcutsq[0], ron0 = s_ronsq[0];

        auto visit = [&](const typename vec4<S>::type& pj)
        {
            S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
            min_image_t<FASTBOX>(a.box, dx, dy, dz);
            const S rsq = dx * dx + dy * dy + dz * dz;
            S fdivr = S(0), e = S(0);
            if (SINGLE)
                {
                if (Pot::eval(rsq, rc0, c0, fdivr, e))
                    {
                    if (SHIFT == 2)
                        xplor(rsq, rc0, ron0, fdivr, e);
                    acc.add(dx, dy, dz, fdivr, fdivr, e);
                    }
                }
            else
                {
                const unsigned int tp = trow + type_of(pj.w);
                const S rcutsq = s_rcutsq[tp];
                if (Pot::eval(rsq, rcutsq, s_coeff[tp], fdivr, e))
                    {
                    if (SHIFT == 2)
                        xplor(rsq, rcutsq, s_ronsq[tp], fdivr, e);
                    acc.add(dx, dy, dz, fdivr, fdivr, e);
                    }
                }
        };

        unsigned int k = lane;
        for (; k + TPP < nn; k += 2 * TPP)
            {
            const unsigned int j0 = __ldg(row + k);
            const unsigned int j1 = __ldg(row + k + TPP);
            const auto p0 = ld4(a.pos + j0);
            const auto p1 = ld4(a.pos + j1);
            visit(p0);
            visit(p1);
            }
        if (k < nn)
            visit(ld4(a.pos + __ldg(row + k)));
        }
    acc.template reduce<TPP>();
    if (active && lane == 0)
        acc.store(a, idx);
    }
// Separation of one list entry with the lazy minimum image. In an orthorhombic box (any mix of periodic and
// non-periodic directions — the latter is what a domain-decomposed rank sees) a pair whose raw separation is inside half
// the shortest PERIODIC box edge (wrap_sq = (L_min / 2)^2; +inf when nothing is periodic) needs no wrap. Returns r^2. The
// wrap itself is a real (not inlined) call so that the compiler cannot turn the rare path into predicated straight-line
// code that every list entry would pay for. LAZY = false (triclinic box): BoxDim::minImage for every entry.
template<class S> struct Sep
    {
    S dx, dy, dz, rsq;
    };
template<class S> __device__ __noinline__ Sep<S> wrapped_separation(S Lx, S Ly, S Lz, S Lxinv, S Lyinv, S Lzinv, unsigned int flags, S dx, S dy, S dz)
    {
    Sep<S> r;
    r.dx = (flags & 1u) ? dx - Lx * rint_small(dx * Lxinv) : dx;
    r.dy = (flags & 2u) ? dy - Ly * rint_small(dy * Lyinv) : dy;
    r.dz = (flags & 4u) ? dz - Lz * rint_small(dz * Lzinv) : dz;
    r.rsq = r.dx * r.dx + r.dy * r.dy + r.dz * r.dz;
    return r;
    }
template<bool LAZY, class S> __device__ __forceinline__ S separation(const Box<S>& box, S wrap_sq, S& dx, S& dy, S& dz)
    {
    if (LAZY)
        {
        S rsq = dx * dx + dy * dy + dz * dz;
        if (rsq > wrap_sq) // only then can a component exceed half a periodic box edge
            {
            const Sep<S> w = wrapped_separation(box.Lx, box.Ly, box.Lz, box.Lxinv, box.Lyinv, box.Lzinv, box.flags, dx, dy, dz);
            dx = w.dx;
            dy = w.dy;
            dz = w.dz;
            rsq = w.rsq;
            }
        return rsq;
        }
    min_image(box, dx, dy, dz);
    return dx * dx + dy * dy + dz * dz;
    }

// One CTA per tile of consecutive particles. Stage: s_pos[s] = pos[uniq[s]] (ascending indices -> coalesced runs).
// Walk: blockDim / TPP groups, each taking tile_size / groups particles. A row is read in blocks of 4 TPP entries:
// every lane fetches its four slots of a block with one 8-byte load (the next block's load is issued before the
// current block is evaluated) and the lanes of a group visit consecutive list entries together (azb200.h, slot rows).
// Stage rows src[uq[0..U)] of a tile into shared memory. The gather is two dependent loads per row (index, then row); eight
// (four) rows per thread are kept in flight so a tile of <= 2048 distinct neighbours costs two (four) memory round trips, not
// 2 * U / 256.
// Warm L2 with the slot rows of a tile (the largest stream the walk reads from global memory: 2 B per list entry) and the
// positions of its own particles. Nothing waits on a prefetch, so the requests overlap the staging gather, and the walk's
// packet loads then cost an L2 hit instead of a DRAM round trip. W = lanes per particle of the slot-row layout.
template<class S>
__device__ __forceinline__ void warm_tile_rows(const PairArgsDev<S>& a, const TilePlanDev& plan, const unsigned int tile, const unsigned int W)
    {
    for (unsigned int p = threadIdx.x; p < plan.tile_size; p += blockDim.x)
        {
        const unsigned int idx = tile * plan.tile_size + p;
        if (idx < a.N)
            {
            const unsigned int nn = __ldg(a.n_neigh + idx);
            const char* r = reinterpret_cast<const char*>(plan.slots + slot_row_begin(plan, a.head_list, idx, W));
            for (unsigned int off = 0; off < nn * 2u; off += 128u)
                prefetch_l2(r + off);
            if (p % (128u / (unsigned int)sizeof(typename vec4<S>::type)) == 0u)
                prefetch_l2(a.pos + idx);
            }
        }
    }

template<class V4>
__device__ __forceinline__ void stage_rows(V4* __restrict__ dst, const V4* __restrict__ src, const unsigned int* __restrict__ uq, const unsigned int U)
    {
    constexpr int B = 128 / sizeof(V4); // 32 registers of rows in flight: 8 float rows, 4 double rows
    const unsigned int step = blockDim.x;
    for (unsigned int base = threadIdx.x; base < U; base += B * step)
        {
        unsigned int j[B];
        V4 v[B];
#pragma unroll
        for (int k = 0; k < B; ++k)
            j[k] = (base + k * step < U) ? __ldg(uq + base + k * step) : 0xffffffffu;
#pragma unroll
        for (int k = 0; k < B; ++k)
            if (j[k] != 0xffffffffu)
                v[k] = ld4(src + j[k]);
#pragma unroll
        for (int k = 0; k < B; ++k)
            if (j[k] != 0xffffffffu)
                dst[base + k * step] = v[k];
        }
    }

// The next block's slot packet, requested where the statement stands: as a plain __ldg the compiler sinks the load to the end of
// a branch-free loop body (its result is only used one iteration later), which turns it into a blocking load — 13 % of the
// kernel's stall samples in the first capture of the interior-tile walk.
__device__ __forceinline__ uint2 ldg_packet_here(const uint2* p)
    {
    uint2 v;
    asm volatile("ld.global.nc.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
    }

// Is the particle at p (inside the box, as HOOMD keeps its local particles) farther than the cutoff from every PERIODIC face?
// Then none of its in-range pairs crosses a face: a pair whose minimum image is shorter than r_cut but whose raw separation
// differs from it by a box edge would need |p| > L / 2 - r_cut in that direction. Thresholds L / 2 - 1.01 r_cut, recomputed per
// call (per row) rather than kept in registers across the walk. rcutsq > 0.
template<class S, class V4> __device__ __forceinline__ bool away_from_periodic_faces(const Box<S>& box, const V4& p, const S rcutsq)
    {
    const S margin = S(1.01) * rcutsq * rsqrt_(rcutsq);
    return ((box.flags & 1u) == 0u || fabs(p.x) < S(0.5) * box.Lx - margin) & ((box.flags & 2u) == 0u || fabs(p.y) < S(0.5) * box.Ly - margin)
           & ((box.flags & 4u) == 0u || fabs(p.z) < S(0.5) * box.Lz - margin);
    }

template<class Pot, class S, int SHIFT, bool VIRIAL, int TPP, bool FASTBOX, bool SINGLE>
__global__ void __launch_bounds__(256) pair_tile_kernel(const PairArgsDev<S> a, const TilePlanDev plan,
                                                        const typename Pot::template Params<S>* __restrict__ params, const S wrap_sq)
    {
    typedef typename Pot::template Coeff<S> coeff_t;
    typedef typename vec4<S>::type V4;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int T2 = a.ntypes * a.ntypes;
    V4* s_pos = reinterpret_cast<V4*>(smem_raw);
    coeff_t* s_coeff = reinterpret_cast<coeff_t*>(s_pos + plan.smem_uniq);
    S* s_rcutsq = reinterpret_cast<S*>(s_coeff + T2);
    S* s_ronsq = s_rcutsq + T2;

    const unsigned int tile = plan.tile_order ? __ldg(plan.tile_order + plan.tile_begin + blockIdx.x) : plan.tile_begin + blockIdx.x;
    const unsigned int U = __ldg(plan.uniq_count + tile);
    const unsigned int* uq = plan.uniq + (size_t)tile * plan.uniq_capacity;
    warm_tile_rows(a, plan, tile, TPP);
    stage_rows(s_pos, a.pos, uq, U);
    for (unsigned int t = threadIdx.x; t < T2; t += blockDim.x)
        {
        const S rc = a.rcutsq[t];
        S ron = S(0);
        bool energy_shift = (SHIFT == 1);
        if (SHIFT == 2)
            {
            ron = a.ronsq[t];
            energy_shift = ron > rc;
            }
        s_coeff[t] = Pot::make(params[t], rc, energy_shift);
        s_rcutsq[t] = Pot::active(s_coeff[t]) ? rc : S(-1); // inactive type pair: no distance passes the cutoff test
        s_ronsq[t] = ron;
        }
    __syncthreads();
    const coeff_t c0 = s_coeff[0];
    const S rc0 = s_rcutsq[0], ron0 = s_ronsq[0];
    const unsigned int pos_base = smem_addr(s_pos);

    constexpr unsigned int M = 4 * TPP; // entries per block of a row
    const unsigned int groups = blockDim.x / TPP;
    const unsigned int gid = threadIdx.x / TPP;
    const unsigned int lane = threadIdx.x % TPP;
    for (unsigned int p = gid; p < plan.tile_size; p += groups) // warp-uniform trip count (host picks blockDim)
        {
        const unsigned int idx = tile * plan.tile_size + p;
        const bool active = idx < a.N;
        PairAccum<S, VIRIAL> acc;
        if (active)
            {
            const unsigned int nn = __ldg(a.n_neigh + idx);
            const uint2* row = reinterpret_cast<const uint2*>(plan.slots + slot_row_begin(plan, a.head_list, idx, TPP)) + lane;
            const V4 pi = ld4(a.pos + idx);
            const unsigned int trow = SINGLE ? 0u : type_of(pi.w) * a.ntypes;

            auto visit = [&](const unsigned int slot, const bool valid)
            {
                const V4 pj = lds4(pos_base, slot, S());
                S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                const S rsq = separation<FASTBOX>(a.box, wrap_sq, dx, dy, dz);
                S fdivr = S(0), e = S(0);
                if (SINGLE)
                    {
                    if (valid && Pot::eval(rsq, rc0, c0, fdivr, e))
                        {
                        if (SHIFT == 2)
                            xplor(rsq, rc0, ron0, fdivr, e);
                        acc.add(dx, dy, dz, fdivr, fdivr, e);
                        }
                    }
                else
                    {
                    const unsigned int tp = trow + type_of(pj.w);
                    const S rcutsq = s_rcutsq[tp];
                    if (valid && Pot::eval(rsq, rcutsq, s_coeff[tp], fdivr, e))
                        {
                        if (SHIFT == 2)
                            xplor(rsq, rcutsq, s_ronsq[tp], fdivr, e);
                        acc.add(dx, dy, dz, fdivr, fdivr, e);
                        }
                    }
            };

            const unsigned int full = nn / M;             // blocks whose 4 TPP entries all exist
            const unsigned int nblk = (nn + M - 1u) / M;  // ... plus the ragged last one
            uint2 next = make_uint2(0u, 0u);
            if (nblk)
                next = __ldg(row);
            for (unsigned int b = 0; b < full; ++b)
                {
                const uint2 s4 = next;
                if (b + 1 < nblk)
                    next = __ldg(row + (b + 1) * TPP);
                visit(s4.x & 0xffffu, true);
                visit(s4.x >> 16, true);
                visit(s4.y & 0xffffu, true);
                visit(s4.y >> 16, true);
                }
            if (full < nblk) // entries full*M + u*TPP + lane, u = 0..3, that exist (padding slots are never evaluated)
                {
                const unsigned int e0 = full * M + lane;
                visit(next.x & 0xffffu, e0 < nn);
                visit(next.x >> 16, e0 + TPP < nn);
                visit(next.y & 0xffffu, e0 + 2 * TPP < nn);
                visit(next.y >> 16, e0 + 3 * TPP < nn);
                }
            }
        acc.template reduce<TPP>();
        if (active && lane == 0)
            acc.store(a, idx);
        }
    }
    } // namespace azb
