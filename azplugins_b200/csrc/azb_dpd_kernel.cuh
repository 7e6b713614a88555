// azb_dpd_kernel.cuh — DPD thermostat pair force kernels (replace HOOMD's gpu_compute_dpd_forces<E> as
// instantiated by /root/reference/src/PotentialPairDPDThermoGPUKernel.cu.inc:21-24 for
// DPDPairEvaluatorGeneralWeight, /root/reference/src/DPDPairEvaluatorGeneralWeight.h:198-255).
//
// A buffered neighbour list holds ~(1 + r_buff/r_cut)^3 entries per in-range pair, and the in-range work
// (Philox4x32-10 + thermostat, ~100 instructions) is ~5x the out-of-range work, so evaluating it under
// `if (rsq < rcutsq)` leaves two thirds of the lanes idle through the expensive block. Both kernels here split the
// walk of a particle's row, done by a group of TPP lanes, into two phases that each run with (nearly) all lanes busy:
//   light: TPP list entries at a time — coalesced index read, one Scalar4 read of pos[j], distance, cutoff test;
//          in-range entries (dx, dy, dz, j) are compacted (ballot + popc) into the group's ring in shared memory;
//   heavy: whenever every group of the warp holds >= TPP queued entries (or has finished its row), each lane pops
//          one entry: reads vel[j], tag[j], draws the pair's random number, evaluates the force.
// dpd_force_kernel_q gathers pos/vel/tag of the neighbours straight from global memory; dpd_tile_kernel takes them
// from a shared-memory copy of the tile's distinct neighbours (azb_tile_plan) addressed by 16-bit slots.
//
// Minimum image: in the fully periodic orthorhombic box a pair whose raw separation is inside half the shortest box
// edge (wrap_sq = (L_min / 2)^2) needs no wrap; only the few pairs straddling the boundary take the wrap path.
// One Philox4x32-10 block per in-range pair, keyed on (seed, timestep) with the round keys precomputed on the host,
// counter = (min tag, max tag).
#pragma once
#include "azb_math.cuh"
#include "azb_launch.cuh"
#include "azb_pair_kernel.cuh"
#include "azb_rng.cuh"

// ring capacity per group in units of TPP (power of two), and the minimum number of resident CTAs per SM asked of the
// tile kernel; build-time switches for occupancy experiments
#ifndef AZB_DPD_QMUL
#define AZB_DPD_QMUL 4
#endif
#ifndef AZB_DPD_LIGHT2
#define AZB_DPD_LIGHT2 true
#endif
#ifdef AZB_DPD_TILE_MINB
#define AZB_DPD_TILE_BOUNDS __launch_bounds__(256, AZB_DPD_TILE_MINB)
#else
#define AZB_DPD_TILE_BOUNDS __launch_bounds__(256)
#endif

namespace azb
    {
template<class S> struct DPDArgsDev
    {
    PairArgsDev<S> pair;
    const typename vec4<S>::type* vel;
    const unsigned int* tag;
    PhiloxRoundKeys rk;
    S deltaT;
    S T;
    };

struct DPDThermo
    {
    template<class S> struct alignas(4 * sizeof(S)) Params
        {
        S A, gamma, s;
        };
    template<class S> struct alignas(4 * sizeof(S)) Coeff
        {
        S A, gamma, sigma, rcutinv; // sigma = rsqrt(dt / (6 kT gamma))
        S rcut, half_s, hA, rcutsq; // hA = A / (2 r_cut): E = hA (r_cut - r)^2
        };
    template<class S> __device__ static Coeff<S> make(const Params<S>& p, S rcutsq, S deltaT, S T)
        {
        Coeff<S> c;
        c.A = p.A;
        c.gamma = p.gamma;
        c.half_s = S(0.5) * p.s;
        c.rcutinv = S(1.0) / sqrt_(rcutsq);
        c.rcut = S(1.0) / c.rcutinv;
        c.sigma = S(1.0) / sqrt_(deltaT / (T * p.gamma * S(6.0)));
        c.hA = S(0.5) * p.A * c.rcutinv;
        c.rcutsq = rcutsq;
        return c;
        }
    };

// queue entry: separation vector + the neighbour's index (global index or tile slot) in the 4th component
template<class S> __device__ __forceinline__ typename vec4<S>::type pack_entry(S dx, S dy, S dz, unsigned int j);
template<> __device__ __forceinline__ float4 pack_entry<float>(float dx, float dy, float dz, unsigned int j)
    {
    return make_float4(dx, dy, dz, __uint_as_float(j));
    }
template<> __device__ __forceinline__ double4a pack_entry<double>(double dx, double dy, double dz, unsigned int j)
    {
    double4a e;
    e.x = dx;
    e.y = dy;
    e.z = dz;
    e.w = __longlong_as_double((long long)j);
    return e;
    }
__device__ __forceinline__ unsigned int entry_index(float w)
    {
    return __float_as_uint(w);
    }
__device__ __forceinline__ unsigned int entry_index(double w)
    {
    return (unsigned int)__double_as_longlong(w);
    }

// One in-range pair: conservative + drag + random force/r, conservative energy and virial
// (DPDPairEvaluatorGeneralWeight.h:198-255; the virial takes the conservative part only, :193-194).
template<class S, bool VIRIAL, class V4>
__device__ __forceinline__ void dpd_pair(const V4& e, const V4& vi, const V4& vj, unsigned int tag_i, unsigned int tag_j,
                                         const DPDThermo::Coeff<S>& c, const PhiloxRoundKeys& rk, PairAccum<S, VIRIAL>& acc)
    {
    const S rsq = e.x * e.x + e.y * e.y + e.z * e.z;
    const S rdotv = e.x * (vi.x - vj.x) + e.y * (vi.y - vj.y) + e.z * (vi.z - vj.z);
    const S rinv = rsqrt_(rsq);
    const S r = rsq * rinv;
    const S alpha = dpd_alpha<S>(tag_i, tag_j, rk);
    const S fcons = c.A * (rinv - c.rcutinv);
    // w = 1 - r/rc can round a hair below 0 right at the cutoff; clamp so pow() never sees a negative base
    S w = fmax(S(0), S(1.0) - r * c.rcutinv);
    S wpow = w;
    if (c.half_s != S(1.0))
        {
        // w^(s/2) with s < 2 is infinitely steep at the cutoff: take r from a correctly rounded sqrt there
        w = fmax(S(0), S(1.0) - sqrt_(rsq) * c.rcutinv);
        wpow = (c.half_s == S(0)) ? S(1.0) : pow_(w, c.half_s);
        }
    const S wR = wpow * rinv;
    const S fdivr = fcons + wR * (c.sigma * alpha - c.gamma * wR * rdotv);
    const S dr = c.rcut - r;
    acc.add(e.x, e.y, e.z, fdivr, fcons, (c.hA * dr) * dr); // hA dr^2 == A (rc - r) - A (rc^2 - r^2) / (2 rc)
    }

// The two-phase walk of one particle's row by a group of TPP lanes. `Src` supplies the neighbour data:
//   index(k)                list entry k of this particle's row (global particle index, or tile slot)
//   pos(j), vel(j), tag(j)  the neighbour's data (global memory, or the tile's shared-memory copy)
//   push / pop              how the neighbour index and the type pair travel through the ring
// All 32 lanes of the warp must call this together (warp votes inside); lanes of idle groups pass nn = 0.
// The ring of a group holds QCAP = 4 TPP entries at shared address q_addr (group rings are QCAP + 1 entries apart so
// that equal positions of different groups fall into different banks).
template<class S, bool VIRIAL, int TPP, bool FASTBOX, bool SINGLE, bool LIGHT2, class Src>
__device__ __forceinline__ void dpd_walk_row(const Src& src, const unsigned int nn, const typename vec4<S>::type& pi,
                                             const typename vec4<S>::type& vi, const unsigned int tag_i, const unsigned int trow,
                                             const Box<S>& box, const S wrap_sq, const unsigned int coeff_addr,
                                             const DPDThermo::Coeff<S>& c0, const unsigned int q_addr, const unsigned int qtp_addr,
                                             const PhiloxRoundKeys& rk, PairAccum<S, VIRIAL>& acc)
    {
    typedef typename vec4<S>::type V4;
    typedef DPDThermo::Coeff<S> coeff_t;
    constexpr unsigned int QCAP = AZB_DPD_QMUL * TPP;
    constexpr unsigned int FULL = 0xffffffffu;
    const unsigned int lane = threadIdx.x % TPP;
    const unsigned int gshift = (threadIdx.x & 31u) & ~(unsigned int)(TPP - 1);
    // lane masks of my group / of the lanes below me in it; pinned in registers (the compiler otherwise rebuilds them
    // from %tid inside the loop)
    unsigned int gmask, ltmask;
    asm volatile("mov.u32 %0, %1;" : "=r"(gmask) : "r"(((TPP == 32) ? FULL : ((1u << TPP) - 1u)) << gshift));
    asm volatile("mov.u32 %0, %1;" : "=r"(ltmask) : "r"(((1u << lane) - 1u) << gshift));
    const unsigned int nn_warp = __reduce_max_sync(FULL, nn); // longest row among the warp's groups

    unsigned int q_head = 0, q_tail = 0; // group-uniform ring state (monotonic; position = counter mod QCAP)
    // LIGHT2: two list entries per lane and light visit (entries k and k + TPP) — two independent chains of
    // load / distance / vote / push per lane, and half as many phase-switch votes. Ring order (hence the lane every
    // in-range pair is evaluated on, hence the summation order) is the same as with one entry per visit.
    constexpr unsigned int STEP = LIGHT2 ? 2u * TPP : (unsigned int)TPP;
    unsigned int k = lane;
    unsigned int j0 = 0, j1 = 0;
    if (k < nn)
        j0 = src.index(k);
    if (LIGHT2 && k + TPP < nn)
        j1 = src.index(k + TPP);
    bool room = true; // warp-uniform: every ring can take another light visit
    for (;;)
        {
        // ---- light phase: skipped by the whole warp once its longest row is exhausted. Straight-line: a lane without
        // an entry re-reads the last valid index (or 0) and is masked out of the cutoff test.
        if (room && k - lane < nn_warp)
            {
            const bool have0 = k < nn, have1 = LIGHT2 && (k + TPP < nn);
            const unsigned int jc0 = j0, jc1 = j1;
            k += STEP;
            if (k < nn)
                j0 = src.index(k);
            if (LIGHT2 && k + TPP < nn)
                j1 = src.index(k + TPP);
            auto light = [&](const unsigned int jc, const bool have, V4& entry, unsigned int& tp_out) -> bool
            {
                const V4 pj = src.pos(jc);
                S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                unsigned int tp = 0;
                S rc = c0.rcutsq;
                if (!SINGLE)
                    {
                    tp = trow + type_of(pj.w);
                    rc = lds_scalar(coeff_addr + tp * (unsigned int)sizeof(coeff_t) + 7u * (unsigned int)sizeof(S), S());
                    }
                const bool in = have & (separation<FASTBOX>(box, wrap_sq, dx, dy, dz) < rc);
                entry = pack_entry<S>(dx, dy, dz, Src::pack(jc, tp));
                tp_out = tp;
                return in;
            };
            V4 e0, e1;
            unsigned int tp0 = 0, tp1 = 0;
            const bool in0 = light(jc0, have0, e0, tp0);
            const bool in1 = LIGHT2 ? light(jc1, have1, e1, tp1) : false;
            const unsigned int b0 = __ballot_sync(FULL, in0);
            const unsigned int b1 = LIGHT2 ? __ballot_sync(FULL, in1) : 0u;
            // branch-free push: every lane computes its ring position, the store is predicated on `in`
            const unsigned int n0 = __popc(b0 & gmask);
            const unsigned int at0 = (q_tail + __popc(b0 & ltmask)) & (QCAP - 1u);
            sts4_if(in0, q_addr + at0 * (unsigned int)sizeof(V4), e0);
            if (!SINGLE && !Src::kPackedTypePair)
                sts32_if(in0, qtp_addr + at0 * 4u, tp0);
            if (LIGHT2)
                {
                const unsigned int at1 = (q_tail + n0 + __popc(b1 & ltmask)) & (QCAP - 1u);
                sts4_if(in1, q_addr + at1 * (unsigned int)sizeof(V4), e1);
                if (!SINGLE && !Src::kPackedTypePair)
                    sts32_if(in1, qtp_addr + at1 * 4u, tp1);
                }
            q_tail += n0 + __popc(b1 & gmask);
            }
        const bool light_done = (k - lane) >= nn;
        const unsigned int cnt = q_tail - q_head;
        const bool all_ready = __all_sync(FULL, (cnt >= TPP) | light_done);

        // ---- heavy phase: one queued in-range pair per lane ------------------------------------------------------
        // A light visit pushes up to STEP entries per group, a heavy pass pops TPP: the next light visit needs
        // cnt <= QCAP - STEP in every group, which one pop may not restore when STEP > TPP (`room`).
        const unsigned int maxcnt = __reduce_max_sync(FULL, cnt);
        room = true;
        if (all_ready || maxcnt > QCAP - STEP)
            {
            room = !LIGHT2 || (maxcnt <= QCAP - STEP + TPP);
            __syncwarp();
            if (lane < cnt)
                {
                const unsigned int at = (q_head + lane) & (QCAP - 1u);
                const V4 e = lds4_at(q_addr + at * (unsigned int)sizeof(V4), S());
                const unsigned int packed = entry_index(e.w);
                const unsigned int jj = Src::index_of(packed);
                const V4 vj = src.vel(jj);
                const unsigned int tag_j = src.tag(jj);
                if (SINGLE)
                    dpd_pair<S, VIRIAL>(e, vi, vj, tag_i, tag_j, c0, rk, acc);
                else
                    {
                    const unsigned int tq = Src::kPackedTypePair ? Src::type_pair_of(packed) : lds32_at(qtp_addr + at * 4u);
                    const unsigned int ca = coeff_addr + tq * (unsigned int)sizeof(coeff_t);
                    const V4 c_lo = lds4_at(ca, S()), c_hi = lds4_at(ca + (unsigned int)sizeof(V4), S());
                    coeff_t c;
                    c.A = c_lo.x, c.gamma = c_lo.y, c.sigma = c_lo.z, c.rcutinv = c_lo.w;
                    c.rcut = c_hi.x, c.half_s = c_hi.y, c.hA = c_hi.z, c.rcutsq = c_hi.w;
                    dpd_pair<S, VIRIAL>(e, vi, vj, tag_i, tag_j, c, rk, acc);
                    }
                }
            q_head += min(cnt, (unsigned int)TPP);
            if (all_ready && __all_sync(FULL, light_done & (q_head == q_tail)))
                break;
            }
        }
    }

// neighbour data straight from global memory; the ring carries the global index, the type pair rides in a side ring
template<class S> struct GlobalSource
    {
    typedef typename vec4<S>::type V4;
    static constexpr bool kPackedTypePair = false;
    const unsigned int* row;
    const V4* pos_;
    const V4* vel_;
    const unsigned int* tag_;
    __device__ __forceinline__ unsigned int index(unsigned int k) const
        {
        return __ldg(row + k);
        }
    __device__ __forceinline__ V4 pos(unsigned int j) const
        {
        return ld4(pos_ + j);
        }
    __device__ __forceinline__ V4 vel(unsigned int j) const
        {
        return ld4(vel_ + j);
        }
    __device__ __forceinline__ unsigned int tag(unsigned int j) const
        {
        return __ldg(tag_ + j);
        }
    __device__ __forceinline__ static unsigned int pack(unsigned int j, unsigned int)
        {
        return j;
        }
    __device__ __forceinline__ static unsigned int index_of(unsigned int packed)
        {
        return packed;
        }
    __device__ __forceinline__ static unsigned int type_pair_of(unsigned int)
        {
        return 0u;
        }
    };

// neighbour data from the tile's shared-memory copy, addressed by 16-bit slots; the ring carries slot | type pair << 16
template<class S> struct TileSource
    {
    typedef typename vec4<S>::type V4;
    static constexpr bool kPackedTypePair = true;
    const uint16_t* row;
    unsigned int pos_addr, vel_addr, tag_addr; // shared addresses of the staged arrays
    __device__ __forceinline__ unsigned int index(unsigned int k) const
        {
        return __ldg(row + k);
        }
    __device__ __forceinline__ V4 pos(unsigned int j) const
        {
        return lds4(pos_addr, j, S());
        }
    __device__ __forceinline__ V4 vel(unsigned int j) const
        {
        return lds4(vel_addr, j, S());
        }
    __device__ __forceinline__ unsigned int tag(unsigned int j) const
        {
        return lds32_at(tag_addr + j * 4u);
        }
    __device__ __forceinline__ static unsigned int pack(unsigned int j, unsigned int tp)
        {
        return j | (tp << 16);
        }
    __device__ __forceinline__ static unsigned int index_of(unsigned int packed)
        {
        return packed & 0xffffu;
        }
    __device__ __forceinline__ static unsigned int type_pair_of(unsigned int packed)
        {
        return packed >> 16;
        }
    };

// shared-memory footprint of the rings of a CTA: one ring of QCAP + 1 entries per group (+ the type-pair side ring)
template<class S, int TPP> __host__ __device__ constexpr unsigned int dpd_ring_stride()
    {
    return ((unsigned int)AZB_DPD_QMUL * TPP + 1u) * (unsigned int)sizeof(typename vec4<S>::type);
    }

template<class S, bool VIRIAL, int TPP, bool FASTBOX, bool SINGLE>
__global__ void __launch_bounds__(256)
    dpd_force_kernel_q(const DPDArgsDev<S> d, const DPDThermo::Params<S>* __restrict__ params, const S wrap_sq)
    {
    typedef DPDThermo::Coeff<S> coeff_t;
    typedef typename vec4<S>::type V4;
    constexpr unsigned int QCAP = AZB_DPD_QMUL * TPP;
    const PairArgsDev<S>& a = d.pair;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int T2 = a.ntypes * a.ntypes;
    const unsigned int groups = blockDim.x / TPP;
    coeff_t* s_coeff = reinterpret_cast<coeff_t*>(smem_raw);
    V4* s_queue = reinterpret_cast<V4*>(s_coeff + T2);
    unsigned int* s_qtp = reinterpret_cast<unsigned int*>(s_queue + groups * (QCAP + 1u));
    for (unsigned int t = threadIdx.x; t < T2; t += blockDim.x)
        s_coeff[t] = DPDThermo::make(params[t], a.rcutsq[t], d.deltaT, d.T);
    __syncthreads();

    const unsigned int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int idx = gtid / TPP;
    const unsigned int gid = threadIdx.x / TPP;
    const bool active = idx < a.N;

    unsigned int nn = 0;
    GlobalSource<S> src = {a.nlist, a.pos, d.vel, d.tag};
    V4 pi = pack_entry<S>(S(0), S(0), S(0), 0u), vi = pi;
    unsigned int tag_i = 0, trow = 0;
    if (active)
        {
        nn = __ldg(a.n_neigh + idx);
        src.row = a.nlist + __ldg(a.head_list + idx);
        pi = ld4(a.pos + idx);
        vi = ld4(d.vel + idx);
        tag_i = __ldg(d.tag + idx);
        trow = SINGLE ? 0u : type_of(pi.w) * a.ntypes;
        }
    const coeff_t c0 = s_coeff[0]; // single-type systems keep the one coefficient set in registers
    PairAccum<S, VIRIAL> acc;
    dpd_walk_row<S, VIRIAL, TPP, FASTBOX, SINGLE, false>(src, nn, pi, vi, tag_i, trow, a.box, wrap_sq, smem_addr(s_coeff), c0,
                                                  smem_addr(s_queue) + gid * dpd_ring_stride<S, TPP>(),
                                                  smem_addr(s_qtp) + gid * (QCAP + 1u) * 4u, d.rk, acc);
    acc.template reduce<TPP>();
    if (active && threadIdx.x % TPP == 0)
        acc.store(a, idx);
    }

// One CTA per tile: stage pos/vel/tag of the tile's distinct neighbours (ascending indices -> coalesced runs; four
// gathers per thread in flight), then every group walks tile_size / (blockDim / TPP) particles against the staged copy.
template<class S, bool VIRIAL, int TPP, bool FASTBOX, bool SINGLE>
__global__ void AZB_DPD_TILE_BOUNDS
    dpd_tile_kernel(const DPDArgsDev<S> d, const TilePlanDev plan, const DPDThermo::Params<S>* __restrict__ params, const S wrap_sq)
    {
    typedef DPDThermo::Coeff<S> coeff_t;
    typedef typename vec4<S>::type V4;
    constexpr unsigned int QCAP = AZB_DPD_QMUL * TPP;
    const PairArgsDev<S>& a = d.pair;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int T2 = a.ntypes * a.ntypes;
    const unsigned int groups = blockDim.x / TPP;
    const unsigned int gid = threadIdx.x / TPP;
    V4* s_pos = reinterpret_cast<V4*>(smem_raw);
    V4* s_vel = s_pos + plan.smem_uniq;
    coeff_t* s_coeff = reinterpret_cast<coeff_t*>(s_vel + plan.smem_uniq);
    V4* s_queue = reinterpret_cast<V4*>(s_coeff + T2);
    unsigned int* s_tag = reinterpret_cast<unsigned int*>(s_queue + groups * (QCAP + 1u));

    const unsigned int tile = plan.tile_order ? __ldg(plan.tile_order + plan.tile_begin + blockIdx.x) : plan.tile_begin + blockIdx.x;
    const unsigned int U = __ldg(plan.uniq_count + tile);
    const unsigned int* uq = plan.uniq + (size_t)tile * plan.uniq_capacity;
    // Warm L2 with what the walk will read from global memory: the slot rows and the row headers / positions /
    // velocities / tags of the tile's own particles. Nothing waits on a prefetch, so the requests overlap the staging
    // gather below, and the walk's loads then cost an L2 hit instead of a DRAM round trip.
    for (unsigned int p = threadIdx.x; p < plan.tile_size; p += blockDim.x)
        {
        const unsigned int idx = tile * plan.tile_size + p;
        if (idx < a.N)
            {
            const unsigned int nn = __ldg(a.n_neigh + idx);
            const char* r = reinterpret_cast<const char*>(plan.slots + slot_row_begin(plan, a.head_list, idx, 1));
            for (unsigned int off = 0; off < nn * 2u; off += 128u)
                prefetch_l2(r + off);
            if (p % 8u == 0u)
                {
                prefetch_l2(a.pos + idx);
                prefetch_l2(d.vel + idx);
                }
            if (p % 32u == 0u)
                prefetch_l2(d.tag + idx);
            }
        }
    for (unsigned int s0 = threadIdx.x; s0 < U; s0 += 4 * blockDim.x)
        {
        unsigned int g[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            g[u] = (s0 + u * blockDim.x < U) ? __ldg(uq + s0 + u * blockDim.x) : 0xffffffffu;
        V4 pv[4], vv[4];
        unsigned int tv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (g[u] != 0xffffffffu)
                {
                pv[u] = ld4(a.pos + g[u]);
                vv[u] = ld4(d.vel + g[u]);
                tv[u] = __ldg(d.tag + g[u]);
                }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (g[u] != 0xffffffffu)
                {
                const unsigned int s = s0 + u * blockDim.x;
                s_pos[s] = pv[u];
                s_vel[s] = vv[u];
                s_tag[s] = tv[u];
                }
        }
    for (unsigned int t = threadIdx.x; t < T2; t += blockDim.x)
        s_coeff[t] = DPDThermo::make(params[t], a.rcutsq[t], d.deltaT, d.T);
    __syncthreads();
    const coeff_t c0 = s_coeff[0];
    const unsigned int coeff_addr = smem_addr(s_coeff);
    const unsigned int q_addr = smem_addr(s_queue) + gid * dpd_ring_stride<S, TPP>();

    TileSource<S> src = {plan.slots, smem_addr(s_pos), smem_addr(s_vel), smem_addr(s_tag)};
    for (unsigned int p = gid; p < plan.tile_size; p += groups) // warp-uniform trip count (host picks blockDim)
        {
        const unsigned int idx = tile * plan.tile_size + p;
        const bool active = idx < a.N;
        unsigned int nn = 0;
        V4 pi = pack_entry<S>(S(0), S(0), S(0), 0u), vi = pi;
        unsigned int tag_i = 0, trow = 0;
        if (active)
            {
            nn = __ldg(a.n_neigh + idx);
            src.row = plan.slots + slot_row_begin(plan, a.head_list, idx, 1);
            pi = ld4(a.pos + idx);
            vi = ld4(d.vel + idx);
            tag_i = __ldg(d.tag + idx);
            trow = SINGLE ? 0u : type_of(pi.w) * a.ntypes;
            }
        PairAccum<S, VIRIAL> acc;
        dpd_walk_row<S, VIRIAL, TPP, FASTBOX, SINGLE, AZB_DPD_LIGHT2>(src, nn, pi, vi, tag_i, trow, a.box, wrap_sq, coeff_addr, c0, q_addr, 0u, d.rk, acc);
        acc.template reduce<TPP>();
        if (active && threadIdx.x % TPP == 0)
            acc.store(a, idx);
        }
    }

// ---- tile kernel with a PRIVATE in-range stack per lane (no ring, no votes) -----------------------------------------------------
// The ring walk above shares a queue between the TPP lanes of a particle and pays for it per visit: two ballots, popc ranks, a
// warp-wide readiness vote, a warp-wide max, a __syncwarp. With the neighbour data in shared memory nothing is gained from several
// lanes reading consecutive list entries, so here every lane walks its own share of the row (entries lane, lane + TPP, ...; TPP = 1:
// the whole row) in plain SIMT: light — four entries at a time, cutoff test, in-range slots pushed (16 bits each) onto the lane's own
// stack in shared memory; heavy — the lane drains its stack, re-reading the neighbour's staged position (cheaper than carrying the
// separation through the stack). The warp reconverges after each phase, so a phase costs its longest lane; there is no other
// coupling between lanes. A lane's pairs are summed in list order, the TPP partial sums by shuffle.
constexpr unsigned int DPD_STACK_CAP = 16u; // in-range entries per lane between two drains

template<class S, bool VIRIAL, int TPP, bool FASTBOX, bool SINGLE>
__global__ void __launch_bounds__(256)
    dpd_stack_tile_kernel(const DPDArgsDev<S> d, const TilePlanDev plan, const DPDThermo::Params<S>* __restrict__ params, const S wrap_sq)
    {
    typedef DPDThermo::Coeff<S> coeff_t;
    typedef typename vec4<S>::type V4;
    const PairArgsDev<S>& a = d.pair;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int T2 = a.ntypes * a.ntypes;
    const unsigned int groups = blockDim.x / TPP;
    const unsigned int gid = threadIdx.x / TPP;
    const unsigned int lane = threadIdx.x % TPP;
    V4* s_pos = reinterpret_cast<V4*>(smem_raw);
    V4* s_vel = s_pos + plan.smem_uniq;
    coeff_t* s_coeff = reinterpret_cast<coeff_t*>(s_vel + plan.smem_uniq);
    unsigned int* s_tag = reinterpret_cast<unsigned int*>(s_coeff + T2);
    uint16_t* s_stack = reinterpret_cast<uint16_t*>(s_tag + plan.smem_uniq); // [DPD_STACK_CAP][blockDim.x]

    const unsigned int tile = plan.tile_order ? __ldg(plan.tile_order + plan.tile_begin + blockIdx.x) : plan.tile_begin + blockIdx.x;
    const unsigned int U = __ldg(plan.uniq_count + tile);
    const unsigned int* uq = plan.uniq + (size_t)tile * plan.uniq_capacity;
    // The group's first particle (the only one when blockDim / TPP == tile_size, the default): its row header, own position /
    // velocity / tag and its first slot packet do not depend on the staged tile — their memory round trips are started here and
    // hide behind the staging gather. The rest of a longer row is warmed in L2.
    unsigned int nn_next = 0, tag_next = 0;
    const uint16_t* row_next = plan.slots;
    V4 pi_next = pack_entry<S>(S(0), S(0), S(0), 0u), vi_next = pi_next;
    uint2 pk_next = make_uint2(0u, 0u);
    auto load_header = [&](const unsigned int p)
    {
        const unsigned int idx = tile * plan.tile_size + p;
        nn_next = 0;
        if (p < plan.tile_size && idx < a.N)
            {
            nn_next = __ldg(a.n_neigh + idx);
            row_next = plan.slots + slot_row_begin(plan, a.head_list, idx, 1);
            pi_next = ld4(a.pos + idx);
            vi_next = ld4(d.vel + idx);
            tag_next = __ldg(d.tag + idx);
            if (4u * lane < nn_next)
                pk_next = __ldg(reinterpret_cast<const uint2*>(row_next + 4u * lane));
            if (lane == 0)
                for (unsigned int off = 128u; off < nn_next * 2u; off += 128u)
                    prefetch_l2(reinterpret_cast<const char*>(row_next) + off);
            }
    };
    // (one lane per particle only: with more lanes the extra live registers cost a resident CTA — C2 0.204 -> 0.228 ms)
    constexpr bool HOIST = TPP == 1;
    if (HOIST)
        load_header(gid);
    else
        for (unsigned int p = threadIdx.x; p < plan.tile_size; p += blockDim.x)
            {
            const unsigned int idx = tile * plan.tile_size + p;
            if (idx < a.N)
                {
                const unsigned int nn = __ldg(a.n_neigh + idx);
                const char* r = reinterpret_cast<const char*>(plan.slots + slot_row_begin(plan, a.head_list, idx, 1));
                for (unsigned int off = 0; off < nn * 2u; off += 128u)
                    prefetch_l2(r + off);
                if (p % 8u == 0u)
                    {
                    prefetch_l2(a.pos + idx);
                    prefetch_l2(d.vel + idx);
                    }
                if (p % 32u == 0u)
                    prefetch_l2(d.tag + idx);
                }
            }
    for (unsigned int s0 = threadIdx.x; s0 < U; s0 += 4 * blockDim.x)
        {
        unsigned int g[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            g[u] = (s0 + u * blockDim.x < U) ? __ldg(uq + s0 + u * blockDim.x) : 0xffffffffu;
        V4 pv[4], vv[4];
        unsigned int tv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (g[u] != 0xffffffffu)
                {
                pv[u] = ld4(a.pos + g[u]);
                vv[u] = ld4(d.vel + g[u]);
                tv[u] = __ldg(d.tag + g[u]);
                }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (g[u] != 0xffffffffu)
                {
                const unsigned int s = s0 + u * blockDim.x;
                s_pos[s] = pv[u];
                s_vel[s] = vv[u];
                s_tag[s] = tv[u];
                }
        }
    for (unsigned int t = threadIdx.x; t < T2; t += blockDim.x)
        s_coeff[t] = DPDThermo::make(params[t], a.rcutsq[t], d.deltaT, d.T);
    __syncthreads();
    const coeff_t c0 = s_coeff[0];
    const unsigned int pos_addr = smem_addr(s_pos), vel_addr = smem_addr(s_vel), tag_addr = smem_addr(s_tag);
    const unsigned int coeff_addr = smem_addr(s_coeff);
    uint16_t* stack = s_stack + threadIdx.x; // entry q at stack[q * blockDim.x]
    const unsigned int sstride = blockDim.x;

    for (unsigned int p = gid; p < plan.tile_size; p += groups)
        {
        const unsigned int idx = tile * plan.tile_size + p;
        const bool active = idx < a.N;
        PairAccum<S, VIRIAL> acc;
        if (!HOIST)
            load_header(p);
        if (active)
            {
            const unsigned int nn = nn_next;
            const uint16_t* row = row_next;
            const V4 pi = pi_next;
            const V4 vi = vi_next;
            const unsigned int tag_i = tag_next;
            uint2 pk = pk_next;
            const unsigned int trow = SINGLE ? 0u : type_of(pi.w) * a.ntypes;
            // my entries: blocks of four consecutive list entries (one 8-byte load: slot rows start 16-byte aligned and are
            // padded to a multiple of four), block b of lane l starting at entry 4 (b TPP + l)
            unsigned int k = 4u * lane, cnt = 0;
            auto eval = [&](const unsigned int jj)
            {
                const V4 pj = lds4(pos_addr, jj, S());
                S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                separation<FASTBOX>(a.box, wrap_sq, dx, dy, dz);
                const V4 e = pack_entry<S>(dx, dy, dz, 0u);
                const V4 vj = lds4(vel_addr, jj, S());
                const unsigned int tag_j = lds32_at(tag_addr + jj * 4u);
                if (SINGLE)
                    dpd_pair<S, VIRIAL>(e, vi, vj, tag_i, tag_j, c0, d.rk, acc);
                else
                    {
                    const unsigned int ca = coeff_addr + (trow + type_of(pj.w)) * (unsigned int)sizeof(coeff_t);
                    const V4 c_lo = lds4_at(ca, S()), c_hi = lds4_at(ca + (unsigned int)sizeof(V4), S());
                    coeff_t c;
                    c.A = c_lo.x, c.gamma = c_lo.y, c.sigma = c_lo.z, c.rcutinv = c_lo.w;
                    c.rcut = c_hi.x, c.half_s = c_hi.y, c.hA = c_hi.z, c.rcutsq = c_hi.w;
                    dpd_pair<S, VIRIAL>(e, vi, vj, tag_i, tag_j, c, d.rk, acc);
                    }
            };
            for (;;)
                {
                // ---- light: four of my entries per pass
                while (k < nn && cnt + 4u <= DPD_STACK_CAP)
                    {
                    const unsigned int sl[4] = {pk.x & 0xffffu, pk.x >> 16, pk.y & 0xffffu, pk.y >> 16};
                    if (k + 4u * TPP < nn) // my next block, one pass ahead (as a plain __ldg the compiler moves the load across the
                                           // back edge to the top of the NEXT pass, next to its use: 15 % of the kernel's stall samples)
                        pk = ldg_packet_here(reinterpret_cast<const uint2*>(row + k + 4u * TPP));
                    bool in[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        {
                        const V4 pj = lds4(pos_addr, sl[u], S());
                        S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                        S rc = c0.rcutsq;
                        if (!SINGLE)
                            rc = lds_scalar(coeff_addr + (trow + type_of(pj.w)) * (unsigned int)sizeof(coeff_t) + 7u * (unsigned int)sizeof(S), S());
                        in[u] = (k + u < nn) & (separation<FASTBOX>(a.box, wrap_sq, dx, dy, dz) < rc);
                        }
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (in[u])
                            {
                            stack[cnt * sstride] = (uint16_t)sl[u];
                            ++cnt;
                            }
                    k += 4u * TPP;
                    }
                // ---- heavy: drain my stack, two pairs per pass (two independent Philox chains in flight)
                unsigned int q = 0;
                for (; q + 2u <= cnt; q += 2u)
                    {
                    const unsigned int ja = stack[q * sstride], jb = stack[(q + 1u) * sstride];
                    eval(ja);
                    eval(jb);
                    }
                if (q < cnt)
                    eval(stack[q * sstride]);
                cnt = 0;
                if (k >= nn)
                    break;
                }
            }
        acc.template reduce<TPP>();
        if (active && lane == 0)
            acc.store(a, idx);
        if (HOIST && p + groups < plan.tile_size)
            load_header(p + groups); // (more particles than groups: not the default launch shape)
        }
    }
    } // namespace azb
