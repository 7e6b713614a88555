// azb_ring_walk.cuh — the two-phase (light / heavy) row walk of the DPD tile kernel (azb_dpd_kernel.cuh, dpd_walk_row), with the
// per-pair work as a functor, for pair evaluators that are expensive next to the cutoff test: TwoPatchMorse (~150 instructions per
// in-range pair, 51 % of C4a's list entries in range — evaluated under `if (in range)` half the lanes idle through the block).
//   light: two list entries per lane and visit — slot, staged position, separation, cutoff test; the in-range ones
//          (dx, dy, dz, slot | type pair << 16) are compacted by ballot + popc into the group's ring in shared memory;
//   heavy: whenever every group of the warp holds >= TPP queued entries (or has finished its row) each lane pops one entry and
//          evaluates it.
// The order in which a particle's in-range pairs are summed is the ring order: lane l of the group takes entries l, l + TPP, ...
// of the row's in-range subsequence, so results do not depend on how visits and passes interleave.
// dpd_walk_row keeps its own copy (its measured behaviour is pinned by profiles/); this one differs only in the functor and in
// the inclusive cutoff test anisotropic evaluators make (AnisoPairEvaluatorTwoPatchMorse.h:135 rejects only rsq > rcutsq).
#pragma once
#include "azb_math.cuh"
#include "azb_pair_kernel.cuh"

namespace azb
    {
template<class S> __device__ __forceinline__ typename vec4<S>::type ring_entry(S dx, S dy, S dz, unsigned int w);
template<> __device__ __forceinline__ float4 ring_entry<float>(float dx, float dy, float dz, unsigned int w)
    {
    return make_float4(dx, dy, dz, __uint_as_float(w));
    }
template<> __device__ __forceinline__ double4a ring_entry<double>(double dx, double dy, double dz, unsigned int w)
    {
    double4a e;
    e.x = dx;
    e.y = dy;
    e.z = dz;
    e.w = __longlong_as_double((long long)w);
    return e;
    }
__device__ __forceinline__ unsigned int ring_word(float w)
    {
    return __float_as_uint(w);
    }
__device__ __forceinline__ unsigned int ring_word(double w)
    {
    return (unsigned int)__double_as_longlong(w);
    }

// ring capacity per group, in entries
template<int TPP> __host__ __device__ constexpr unsigned int ring_capacity()
    {
    return 4u * TPP;
    }
// distance between the rings of neighbouring groups (one entry of padding: equal positions fall into different banks)
template<class S, int TPP> __host__ __device__ constexpr unsigned int ring_stride()
    {
    return (ring_capacity<TPP>() + 1u) * (unsigned int)sizeof(typename vec4<S>::type);
    }

// All 32 lanes of the warp call this together (warp votes inside); lanes of idle groups pass nn = 0.
//   row        the particle's slot row (plain order, lane_width 1)
//   pos_addr   shared address of the staged positions
//   rc_addr, rc_stride   shared address / byte stride of the per-type-pair cutoff^2 table (unused when SINGLE: rc0)
//   heavy(e, slot, tp)   evaluates one in-range pair: e = (dx, dy, dz, .) with the minimum image applied
template<class S, int TPP, bool FASTBOX, bool SINGLE, bool STRICT, class Heavy>
__device__ __forceinline__ void ring_walk_row(const uint16_t* __restrict__ row, const unsigned int nn, const typename vec4<S>::type& pi,
                                              const unsigned int trow, const Box<S>& box, const S wrap_sq, const unsigned int pos_addr,
                                              const unsigned int rc_addr, const unsigned int rc_stride, const S rc0, const unsigned int q_addr,
                                              Heavy&& heavy)
    {
    typedef typename vec4<S>::type V4;
    constexpr unsigned int QCAP = ring_capacity<TPP>();
    constexpr unsigned int FULL = 0xffffffffu;
    constexpr unsigned int STEP = 2u * TPP;
    const unsigned int lane = threadIdx.x % TPP;
    const unsigned int gshift = (threadIdx.x & 31u) & ~(unsigned int)(TPP - 1);
    unsigned int gmask, ltmask;
    asm volatile("mov.u32 %0, %1;" : "=r"(gmask) : "r"(((TPP == 32) ? FULL : ((1u << TPP) - 1u)) << gshift));
    asm volatile("mov.u32 %0, %1;" : "=r"(ltmask) : "r"(((1u << lane) - 1u) << gshift));
    const unsigned int nn_warp = __reduce_max_sync(FULL, nn); // longest row among the warp's groups

    unsigned int q_head = 0, q_tail = 0; // group-uniform ring state (monotonic; position = counter mod QCAP)
    unsigned int k = lane;
    unsigned int j0 = 0, j1 = 0;
    if (k < nn)
        j0 = __ldg(row + k);
    if (k + TPP < nn)
        j1 = __ldg(row + k + TPP);
    bool room = true; // warp-uniform: every ring can take another light visit
    for (;;)
        {
        if (room && k - lane < nn_warp)
            {
            const bool have0 = k < nn, have1 = k + TPP < nn;
            const unsigned int jc0 = j0, jc1 = j1;
            k += STEP;
            if (k < nn)
                j0 = __ldg(row + k);
            if (k + TPP < nn)
                j1 = __ldg(row + k + TPP);
            auto light = [&](const unsigned int jc, const bool have, V4& entry) -> bool
            {
                const V4 pj = lds4(pos_addr, jc, S());
                S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                unsigned int tp = 0;
                S rc = rc0;
                if (!SINGLE)
                    {
                    tp = trow + type_of(pj.w);
                    rc = lds_scalar(rc_addr + tp * rc_stride, S());
                    }
                const S rsq = separation<FASTBOX>(box, wrap_sq, dx, dy, dz);
                const bool in = have & (STRICT ? (rsq < rc) : !(rsq > rc));
                entry = ring_entry<S>(dx, dy, dz, jc | (tp << 16));
                return in;
            };
            V4 e0, e1;
            const bool in0 = light(jc0, have0, e0);
            const bool in1 = light(jc1, have1, e1);
            const unsigned int b0 = __ballot_sync(FULL, in0);
            const unsigned int b1 = __ballot_sync(FULL, in1);
            const unsigned int n0 = __popc(b0 & gmask);
            const unsigned int at0 = (q_tail + __popc(b0 & ltmask)) & (QCAP - 1u);
            sts4_if(in0, q_addr + at0 * (unsigned int)sizeof(V4), e0);
            const unsigned int at1 = (q_tail + n0 + __popc(b1 & ltmask)) & (QCAP - 1u);
            sts4_if(in1, q_addr + at1 * (unsigned int)sizeof(V4), e1);
            q_tail += n0 + __popc(b1 & gmask);
            }
        const bool light_done = (k - lane) >= nn;
        const unsigned int cnt = q_tail - q_head;
        const bool all_ready = __all_sync(FULL, (cnt >= TPP) | light_done);
        // a light visit pushes up to STEP entries per group, a heavy pass pops TPP: the next light visit needs
        // cnt <= QCAP - STEP in every group
        const unsigned int maxcnt = __reduce_max_sync(FULL, cnt);
        room = true;
        if (all_ready || maxcnt > QCAP - STEP)
            {
            room = maxcnt <= QCAP - STEP + TPP;
            __syncwarp();
            if (lane < cnt)
                {
                const unsigned int at = (q_head + lane) & (QCAP - 1u);
                const V4 e = lds4_at(q_addr + at * (unsigned int)sizeof(V4), S());
                const unsigned int packed = ring_word(e.w);
                heavy(e, packed & 0xffffu, packed >> 16);
                }
            q_head += min(cnt, (unsigned int)TPP);
            if (all_ready && __all_sync(FULL, light_done & (q_head == q_tail)))
                break;
            }
        }
    }
    } // namespace azb
