// azb_plj_packed.cuh — PerturbedLennardJones tile kernel with packed FP32 arithmetic (sm_100 FFMA2 / FMUL2 / FADD2).
//
// The generic tile kernel (azb_pair_kernel.cuh) is issue-bound on the 4 M-particle LJ workload: ~44 instructions per
// list entry, ~30 of them scalar FP32. Scalar FFMA already issues at the full 128 lanes/clk/SM on B200
// (tools/microbench/ffma2.cu), so the FMA pipe is not the limit — the issue slots are. Here every lane evaluates two
// neighbours at a time in the two halves of 64-bit register pairs (`fma.rn.f32x2`), which halves the FP issue slots.
// Out-of-range and padding entries are not branched around: their r^2 is replaced by a huge value, which drives
// r^-6 to zero (flush-to-zero), so force and energy vanish and the accumulators are untouched. Scope: float, one
// particle type, orthorhombic box, modes none / shift; everything else takes the generic kernel.
// Per-pair math: /root/reference/src/PairEvaluatorPerturbedLennardJones.h:117-155.
#pragma once
#include "azb_pair_kernel.cuh"

namespace azb
    {
__device__ __forceinline__ float2 f2(float a, float b)
    {
    return make_float2(a, b);
    }

template<int SHIFT, bool VIRIAL, int TPP>
__global__ void __launch_bounds__(256, 4) plj_packed_tile_kernel(const PairArgsDev<float> a, const TilePlanDev plan,
                                                              const PerturbedLJ::Params<float>* __restrict__ params, const float wrap_sq)
    {
    extern __shared__ __align__(32) unsigned char smem_raw[];
    float4* s_pos = reinterpret_cast<float4*>(smem_raw);

    const unsigned int tile = plan.tile_order ? __ldg(plan.tile_order + plan.tile_begin + blockIdx.x) : plan.tile_begin + blockIdx.x;
    const unsigned int U = __ldg(plan.uniq_count + tile);
    const unsigned int* uq = plan.uniq + (size_t)tile * plan.uniq_capacity;
    // the group's first particle: its row header and first slot packet do not depend on the staged tile, so their two memory
    // round trips are started before the staging gather and hide behind it
    constexpr unsigned int M = 4 * TPP;
    const unsigned int groups = blockDim.x / TPP;
    const unsigned int gid = threadIdx.x / TPP;
    const unsigned int lane = threadIdx.x % TPP;
    unsigned int nn_next = 0;
    const uint2* row_next = nullptr;
    float4 pi_next = make_float4(0.f, 0.f, 0.f, 0.f);
    uint2 pkt_next = make_uint2(0u, 0u);
    auto load_header = [&](const unsigned int p)
    {
        const unsigned int idx = tile * plan.tile_size + p;
        nn_next = 0;
        if (p < plan.tile_size && idx < a.N)
            {
            nn_next = __ldg(a.n_neigh + idx);
            row_next = reinterpret_cast<const uint2*>(plan.slots + slot_row_begin(plan, a.head_list, idx, TPP)) + lane;
            pi_next = ld4(a.pos + idx);
            if (nn_next)
                pkt_next = __ldg(row_next);
            }
    };
    load_header(gid);
    warm_tile_rows(a, plan, tile, TPP);
    stage_rows(s_pos, a.pos, uq, U);
    // one particle type: every thread derives the coefficients itself (same expressions as the generic kernel's table)
    const float rc = __ldg(a.rcutsq);
    const PerturbedLJ::Coeff<float> c = PerturbedLJ::make(params[0], rc, SHIFT == 1);
    const float rc0 = PerturbedLJ::active(c) ? rc : -1.0f;
    __syncthreads();
    const unsigned int pos_base = smem_addr(s_pos);

    const float2 k12lj1 = f2(12.0f * c.lj1, 12.0f * c.lj1), km6lj2 = f2(-6.0f * c.lj2, -6.0f * c.lj2);
    const float2 klj1 = f2(c.lj1, c.lj1), kmlj2 = f2(-c.lj2, -c.lj2), kminus1 = f2(-1.0f, -1.0f);
    // energy added on top of m * u: inside the WCA range (m = 1) the WCA shift, outside (m = lambda) nothing; minus the
    // cutoff shift in mode shift
    const float add_wca = c.wca_shift - c.e_shift, add_far = -c.e_shift;
    const float BIG = 1.0e30f;

    for (unsigned int p = gid; p < plan.tile_size; p += groups)
        {
        const unsigned int idx = tile * plan.tile_size + p;
        const bool active = idx < a.N;
        float2 fx = f2(0, 0), fy = fx, fz = fx, pe = fx, vxx = fx, vxy = fx, vxz = fx, vyy = fx, vyz = fx, vzz = fx;
        if (active)
            {
            const unsigned int nn = nn_next;
            const uint2* row = row_next;
            const float4 pi = pi_next;
            const float2 pix = f2(pi.x, pi.x), piy = f2(pi.y, pi.y), piz = f2(pi.z, pi.z);

            auto visit2 = [&](const unsigned int sA, const unsigned int sB, const bool vA, const bool vB)
            {
                const float4 pA = lds4(pos_base, sA, float()), pB = lds4(pos_base, sB, float());
                float2 dx = __ffma2_rn(f2(pA.x, pB.x), kminus1, pix);
                float2 dy = __ffma2_rn(f2(pA.y, pB.y), kminus1, piy);
                float2 dz = __ffma2_rn(f2(pA.z, pB.z), kminus1, piz);
                float2 rsq = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
                if (fmaxf(rsq.x, rsq.y) > wrap_sq) // a pair straddling the periodic box: wrap (rare, real call)
                    {
                    const Sep<float> wA = wrapped_separation(a.box.Lx, a.box.Ly, a.box.Lz, a.box.Lxinv, a.box.Lyinv, a.box.Lzinv, a.box.flags,
                                                             dx.x, dy.x, dz.x);
                    const Sep<float> wB = wrapped_separation(a.box.Lx, a.box.Ly, a.box.Lz, a.box.Lxinv, a.box.Lyinv, a.box.Lzinv, a.box.flags,
                                                             dx.y, dy.y, dz.y);
                    dx = f2(wA.dx, wB.dx);
                    dy = f2(wA.dy, wB.dy);
                    dz = f2(wA.dz, wB.dz);
                    // same operation sequence as the unwrapped path: an entry's r^2 must not depend on whether its partner in
                    // the register pair (another entry, or row padding — which differs between plans) needed the wrap
                    rsq = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
                    }
                // entries that do not count (beyond the cutoff, padding) get a huge r^2: r^-6 flushes to zero
                const bool inA = vA & (rsq.x < rc0), inB = vB & (rsq.y < rc0);
                const float2 rq = f2(inA ? rsq.x : BIG, inB ? rsq.y : BIG);
                const bool wcaA = rq.x < c.rwcasq, wcaB = rq.y < c.rwcasq;
                const float2 m = f2(wcaA ? 1.0f : c.lam, wcaB ? 1.0f : c.lam);
                float2 add;
                if (SHIFT == 1)
                    add = f2(inA ? (wcaA ? add_wca : add_far) : 0.0f, inB ? (wcaB ? add_wca : add_far) : 0.0f);
                else
                    add = f2(wcaA ? add_wca : 0.0f, wcaB ? add_wca : 0.0f);
                const float2 r2inv = f2(rcp(rq.x), rcp(rq.y));
                const float2 r6inv = __fmul2_rn(__fmul2_rn(r2inv, r2inv), r2inv);
                float2 f = __fmul2_rn(__fmul2_rn(r2inv, r6inv), __ffma2_rn(r6inv, k12lj1, km6lj2));
                const float2 u = __fmul2_rn(r6inv, __ffma2_rn(r6inv, klj1, kmlj2));
                f = __fmul2_rn(f, m);
                pe = __fadd2_rn(pe, __ffma2_rn(u, m, add));
                fx = __ffma2_rn(dx, f, fx);
                fy = __ffma2_rn(dy, f, fy);
                fz = __ffma2_rn(dz, f, fz);
                if (VIRIAL)
                    {
                    const float2 hx = __fmul2_rn(f, dx), hy = __fmul2_rn(f, dy), hz = __fmul2_rn(f, dz);
                    vxx = __ffma2_rn(hx, dx, vxx);
                    vxy = __ffma2_rn(hx, dy, vxy);
                    vxz = __ffma2_rn(hx, dz, vxz);
                    vyy = __ffma2_rn(hy, dy, vyy);
                    vyz = __ffma2_rn(hy, dz, vyz);
                    vzz = __ffma2_rn(hz, dz, vzz);
                    }
            };

            const unsigned int full = nn / M;
            const unsigned int nblk = (nn + M - 1u) / M;
            uint2 next = pkt_next;
            load_header(p + groups); // the group's next particle, if the tile has more particles than the CTA has groups
            for (unsigned int b = 0; b < full; ++b)
                {
                const uint2 s4 = next;
                if (b + 1 < nblk)
                    next = __ldg(row + (b + 1) * TPP);
                visit2(s4.x & 0xffffu, s4.x >> 16, true, true);
                visit2(s4.y & 0xffffu, s4.y >> 16, true, true);
                }
            if (full < nblk) // the ragged last block: padding slots are 0 (a valid address) and masked out
                {
                const unsigned int e0 = full * M + lane;
                visit2(next.x & 0xffffu, next.x >> 16, e0 < nn, e0 + TPP < nn);
                visit2(next.y & 0xffffu, next.y >> 16, e0 + 2 * TPP < nn, e0 + 3 * TPP < nn);
                }
            }
        PairAccum<float, VIRIAL> acc;
        acc.fx = fx.x + fx.y;
        acc.fy = fy.x + fy.y;
        acc.fz = fz.x + fz.y;
        acc.pe = pe.x + pe.y;
        if (VIRIAL)
            {
            acc.vxx = vxx.x + vxx.y;
            acc.vxy = vxy.x + vxy.y;
            acc.vxz = vxz.x + vxz.y;
            acc.vyy = vyy.x + vyy.y;
            acc.vyz = vyz.x + vyz.y;
            acc.vzz = vzz.x + vzz.y;
            }
        acc.template reduce<TPP>();
        if (active && lane == 0)
            acc.store(a, idx);
        }
    }
    } // namespace azb
