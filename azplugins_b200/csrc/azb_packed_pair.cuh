// azb_packed_pair.cuh — tile kernel with packed FP32 arithmetic (sm_100 FFMA2 / FMUL2 / FADD2) for isotropic pair potentials
// other than PerturbedLennardJones (which has its own, azb_plj_packed.cuh — same walk, kept separate so that its measured
// behaviour does not move). Written for ExpandedYukawa, the dominant kernel of C5 (8 M particles, 57 list entries per
// particle, 58 % of them inside r_cut = 2.0): the generic tile kernel spends 48 thread-instructions per list entry at 82 %
// issue-active and 23 of 32 lanes (the in-range branch), i.e. it is issue-bound. Here every lane evaluates TWO neighbours at a
// time in the halves of 64-bit register pairs, and nothing branches on the cutoff: an entry that does not count (beyond the
// cutoff, row padding) is evaluated at r = r_cut and multiplied by a 0 / 1 mask.
// Scope: float, one particle type, orthorhombic box, modes none / shift; everything else takes the generic kernel.
//
// A policy `PK` supplies the per-pair math on pairs of entries:
//   typedef Pot;                          the potential (Params / Coeff / make / active)
//   struct K;  static K consts(coeff)     packed constants
//   static void eval2(K, rsq2, inA, inB, f2&, u2&)   force / r and pair energy (unshifted) of two entries, multiplied by the
//                                                    entries' 0 / 1 masks (inA, inB); rsq2 is finite and positive
//   static bool finite_beyond_cutoff(coeff, rcutsq)   eval2 is finite (no inf, no NaN) for every r^2 >= rcutsq
#pragma once
#include "azb_pair_kernel.cuh"
#include "azb_pair_potentials.cuh"
#include <type_traits>

namespace azb
    {
__device__ __forceinline__ float2 pk2(float a, float b)
    {
    return make_float2(a, b);
    }

// ExpandedYukawa — /root/reference/src/PairEvaluatorExpandedYukawa.h:92-115:
//   u = eps exp(-kappa (r - delta)) / (r - delta),  f / r = u (kappa + 1 / (r - delta)) / r
struct PackedExpandedYukawa
    {
    typedef ExpandedYukawa Pot;
    struct K
        {
        float2 mdelta, mkl2e, kappa;
        float eps;
        };
    // The kernel is bound by the FMA pipe (an FFMA2 holds it for two cycles: tools/microbench/mufu.cu; 84 % busy in the capture
    // of the first version), so what counts is the number of FP32 operations per entry, packed or not:
    //  * exp(-kappa (r - delta)) = ex2((r - delta) * (-kappa log2 e)) with the constant folded once per thread — HOOMD's own
    //    fast::exp on the GPU is __expf, the same single MUFU.EX2 on a rounded product. The argument's rounding error,
    //    |t| 6e-8 relative on the exponential, is below one float ulp of the pair force for |t| < 1 and always far below the
    //    1e-5 parity tolerance (a pair with |t| >> 1 contributes e^-|t| of the particle's force). exp_'s two-term correction
    //    (five more FP32 operations per entry) is left to the generic kernel;
    //  * epsilon rides on the 0 / 1 cutoff mask (a select on the ALU pipe instead of a multiplication), and the mask is
    //    applied once, to u: the force inherits it.
    __device__ __forceinline__ static K consts(const ExpandedYukawa::Coeff<float>& c)
        {
        K k;
        const float kl = (float)(-(double)c.kappa * 1.4426950408889634);
        k.mdelta = pk2(-c.delta, -c.delta);
        k.mkl2e = pk2(kl, kl);
        k.kappa = pk2(c.kappa, c.kappa);
        k.eps = c.epsilon;
        return k;
        }
    // eval2 stays finite for every r^2 >= r_cut^2 when delta < r_cut (r - delta > 0, the exponential only underflows): a listed
    // entry beyond the cutoff can then be evaluated where it is and multiplied by the zero mask, without the r_cut substitute
    __device__ __forceinline__ static bool finite_beyond_cutoff(const ExpandedYukawa::Coeff<float>& c, const float rcutsq)
        {
        return c.delta >= 0.0f && c.delta * c.delta < rcutsq;
        }
    // f = force / r and u = pair energy (unshifted) of two entries, both already multiplied by the entries' masks
    __device__ __forceinline__ static void eval2(const K& k, const float2 rsq, const bool inA, const bool inB, float2& f, float2& u)
        {
        const float2 rinv = pk2(rsqrt_(rsq.x), rsqrt_(rsq.y));
        const float2 rd = __ffma2_rn(rsq, rinv, k.mdelta);
        const float2 rdinv = pk2(rcp(rd.x), rcp(rd.y));
        const float2 t = __fmul2_rn(rd, k.mkl2e);
        float ea, eb;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ea) : "f"(t.x));
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(eb) : "f"(t.y));
        const float2 em = pk2(inA ? k.eps : 0.0f, inB ? k.eps : 0.0f);
        u = __fmul2_rn(__fmul2_rn(pk2(ea, eb), rdinv), em);
        f = __fmul2_rn(__fmul2_rn(u, __fadd2_rn(k.kappa, rdinv)), rinv);
        }
    };

template<class PK, int SHIFT, bool VIRIAL, int TPP>
__global__ void __launch_bounds__(128, 8) packed_pair_tile_kernel(const PairArgsDev<float> a, const TilePlanDev plan,
                                                                 const typename PK::Pot::template Params<float>* __restrict__ params,
                                                                 const float wrap_sq)
    {
    typedef typename PK::Pot Pot;
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
    const typename Pot::template Coeff<float> c = Pot::make(params[0], rc, SHIFT == 1);
    const float rc0 = Pot::active(c) ? rc : -1.0f; // an inactive pair (vanishing prefactor) never passes the cutoff test
    const float r_safe = rc > 0.0f ? rc : 1.0f;    // where masked-out entries are evaluated: any finite, harmless distance
    const bool no_subst = rc0 > 0.0f && PK::finite_beyond_cutoff(c, rc);
    typedef std::integral_constant<bool, true> true_c;
    typedef std::integral_constant<bool, false> false_c;
    const typename PK::K k = PK::consts(c);
    const float e_shift = SHIFT == 1 ? c.e_shift : 0.0f;
    __syncthreads();
    const bool never = wrap_sq < 0.0f; // an opaque, always false, warp-uniform predicate (wrap_sq is a squared length)
    const unsigned int pos_base = smem_addr(s_pos);
    const float2 kminus1 = pk2(-1.0f, -1.0f);

    for (unsigned int p = gid; p < plan.tile_size; p += groups)
        {
        const unsigned int idx = tile * plan.tile_size + p;
        const bool active = idx < a.N;
        float2 fx = pk2(0, 0), fy = fx, fz = fx, pe = fx, vxx = fx, vxy = fx, vxz = fx, vyy = fx, vyz = fx, vzz = fx;
        // Rows away from the periodic faces: a particle inside the box and farther than r_cut from every periodic face has no
        // in-range pair across a face (the raw separation of such a pair would exceed L - r_cut in that direction), so its walk
        // skips the per-visit wrap test; a listed pair that does straddle is beyond the cutoff and fails the cutoff test with
        // its raw separation as it does with the wrapped one (away_from_periodic_faces, azb_pair_kernel.cuh). Decided per warp (the rows of a warp walk in lockstep): ~86 % of the rows
        // of the 8 M system.
        const bool row_inner = away_from_periodic_faces(a.box, pi_next, r_safe);
        const bool nowrap = __all_sync(0xffffffffu, !active || row_inner);
        if (active)
            {
            const unsigned int nn = nn_next;
            const uint2* row = row_next;
            const float4 pi = pi_next;
            const float2 pix = pk2(pi.x, pi.x), piy = pk2(pi.y, pi.y), piz = pk2(pi.z, pi.z);

            // SUBST: entries that do not count are evaluated at r_safe; without it (full blocks of a potential that is finite
            // beyond the cutoff: every entry is a real list entry, r > 0) they are evaluated where they are — four selects less
            // WRAP: the lazy minimum-image test; without it (interior tile and own particle near the tile's reference) dx is final
            auto visit2 = [&](const unsigned int sA, const unsigned int sB, const bool vA, const bool vB, auto subst, auto wrap)
            {
                constexpr bool SUBST = decltype(subst)::value;
                constexpr bool WRAP = decltype(wrap)::value;
                const float4 pA = lds4(pos_base, sA, float()), pB = lds4(pos_base, sB, float());
                float2 dx = __ffma2_rn(pk2(pA.x, pB.x), kminus1, pix);
                float2 dy = __ffma2_rn(pk2(pA.y, pB.y), kminus1, piy);
                float2 dz = __ffma2_rn(pk2(pA.z, pB.z), kminus1, piz);
                float2 rsq = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
                if (WRAP && fmaxf(rsq.x, rsq.y) > wrap_sq) // a pair straddling the periodic box: wrap (rare, real call)
                    {
                    const Sep<float> wA = wrapped_separation(a.box.Lx, a.box.Ly, a.box.Lz, a.box.Lxinv, a.box.Lyinv, a.box.Lzinv, a.box.flags,
                                                             dx.x, dy.x, dz.x);
                    const Sep<float> wB = wrapped_separation(a.box.Lx, a.box.Ly, a.box.Lz, a.box.Lxinv, a.box.Lyinv, a.box.Lzinv, a.box.flags,
                                                             dx.y, dy.y, dz.y);
                    dx = pk2(wA.dx, wB.dx);
                    dy = pk2(wA.dy, wB.dy);
                    dz = pk2(wA.dz, wB.dz);
                    // same operation sequence as the unwrapped path: an entry's r^2 must not depend on whether its partner in
                    // the register pair (another entry, or row padding — which differs between plans) needed the wrap
                    rsq = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
                    }
                // entries that do not count (beyond the cutoff, padding) are evaluated at r_cut and masked to zero
                const bool inA = vA & (rsq.x < rc0), inB = vB & (rsq.y < rc0);
                const float2 rq = SUBST ? pk2(inA ? rsq.x : r_safe, inB ? rsq.y : r_safe) : rsq;
                float2 f, u;
                PK::eval2(k, rq, inA, inB, f, u);
                pe = __fadd2_rn(pe, u);
                if (SHIFT == 1)
                    pe = __fadd2_rn(pe, pk2(inA ? -e_shift : 0.0f, inB ? -e_shift : 0.0f));
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
            auto full_blocks = [&](auto subst, auto wrap)
            {
                for (unsigned int b = 0; b < full; ++b)
                    {
                    const uint2 s4 = next;
                    if (b + 1 < nblk)
                        next = ldg_packet_here(row + (b + 1) * TPP);
                    if (!decltype(wrap)::value && never) // (never taken: ends the basic block, so that the load above is issued here)
                        asm volatile("trap;");
                    visit2(s4.x & 0xffffu, s4.x >> 16, true, true, subst, wrap);
                    visit2(s4.y & 0xffffu, s4.y >> 16, true, true, subst, wrap);
                    }
            };
            if (nowrap)
                {
                if (no_subst)
                    full_blocks(false_c(), false_c());
                else
                    full_blocks(true_c(), false_c());
                }
            else
                {
                if (no_subst)
                    full_blocks(false_c(), true_c());
                else
                    full_blocks(true_c(), true_c());
                }
            if (full < nblk) // the ragged last block: padding slots are 0 (a valid address) and masked out
                {
                const unsigned int e0 = full * M + lane;
                visit2(next.x & 0xffffu, next.x >> 16, e0 < nn, e0 + TPP < nn, true_c(), true_c());
                visit2(next.y & 0xffffu, next.y >> 16, e0 + 2 * TPP < nn, e0 + 3 * TPP < nn, true_c(), true_c());
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

// SpecialTileKernel hook (azb_pair_launch.cuh) for a packed policy: modes none and shift, float, one type, lazy-wrap boxes
template<class PK, int SHIFT, bool VIRIAL> struct PackedPairLaunch
    {
    static bool try_launch(const PairArgsDev<float>& d, const TilePlanDev& plan, const void* d_params, float wrap_sq, bool lazy, bool single,
                           unsigned int tpp, unsigned int grid, unsigned int block, cudaStream_t stream)
        {
        static const bool disabled = getenv("AZB_NO_PACKED") != nullptr; // A/B switch against the generic tile kernel
        if (disabled || !lazy || !single || block > 128u)
            return false;
        typedef typename PK::Pot::template Params<float> param_t;
        const size_t smem = (size_t)plan.smem_uniq * sizeof(float4);
        if (smem > AZB_MAX_SMEM)
            return false;
#define AZB_PACKED_PAIR(TPP)                                                                                                   \
    do                                                                                                                         \
        {                                                                                                                      \
        if (ensure_smem<packed_pair_tile_kernel<PK, SHIFT, VIRIAL, TPP>>(smem) != 0)                                           \
            return false;                                                                                                      \
        packed_pair_tile_kernel<PK, SHIFT, VIRIAL, TPP><<<grid, block, smem, stream>>>(d, plan, (const param_t*)d_params, wrap_sq); \
        } while (0)
        switch (tpp)
            {
        case 2:
            AZB_PACKED_PAIR(2);
            break;
        case 4:
            AZB_PACKED_PAIR(4);
            break;
        default:
            return false;
            }
#undef AZB_PACKED_PAIR
        return true;
        }
    };
    } // namespace azb
