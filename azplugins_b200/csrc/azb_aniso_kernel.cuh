// azb_aniso_kernel.cuh — anisotropic pair force + torque kernel for TwoPatchMorse (replaces HOOMD's
// gpu_compute_pair_aniso_forces<E> as instantiated by /root/reference/src/AnisoPotentialPairGPUKernel.cu.inc:21-25;
// per-pair math: /root/reference/src/AnisoPairEvaluatorTwoPatchMorse.h:127-216).
//
// The patch director n = rotate(q, e_x) depends on one particle only: n_i is computed once per particle, n_j once
// per neighbour from the gathered quaternion. Energy shift U_M(r_cut) is hoisted into the per-type-pair table.
#pragma once
#include "azb_math.cuh"
#include "azb_pair_kernel.cuh"
#include "azb_ring_walk.cuh"

namespace azb
    {
template<class S> struct AnisoArgsDev
    {
    PairArgsDev<S> pair;
    const typename vec4<S>::type* orientation;
    typename vec4<S>::type* torque;
    };

struct TwoPatchMorse
    {
    // AnisoPairEvaluatorTwoPatchMorse.h:63-68 — no alignment attribute: 5 Scalars + bool, padded to a Scalar multiple
    template<class S> struct Params
        {
        S M_d, M_rinv, r_eq, omega, alpha;
        bool repulsion;
        };
    template<class S> struct Coeff
        {
        S M_d, M_rinv, r_eq, omega, alpha, U_shift;
        int repulsion;
        };
    template<class S> __device__ static Coeff<S> make(const Params<S>& p, S rcutsq, bool energy_shift)
        {
        Coeff<S> c;
        c.M_d = p.M_d;
        c.M_rinv = p.M_rinv;
        c.r_eq = p.r_eq;
        c.omega = p.omega;
        c.alpha = p.alpha;
        c.repulsion = p.repulsion ? 1 : 0;
        c.U_shift = S(0);
        if (energy_shift)
            {
            const S rcut = sqrt_(rcutsq);
            const S ex = exp_(-(rcut - c.r_eq) * c.M_rinv);
            const S om = S(1.0) - ex;
            c.U_shift = c.M_d * (om * om - S(1.0));
            }
        return c;
        }
    };

// rotate(quat(s, v), (1,0,0))
template<class S, class Q> __device__ __forceinline__ void patch_director(const Q& q, S& nx, S& ny, S& nz)
    {
    const S s = q.x, vx = q.y, vy = q.z, vz = q.w;
    nx = s * s - (vx * vx + vy * vy + vz * vz) + S(2) * vx * vx;
    ny = S(2) * (s * vz + vx * vy);
    nz = S(2) * (vx * vz - s * vy);
    }

// per-lane accumulators of one particle: force, energy, torque, virial
template<class S, bool VIRIAL> struct AnisoAccum
    {
    S fx = S(0), fy = S(0), fz = S(0), pe = S(0), tx = S(0), ty = S(0), tz = S(0);
    S vxx = S(0), vxy = S(0), vxz = S(0), vyy = S(0), vyz = S(0), vzz = S(0);
    template<int TPP> __device__ __forceinline__ void reduce()
        {
        if (TPP > 1)
            {
            fx = group_sum<TPP>(fx);
            fy = group_sum<TPP>(fy);
            fz = group_sum<TPP>(fz);
            pe = group_sum<TPP>(pe);
            tx = group_sum<TPP>(tx);
            ty = group_sum<TPP>(ty);
            tz = group_sum<TPP>(tz);
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
    __device__ __forceinline__ void store(const AnisoArgsDev<S>& d, unsigned int idx) const
        {
        const PairArgsDev<S>& a = d.pair;
        st4(a.force + idx, fx, fy, fz, S(0.5) * pe);
        st4(d.torque + idx, tx, ty, tz, S(0));
        if (VIRIAL)
            {
            a.virial[0 * a.virial_pitch + idx] = vxx;
            a.virial[1 * a.virial_pitch + idx] = vxy;
            a.virial[2 * a.virial_pitch + idx] = vxz;
            a.virial[3 * a.virial_pitch + idx] = vyy;
            a.virial[4 * a.virial_pitch + idx] = vyz;
            a.virial[5 * a.virial_pitch + idx] = vzz;
            }
        }
    };

// One in-range pair (AnisoPairEvaluatorTwoPatchMorse.h:127-216): separation d = r_i - r_j, patch directors n_i, n_j.
template<class S, bool VIRIAL>
__device__ __forceinline__ void tpm_pair(S dx, S dy, S dz, S rsq, S nix, S niy, S niz, S njx, S njy, S njz, const TwoPatchMorse::Coeff<S>& c,
                                         AnisoAccum<S, VIRIAL>& acc)
    {
    const S rinv = rsqrt_(rsq);
    const S r = rsq * rinv;
    const S ux = dx * rinv, uy = dy * rinv, uz = dz * rinv;

    S UMorse = -c.M_d, dUMorse_dr = S(0);
    if (r > c.r_eq || c.repulsion)
        {
        const S ex = exp_(-(r - c.r_eq) * c.M_rinv);
        const S om = S(1.0) - ex;
        UMorse = c.M_d * (om * om - S(1.0));
        dUMorse_dr = S(2.0) * c.M_d * c.M_rinv * ex * om;
        }
    const S gi = ux * nix + uy * niy + uz * niz;
    const S gi_exp = exp_(-c.omega * (gi * gi - c.alpha));
    const S Oi = rcp(S(1.0) + gi_exp);
    const S gj = ux * njx + uy * njy + uz * njz;
    const S gj_exp = exp_(-c.omega * (gj * gj - c.alpha));
    const S Oj = rcp(S(1.0) + gj_exp);

    const S e = (UMorse - c.U_shift) * Oi * Oj;
    const S dU_dr = dUMorse_dr * Oi * Oj;
    const S dU_dgi = S(2.0) * c.omega * gi * gi_exp * Oi * Oi * UMorse * Oj;
    const S dU_dgj = S(2.0) * c.omega * gj * gj_exp * Oj * Oj * UMorse * Oi;

    // n_perp = -u x (u x n) = n - (u.n) u   (|u| = 1)
    const S pix = nix - gi * ux, piy = niy - gi * uy, piz = niz - gi * uz;
    const S pjx = njx - gj * ux, pjy = njy - gj * uy, pjz = njz - gj * uz;
    const S Fx = -dU_dr * ux - rinv * (dU_dgi * pix + dU_dgj * pjx);
    const S Fy = -dU_dr * uy - rinv * (dU_dgi * piy + dU_dgj * pjy);
    const S Fz = -dU_dr * uz - rinv * (dU_dgi * piz + dU_dgj * pjz);
    // t_i = dU_dgi (u x n_i)
    acc.tx += dU_dgi * (uy * niz - uz * niy);
    acc.ty += dU_dgi * (uz * nix - ux * niz);
    acc.tz += dU_dgi * (ux * niy - uy * nix);
    acc.fx += Fx;
    acc.fy += Fy;
    acc.fz += Fz;
    acc.pe += e;
    if (VIRIAL)
        {
        acc.vxx += S(0.5) * dx * Fx;
        acc.vxy += S(0.5) * dy * Fx;
        acc.vxz += S(0.5) * dz * Fx;
        acc.vyy += S(0.5) * dy * Fy;
        acc.vyz += S(0.5) * dz * Fy;
        acc.vzz += S(0.5) * dz * Fz;
        }
    }

template<class S, int SHIFT, bool VIRIAL, int TPP, bool FASTBOX>
__global__ void __launch_bounds__(256) aniso_force_kernel(const AnisoArgsDev<S> d, const TwoPatchMorse::Params<S>* __restrict__ params)
    {
    typedef TwoPatchMorse::Coeff<S> coeff_t;
    const PairArgsDev<S>& a = d.pair;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int T2 = a.ntypes * a.ntypes;
    coeff_t* s_coeff = reinterpret_cast<coeff_t*>(smem_raw);
    S* s_rcutsq = reinterpret_cast<S*>(s_coeff + T2);
    for (unsigned int t = threadIdx.x; t < T2; t += blockDim.x)
        {
        const S rc = a.rcutsq[t];
        s_coeff[t] = TwoPatchMorse::make(params[t], rc, SHIFT == 1);
        s_rcutsq[t] = rc;
        }
    __syncthreads();

    const unsigned int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int idx = gtid / TPP;
    const unsigned int lane = threadIdx.x % TPP;
    const bool active = idx < a.N;

    AnisoAccum<S, VIRIAL> acc;
    if (active)
        {
        const unsigned int nn = __ldg(a.n_neigh + idx);
        const unsigned int* row = a.nlist + __ldg(a.head_list + idx);
        const auto pi = ld4(a.pos + idx);
        const auto qi = ld4(d.orientation + idx);
        S nix, niy, niz;
        patch_director<S>(qi, nix, niy, niz);
        const unsigned int trow = type_of(pi.w) * a.ntypes;

        for (unsigned int k = lane; k < nn; k += TPP)
            {
            const unsigned int j = __ldg(row + k);
            const auto pj = ld4(a.pos + j);
            S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
            min_image_t<FASTBOX>(a.box, dx, dy, dz);
            const S rsq = dx * dx + dy * dy + dz * dz;
            const unsigned int tp = trow + type_of(pj.w);
            if (!(rsq > s_rcutsq[tp])) // the evaluator rejects only rsq > rcutsq (strict), :135
                {
                const auto qj = ld4(d.orientation + j);
                S njx, njy, njz;
                patch_director<S>(qj, njx, njy, njz);
                tpm_pair<S, VIRIAL>(dx, dy, dz, rsq, nix, niy, niz, njx, njy, njz, s_coeff[tp], acc);
                }
            }
        }
    acc.template reduce<TPP>();
    if (active && lane == 0)
        acc.store(d, idx);
    }

// Tile-staged flavour: one CTA per tile stages, per distinct neighbour, the position and the patch director
// n_j = rotate(q_j, e_x) (computed once per tile instead of once per pair); rows are walked through the lane-transposed
// 16-bit slots like the isotropic tile kernel.
template<class S, int SHIFT, bool VIRIAL, int TPP, bool FASTBOX>
__global__ void __launch_bounds__(256)
    aniso_tile_kernel(const AnisoArgsDev<S> d, const TilePlanDev plan, const TwoPatchMorse::Params<S>* __restrict__ params, const S wrap_sq)
    {
    typedef TwoPatchMorse::Coeff<S> coeff_t;
    typedef typename vec4<S>::type V4;
    const PairArgsDev<S>& a = d.pair;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int T2 = a.ntypes * a.ntypes;
    V4* s_pos = reinterpret_cast<V4*>(smem_raw);
    V4* s_dir = s_pos + plan.smem_uniq;
    coeff_t* s_coeff = reinterpret_cast<coeff_t*>(s_dir + plan.smem_uniq);
    S* s_rcutsq = reinterpret_cast<S*>(s_coeff + T2);

    const unsigned int tile = plan.tile_order ? __ldg(plan.tile_order + plan.tile_begin + blockIdx.x) : plan.tile_begin + blockIdx.x;
    const unsigned int U = __ldg(plan.uniq_count + tile);
    const unsigned int* uq = plan.uniq + (size_t)tile * plan.uniq_capacity;
    warm_tile_rows(a, plan, tile, TPP);
    // (64 / sizeof(V4)) rows per thread in flight: the gather is two dependent loads per row
    constexpr int B = 64 / sizeof(V4);
    for (unsigned int s0 = threadIdx.x; s0 < U; s0 += B * blockDim.x)
        {
        unsigned int g[B];
        V4 pv[B], qv[B];
#pragma unroll
        for (int u = 0; u < B; ++u)
            g[u] = (s0 + u * blockDim.x < U) ? __ldg(uq + s0 + u * blockDim.x) : 0xffffffffu;
#pragma unroll
        for (int u = 0; u < B; ++u)
            if (g[u] != 0xffffffffu)
                {
                pv[u] = ld4(a.pos + g[u]);
                qv[u] = ld4(d.orientation + g[u]);
                }
#pragma unroll
        for (int u = 0; u < B; ++u)
            if (g[u] != 0xffffffffu)
                {
                const unsigned int s = s0 + u * blockDim.x;
                s_pos[s] = pv[u];
                V4 n = qv[u];
                patch_director<S>(qv[u], n.x, n.y, n.z);
                s_dir[s] = n;
                }
        }
    for (unsigned int t = threadIdx.x; t < T2; t += blockDim.x)
        {
        const S rc = a.rcutsq[t];
        s_coeff[t] = TwoPatchMorse::make(params[t], rc, SHIFT == 1);
        s_rcutsq[t] = rc;
        }
    __syncthreads();
    const unsigned int pos_base = smem_addr(s_pos), dir_base = smem_addr(s_dir);

    constexpr unsigned int M = 4 * TPP;
    const unsigned int groups = blockDim.x / TPP;
    const unsigned int gid = threadIdx.x / TPP;
    const unsigned int lane = threadIdx.x % TPP;
    for (unsigned int p = gid; p < plan.tile_size; p += groups)
        {
        const unsigned int idx = tile * plan.tile_size + p;
        const bool active = idx < a.N;
        AnisoAccum<S, VIRIAL> acc;
        if (active)
            {
            const unsigned int nn = __ldg(a.n_neigh + idx);
            const uint2* row = reinterpret_cast<const uint2*>(plan.slots + slot_row_begin(plan, a.head_list, idx, TPP)) + lane;
            const V4 pi = ld4(a.pos + idx);
            const V4 qi = ld4(d.orientation + idx);
            S nix, niy, niz;
            patch_director<S>(qi, nix, niy, niz);
            const unsigned int trow = type_of(pi.w) * a.ntypes;

            auto visit = [&](const unsigned int slot, const bool valid)
            {
                const V4 pj = lds4(pos_base, slot, S());
                S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                const S rsq = separation<FASTBOX>(a.box, wrap_sq, dx, dy, dz);
                const unsigned int tp = trow + type_of(pj.w);
                if (valid && !(rsq > s_rcutsq[tp]))
                    {
                    const V4 nj = lds4(dir_base, slot, S());
                    tpm_pair<S, VIRIAL>(dx, dy, dz, rsq, nix, niy, niz, nj.x, nj.y, nj.z, s_coeff[tp], acc);
                    }
            };

            const unsigned int full = nn / M;
            const unsigned int nblk = (nn + M - 1u) / M;
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
            if (full < nblk)
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
            acc.store(d, idx);
        }
    }

// Tile-staged flavour with in-range compaction (azb_ring_walk.cuh): same staging as aniso_tile_kernel, but the ~150-instruction
// evaluator runs on queued in-range pairs only, one per lane, instead of under a branch that 49 % of C4a's list entries skip.
// Slot rows in plain order (plan lane_width 1), TPP lanes per particle.
template<class S, int SHIFT, bool VIRIAL, int TPP, bool FASTBOX, bool SINGLE>
__global__ void __launch_bounds__(256)
    aniso_ring_tile_kernel(const AnisoArgsDev<S> d, const TilePlanDev plan, const TwoPatchMorse::Params<S>* __restrict__ params, const S wrap_sq)
    {
    typedef TwoPatchMorse::Coeff<S> coeff_t;
    typedef typename vec4<S>::type V4;
    const PairArgsDev<S>& a = d.pair;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int T2 = a.ntypes * a.ntypes;
    const unsigned int groups = blockDim.x / TPP;
    const unsigned int gid = threadIdx.x / TPP;
    V4* s_pos = reinterpret_cast<V4*>(smem_raw);
    V4* s_dir = s_pos + plan.smem_uniq;
    V4* s_queue = s_dir + plan.smem_uniq;
    coeff_t* s_coeff = reinterpret_cast<coeff_t*>(reinterpret_cast<unsigned char*>(s_queue) + groups * ring_stride<S, TPP>());
    S* s_rcutsq = reinterpret_cast<S*>(s_coeff + T2);

    const unsigned int tile = plan.tile_order ? __ldg(plan.tile_order + plan.tile_begin + blockIdx.x) : plan.tile_begin + blockIdx.x;
    const unsigned int U = __ldg(plan.uniq_count + tile);
    const unsigned int* uq = plan.uniq + (size_t)tile * plan.uniq_capacity;
    warm_tile_rows(a, plan, tile, 1u);
    constexpr int B = 64 / sizeof(V4);
    for (unsigned int s0 = threadIdx.x; s0 < U; s0 += B * blockDim.x)
        {
        unsigned int g[B];
        V4 pv[B], qv[B];
#pragma unroll
        for (int u = 0; u < B; ++u)
            g[u] = (s0 + u * blockDim.x < U) ? __ldg(uq + s0 + u * blockDim.x) : 0xffffffffu;
#pragma unroll
        for (int u = 0; u < B; ++u)
            if (g[u] != 0xffffffffu)
                {
                pv[u] = ld4(a.pos + g[u]);
                qv[u] = ld4(d.orientation + g[u]);
                }
#pragma unroll
        for (int u = 0; u < B; ++u)
            if (g[u] != 0xffffffffu)
                {
                const unsigned int s = s0 + u * blockDim.x;
                s_pos[s] = pv[u];
                V4 n = qv[u];
                patch_director<S>(qv[u], n.x, n.y, n.z);
                s_dir[s] = n;
                }
        }
    for (unsigned int t = threadIdx.x; t < T2; t += blockDim.x)
        {
        const S rc = a.rcutsq[t];
        s_coeff[t] = TwoPatchMorse::make(params[t], rc, SHIFT == 1);
        s_rcutsq[t] = rc;
        }
    __syncthreads();
    const unsigned int pos_base = smem_addr(s_pos), dir_base = smem_addr(s_dir);
    const unsigned int q_addr = smem_addr(s_queue) + gid * ring_stride<S, TPP>();
    const unsigned int rc_addr = smem_addr(s_rcutsq);
    const coeff_t c0 = s_coeff[0]; // one particle type: the coefficient set stays in registers
    const S rc0 = s_rcutsq[0];

    for (unsigned int p = gid; p < plan.tile_size; p += groups) // warp-uniform trip count
        {
        const unsigned int idx = tile * plan.tile_size + p;
        const bool active = idx < a.N;
        unsigned int nn = 0, trow = 0;
        const uint16_t* row = plan.slots;
        V4 pi = ring_entry<S>(S(0), S(0), S(0), 0u);
        S nix = S(0), niy = S(0), niz = S(0);
        if (active)
            {
            nn = __ldg(a.n_neigh + idx);
            row = plan.slots + slot_row_begin(plan, a.head_list, idx, 1);
            pi = ld4(a.pos + idx);
            const V4 qi = ld4(d.orientation + idx);
            patch_director<S>(qi, nix, niy, niz);
            trow = SINGLE ? 0u : type_of(pi.w) * a.ntypes;
            }
        AnisoAccum<S, VIRIAL> acc;
        ring_walk_row<S, TPP, FASTBOX, SINGLE, false>(row, nn, pi, trow, a.box, wrap_sq, pos_base, rc_addr, (unsigned int)sizeof(S), rc0, q_addr,
                                                      [&](const V4& e, const unsigned int slot, const unsigned int tp)
                                                      {
                                                          const V4 nj = lds4(dir_base, slot, S());
                                                          const S rsq = e.x * e.x + e.y * e.y + e.z * e.z;
                                                          if (SINGLE)
                                                              tpm_pair<S, VIRIAL>(e.x, e.y, e.z, rsq, nix, niy, niz, nj.x, nj.y, nj.z, c0, acc);
                                                          else
                                                              tpm_pair<S, VIRIAL>(e.x, e.y, e.z, rsq, nix, niy, niz, nj.x, nj.y, nj.z, s_coeff[tp], acc);
                                                      });
        acc.template reduce<TPP>();
        if (active && threadIdx.x % TPP == 0)
            acc.store(d, idx);
        }
    }
    } // namespace azb
