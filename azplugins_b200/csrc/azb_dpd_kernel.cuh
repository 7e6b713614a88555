// azb_dpd_kernel.cuh — DPD thermostat pair force kernel (replaces HOOMD's gpu_compute_dpd_forces<E> as
// instantiated by /root/reference/src/PotentialPairDPDThermoGPUKernel.cu.inc:21-24 for
// DPDPairEvaluatorGeneralWeight, /root/reference/src/DPDPairEvaluatorGeneralWeight.h:198-255).
//
// Same TPP-lanes-per-particle neighbour walk as the isotropic kernel. vel[j] and tag[j] are gathered only for
// neighbours inside the cutoff (the reference's driver gathers them for every list entry; ~2/3 of the entries of a
// buffered list are outside r_cut). One Philox4x32-10 block per in-range pair, keyed on (seed, timestep) with the
// round keys precomputed on the host, counter = (min tag, max tag).
#pragma once
#include "azb_math.cuh"
#include "azb_pair_kernel.cuh"
#include "azb_rng.cuh"

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
    template<class S> struct Coeff
        {
        S A, gamma, half_s, rcutinv, rcut, sigma; // sigma = rsqrt(dt / (6 kT gamma))
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
        return c;
        }
    };

template<class S, bool VIRIAL, int TPP, bool FASTBOX>
__global__ void __launch_bounds__(256) dpd_force_kernel(const DPDArgsDev<S> d, const DPDThermo::Params<S>* __restrict__ params)
    {
    typedef DPDThermo::Coeff<S> coeff_t;
    const PairArgsDev<S>& a = d.pair;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int T2 = a.ntypes * a.ntypes;
    coeff_t* s_coeff = reinterpret_cast<coeff_t*>(smem_raw);
    S* s_rcutsq = reinterpret_cast<S*>(s_coeff + T2);
    for (unsigned int t = threadIdx.x; t < T2; t += blockDim.x)
        {
        const S rc = a.rcutsq[t];
        s_coeff[t] = DPDThermo::make(params[t], rc, d.deltaT, d.T);
        s_rcutsq[t] = rc;
        }
    __syncthreads();

    const unsigned int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int idx = gtid / TPP;
    const unsigned int lane = threadIdx.x % TPP;
    const bool active = idx < a.N;

    S fx = S(0), fy = S(0), fz = S(0), pe = S(0);
    S vxx = S(0), vxy = S(0), vxz = S(0), vyy = S(0), vyz = S(0), vzz = S(0);
    if (active)
        {
        const unsigned int nn = __ldg(a.n_neigh + idx);
        const unsigned int* row = a.nlist + __ldg(a.head_list + idx);
        const auto pi = ld4(a.pos + idx);
        const auto vi = ld4(d.vel + idx);
        const unsigned int tag_i = __ldg(d.tag + idx);
        const unsigned int trow = type_of(pi.w) * a.ntypes;

        for (unsigned int k = lane; k < nn; k += TPP)
            {
            const unsigned int j = __ldg(row + k);
            const auto pj = ld4(a.pos + j);
            S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
            min_image_t<FASTBOX>(a.box, dx, dy, dz);
            const S rsq = dx * dx + dy * dy + dz * dz;
            const unsigned int tp = trow + type_of(pj.w);
            const S rcutsq = s_rcutsq[tp];
            if (rsq < rcutsq)
                {
                const auto vj = ld4(d.vel + j);
                const unsigned int tag_j = __ldg(d.tag + j);
                const coeff_t c = s_coeff[tp];
                const S rdotv = dx * (vi.x - vj.x) + dy * (vi.y - vj.y) + dz * (vi.z - vj.z);
                const S rinv = rsqrt_(rsq);
                const S r = rsq * rinv;
                const S alpha = dpd_alpha<S>(tag_i, tag_j, d.rk);
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
                S fdivr = fcons - c.gamma * wR * wR * rdotv;
                fdivr += c.sigma * wR * alpha;
                const S e = c.A * (c.rcut - r) - S(0.5) * c.A * c.rcutinv * (rcutsq - rsq);
                fx += dx * fdivr;
                fy += dy * fdivr;
                fz += dz * fdivr;
                pe += e;
                if (VIRIAL)
                    {
                    const S h = S(0.5) * fcons; // conservative part only (DPDPairEvaluatorGeneralWeight.h:193-194)
                    vxx += h * dx * dx;
                    vxy += h * dx * dy;
                    vxz += h * dx * dz;
                    vyy += h * dy * dy;
                    vyz += h * dy * dz;
                    vzz += h * dz * dz;
                    }
                }
            }
        }
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
    if (active && lane == 0)
        {
        st4(a.force + idx, fx, fy, fz, S(0.5) * pe);
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
    }
// ---------------------------------------------------------------------------------------------------------
// Queue kernel (the production DPD path). A buffered neighbour list holds ~(1 + r_buff/r_cut)^3 entries per
// in-range pair, and the in-range work (Philox4x32-10 + thermostat) is ~6x the out-of-range work, so evaluating
// it under `if (rsq < rcutsq)` leaves two thirds of the lanes idle through the expensive block. Here every group
// of TPP lanes splits the walk in two phases that each run with (nearly) all lanes busy:
//   light: TPP list entries at a time — coalesced index read, one 16/32 B gather of pos[j], distance, cutoff test;
//          in-range entries (dx, dy, dz, j) are compacted (ballot + popc) into the group's ring in shared memory;
//   heavy: whenever every group of the warp holds >= TPP queued entries (or has finished its row), each lane pops
//          one entry: gathers vel[j], tag[j], draws the pair's random number, evaluates the force.
// Minimum image: in the fully periodic orthorhombic box a pair whose raw separation is inside half the shortest box
// edge (wrap_sq = (L_min / 2)^2) needs no wrap; only the few pairs straddling the boundary take the wrap path.
// ---------------------------------------------------------------------------------------------------------
template<class S> struct alignas(4 * sizeof(S)) DPDCoeffQ
    {
    S A, gamma, sigma, rcutinv; // sigma = rsqrt(dt / (6 kT gamma))
    S rcut, half_s, hA, rcutsq; // hA = A / (2 r_cut): E = hA (r_cut - r)^2
    };

template<class S> __device__ __forceinline__ DPDCoeffQ<S> make_dpd_coeff_q(const DPDThermo::Params<S>& p, S rcutsq, S deltaT, S T)
    {
    DPDCoeffQ<S> c;
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

template<class S, bool VIRIAL, int TPP, bool FASTBOX, bool SINGLE>
__global__ void __launch_bounds__(256)
    dpd_force_kernel_q(const DPDArgsDev<S> d, const DPDThermo::Params<S>* __restrict__ params, const S wrap_sq)
    {
    typedef DPDCoeffQ<S> coeff_t;
    typedef typename vec4<S>::type V4;
    constexpr unsigned int QCAP = 4 * TPP;
    constexpr unsigned int FULL = 0xffffffffu;
    const PairArgsDev<S>& a = d.pair;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int T2 = a.ntypes * a.ntypes;
    V4* s_queue = reinterpret_cast<V4*>(smem_raw);
    unsigned int* s_qtp = reinterpret_cast<unsigned int*>(s_queue + 4 * blockDim.x);
    coeff_t* s_coeff = reinterpret_cast<coeff_t*>(s_qtp + (SINGLE ? 0u : 4u * blockDim.x));
    for (unsigned int t = threadIdx.x; t < T2; t += blockDim.x)
        s_coeff[t] = make_dpd_coeff_q(params[t], a.rcutsq[t], d.deltaT, d.T);
    __syncthreads();

    const unsigned int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int idx = gtid / TPP;
    const unsigned int lane = threadIdx.x % TPP;
    const unsigned int gshift = (threadIdx.x & 31u) & ~(unsigned int)(TPP - 1);
    const unsigned int gbits = (TPP == 32) ? FULL : ((1u << TPP) - 1u);
    const bool active = idx < a.N;
    V4* const q = s_queue + (threadIdx.x / TPP) * QCAP;
    unsigned int* const qtp = s_qtp + (threadIdx.x / TPP) * QCAP;

    unsigned int nn = 0;
    const unsigned int* row = a.nlist;
    V4 pi = pack_entry<S>(S(0), S(0), S(0), 0u), vi = pi;
    unsigned int tag_i = 0, trow = 0;
    if (active)
        {
        nn = __ldg(a.n_neigh + idx);
        row = a.nlist + __ldg(a.head_list + idx);
        pi = ld4(a.pos + idx);
        vi = ld4(d.vel + idx);
        tag_i = __ldg(d.tag + idx);
        trow = SINGLE ? 0u : type_of(pi.w) * a.ntypes;
        }
    const coeff_t c0 = s_coeff[0]; // single-type systems keep the one coefficient set in registers

    S fx = S(0), fy = S(0), fz = S(0), pe = S(0);
    S vxx = S(0), vxy = S(0), vxz = S(0), vyy = S(0), vyz = S(0), vzz = S(0);

    unsigned int q_head = 0, q_cnt = 0; // group-uniform ring state
    unsigned int k = lane;
    unsigned int j_next = (k < nn) ? __ldg(row + k) : 0u;
    for (;;)
        {
        // ---- light phase: one list entry per lane ----------------------------------------------------------------
        bool in = false;
        S dx = S(0), dy = S(0), dz = S(0);
        unsigned int tp = 0;
        const unsigned int j = j_next;
        const bool have = k < nn;
        k += TPP;
        if (k < nn)
            j_next = __ldg(row + k);
        if (have)
            {
            const V4 pj = ld4(a.pos + j);
            dx = pi.x - pj.x;
            dy = pi.y - pj.y;
            dz = pi.z - pj.z;
            S rc;
            if (SINGLE)
                rc = c0.rcutsq;
            else
                {
                tp = trow + type_of(pj.w);
                rc = s_coeff[tp].rcutsq;
                }
            S rsq;
            if (FASTBOX)
                {
                rsq = dx * dx + dy * dy + dz * dz;
                if (rsq > wrap_sq) // only then can a component exceed half a box edge
                    {
                    min_image_ortho(a.box, dx, dy, dz);
                    rsq = dx * dx + dy * dy + dz * dz;
                    }
                }
            else
                {
                min_image(a.box, dx, dy, dz);
                rsq = dx * dx + dy * dy + dz * dz;
                }
            in = rsq < rc;
            }
        const unsigned int b = (__ballot_sync(FULL, in) >> gshift) & gbits;
        if (in)
            {
            const unsigned int slot = (q_head + q_cnt + __popc(b & ((1u << lane) - 1u))) & (QCAP - 1u);
            q[slot] = pack_entry<S>(dx, dy, dz, j);
            if (!SINGLE)
                qtp[slot] = tp;
            }
        q_cnt += __popc(b);
        const bool light_done = (k - lane) >= nn;
        const bool go = __all_sync(FULL, q_cnt >= TPP || light_done) || __any_sync(FULL, q_cnt > QCAP - TPP);

        // ---- heavy phase: one queued in-range pair per lane ------------------------------------------------------
        if (go)
            {
            __syncwarp();
            if (lane < q_cnt)
                {
                const unsigned int slot = (q_head + lane) & (QCAP - 1u);
                const V4 e = q[slot];
                const unsigned int jj = entry_index(e.w);
                const V4 vj = ld4(d.vel + jj);
                const unsigned int tag_j = __ldg(d.tag + jj);
                coeff_t c = c0;
                if (!SINGLE)
                    c = s_coeff[qtp[slot]];
                const S rsq = e.x * e.x + e.y * e.y + e.z * e.z;
                const S rdotv = e.x * (vi.x - vj.x) + e.y * (vi.y - vj.y) + e.z * (vi.z - vj.z);
                const S rinv = rsqrt_(rsq);
                const S r = rsq * rinv;
                const S alpha = dpd_alpha<S>(tag_i, tag_j, d.rk);
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
                fx += e.x * fdivr;
                fy += e.y * fdivr;
                fz += e.z * fdivr;
                pe += (c.hA * dr) * dr; // == A (rc - r) - A (rc^2 - r^2) / (2 rc)
                if (VIRIAL)
                    {
                    const S h = S(0.5) * fcons; // conservative part only (DPDPairEvaluatorGeneralWeight.h:193-194)
                    const S hx = h * e.x, hy = h * e.y, hz = h * e.z;
                    vxx += hx * e.x;
                    vxy += hx * e.y;
                    vxz += hx * e.z;
                    vyy += hy * e.y;
                    vyz += hy * e.z;
                    vzz += hz * e.z;
                    }
                }
            const unsigned int n = min(q_cnt, (unsigned int)TPP);
            q_head += n;
            q_cnt -= n;
            }
        if (__all_sync(FULL, light_done && q_cnt == 0u))
            break;
        }
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
    if (active && lane == 0)
        {
        st4(a.force + idx, fx, fy, fz, S(0.5) * pe);
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
    }
    } // namespace azb
