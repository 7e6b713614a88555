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
    } // namespace azb
