// azb_bond_kernel.cuh — bond force kernel (replaces HOOMD's gpu_compute_bond_forces<E, 2> as instantiated by
// /root/reference/src/PotentialBondGPUKernel.cu.inc:25-29). One thread per particle walks its column of the
// per-particle bond table (entry k of particle i at [k*pitch + i]: coalesced across the warp), gathers the partner
// position, evaluates DoubleWell (/root/reference/src/BondEvaluatorDoubleWell.h:90-113) or Quartic
// (/root/reference/src/BondEvaluatorQuartic.h:116-200) and writes force, half the bond energy and the virial.
// An evaluator that rejects its parameters raises *flags (host class throws, as HOOMD's PotentialBondGPU does).
#pragma once
#include "azb_math.cuh"

namespace azb
    {
struct BondGroup
    {
    unsigned int idx, type;
    };

template<class S> struct BondArgsDev
    {
    typename vec4<S>::type* force;
    S* virial;
    size_t virial_pitch;
    unsigned int N;
    const typename vec4<S>::type* pos;
    Box<S> box;
    const uint2* bondlist;
    const unsigned int* n_bonds;
    unsigned int pitch;
    unsigned int n_bond_types;
    unsigned int compute_virial;
    };

struct DoubleWell
    {
    template<class S> struct alignas(4 * sizeof(S)) Params
        {
        S r_1, r_diff, U_1, U_tilt;
        };
    template<class S> __device__ __forceinline__ static bool eval(S rsq, const Params<S>& p, S& fdivr, S& e)
        {
        if (p.r_diff == S(0.0))
            return false;
        const S rinv = rsqrt_(rsq);
        const S r = rsq * rinv;
        const S rdinv = rcp(p.r_diff);
        const S x = (p.r_1 - r) * rdinv;
        const S y = S(1.0) - x * x;
        const S y2 = y * y;
        e = p.U_1 * y2 + p.U_tilt * (S(1.0) - x - y2);
        fdivr = (S(4.0) * x * y * (p.U_tilt - p.U_1) - p.U_tilt) * rdinv * rinv;
        return true;
        }
    };

struct Quartic
    {
    template<class S> struct alignas(4 * sizeof(S)) Params
        {
        S k, r_0, b_1, b_2, U_0, sigma_6, epsilon_x_4, delta;
        };
    template<class S> __device__ __forceinline__ static bool eval(S rsq, const Params<S>& p, S& fdivr, S& e)
        {
        if (p.r_0 == S(0.0))
            return false;
        const S lj1 = p.epsilon_x_4 * p.sigma_6 * p.sigma_6;
        const S lj2 = p.epsilon_x_4 * p.sigma_6;
        S f = S(0), u = S(0);
        S r_red = S(1.0);
        if (p.delta == S(0.0))
            {
            const S r2inv = rcp(rsq);
            const S r6inv = r2inv * r2inv * r2inv;
            // WCA on when r^-6 > sigma^-6 / 2; (lj2 / lj1) is only formed when lj1 != 0
            if (lj1 != S(0) && r6inv > (lj2 / lj1) / S(2.0))
                {
                const S epsilon = lj2 * lj2 / S(4.0) / lj1; // the local that shadows the member (:150)
                f += r2inv * r6inv * (S(12.0) * lj1 * r6inv - S(6.0) * lj2);
                u += r6inv * (lj1 * r6inv - lj2) + epsilon;
                }
            if (rsq < p.r_0 * p.r_0)
                r_red = sqrt_(rsq) - p.r_0;
            }
        else
            {
            const S r = sqrt_(rsq) - p.delta;
            const S rinv = rcp(r);
            const S r2inv = rinv * rinv;
            const S r6inv = r2inv * r2inv * r2inv;
            if (lj1 != S(0) && r6inv > (lj2 / lj1) / S(2.0))
                {
                const S epsilon = p.epsilon_x_4 / S(4.); // the member (:126,175)
                f += r6inv * (S(12.0) * lj1 * r6inv - S(6.0) * lj2) * rinv * rcp(r + p.delta);
                u += r6inv * (lj1 * r6inv - lj2) + epsilon;
                }
            if (r < p.r_0)
                r_red = r - p.r_0;
            }
        if (r_red < S(0))
            {
            f += S(-1) * p.k * r_red * (S(4) * r_red * r_red - S(3) * (p.b_1 + p.b_2) * r_red + S(2) * p.b_1 * p.b_2)
                 * rcp(r_red + p.r_0 + p.delta);
            u += p.k * (r_red - p.b_1) * (r_red - p.b_2) * r_red * r_red + p.U_0;
            }
        else
            u += p.U_0;
        fdivr = f;
        e = u;
        return true;
        }
    };

template<class Pot, class S> __global__ void __launch_bounds__(256) bond_force_kernel(const BondArgsDev<S> a,
                                                                                      const typename Pot::template Params<S>* __restrict__ params,
                                                                                      unsigned int* __restrict__ flags)
    {
    typedef typename Pot::template Params<S> param_t;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    param_t* s_params = reinterpret_cast<param_t*>(smem_raw);
    for (unsigned int t = threadIdx.x; t < a.n_bond_types; t += blockDim.x)
        s_params[t] = params[t];
    __syncthreads();

    const unsigned int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.N)
        return;
    const unsigned int nb = __ldg(a.n_bonds + idx);
    const auto pi = ld4(a.pos + idx);
    S fx = S(0), fy = S(0), fz = S(0), pe = S(0);
    S vxx = S(0), vxy = S(0), vxz = S(0), vyy = S(0), vyz = S(0), vzz = S(0);
    for (unsigned int k = 0; k < nb; ++k)
        {
        const uint2 g = __ldg(a.bondlist + (size_t)k * a.pitch + idx);
        const auto pj = ld4(a.pos + g.x);
        S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
        if (a.box.flags == 7u)
            min_image_ortho(a.box, dx, dy, dz);
        else
            min_image(a.box, dx, dy, dz);
        const S rsq = dx * dx + dy * dy + dz * dz;
        S fdivr = S(0), e = S(0);
        if (Pot::eval(rsq, s_params[g.y], fdivr, e))
            {
            fx += dx * fdivr;
            fy += dy * fdivr;
            fz += dz * fdivr;
            pe += e;
            const S h = S(0.5) * fdivr;
            vxx += h * dx * dx;
            vxy += h * dx * dy;
            vxz += h * dx * dz;
            vyy += h * dy * dy;
            vyz += h * dy * dz;
            vzz += h * dz * dz;
            }
        else
            *flags = 1u;
        }
    st4(a.force + idx, fx, fy, fz, S(0.5) * pe);
    if (a.compute_virial)
        {
        a.virial[0 * a.virial_pitch + idx] = vxx;
        a.virial[1 * a.virial_pitch + idx] = vxy;
        a.virial[2 * a.virial_pitch + idx] = vxz;
        a.virial[3 * a.virial_pitch + idx] = vyy;
        a.virial[4 * a.virial_pitch + idx] = vyz;
        a.virial[5 * a.virial_pitch + idx] = vzz;
        }
    }
    } // namespace azb
