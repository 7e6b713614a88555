// azb_pair_potentials.cuh — device-side isotropic pair potentials of hoomd.azplugins.
//
// Each potential has
//   Params<S> : the reference's param_type, bit-for-bit (field order, size, alignment) — what HOOMD's host class
//               hands over in d_params;
//   Coeff<S>  : what the kernel keeps per type pair in shared memory — everything the reference's evaluator
//               constructor recomputes for every pair (lj1, lj2, wca_shift, sigma_6, ...) hoisted to once per
//               block, plus the energy shift at r_cut, which depends only on the type pair;
//   make()    : Params -> Coeff (same expressions as the reference ctor, same precision);
//   eval()    : per-pair force/r and energy; returns false exactly when the reference's evalForceAndEnergy does.
#pragma once
#include "azb_math.cuh"

namespace azb
    {
// ---------------------------------------------------------------------------------------------------------
// PerturbedLennardJones — /root/reference/src/PairEvaluatorPerturbedLennardJones.h:25-66 (params),
// :97-104 (ctor), :117-155 (evalForceAndEnergy)
// ---------------------------------------------------------------------------------------------------------
struct PerturbedLJ
    {
    template<class S> struct alignas(4 * sizeof(S)) Params
        {
        S sigma_6, epsilon_x_4, attraction_scale_factor, rwcasq;
        };
    template<class S> struct Coeff
        {
        S lj1, lj2, lam, rwcasq, wca_shift, e_shift;
        };
    template<class S> __device__ static Coeff<S> make(const Params<S>& p, S rcutsq, bool energy_shift)
        {
        Coeff<S> c;
        c.lj1 = p.epsilon_x_4 * p.sigma_6 * p.sigma_6;
        c.lj2 = p.epsilon_x_4 * p.sigma_6;
        c.lam = p.attraction_scale_factor;
        c.rwcasq = p.rwcasq;
        c.wca_shift = p.epsilon_x_4 * (S(1.0) - c.lam) / S(4.0);
        c.e_shift = S(0);
        if (energy_shift)
            {
            const S rcut2inv = S(1.0) / rcutsq;
            const S rcut6inv = rcut2inv * rcut2inv * rcut2inv;
            S sh = rcut6inv * (c.lj1 * rcut6inv - c.lj2);
            if (rcutsq < c.rwcasq)
                sh += c.wca_shift;
            else
                sh *= c.lam;
            c.e_shift = sh;
            }
        return c;
        }
    // The reference's evaluators return false for a type pair with a vanishing prefactor; the kernels fold that into the
    // cutoff they test against (r_cut^2 = -1 for such pairs), so eval() itself only tests the distance.
    template<class S> __device__ __forceinline__ static bool active(const Coeff<S>& c)
        {
        return c.lj1 != S(0);
        }
    template<class S> __device__ __forceinline__ static bool eval(S rsq, S rcutsq, const Coeff<S>& c, S& fdivr, S& e)
        {
        if (rsq < rcutsq)
            {
            const S r2inv = rcp(rsq);
            const S r6inv = r2inv * r2inv * r2inv;
            S f = r2inv * r6inv * (S(12.0) * c.lj1 * r6inv - S(6.0) * c.lj2);
            S u = r6inv * (c.lj1 * r6inv - c.lj2);
            if (rsq < c.rwcasq)
                u += c.wca_shift;
            else
                {
                f *= c.lam;
                u *= c.lam;
                }
            fdivr = f;
            e = u - c.e_shift;
            return true;
            }
        return false;
        }
    };

// ---------------------------------------------------------------------------------------------------------
// Colloid — /root/reference/src/PairEvaluatorColloid.h:23-57 (params), :81-89 (ctor), :101-113 (solvent-solvent),
// :125-152 (colloid-solvent), :164-220 (colloid-colloid), :233-269 (dispatch + shift)
// ---------------------------------------------------------------------------------------------------------
struct Colloid
    {
    template<class S> struct alignas(4 * sizeof(S)) Params
        {
        S A, a_1, a_2, sigma_3;
        };
    template<class S> struct Coeff
        {
        S A, sigma_3, sigma_6, ai, aj, e_shift;
        };

    template<bool force, class S> __device__ __forceinline__ static S solvent_solvent(const Coeff<S>& c, S& fdivr, S rsq)
        {
        const S r2inv = rcp(rsq);
        const S r6inv = r2inv * r2inv * r2inv;
        const S c1 = c.A * c.sigma_6 / S(36.0);
        if (force)
            fdivr = S(6.0) * c1 * r2inv * r6inv * (S(2.0) * c.sigma_6 * r6inv - S(1.0));
        return c1 * r6inv * (c.sigma_6 * r6inv - S(1.0));
        }
    template<bool force, class S> __device__ __forceinline__ static S colloid_solvent(const Coeff<S>& c, S& fdivr, S rsq)
        {
        const S a = (c.ai > c.aj) ? c.ai : c.aj;
        const S asq = a * a;
        const S d = asq - rsq;
        const S rsqsq = rsq * rsq;
        const S d3 = d * d * d;
        const S d6 = d3 * d3;
        const S d6inv = rcp(d6);
        const S fR = c.sigma_3 * c.A * a * asq * rcp(d3);
        if (force)
            fdivr = S(4.0 / 15.0) * fR
                    * (S(2.0) * (asq + rsq) * (asq * (S(5.0) * asq + S(22.0) * rsq) + S(5.0) * rsqsq) * c.sigma_6 * d6inv - S(5.0))
                    * rcp(d);
        return S(2.0 / 9.0) * fR
               * (S(1.0) - (asq * (asq * (asq / S(3.0) + S(3.0) * rsq) + S(4.2) * rsqsq) + rsq * rsqsq) * c.sigma_6 * d6inv);
        }
    template<bool force, class S> __device__ static S colloid_colloid(const Coeff<S>& c, S& fdivr, S rsq)
        {
        // The repulsive part sums +-h terms that cancel to many digits near contact; keep exact division/sqrt here.
        const S r = sqrt_(rsq);
        const S k0 = c.ai * c.aj, k1 = c.ai + c.aj, k2 = c.ai - c.aj;
        const S k3 = k1 + r, k4 = k1 - r, k5 = k2 + r, k6 = k2 - r;
        const S k7 = S(1.0) / (k3 * k4);
        const S k8 = S(1.0) / (k5 * k6);
        const S k3inv = S(1.0) / k3;
        S g0 = k3inv * k3inv;
        g0 *= g0 * g0;
        g0 *= k3inv;
        const S k4inv = S(1.0) / k4;
        S g1 = k4inv * k4inv;
        g1 *= g1 * g1;
        g1 *= k4inv;
        const S k5inv = S(1.0) / k5;
        S g2 = k5inv * k5inv;
        g2 *= g2 * g2;
        g2 *= k5inv;
        const S k6inv = S(1.0) / k6;
        S g3 = k6inv * k6inv;
        g3 *= g3 * g3;
        g3 *= k6inv;
        const S h0 = ((k3 + S(5.0) * k1) * k3 + S(30.0) * k0) * g0;
        const S h1 = ((k4 + S(5.0) * k1) * k4 + S(30.0) * k0) * g1;
        const S h2 = ((k5 + S(5.0) * k2) * k5 - S(30.0) * k0) * g2;
        const S h3 = ((k6 + S(5.0) * k2) * k6 - S(30.0) * k0) * g3;
        g0 *= S(42.0) * k0 * k3inv + S(6.0) * k1 + k3;
        g1 *= S(42.0) * k0 * k4inv + S(6.0) * k1 + k4;
        g2 *= S(-42.0) * k0 * k5inv + S(6.0) * k2 + k5;
        g3 *= S(-42.0) * k0 * k6inv + S(6.0) * k2 + k6;
        const S rinv = S(1.0) / r;
        const S fR = c.A * c.sigma_6 * rinv / S(37800.0);
        S pair_eng = fR * (h0 - h1 - h2 + h3);
        if (force)
            {
            const S dUR = pair_eng * rinv + S(5.0) * fR * (g0 + g1 - g2 - g3);
            const S dUA
                = -c.A / S(3.0) * r * ((S(2.0) * k0 * k7 + S(1.0)) * k7 + (S(2.0) * k0 * k8 - S(1.0)) * k8);
            fdivr = (dUR + dUA) * rinv;
            }
        pair_eng += c.A / S(6.0) * (S(2.0) * k0 * (k7 + k8) - log_(k8 / k7));
        return pair_eng;
        }
    template<bool force, class S> __device__ __forceinline__ static S branch(const Coeff<S>& c, S& fdivr, S rsq)
        {
        if (c.ai == S(0) && c.aj == S(0))
            return solvent_solvent<force>(c, fdivr, rsq);
        else if (c.ai != S(0) && c.aj != S(0))
            return colloid_colloid<force>(c, fdivr, rsq);
        else
            return colloid_solvent<force>(c, fdivr, rsq);
        }
    template<class S> __device__ static Coeff<S> make(const Params<S>& p, S rcutsq, bool energy_shift)
        {
        Coeff<S> c;
        c.A = p.A;
        c.sigma_3 = p.sigma_3;
        c.sigma_6 = p.sigma_3 * p.sigma_3;
        c.ai = p.a_1;
        c.aj = p.a_2;
        c.e_shift = S(0);
        if (energy_shift && c.A != S(0))
            {
            S dummy;
            c.e_shift = branch<false>(c, dummy, rcutsq);
            }
        return c;
        }
    // The reference's evaluators return false for a type pair with a vanishing prefactor; the kernels fold that into the
    // cutoff they test against (r_cut^2 = -1 for such pairs), so eval() itself only tests the distance.
    template<class S> __device__ __forceinline__ static bool active(const Coeff<S>& c)
        {
        return c.A != S(0);
        }
    template<class S> __device__ __forceinline__ static bool eval(S rsq, S rcutsq, const Coeff<S>& c, S& fdivr, S& e)
        {
        if (rsq < rcutsq)
            {
            e = branch<true>(c, fdivr, rsq) - c.e_shift;
            return true;
            }
        return false;
        }
    };

// ---------------------------------------------------------------------------------------------------------
// ExpandedYukawa — /root/reference/src/PairEvaluatorExpandedYukawa.h:23-53 (params), :73-79, :92-115
// ---------------------------------------------------------------------------------------------------------
struct ExpandedYukawa
    {
    template<class S> struct alignas(4 * sizeof(S)) Params
        {
        S epsilon, kappa, delta;
        };
    template<class S> struct Coeff
        {
        S epsilon, kappa, delta, e_shift;
        };
    template<class S> __device__ static Coeff<S> make(const Params<S>& p, S rcutsq, bool energy_shift)
        {
        Coeff<S> c;
        c.epsilon = p.epsilon;
        c.kappa = p.kappa;
        c.delta = p.delta;
        c.e_shift = S(0);
        if (energy_shift)
            {
            const S rcut = sqrt_(rcutsq);
            const S rcut_delta = rcut - c.delta;
            c.e_shift = c.epsilon * exp_(-c.kappa * rcut_delta) / rcut_delta;
            }
        return c;
        }
    // The reference's evaluators return false for a type pair with a vanishing prefactor; the kernels fold that into the
    // cutoff they test against (r_cut^2 = -1 for such pairs), so eval() itself only tests the distance.
    template<class S> __device__ __forceinline__ static bool active(const Coeff<S>& c)
        {
        return c.epsilon != S(0);
        }
    template<class S> __device__ __forceinline__ static bool eval(S rsq, S rcutsq, const Coeff<S>& c, S& fdivr, S& e)
        {
        if (rsq < rcutsq)
            {
            const S rinv = rsqrt_(rsq);
            const S r = rsq * rinv;
            const S r_delta = r - c.delta;
            const S r_delta_inv = rcp(r_delta);
            const S u = c.epsilon * exp_(-c.kappa * r_delta) * r_delta_inv;
            fdivr = u * (c.kappa + r_delta_inv) * rinv;
            e = u - c.e_shift;
            return true;
            }
        return false;
        }
    };

// ---------------------------------------------------------------------------------------------------------
// Hertz — /root/reference/src/PairEvaluatorHertz.h:23-47 (params), :77-80, :93-110. energy_shift is ignored.
// ---------------------------------------------------------------------------------------------------------
struct Hertz
    {
    template<class S> struct Params
        {
        S epsilon;
        };
    template<class S> struct Coeff
        {
        S epsilon, rcutinv;
        };
    template<class S> __device__ static Coeff<S> make(const Params<S>& p, S rcutsq, bool)
        {
        Coeff<S> c;
        c.epsilon = p.epsilon;
        c.rcutinv = S(1.0) / sqrt_(rcutsq);
        return c;
        }
    // The reference's evaluators return false for a type pair with a vanishing prefactor; the kernels fold that into the
    // cutoff they test against (r_cut^2 = -1 for such pairs), so eval() itself only tests the distance.
    template<class S> __device__ __forceinline__ static bool active(const Coeff<S>& c)
        {
        return c.epsilon != S(0);
        }
    template<class S> __device__ __forceinline__ static bool eval(S rsq, S rcutsq, const Coeff<S>& c, S& fdivr, S& e)
        {
        if (rsq < rcutsq)
            {
            const S rinv = rsqrt_(rsq);
            const S r = rsq * rinv;
            const S x = S(1.0) - r * c.rcutinv;
            const S xsqrt = sqrt_(x);
            const S ex3p2 = c.epsilon * x * xsqrt;
            fdivr = S(2.5) * ex3p2 * rinv * c.rcutinv;
            e = ex3p2 * x;
            return true;
            }
        return false;
        }
    };

// ---------------------------------------------------------------------------------------------------------
// DPD conservative part — /root/reference/src/DPDPairEvaluatorGeneralWeight.h:32-62 (params), :165-183
// (class PotentialPairConservativeGeneralWeight, export_PotentialPairDPDThermo.cc.inc:33-35)
// ---------------------------------------------------------------------------------------------------------
struct DPDConservative
    {
    template<class S> struct alignas(4 * sizeof(S)) Params
        {
        S A, gamma, s;
        };
    template<class S> struct Coeff
        {
        S A, rcutinv, rcut;
        };
    template<class S> __device__ static Coeff<S> make(const Params<S>& p, S rcutsq, bool)
        {
        Coeff<S> c;
        c.A = p.A;
        c.rcutinv = S(1.0) / sqrt_(rcutsq);
        c.rcut = S(1.0) / c.rcutinv;
        return c;
        }
    // The reference's evaluators return false for a type pair with a vanishing prefactor; the kernels fold that into the
    // cutoff they test against (r_cut^2 = -1 for such pairs), so eval() itself only tests the distance.
    template<class S> __device__ __forceinline__ static bool active(const Coeff<S>& c)
        {
        return true;
        }
    template<class S> __device__ __forceinline__ static bool eval(S rsq, S rcutsq, const Coeff<S>& c, S& fdivr, S& e)
        {
        if (rsq < rcutsq)
            {
            const S rinv = rsqrt_(rsq);
            const S r = rsq * rinv;
            fdivr = c.A * (rinv - c.rcutinv);
            e = c.A * (c.rcut - r) - S(0.5) * c.A * c.rcutinv * (rcutsq - rsq);
            return true;
            }
        return false;
        }
    };
    } // namespace azb
