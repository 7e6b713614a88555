// DPD thermostat launcher (float and double; virial on/off; tpp 1..32).
#include <cstdlib>

#include "azb_dpd_kernel.cuh"
#include "azb_launch.cuh"

namespace azb
    {
template<class S, bool VIRIAL> static int launch_dpd_variant(const azb_dpd_args& a, const void* d_params, cudaStream_t stream)
    {
    typedef DPDThermo::Params<S> param_t;
    typedef DPDThermo::Coeff<S> coeff_t;
    static const bool use_v1 = getenv("AZB_DPD_V1") != nullptr; // A/B switch while the queue kernel is being tuned
    DPDArgsDev<S> d;
    d.pair = make_pair_args<S>(a.pair);
    d.vel = (const typename vec4<S>::type*)a.d_vel;
    d.tag = a.d_tag;
    // the evaluator stores the timestep in an `unsigned int` (DPDPairEvaluatorGeneralWeight.h:122,273)
    d.rk = make_round_keys(200, (uint64_t)(uint32_t)a.timestep, a.seed);
    d.deltaT = (S)a.deltaT;
    d.T = (S)a.T;
    // the queue kernel balances in-range work inside a group, so narrow groups (less idle tail, cheaper reductions) win
    unsigned int tpp = a.pair.threads_per_particle;
    if (tpp == 0)
        {
        const double nbar = a.pair.N ? (double)a.pair.size_nlist / (double)a.pair.N : 0.0;
        tpp = use_v1 ? pick_tpp(a.pair) : (nbar <= 8 ? 1 : (nbar <= 16 ? 2 : (nbar <= 96 ? 4 : (nbar <= 256 ? 8 : 16))));
        }
    if (!valid_tpp(tpp))
        return AZB_ERR_BAD_ARGUMENT;
    const unsigned int block = pick_block(a.pair.block_size);
    const unsigned int per_block = block / tpp;
    const unsigned int grid = (a.pair.N + per_block - 1) / per_block;
    const size_t T2 = (size_t)a.pair.ntypes * a.pair.ntypes;
    const size_t smem = T2 * (sizeof(coeff_t) + sizeof(S));
    if (smem > 48 * 1024)
        return AZB_ERR_BAD_ARGUMENT;
    if (use_v1)
        {
#define AZB_CALL(TPP)                                                                           \
    do                                                                                          \
        {                                                                                       \
        if (d.pair.box.flags == 7u)                                                             \
            dpd_force_kernel<S, VIRIAL, TPP, true><<<grid, block, smem, stream>>>(d, (const param_t*)d_params);  \
        else                                                                                    \
            dpd_force_kernel<S, VIRIAL, TPP, false><<<grid, block, smem, stream>>>(d, (const param_t*)d_params); \
        } while (0)
        AZB_DISPATCH_TPP(tpp, AZB_CALL)
#undef AZB_CALL
        return (int)cudaGetLastError();
        }
    // queue kernel: per-group ring of in-range entries + coefficient table in shared memory
    const bool fastbox = d.pair.box.flags == 7u;
    const bool single = a.pair.ntypes == 1;
    const S lmin = fmin(d.pair.box.Lx, fmin(d.pair.box.Ly, d.pair.box.Lz));
    const S wrap_sq = S(0.25) * lmin * lmin;
    const size_t smem_q = (size_t)block * 4 * sizeof(typename vec4<S>::type) + (single ? 0 : (size_t)block * 4 * sizeof(unsigned int))
                          + T2 * sizeof(DPDCoeffQ<S>);
    if (smem_q > 200 * 1024)
        return AZB_ERR_BAD_ARGUMENT;
#define AZB_LAUNCH_Q(TPP, FB, SG)                                                                   \
    do                                                                                              \
        {                                                                                           \
        auto kern = dpd_force_kernel_q<S, VIRIAL, TPP, FB, SG>;                                     \
        if (smem_q > 48 * 1024)                                                                     \
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_q);   \
        kern<<<grid, block, smem_q, stream>>>(d, (const param_t*)d_params, wrap_sq);                \
        } while (0)
#define AZB_CALL(TPP)                            \
    do                                           \
        {                                        \
        if (fastbox && single)                   \
            AZB_LAUNCH_Q(TPP, true, true);       \
        else if (fastbox)                        \
            AZB_LAUNCH_Q(TPP, true, false);      \
        else if (single)                         \
            AZB_LAUNCH_Q(TPP, false, true);      \
        else                                     \
            AZB_LAUNCH_Q(TPP, false, false);     \
        } while (0)
    AZB_DISPATCH_TPP(tpp, AZB_CALL)
#undef AZB_CALL
#undef AZB_LAUNCH_Q
    return (int)cudaGetLastError();
    }

template<class S> int launch_dpd(const azb_dpd_args& a, const void* d_params, cudaStream_t stream)
    {
    if (a.pair.N == 0)
        return 0;
    return a.pair.compute_virial ? launch_dpd_variant<S, true>(a, d_params, stream)
                                 : launch_dpd_variant<S, false>(a, d_params, stream);
    }
template int launch_dpd<float>(const azb_dpd_args&, const void*, cudaStream_t);
template int launch_dpd<double>(const azb_dpd_args&, const void*, cudaStream_t);
    } // namespace azb
