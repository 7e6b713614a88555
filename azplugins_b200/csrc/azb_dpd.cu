// DPD thermostat launcher (float and double; virial on/off; tpp 1..32).
#include "azb_dpd_kernel.cuh"
#include "azb_launch.cuh"

namespace azb
    {
template<class S, bool VIRIAL> static int launch_dpd_variant(const azb_dpd_args& a, const void* d_params, cudaStream_t stream)
    {
    typedef DPDThermo::Params<S> param_t;
    typedef DPDThermo::Coeff<S> coeff_t;
    DPDArgsDev<S> d;
    d.pair = make_pair_args<S>(a.pair);
    d.vel = (const typename vec4<S>::type*)a.d_vel;
    d.tag = a.d_tag;
    // the evaluator stores the timestep in an `unsigned int` (DPDPairEvaluatorGeneralWeight.h:122,273)
    d.rk = make_round_keys(200, (uint64_t)(uint32_t)a.timestep, a.seed);
    d.deltaT = (S)a.deltaT;
    d.T = (S)a.T;
    const unsigned int tpp = pick_tpp(a.pair);
    if (!valid_tpp(tpp))
        return AZB_ERR_BAD_ARGUMENT;
    const unsigned int block = pick_block(a.pair.block_size);
    const unsigned int per_block = block / tpp;
    const unsigned int grid = (a.pair.N + per_block - 1) / per_block;
    const size_t T2 = (size_t)a.pair.ntypes * a.pair.ntypes;
    const size_t smem = T2 * (sizeof(coeff_t) + sizeof(S));
    if (smem > 48 * 1024)
        return AZB_ERR_BAD_ARGUMENT;
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
