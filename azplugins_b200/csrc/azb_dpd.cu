// DPD thermostat launcher (float and double; virial on/off; tpp 1..32; direct or tile-staged neighbour data).
#include <cstdlib>

#include "azb_dpd_kernel.cuh"
#include "azb_launch.cuh"

namespace azb
    {
template<class S, bool VIRIAL> static int launch_dpd_variant(const azb_dpd_args& a, const void* d_params, cudaStream_t stream)
    {
    typedef DPDThermo::Params<S> param_t;
    typedef DPDThermo::Coeff<S> coeff_t;
    typedef typename vec4<S>::type V4;
    DPDArgsDev<S> d;
    d.pair = make_pair_args<S>(a.pair);
    d.vel = (const V4*)a.d_vel;
    d.tag = a.d_tag;
    // the evaluator stores the timestep in an `unsigned int` (DPDPairEvaluatorGeneralWeight.h:122,273)
    d.rk = make_round_keys(200, (uint64_t)(uint32_t)a.timestep, a.seed);
    d.deltaT = (S)a.deltaT;
    d.T = (S)a.T;
    // the queue balances in-range work inside a group, so narrow groups (less idle tail, cheaper reductions) win
    unsigned int tpp = a.pair.threads_per_particle;
    if (tpp == 0)
        {
        // rows of a pruned plan are shorter than the list's
        const size_t entries = (a.pair.tiles && a.pair.tiles->n_entries) ? a.pair.tiles->n_entries : a.pair.size_nlist;
        const double nbar = a.pair.N ? (double)entries / (double)a.pair.N : 0.0;
        tpp = nbar <= 8 ? 1 : (nbar <= 16 ? 2 : (nbar <= 96 ? 4 : (nbar <= 256 ? 8 : 16)));
        }
    if (a.pair.tiles && a.pair.threads_per_particle == 0)
        tpp = tpp < 2 ? 2 : (tpp > 16 ? 16 : tpp); // tile kernels exist for 2..16 lanes per particle
    if (!valid_tpp(tpp))
        return AZB_ERR_BAD_ARGUMENT;
    const size_t T2 = (size_t)a.pair.ntypes * a.pair.ntypes;
    S wrap_sq;
    const bool fastbox = lazy_wrap(d.pair.box, wrap_sq);
    const bool single = a.pair.ntypes == 1;

#define AZB_BY_FLAVOUR(LAUNCH, TPP)        \
    do                                     \
        {                                  \
        if (fastbox && single)             \
            LAUNCH(TPP, true, true);       \
        else if (fastbox)                  \
            LAUNCH(TPP, true, false);      \
        else if (single)                   \
            LAUNCH(TPP, false, true);      \
        else                               \
            LAUNCH(TPP, false, false);     \
        } while (0)

    // Tile kernels, lanes per particle left to the engine: lane-private stacks (dpd_stack_tile_kernel) for short and medium rows —
    // one lane per particle up to 20 entries per row (C5's pruned view: 0.726 -> 0.537 ms), two up to 64 (C2: 0.221 -> 0.204 ms) —
    // and the shared-ring walk beyond. AZB_DPD_STACK = 1 | 2 | 4 forces a stack width, 0 the ring (A/B).
    static const char* stack_env = getenv("AZB_DPD_STACK");
    unsigned int stack_tpp = 0u;
    if (a.pair.tiles && a.pair.threads_per_particle == 0)
        {
        const size_t entries = a.pair.tiles->n_entries ? a.pair.tiles->n_entries : a.pair.size_nlist;
        const double nbar = a.pair.N ? (double)entries / (double)a.pair.N : 0.0;
        stack_tpp = stack_env ? (unsigned int)atoi(stack_env) : (nbar <= 20.0 ? 1u : (nbar <= 64.0 ? 2u : 0u));
        }
    if (a.pair.tiles && (stack_tpp == 1u || stack_tpp == 2u || stack_tpp == 4u))
        {
        TilePlanDev plan;
        unsigned int block = 0, grid = 0;
        const int bad = make_tile_plan(a.pair, stack_tpp, 1u, plan, block, grid);
        if (bad)
            return bad;
        const size_t smem = (size_t)plan.smem_uniq * (2 * sizeof(V4) + sizeof(unsigned int)) + T2 * sizeof(coeff_t)
                            + (size_t)DPD_STACK_CAP * block * sizeof(uint16_t);
        if (smem > AZB_MAX_SMEM)
            return AZB_ERR_BAD_ARGUMENT;
#define AZB_LAUNCH_S(TPP, FB, SG)                                                                                                 \
    do                                                                                                                            \
        {                                                                                                                         \
        if (const int e_ = ensure_smem<dpd_stack_tile_kernel<S, VIRIAL, TPP, FB, SG>>(smem)) return e_;                           \
        dpd_stack_tile_kernel<S, VIRIAL, TPP, FB, SG><<<grid, block, smem, stream>>>(d, plan, (const param_t*)d_params, wrap_sq);  \
        } while (0)
        if (stack_tpp == 1u)
            AZB_BY_FLAVOUR(AZB_LAUNCH_S, 1);
        else if (stack_tpp == 2u)
            AZB_BY_FLAVOUR(AZB_LAUNCH_S, 2);
        else
            AZB_BY_FLAVOUR(AZB_LAUNCH_S, 4);
#undef AZB_LAUNCH_S
        return (int)cudaGetLastError();
        }
    if (a.pair.tiles)
        {
        TilePlanDev plan;
        unsigned int block = 0, grid = 0;
        const int bad = make_tile_plan(a.pair, tpp, 1u, plan, block, grid);
        if (bad)
            return bad;
        if (a.pair.ntypes > 255)
            return AZB_ERR_BAD_ARGUMENT; // the ring packs slot | type pair << 16
        const size_t smem = (size_t)plan.smem_uniq * (2 * sizeof(V4) + sizeof(unsigned int)) + (size_t)(block / tpp) * (AZB_DPD_QMUL * tpp + 1) * sizeof(V4)
                            + T2 * sizeof(coeff_t);
        if (smem > AZB_MAX_SMEM)
            return AZB_ERR_BAD_ARGUMENT;
#define AZB_LAUNCH_T(TPP, FB, SG)                                                            \
    do                                                                                       \
        {                                                                                    \
        if (const int e_ = ensure_smem<dpd_tile_kernel<S, VIRIAL, TPP, FB, SG>>(smem)) return e_;                          \
        dpd_tile_kernel<S, VIRIAL, TPP, FB, SG><<<grid, block, smem, stream>>>(d, plan, (const param_t*)d_params, wrap_sq); \
        } while (0)
        switch (tpp)
            {
        case 2:
            AZB_BY_FLAVOUR(AZB_LAUNCH_T, 2);
            break;
        case 4:
            AZB_BY_FLAVOUR(AZB_LAUNCH_T, 4);
            break;
        case 8:
            AZB_BY_FLAVOUR(AZB_LAUNCH_T, 8);
            break;
        case 16:
            AZB_BY_FLAVOUR(AZB_LAUNCH_T, 16);
            break;
        default:
            return AZB_ERR_BAD_ARGUMENT; // tile kernels exist for 2..16 lanes per particle
            }
#undef AZB_LAUNCH_T
        return (int)cudaGetLastError();
        }

    const unsigned int block = pick_block(a.pair.block_size);
    const unsigned int per_block = block / tpp;
    const unsigned int grid = (a.pair.N + per_block - 1) / per_block;
    const size_t smem = (size_t)(block / tpp) * (AZB_DPD_QMUL * tpp + 1) * (sizeof(V4) + sizeof(unsigned int)) + T2 * sizeof(coeff_t);
    if (smem > AZB_MAX_SMEM)
        return AZB_ERR_BAD_ARGUMENT;
#define AZB_LAUNCH_Q(TPP, FB, SG)                                                            \
    do                                                                                       \
        {                                                                                    \
        if (const int e_ = ensure_smem<dpd_force_kernel_q<S, VIRIAL, TPP, FB, SG>>(smem)) return e_;                       \
        dpd_force_kernel_q<S, VIRIAL, TPP, FB, SG><<<grid, block, smem, stream>>>(d, (const param_t*)d_params, wrap_sq); \
        } while (0)
#define AZB_CALL(TPP) AZB_BY_FLAVOUR(AZB_LAUNCH_Q, TPP)
    AZB_DISPATCH_TPP(tpp, AZB_CALL)
#undef AZB_CALL
#undef AZB_LAUNCH_Q
#undef AZB_BY_FLAVOUR
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
