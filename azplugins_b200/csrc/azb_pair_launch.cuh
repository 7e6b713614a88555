// azb_pair_launch.cuh — definition of launch_pair<Pot, S>; included by the per-potential translation units so the
// (shift mode x virial x tpp) kernel variants of different potentials compile in parallel.
#pragma once
#include "azb_launch.cuh"

namespace azb
    {
// Hook for potential-specific tile kernels (e.g. the packed-FP32 PerturbedLJ kernel, azb_plj_packed.cuh): a translation unit
// specialises this before it instantiates launch_pair. try_launch returns true when it has issued the launch.
template<class Pot, class S, int SHIFT, bool VIRIAL> struct SpecialTileKernel
    {
    static bool try_launch(const PairArgsDev<S>&, const TilePlanDev&, const void*, S, bool, bool, unsigned int, unsigned int, unsigned int,
                           cudaStream_t)
        {
        return false;
        }
    };

template<class Pot, class S, int SHIFT, bool VIRIAL>
static int launch_pair_variant(const azb_pair_args& a, const void* d_params, cudaStream_t stream)
    {
    typedef typename Pot::template Params<S> param_t;
    typedef typename Pot::template Coeff<S> coeff_t;
    const PairArgsDev<S> d = make_pair_args<S>(a);
    const size_t T2 = (size_t)a.ntypes * a.ntypes;
    const bool fastbox = d.box.flags == 7u; // direct kernels: straight-line minimum image in the fully periodic box
    if (a.tiles)
        {
        // tile-staged flavour: neighbour positions come from shared memory, so narrow groups lose nothing on
        // coalescing and save on the reduction. The slot rows are laid out for one group width: that is the width.
        const unsigned int tpp = a.tiles->lane_width;
        if (!(tpp == 2 || tpp == 4 || tpp == 8 || tpp == 16) || (a.threads_per_particle && a.threads_per_particle != tpp))
            return AZB_ERR_BAD_ARGUMENT;
        TilePlanDev plan;
        unsigned int block = 0, grid = 0;
        // 128-thread CTAs: eight resident per SM instead of four, so staging phases interleave more finely with walks
        // (4 M LJ: 0.931 ms against 0.975 ms at 256 threads; TwoPatchMorse 2 M: 0.290 against 0.306 ms)
        const int bad = make_tile_plan(a, tpp, tpp, plan, block, grid, 128u);
        if (bad)
            return bad;
        const size_t smem = (size_t)plan.smem_uniq * sizeof(typename vec4<S>::type) + T2 * (sizeof(coeff_t) + 2 * sizeof(S));
        if (smem > AZB_MAX_SMEM)
            return AZB_ERR_BAD_ARGUMENT;
        S wrap_sq;
        const bool lazy = lazy_wrap(d.box, wrap_sq);
        const bool single = a.ntypes == 1;
        if (SpecialTileKernel<Pot, S, SHIFT, VIRIAL>::try_launch(d, plan, d_params, wrap_sq, lazy, single, tpp, grid, block, stream))
            return (int)cudaGetLastError();
#define AZB_LAUNCH_T(TPP, FB, SG)                                                                                         \
    do                                                                                                                    \
        {                                                                                                                 \
        if (const int e_ = ensure_smem<pair_tile_kernel<Pot, S, SHIFT, VIRIAL, TPP, FB, SG>>(smem)) return e_;                                          \
        pair_tile_kernel<Pot, S, SHIFT, VIRIAL, TPP, FB, SG><<<grid, block, smem, stream>>>(d, plan, (const param_t*)d_params, wrap_sq); \
        } while (0)
#define AZB_CALL_T(TPP)                         \
    do                                          \
        {                                       \
        if (lazy && single)                     \
            AZB_LAUNCH_T(TPP, true, true);      \
        else if (lazy)                          \
            AZB_LAUNCH_T(TPP, true, false);     \
        else                                    \
            AZB_LAUNCH_T(TPP, false, false);    \
        } while (0)
        switch (tpp)
            {
        case 2:
            AZB_CALL_T(2);
            break;
        case 4:
            AZB_CALL_T(4);
            break;
        case 8:
            AZB_CALL_T(8);
            break;
        default:
            AZB_CALL_T(16);
            break;
            }
#undef AZB_CALL_T
#undef AZB_LAUNCH_T
        return (int)cudaGetLastError();
        }
    const unsigned int tpp = pick_tpp(a);
    if (!valid_tpp(tpp))
        return AZB_ERR_BAD_ARGUMENT;
    const unsigned int block = pick_block(a.block_size);
    const unsigned int per_block = block / tpp;
    const unsigned int grid = (a.N + per_block - 1) / per_block;
    const size_t smem = T2 * (sizeof(coeff_t) + 2 * sizeof(S));
    if (smem > 200 * 1024)
        return AZB_ERR_BAD_ARGUMENT;
    // kernel flavour: fully periodic orthorhombic box -> straight-line minimum image; one particle type -> coefficients
    // in registers. Anything else takes the general kernel.
    const int flavour = fastbox ? (a.ntypes == 1 ? 2 : 1) : 0;
#define AZB_LAUNCH(TPP, FB, SG)                                                                                            \
    do                                                                                                                     \
        {                                                                                                                  \
        auto kern = pair_force_kernel<Pot, S, SHIFT, VIRIAL, TPP, FB, SG>;                                                 \
        if (smem > 48 * 1024)                                                                                              \
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);                            \
        kern<<<grid, block, smem, stream>>>(d, (const param_t*)d_params);                                                  \
        } while (0)
#define AZB_CALL(TPP)                          \
    do                                         \
        {                                      \
        if (flavour == 2)                      \
            AZB_LAUNCH(TPP, true, true);       \
        else if (flavour == 1)                 \
            AZB_LAUNCH(TPP, true, false);      \
        else                                   \
            AZB_LAUNCH(TPP, false, false);     \
        } while (0)
    AZB_DISPATCH_TPP(tpp, AZB_CALL)
#undef AZB_CALL
#undef AZB_LAUNCH
    return (int)cudaGetLastError();
    }

template<class Pot, class S> int launch_pair(const azb_pair_args& a, const void* d_params, cudaStream_t stream)
    {
    if (a.N == 0)
        return 0;
    const bool v = a.compute_virial != 0;
    switch (a.shift_mode)
        {
    case AZB_SHIFT_NONE:
        return v ? launch_pair_variant<Pot, S, 0, true>(a, d_params, stream) : launch_pair_variant<Pot, S, 0, false>(a, d_params, stream);
    case AZB_SHIFT_SHIFT:
        return v ? launch_pair_variant<Pot, S, 1, true>(a, d_params, stream) : launch_pair_variant<Pot, S, 1, false>(a, d_params, stream);
    default:
        return v ? launch_pair_variant<Pot, S, 2, true>(a, d_params, stream) : launch_pair_variant<Pot, S, 2, false>(a, d_params, stream);
        }
    }
    } // namespace azb

#define AZB_INSTANTIATE_PAIR(Pot)                                                                              \
    template int azb::launch_pair<azb::Pot, float>(const azb_pair_args&, const void*, cudaStream_t);         \
    template int azb::launch_pair<azb::Pot, double>(const azb_pair_args&, const void*, cudaStream_t);
