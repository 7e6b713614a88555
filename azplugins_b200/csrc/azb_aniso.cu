// TwoPatchMorse launcher (float and double; shift none/shift; virial on/off; tpp 1..32).
#include "azb_aniso_kernel.cuh"
#include "azb_launch.cuh"

namespace azb
    {
template<class S, int SHIFT, bool VIRIAL> static int launch_aniso_variant(const azb_aniso_args& a, const void* d_params, cudaStream_t stream)
    {
    typedef TwoPatchMorse::Params<S> param_t;
    typedef TwoPatchMorse::Coeff<S> coeff_t;
    AnisoArgsDev<S> d;
    d.pair = make_pair_args<S>(a.pair);
    d.orientation = (const typename vec4<S>::type*)a.d_orientation;
    d.torque = (typename vec4<S>::type*)a.d_torque;
    const size_t T2t = (size_t)a.pair.ntypes * a.pair.ntypes;
    if (a.pair.tiles)
        {
        if (a.pair.tiles->lane_width == 1u)
            {
            // plain-order slot rows: the in-range-compacting walk (azb_ring_walk.cuh), 4 lanes per particle unless asked otherwise
            const unsigned int tpp = a.pair.threads_per_particle ? a.pair.threads_per_particle : 4u;
            TilePlanDev plan;
            unsigned int block = 0, grid = 0;
            const int bad = make_tile_plan(a.pair, tpp, 1u, plan, block, grid);
            if (bad)
                return bad;
            if (a.pair.ntypes > 255)
                return AZB_ERR_BAD_ARGUMENT; // the ring packs slot | type pair << 16
            S wrap_sq;
            const bool lazy = lazy_wrap(d.pair.box, wrap_sq);
            const bool single = a.pair.ntypes == 1;
#define AZB_RING_T(TPP, FB, SG)                                                                                                              \
    do                                                                                                                                       \
        {                                                                                                                                    \
        const size_t smem = (size_t)plan.smem_uniq * 2 * sizeof(typename vec4<S>::type) + (size_t)(block / TPP) * ring_stride<S, TPP>()      \
                            + T2t * (sizeof(coeff_t) + sizeof(S));                                                                           \
        if (smem > AZB_MAX_SMEM)                                                                                                             \
            return AZB_ERR_BAD_ARGUMENT;                                                                                                     \
        if (const int e_ = ensure_smem<aniso_ring_tile_kernel<S, SHIFT, VIRIAL, TPP, FB, SG>>(smem)) return e_;                              \
        aniso_ring_tile_kernel<S, SHIFT, VIRIAL, TPP, FB, SG><<<grid, block, smem, stream>>>(d, plan, (const param_t*)d_params, wrap_sq);     \
        } while (0)
#define AZB_RING_F(TPP)                    \
    do                                     \
        {                                  \
        if (lazy && single)                \
            AZB_RING_T(TPP, true, true);   \
        else if (lazy)                     \
            AZB_RING_T(TPP, true, false);  \
        else                               \
            AZB_RING_T(TPP, false, false); \
        } while (0)
            switch (tpp)
                {
            case 2:
                AZB_RING_F(2);
                break;
            case 4:
                AZB_RING_F(4);
                break;
            case 8:
                AZB_RING_F(8);
                break;
            default:
                return AZB_ERR_BAD_ARGUMENT;
                }
#undef AZB_RING_F
#undef AZB_RING_T
            return (int)cudaGetLastError();
            }
        const unsigned int tpp = a.pair.tiles->lane_width; // the slot rows are laid out for one group width
        if (!(tpp == 2 || tpp == 4 || tpp == 8 || tpp == 16) || (a.pair.threads_per_particle && a.pair.threads_per_particle != tpp))
            return AZB_ERR_BAD_ARGUMENT;
        TilePlanDev plan;
        unsigned int block = 0, grid = 0;
        const int bad = make_tile_plan(a.pair, tpp, tpp, plan, block, grid, 128u);
        if (bad)
            return bad;
        const size_t smem = (size_t)plan.smem_uniq * 2 * sizeof(typename vec4<S>::type) + T2t * (sizeof(coeff_t) + sizeof(S));
        if (smem > AZB_MAX_SMEM)
            return AZB_ERR_BAD_ARGUMENT;
        S wrap_sq;
        const bool lazy = lazy_wrap(d.pair.box, wrap_sq);
#define AZB_CALL_T(TPP)                                                                                                        \
    do                                                                                                                         \
        {                                                                                                                      \
        if (lazy)                                                                                                              \
            {                                                                                                                  \
            if (const int e_ = ensure_smem<aniso_tile_kernel<S, SHIFT, VIRIAL, TPP, true>>(smem)) return e_;                                                 \
            aniso_tile_kernel<S, SHIFT, VIRIAL, TPP, true><<<grid, block, smem, stream>>>(d, plan, (const param_t*)d_params, wrap_sq);  \
            }                                                                                                                  \
        else                                                                                                                   \
            {                                                                                                                  \
            if (const int e_ = ensure_smem<aniso_tile_kernel<S, SHIFT, VIRIAL, TPP, false>>(smem)) return e_;                                                \
            aniso_tile_kernel<S, SHIFT, VIRIAL, TPP, false><<<grid, block, smem, stream>>>(d, plan, (const param_t*)d_params, wrap_sq); \
            }                                                                                                                  \
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
        return (int)cudaGetLastError();
        }
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
            aniso_force_kernel<S, SHIFT, VIRIAL, TPP, true><<<grid, block, smem, stream>>>(d, (const param_t*)d_params);  \
        else                                                                                    \
            aniso_force_kernel<S, SHIFT, VIRIAL, TPP, false><<<grid, block, smem, stream>>>(d, (const param_t*)d_params); \
        } while (0)
    AZB_DISPATCH_TPP(tpp, AZB_CALL)
#undef AZB_CALL
    return (int)cudaGetLastError();
    }

template<class S> int launch_aniso(const azb_aniso_args& a, const void* d_params, cudaStream_t stream)
    {
    if (a.pair.N == 0)
        return 0;
    if (!a.d_orientation || !a.d_torque)
        return AZB_ERR_BAD_ARGUMENT;
    const bool v = a.pair.compute_virial != 0;
    // HOOMD's AnisoPotentialPair knows modes none and shift; TwoPatchMorse implements the shift (:122-125)
    if (a.pair.shift_mode == AZB_SHIFT_SHIFT)
        return v ? launch_aniso_variant<S, 1, true>(a, d_params, stream) : launch_aniso_variant<S, 1, false>(a, d_params, stream);
    if (a.pair.shift_mode == AZB_SHIFT_NONE)
        return v ? launch_aniso_variant<S, 0, true>(a, d_params, stream) : launch_aniso_variant<S, 0, false>(a, d_params, stream);
    return AZB_ERR_BAD_ARGUMENT;
    }
template int launch_aniso<float>(const azb_aniso_args&, const void*, cudaStream_t);
template int launch_aniso<double>(const azb_aniso_args&, const void*, cudaStream_t);
    } // namespace azb
