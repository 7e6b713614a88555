// PerturbedLJ kernel variants (all shift modes, virial on/off, tpp 1..32, float and double) + the packed-FP32 tile kernel
// for the single-type float case.
#include <cstdlib>

#include "azb_pair_launch.cuh"
#include "azb_plj_packed.cuh"

namespace azb
    {
template<int SHIFT, bool VIRIAL> struct PackedPLJ
    {
    static bool try_launch(const PairArgsDev<float>& d, const TilePlanDev& plan, const void* d_params, float wrap_sq, bool lazy, bool single,
                           unsigned int tpp, unsigned int grid, unsigned int block, cudaStream_t stream)
        {
        static const bool disabled = getenv("AZB_NO_PACKED") != nullptr; // A/B switch against the generic tile kernel
        if (disabled || !lazy || !single)
            return false;
        typedef PerturbedLJ::Params<float> param_t;
        const size_t smem = (size_t)plan.smem_uniq * sizeof(float4);
        if (smem > AZB_MAX_SMEM)
            return false;
#define AZB_PACKED(TPP)                                                                                                      \
    do                                                                                                                       \
        {                                                                                                                    \
        if (ensure_smem<plj_packed_tile_kernel<SHIFT, VIRIAL, TPP>>(smem) != 0) return false;                                                       \
        plj_packed_tile_kernel<SHIFT, VIRIAL, TPP><<<grid, block, smem, stream>>>(d, plan, (const param_t*)d_params, wrap_sq); \
        } while (0)
        switch (tpp)
            {
        case 2:
            AZB_PACKED(2);
            break;
        case 4:
            AZB_PACKED(4);
            break;
        case 8:
            AZB_PACKED(8);
            break;
        default:
            AZB_PACKED(16);
            break;
            }
#undef AZB_PACKED
        return true;
        }
    };
// modes none and shift; xplor keeps the generic kernel
template<bool VIRIAL> struct SpecialTileKernel<PerturbedLJ, float, 0, VIRIAL> : PackedPLJ<0, VIRIAL>
    {
    };
template<bool VIRIAL> struct SpecialTileKernel<PerturbedLJ, float, 1, VIRIAL> : PackedPLJ<1, VIRIAL>
    {
    };
    } // namespace azb

AZB_INSTANTIATE_PAIR(PerturbedLJ)
