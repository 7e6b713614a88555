// ExpandedYukawa kernel variants (all shift modes, virial on/off, tpp 1..32, float and double) + the packed-FP32 tile kernel
// for the single-type float case (azb_packed_pair.cuh).
#include <cstdlib>

#include "azb_pair_launch.cuh"
#include "azb_packed_pair.cuh"

namespace azb
    {
// modes none and shift; xplor keeps the generic kernel
template<bool VIRIAL> struct SpecialTileKernel<ExpandedYukawa, float, 0, VIRIAL> : PackedPairLaunch<PackedExpandedYukawa, 0, VIRIAL>
    {
    };
template<bool VIRIAL> struct SpecialTileKernel<ExpandedYukawa, float, 1, VIRIAL> : PackedPairLaunch<PackedExpandedYukawa, 1, VIRIAL>
    {
    };
    } // namespace azb

AZB_INSTANTIATE_PAIR(ExpandedYukawa)
