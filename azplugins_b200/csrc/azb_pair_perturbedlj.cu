// PerturbedLJ kernel variants (all shift modes, virial on/off, tpp 1..32, float and double).
#include "azb_pair_launch.cuh"
AZB_INSTANTIATE_PAIR(PerturbedLJ)
