// azb_abi.cu — the extern "C" boundary declared in include/azb200.h. No torch types, no exceptions: plain pointers
// and sizes in, error code out. Each entry point forwards to the typed launcher of one kernel family.
#include <cuda_runtime.h>

#include "azb200.h"
#include "azb_aniso_kernel.cuh"
#include "azb_bond_kernel.cuh"
#include "azb_dpd_kernel.cuh"
#include "azb_launch.cuh"

namespace azb
    {
template<class S> int launch_dpd(const azb_dpd_args&, const void*, cudaStream_t);
template<class S> int launch_aniso(const azb_aniso_args&, const void*, cudaStream_t);
template<class Pot, class S> int launch_bond(const azb_bond_args&, const void*, unsigned int*, cudaStream_t);

template<class Pot> static int pair_by_scalar(int scalar_bits, const azb_pair_args& a, const void* p, cudaStream_t s)
    {
    if (scalar_bits == 32)
        return launch_pair<Pot, float>(a, p, s);
    if (scalar_bits == 64)
        return launch_pair<Pot, double>(a, p, s);
    return AZB_ERR_BAD_SCALAR;
    }

template<template<class> class P> static size_t psize(int bits)
    {
    return bits == 32 ? sizeof(P<float>) : (bits == 64 ? sizeof(P<double>) : 0);
    }
template<template<class> class P> static size_t palign(int bits)
    {
    return bits == 32 ? alignof(P<float>) : (bits == 64 ? alignof(P<double>) : 0);
    }

static size_t param_size_or_align(int kind, int potential, int bits, bool want_align)
    {
#define AZB_PS(P) (want_align ? palign<P>(bits) : psize<P>(bits))
    switch (kind)
        {
    case AZB_KIND_PAIR:
        switch (potential)
            {
        case AZB_PAIR_PERTURBED_LENNARD_JONES:
            return AZB_PS(PerturbedLJ::Params);
        case AZB_PAIR_COLLOID:
            return AZB_PS(Colloid::Params);
        case AZB_PAIR_EXPANDED_YUKAWA:
            return AZB_PS(ExpandedYukawa::Params);
        case AZB_PAIR_HERTZ:
            return AZB_PS(Hertz::Params);
        case AZB_PAIR_DPD_CONSERVATIVE:
            return AZB_PS(DPDConservative::Params);
            }
        return 0;
    case AZB_KIND_DPD:
        return potential == AZB_DPD_GENERAL_WEIGHT ? AZB_PS(DPDThermo::Params) : 0;
    case AZB_KIND_ANISO:
        return potential == AZB_ANISO_TWO_PATCH_MORSE ? AZB_PS(TwoPatchMorse::Params) : 0;
    case AZB_KIND_BOND:
        if (potential == AZB_BOND_DOUBLE_WELL)
            return AZB_PS(DoubleWell::Params);
        if (potential == AZB_BOND_QUARTIC)
            return AZB_PS(Quartic::Params);
        return 0;
        }
#undef AZB_PS
    return 0;
    }
    } // namespace azb

extern "C"
    {
    int azb_abi_version(void)
        {
        return AZB200_ABI_VERSION;
        }

    size_t azb_param_size(int kind, int potential, int scalar_bits)
        {
        return azb::param_size_or_align(kind, potential, scalar_bits, false);
        }

    size_t azb_param_align(int kind, int potential, int scalar_bits)
        {
        return azb::param_size_or_align(kind, potential, scalar_bits, true);
        }

    const char* azb_potential_name(int kind, int potential)
        {
        static const char* pair_names[] = {"PerturbedLennardJones", "colloid", "ExpandedYukawa", "hertz", "dpd_gen"};
        static const char* bond_names[] = {"DoubleWell", "Quartic"};
        if (kind == AZB_KIND_PAIR && potential >= 0 && potential < AZB_PAIR_COUNT_)
            return pair_names[potential];
        if (kind == AZB_KIND_DPD && potential == AZB_DPD_GENERAL_WEIGHT)
            return "dpd_gen";
        if (kind == AZB_KIND_ANISO && potential == AZB_ANISO_TWO_PATCH_MORSE)
            return "TwoPatchMorse";
        if (kind == AZB_KIND_BOND && potential >= 0 && potential < AZB_BOND_COUNT_)
            return bond_names[potential];
        return NULL;
        }

    int azb_pair_forces(int potential, int scalar_bits, const azb_pair_args* args, const void* d_params, void* stream)
        {
        const int bad = azb::check_pair_args(args);
        if (bad)
            return bad;
        if (!d_params)
            return AZB_ERR_BAD_ARGUMENT;
        cudaStream_t s = (cudaStream_t)stream;
        switch (potential)
            {
        case AZB_PAIR_PERTURBED_LENNARD_JONES:
            return azb::pair_by_scalar<azb::PerturbedLJ>(scalar_bits, *args, d_params, s);
        case AZB_PAIR_COLLOID:
            return azb::pair_by_scalar<azb::Colloid>(scalar_bits, *args, d_params, s);
        case AZB_PAIR_EXPANDED_YUKAWA:
            return azb::pair_by_scalar<azb::ExpandedYukawa>(scalar_bits, *args, d_params, s);
        case AZB_PAIR_HERTZ:
            return azb::pair_by_scalar<azb::Hertz>(scalar_bits, *args, d_params, s);
        case AZB_PAIR_DPD_CONSERVATIVE:
            return azb::pair_by_scalar<azb::DPDConservative>(scalar_bits, *args, d_params, s);
            }
        return AZB_ERR_BAD_POTENTIAL;
        }

    int azb_dpd_forces(int potential, int scalar_bits, const azb_dpd_args* args, const void* d_params, void* stream)
        {
        if (!args)
            return AZB_ERR_BAD_ARGUMENT;
        const int bad = azb::check_pair_args(&args->pair);
        if (bad)
            return bad;
        if (!d_params || (args->pair.N && (!args->d_vel || !args->d_tag)) || args->pair.shift_mode != AZB_SHIFT_NONE)
            return AZB_ERR_BAD_ARGUMENT;
        if (potential != AZB_DPD_GENERAL_WEIGHT)
            return AZB_ERR_BAD_POTENTIAL;
        if (scalar_bits == 32)
            return azb::launch_dpd<float>(*args, d_params, (cudaStream_t)stream);
        if (scalar_bits == 64)
            return azb::launch_dpd<double>(*args, d_params, (cudaStream_t)stream);
        return AZB_ERR_BAD_SCALAR;
        }

    int azb_aniso_forces(int potential, int scalar_bits, const azb_aniso_args* args, const void* d_params, void* stream)
        {
        if (!args)
            return AZB_ERR_BAD_ARGUMENT;
        const int bad = azb::check_pair_args(&args->pair);
        if (bad)
            return bad;
        if (!d_params)
            return AZB_ERR_BAD_ARGUMENT;
        if (potential != AZB_ANISO_TWO_PATCH_MORSE)
            return AZB_ERR_BAD_POTENTIAL;
        if (scalar_bits == 32)
            return azb::launch_aniso<float>(*args, d_params, (cudaStream_t)stream);
        if (scalar_bits == 64)
            return azb::launch_aniso<double>(*args, d_params, (cudaStream_t)stream);
        return AZB_ERR_BAD_SCALAR;
        }

    int azb_bond_forces(int potential, int scalar_bits, const azb_bond_args* args, const void* d_params, unsigned int* d_flags, void* stream)
        {
        if (!args || !d_params)
            return AZB_ERR_BAD_ARGUMENT;
        if (scalar_bits != 32 && scalar_bits != 64)
            return AZB_ERR_BAD_SCALAR;
        cudaStream_t s = (cudaStream_t)stream;
        switch (potential)
            {
        case AZB_BOND_DOUBLE_WELL:
            return scalar_bits == 32 ? azb::launch_bond<azb::DoubleWell, float>(*args, d_params, d_flags, s)
                                     : azb::launch_bond<azb::DoubleWell, double>(*args, d_params, d_flags, s);
        case AZB_BOND_QUARTIC:
            return scalar_bits == 32 ? azb::launch_bond<azb::Quartic, float>(*args, d_params, d_flags, s)
                                     : azb::launch_bond<azb::Quartic, double>(*args, d_params, d_flags, s);
            }
        return AZB_ERR_BAD_POTENTIAL;
        }
    }
