// MOCK (compile-only guard) of hoomd/md/PotentialPairGPU.cuh: pair_args_t as recalled from HOOMD-blue v5 and the declaration
// of the kernel driver the reference instantiates (/root/reference/src/PotentialPairGPUKernel.cu.inc:25-28).
#pragma once
#include "hoomd/BoxDim.h"
#include "hoomd/HOOMDMath.h"
namespace hoomd
    {
namespace md
    {
namespace kernel
    {
struct pair_args_t
    {
    Scalar4* d_force;
    Scalar* d_virial;
    const size_t virial_pitch;
    const unsigned int N;
    const unsigned int n_max;
    const Scalar4* d_pos;
    const Scalar* d_charge;
    const BoxDim box;
    const unsigned int* d_n_neigh;
    const unsigned int* d_nlist;
    const size_t* d_head_list;
    const Scalar* d_rcutsq;
    const Scalar* d_ronsq;
    const size_t size_neigh_list;
    const unsigned int ntypes;
    const unsigned int block_size;
    const unsigned int shift_mode;
    const unsigned int compute_virial;
    const unsigned int threads_per_particle;
    const hipDeviceProp_t& devprop;
    };
template<class evaluator> hipError_t gpu_compute_pair_forces(const pair_args_t& pair_args, const typename evaluator::param_type* d_params);
    } // namespace kernel
    } // namespace md
    } // namespace hoomd
