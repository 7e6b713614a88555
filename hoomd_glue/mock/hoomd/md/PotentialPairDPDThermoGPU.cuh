// MOCK (compile-only guard) of hoomd/md/PotentialPairDPDThermoGPU.cuh: dpd_pair_args_t as recalled from HOOMD-blue v5.
#pragma once
#include "hoomd/BoxDim.h"
#include "hoomd/HOOMDMath.h"
namespace hoomd
    {
namespace md
    {
namespace kernel
    {
struct dpd_pair_args_t
    {
    Scalar4* d_force;
    Scalar* d_virial;
    const size_t virial_pitch;
    const unsigned int N;
    const unsigned int n_max;
    const Scalar4* d_pos;
    const Scalar4* d_vel;
    const unsigned int* d_tag;
    const BoxDim box;
    const unsigned int* d_n_neigh;
    const unsigned int* d_nlist;
    const size_t* d_head_list;
    const Scalar* d_rcutsq;
    const size_t size_neigh_list;
    const unsigned int ntypes;
    const unsigned int block_size;
    const uint16_t seed;
    const uint64_t timestep;
    const Scalar deltaT;
    const Scalar T;
    const unsigned int shift_mode;
    const unsigned int compute_virial;
    const unsigned int threads_per_particle;
    const hipDeviceProp_t& devprop;
    };
template<class evaluator> hipError_t gpu_compute_dpd_forces(const dpd_pair_args_t& args, const typename evaluator::param_type* d_params);
    } // namespace kernel
    } // namespace md
    } // namespace hoomd
