// MOCK (compile-only guard) of hoomd/md/AnisoPotentialPairGPU.cuh: a_pair_args_t as recalled from HOOMD-blue v5
// (it carries no size_neigh_list).
#pragma once
#include "hoomd/BoxDim.h"
#include "hoomd/HOOMDMath.h"
namespace hoomd
    {
namespace md
    {
namespace kernel
    {
struct a_pair_args_t
    {
    Scalar4* d_force;
    Scalar4* d_torque;
    Scalar* d_virial;
    const size_t virial_pitch;
    const unsigned int N;
    const unsigned int n_max;
    const Scalar4* d_pos;
    const Scalar* d_charge;
    const Scalar4* d_orientation;
    const unsigned int* d_tag;
    const BoxDim box;
    const unsigned int* d_n_neigh;
    const unsigned int* d_nlist;
    const size_t* d_head_list;
    const Scalar* d_rcutsq;
    const unsigned int ntypes;
    const unsigned int block_size;
    const unsigned int shift_mode;
    const unsigned int compute_virial;
    const unsigned int threads_per_particle;
    const hipDeviceProp_t& devprop;
    };
template<class evaluator>
hipError_t gpu_compute_pair_aniso_forces(const a_pair_args_t& args, const typename evaluator::param_type* d_params,
                                         const typename evaluator::shape_type* d_shape_params);
    } // namespace kernel
    } // namespace md
    } // namespace hoomd
