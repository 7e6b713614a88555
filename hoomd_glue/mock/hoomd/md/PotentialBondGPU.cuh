// MOCK (compile-only guard) of hoomd/md/PotentialBondGPU.cuh (+ the bits of hoomd/BondedGroupData.cuh and hoomd/Index1D.h it
// pulls in): bond_args_t<2>, group_storage<2>, Index2D as recalled from HOOMD-blue v5.
#pragma once
#include "hoomd/BoxDim.h"
#include "hoomd/HOOMDMath.h"
namespace hoomd
    {
template<unsigned int group_size> union group_storage
    {
    unsigned int tag[group_size];
    unsigned int idx[group_size];
    };
class Index2D
    {
    public:
    Index2D(unsigned int w = 0, unsigned int h = 0) : m_w(w), m_h(h) { }
    unsigned int getW() const
        {
        return m_w;
        }
    unsigned int getH() const
        {
        return m_h;
        }

    private:
    unsigned int m_w, m_h;
    };
namespace md
    {
namespace kernel
    {
template<int group_size> struct bond_args_t
    {
    Scalar4* d_force;
    Scalar* d_virial;
    const size_t virial_pitch;
    unsigned int N;
    unsigned int n_max;
    const Scalar4* d_pos;
    const Scalar* d_charge;
    const BoxDim box;
    const group_storage<group_size>* d_gpu_bondlist;
    const Index2D gpu_table_indexer;
    const unsigned int* d_gpu_bond_pos;
    const unsigned int* d_gpu_n_bonds;
    const unsigned int n_bond_types;
    const unsigned int block_size;
    const hipDeviceProp_t& devprop;
    };
template<class evaluator, int group_size>
hipError_t gpu_compute_bond_forces(const bond_args_t<group_size>& bond_args, const typename evaluator::param_type* d_params,
                                   unsigned int* d_flags);
    } // namespace kernel
    } // namespace md
    } // namespace hoomd
