// Bond launcher (DoubleWell, Quartic; float and double).
#include "azb_bond_kernel.cuh"
#include "azb_launch.cuh"

namespace azb
    {
template<class Pot, class S> int launch_bond(const azb_bond_args& a, const void* d_params, unsigned int* d_flags, cudaStream_t stream)
    {
    typedef typename Pot::template Params<S> param_t;
    if (a.N == 0)
        return 0;
    if (!a.d_force || !a.d_pos || !a.d_gpu_bondlist || !a.d_gpu_n_bonds || !d_flags || a.pitch < a.N)
        return AZB_ERR_BAD_ARGUMENT;
    if (a.compute_virial && (!a.d_virial || a.virial_pitch < a.N))
        return AZB_ERR_BAD_ARGUMENT;
    BondArgsDev<S> d;
    d.force = (typename vec4<S>::type*)a.d_force;
    d.virial = (S*)a.d_virial;
    d.virial_pitch = a.virial_pitch;
    d.N = a.N;
    d.pos = (const typename vec4<S>::type*)a.d_pos;
    d.box = make_box<S>(a.box);
    d.bondlist = (const uint2*)a.d_gpu_bondlist;
    d.n_bonds = a.d_gpu_n_bonds;
    d.pitch = a.pitch;
    d.n_bond_types = a.n_bond_types;
    d.compute_virial = a.compute_virial;
    const unsigned int block = pick_block(a.block_size ? a.block_size : 256);
    const unsigned int grid = (a.N + block - 1) / block;
    const size_t smem = (size_t)a.n_bond_types * sizeof(param_t);
    if (smem > 200 * 1024)
        return AZB_ERR_BAD_ARGUMENT;
    auto kern = bond_force_kernel<Pot, S>;
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<grid, block, smem, stream>>>(d, (const param_t*)d_params, d_flags);
    return (int)cudaGetLastError();
    }
template int launch_bond<DoubleWell, float>(const azb_bond_args&, const void*, unsigned int*, cudaStream_t);
template int launch_bond<DoubleWell, double>(const azb_bond_args&, const void*, unsigned int*, cudaStream_t);
template int launch_bond<Quartic, float>(const azb_bond_args&, const void*, unsigned int*, cudaStream_t);
template int launch_bond<Quartic, double>(const azb_bond_args&, const void*, unsigned int*, cudaStream_t);
    } // namespace azb
