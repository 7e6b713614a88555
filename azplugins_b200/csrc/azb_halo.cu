// azb_halo.cu — packing of boundary particles for the ghost halo exchange (SURVEY.md §8 e). HOOMD's CommunicatorGPU does
// this step over MPI; here the packed buffer is handed to NCCL send/recv by the host layer (azplugins_b200/domain.py).
#include <cuda_runtime.h>

#include "azb200.h"
#include "azb_math.cuh"

namespace azb
    {
template<class S>
__global__ void gather_scalar4_kernel(const typename vec4<S>::type* __restrict__ src, const unsigned int* __restrict__ idx,
                                      const typename vec4<S>::type* __restrict__ shift, typename vec4<S>::type* __restrict__ out, size_t n)
    {
    const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n)
        return;
    auto v = ld4(src + __ldg(idx + e));
    if (shift)
        {
        const auto s = ld4(shift + e);
        v.x += s.x;
        v.y += s.y;
        v.z += s.z;
        }
    st4(out + e, v.x, v.y, v.z, v.w);
    }
    } // namespace azb

extern "C" int azb_gather_scalar4(int scalar_bits, const void* d_src, const unsigned int* d_idx, const void* d_shift, void* d_out, size_t n,
                                  void* stream)
    {
    if (n == 0)
        return 0;
    if (!d_src || !d_idx || !d_out)
        return AZB_ERR_BAD_ARGUMENT;
    const unsigned int block = 256;
    const unsigned int grid = (unsigned int)((n + block - 1) / block);
    cudaStream_t s = (cudaStream_t)stream;
    if (scalar_bits == 32)
        azb::gather_scalar4_kernel<float><<<grid, block, 0, s>>>((const float4*)d_src, d_idx, (const float4*)d_shift, (float4*)d_out, n);
    else if (scalar_bits == 64)
        azb::gather_scalar4_kernel<double><<<grid, block, 0, s>>>((const azb::double4a*)d_src, d_idx, (const azb::double4a*)d_shift,
                                                                  (azb::double4a*)d_out, n);
    else
        return AZB_ERR_BAD_SCALAR;
    return (int)cudaGetLastError();
    }
