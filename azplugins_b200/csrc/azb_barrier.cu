// azb_barrier.cu — harmonic barrier forces: the B200 form of azplugins::gpu::compute_harmonic_barrier
// (/root/reference/src/HarmonicBarrierGPU.cuh:49-133). Streaming work: 16 (32) bytes in, 16 (32) bytes out per particle plus
// the zeroed virial rows, so the kernel is written for coalesced 128-bit traffic and nothing else: one thread per particle, a
// grid-stride loop sized to the SM count, the virial cleared by the same threads instead of a second memset pass.
#include <cuda_runtime.h>

#include "azb200.h"
#include "azb_launch.cuh"
#include <cmath>

#include "azb_math.cuh"

namespace azb
    {
template<class S> struct BarrierArgsDev
    {
    typename vec4<S>::type* force;
    S* virial;
    size_t virial_pitch;
    unsigned int N, ntypes;
    const typename vec4<S>::type* pos;
    const S* params; // k, offset interleaved
    Box<S> box;
    S location;
    };

// BoxDim::wrap for a position (HarmonicBarrierGPU.cuh:76-78): shift by ONE image per periodic direction, z then y then x,
// with the tilt terms of a triclinic box (HOOMD-blue v5 BoxDim.h; restated, HOOMD is not under /root/reference).
template<class S> __device__ __forceinline__ void wrap_once(const Box<S>& b, S& x, S& y, S& z)
    {
    const S hx = b.Lx / S(2), hy = b.Ly / S(2), hz = b.Lz / S(2);
    if (b.flags & 4u)
        {
        if (z >= hz)
            {
            z -= b.Lz;
            y -= b.Lz * b.yz;
            x -= b.Lz * b.xz;
            }
        else if (z < -hz)
            {
            z += b.Lz;
            y += b.Lz * b.yz;
            x += b.Lz * b.xz;
            }
        }
    if (b.flags & 2u)
        {
        const S tilt_y = b.yz * z;
        if (y >= hy + tilt_y)
            {
            y -= b.Ly;
            x -= b.Ly * b.xy;
            }
        else if (y < -hy + tilt_y)
            {
            y += b.Ly;
            x += b.Ly * b.xy;
            }
        }
    if (b.flags & 1u)
        {
        const S tilt_x = (b.xz - b.xy * b.yz) * z + b.xy * y;
        if (x >= hx + tilt_x)
            x -= b.Lx;
        else if (x < -hx + tilt_x)
            x += b.Lx;
        }
    }

template<class S, int GEOMETRY> __global__ void __launch_bounds__(256) barrier_kernel(const BarrierArgsDev<S> a)
    {
    extern __shared__ unsigned char s_raw[];
    S* s_params = (S*)s_raw;
    for (unsigned int t = threadIdx.x; t < 2u * a.ntypes; t += blockDim.x)
        s_params[t] = a.params[t];
    __syncthreads();
    const unsigned int stride = gridDim.x * blockDim.x;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.N; i += stride)
        {
        const auto p = ld4(a.pos + i);
        S x = p.x, y = p.y, z = p.z;
        const unsigned int type = type_of(p.w);
        const S k = s_params[2u * type], offset = s_params[2u * type + 1u];
        wrap_once(a.box, x, y, z);
        S fx = S(0), fy = S(0), fz = S(0), e = S(0);
        if (GEOMETRY == AZB_BARRIER_PLANAR)
            {
            // PlanarBarrierEvaluator.h:36-48
            const S dy = y - (a.location + offset);
            if (dy > S(0))
                {
                fy = -k * dy;
                e = S(-0.5) * fy * dy;
                }
            }
        else
            {
            // SphericalBarrierEvaluator.h:36-51
            const S r = sqrt_(x * x + y * y + z * z);
            const S dr = r - (a.location + offset);
            if (dr > S(0))
                {
                const S k_dr = k * dr;
                const S c = -(k_dr / r);
                fx = c * x;
                fy = c * y;
                fz = c * z;
                e = S(0.5) * k_dr * dr;
                }
            }
        st4(a.force + i, fx, fy, fz, e);
        if (a.virial)
            {
#pragma unroll
            for (int c = 0; c < 6; ++c)
                a.virial[(size_t)c * a.virial_pitch + i] = S(0);
            }
        }
    }

// BarrierEvaluatorT::valid(global_box), in Scalar arithmetic like the reference's host call (HarmonicBarrier.h:125-133)
template<class S> static bool barrier_valid(int geometry, const azb_box& b, S location)
    {
    const S Lx = (S)b.Lx, Ly = (S)b.Ly, Lz = (S)b.Lz, xy = (S)b.xy, xz = (S)b.xz, yz = (S)b.yz;
    if (geometry == AZB_BARRIER_PLANAR)
        {
        // lo = makeCoordinates(0,0,0), hi = makeCoordinates(1,1,1): only y is needed (PlanarBarrierEvaluator.h:51-56)
        const S lo_z = -Lz / S(2), hi_z = lo_z + Lz;
        const S lo_y = -Ly / S(2) + yz * lo_z;
        const S hi_y = (-Ly / S(2) + Ly) + yz * hi_z;
        return location >= lo_y && location < hi_y;
        }
    // BoxDim::getNearestPlaneDistance (SphericalBarrierEvaluator.h:54-60)
    const S term = xy * yz - xz;
    const S dx = Lx / std::sqrt(S(1) + xy * xy + term * term);
    const S dy = Ly / std::sqrt(S(1) + yz * yz);
    const S dz = Lz;
    const S two_R = S(2) * location;
    return location >= S(0) && dx >= two_R && dy >= two_R && dz >= two_R;
    }

template<class S> static int launch_barrier(int geometry, const azb_barrier_args& a, cudaStream_t stream)
    {
    if (!barrier_valid<S>(geometry, a.global_box, (S)a.location))
        return AZB_ERR_BARRIER_INVALID;
    if (a.N == 0)
        return 0;
    if (!a.d_force || !a.d_pos || !a.d_params || a.ntypes == 0 || (a.d_virial && a.virial_pitch < a.N))
        return AZB_ERR_BAD_ARGUMENT;
    BarrierArgsDev<S> d;
    d.force = (typename vec4<S>::type*)a.d_force;
    d.virial = (S*)a.d_virial;
    d.virial_pitch = a.virial_pitch;
    d.N = a.N;
    d.ntypes = a.ntypes;
    d.pos = (const typename vec4<S>::type*)a.d_pos;
    d.params = (const S*)a.d_params;
    d.box = make_box<S>(a.global_box);
    d.location = (S)a.location;
    const size_t smem = 2 * sizeof(S) * a.ntypes;
    if (smem > 48 * 1024)
        return AZB_ERR_BAD_ARGUMENT;
    unsigned int block = a.block_size ? a.block_size : 256u;
    if (block > 256u || (block & 31u))
        return AZB_ERR_BAD_ARGUMENT;
    // a few CTAs per SM in flight cover the HBM latency; more rows per thread beyond that only lengthen the tail
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned int need = (a.N + block - 1) / block;
    const unsigned int cap = (unsigned int)sms * 8u * 4u;
    const unsigned int grid = need < cap ? need : cap;
    if (geometry == AZB_BARRIER_PLANAR)
        barrier_kernel<S, AZB_BARRIER_PLANAR><<<grid, block, smem, stream>>>(d);
    else
        barrier_kernel<S, AZB_BARRIER_SPHERICAL><<<grid, block, smem, stream>>>(d);
    return (int)cudaGetLastError();
    }
    } // namespace azb

extern "C" int azb_barrier_forces(int geometry, int scalar_bits, const azb_barrier_args* args, void* stream)
    {
    if (!args)
        return AZB_ERR_BAD_ARGUMENT;
    if (geometry < 0 || geometry >= AZB_BARRIER_COUNT_)
        return AZB_ERR_BAD_POTENTIAL;
    if (scalar_bits == 32)
        return azb::launch_barrier<float>(geometry, *args, (cudaStream_t)stream);
    if (scalar_bits == 64)
        return azb::launch_barrier<double>(geometry, *args, (cudaStream_t)stream);
    return AZB_ERR_BAD_SCALAR;
    }
