// azb_integrate.cu — the two half steps of velocity Verlet around the force path and the neighbour-list rebuild trigger
// (SURVEY.md §8 f2). Pure streaming kernels: one thread per particle, 128-bit loads and stores, the net-force sum over the
// attached force objects folded into the second half step instead of a separate pass over N.
#include <cuda_runtime.h>

#include "azb200.h"
#include "azb_launch.cuh"
#include "azb_math.cuh"

namespace azb
    {
constexpr unsigned int NVE_MAX_FORCES = 8;

template<class S> struct NveDev
    {
    typedef typename vec4<S>::type S4;
    S4* pos;
    S4* vel;
    S4* accel;
    int* image;
    unsigned int N;
    Box<S> box;
    S dt;
    const S4* forces[NVE_MAX_FORCES];
    unsigned int n_forces;
    S4* net_force;
    };

// BoxDim::wrap with image counting (one image per periodic direction; z, y, x with the tilt terms)
template<class S> __device__ __forceinline__ void wrap_count(const Box<S>& b, S& x, S& y, S& z, int& ix, int& iy, int& iz)
    {
    const S hx = b.Lx / S(2), hy = b.Ly / S(2), hz = b.Lz / S(2);
    if (b.flags & 4u)
        {
        if (z >= hz)
            {
            z -= b.Lz;
            y -= b.Lz * b.yz;
            x -= b.Lz * b.xz;
            ++iz;
            }
        else if (z < -hz)
            {
            z += b.Lz;
            y += b.Lz * b.yz;
            x += b.Lz * b.xz;
            --iz;
            }
        }
    if (b.flags & 2u)
        {
        const S tilt_y = b.yz * z;
        if (y >= hy + tilt_y)
            {
            y -= b.Ly;
            x -= b.Ly * b.xy;
            ++iy;
            }
        else if (y < -hy + tilt_y)
            {
            y += b.Ly;
            x += b.Ly * b.xy;
            --iy;
            }
        }
    if (b.flags & 1u)
        {
        const S tilt_x = (b.xz - b.xy * b.yz) * z + b.xy * y;
        if (x >= hx + tilt_x)
            {
            x -= b.Lx;
            ++ix;
            }
        else if (x < -hx + tilt_x)
            {
            x += b.Lx;
            --ix;
            }
        }
    }

template<class S> __global__ void __launch_bounds__(256) nve_step_one_kernel(const NveDev<S> a)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.N)
        return;
    auto p = ld4(a.pos + i);
    auto v = ld4(a.vel + i);
    const auto acc = ld4(a.accel + i);
    const S half_dt = S(0.5) * a.dt;
    v.x += half_dt * acc.x;
    v.y += half_dt * acc.y;
    v.z += half_dt * acc.z;
    S x = p.x + a.dt * v.x, y = p.y + a.dt * v.y, z = p.z + a.dt * v.z;
    int ix = 0, iy = 0, iz = 0;
    wrap_count(a.box, x, y, z, ix, iy, iz);
    st4(a.pos + i, x, y, z, p.w);
    st4(a.vel + i, v.x, v.y, v.z, v.w);
    if (a.image && (ix | iy | iz))
        {
        a.image[3u * i] += ix;
        a.image[3u * i + 1u] += iy;
        a.image[3u * i + 2u] += iz;
        }
    }

template<class S> __global__ void __launch_bounds__(256) nve_step_two_kernel(const NveDev<S> a)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.N)
        return;
    S fx = S(0), fy = S(0), fz = S(0), e = S(0);
#pragma unroll
    for (unsigned int k = 0; k < NVE_MAX_FORCES; ++k)
        if (k < a.n_forces)
            {
            const auto f = ld4(a.forces[k] + i);
            fx += f.x;
            fy += f.y;
            fz += f.z;
            e += f.w;
            }
    auto v = ld4(a.vel + i);
    const S minv = S(1) / v.w;
    const S ax = fx * minv, ay = fy * minv, az = fz * minv;
    const S half_dt = S(0.5) * a.dt;
    st4(a.vel + i, v.x + half_dt * ax, v.y + half_dt * ay, v.z + half_dt * az, v.w);
    st4(a.accel + i, ax, ay, az, S(0));
    if (a.net_force)
        st4(a.net_force + i, fx, fy, fz, e);
    }

template<class S>
__global__ void __launch_bounds__(256) displacement_kernel(const typename vec4<S>::type* pos, const typename vec4<S>::type* last, unsigned int N,
                                                           const Box<S> box, S max_sq, unsigned int* flag)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool moved = false;
    if (i < N)
        {
        const auto p = ld4(pos + i);
        const auto q = ld4(last + i);
        S dx = p.x - q.x, dy = p.y - q.y, dz = p.z - q.z;
        min_image(box, dx, dy, dz);
        moved = dx * dx + dy * dy + dz * dz > max_sq;
        }
    if (__any_sync(0xffffffffu, moved) && (threadIdx.x & 31u) == 0u)
        atomicOr(flag, 1u);
    }

template<class S> static int fill_nve(const azb_nve_args& a, NveDev<S>& d, bool step_two)
    {
    typedef typename vec4<S>::type S4;
    if (!a.d_pos || !a.d_vel || !a.d_accel)
        return AZB_ERR_BAD_ARGUMENT;
    d.pos = (S4*)a.d_pos;
    d.vel = (S4*)a.d_vel;
    d.accel = (S4*)a.d_accel;
    d.image = a.d_image;
    d.N = a.N;
    d.box = make_box<S>(a.box);
    d.dt = (S)a.deltaT;
    d.n_forces = 0;
    d.net_force = (S4*)a.d_net_force;
    for (unsigned int k = 0; k < NVE_MAX_FORCES; ++k)
        d.forces[k] = nullptr;
    if (step_two)
        {
        if (a.n_forces > NVE_MAX_FORCES || (a.n_forces && !a.d_forces))
            return AZB_ERR_BAD_ARGUMENT;
        for (unsigned int k = 0; k < a.n_forces; ++k)
            {
            if (!a.d_forces[k])
                return AZB_ERR_BAD_ARGUMENT;
            d.forces[k] = (const S4*)a.d_forces[k];
            }
        d.n_forces = a.n_forces;
        }
    return 0;
    }

template<class S> static int launch_nve(const azb_nve_args& a, bool step_two, cudaStream_t stream)
    {
    if (a.N == 0)
        return 0;
    NveDev<S> d;
    const int bad = fill_nve<S>(a, d, step_two);
    if (bad)
        return bad;
    const unsigned int block = 256, grid = (a.N + block - 1) / block;
    if (step_two)
        nve_step_two_kernel<S><<<grid, block, 0, stream>>>(d);
    else
        nve_step_one_kernel<S><<<grid, block, 0, stream>>>(d);
    return (int)cudaGetLastError();
    }
    } // namespace azb

extern "C" int azb_nve_step_one(int scalar_bits, const azb_nve_args* args, void* stream)
    {
    if (!args)
        return AZB_ERR_BAD_ARGUMENT;
    if (scalar_bits == 32)
        return azb::launch_nve<float>(*args, false, (cudaStream_t)stream);
    if (scalar_bits == 64)
        return azb::launch_nve<double>(*args, false, (cudaStream_t)stream);
    return AZB_ERR_BAD_SCALAR;
    }

extern "C" int azb_nve_step_two(int scalar_bits, const azb_nve_args* args, void* stream)
    {
    if (!args)
        return AZB_ERR_BAD_ARGUMENT;
    if (scalar_bits == 32)
        return azb::launch_nve<float>(*args, true, (cudaStream_t)stream);
    if (scalar_bits == 64)
        return azb::launch_nve<double>(*args, true, (cudaStream_t)stream);
    return AZB_ERR_BAD_SCALAR;
    }

extern "C" int azb_displacement_check(int scalar_bits, const void* d_pos, const void* d_last_pos, unsigned int N, const azb_box* box,
                                      double max_disp_sq, unsigned int* d_flag, void* stream)
    {
    if (N == 0)
        return 0;
    if (!d_pos || !d_last_pos || !box || !d_flag)
        return AZB_ERR_BAD_ARGUMENT;
    const unsigned int block = 256, grid = (N + block - 1) / block;
    cudaStream_t s = (cudaStream_t)stream;
    if (scalar_bits == 32)
        azb::displacement_kernel<float><<<grid, block, 0, s>>>((const float4*)d_pos, (const float4*)d_last_pos, N, azb::make_box<float>(*box),
                                                                (float)max_disp_sq, d_flag);
    else if (scalar_bits == 64)
        azb::displacement_kernel<double><<<grid, block, 0, s>>>((const azb::double4a*)d_pos, (const azb::double4a*)d_last_pos, N,
                                                                 azb::make_box<double>(*box), max_disp_sq, d_flag);
    else
        return AZB_ERR_BAD_SCALAR;
    return (int)cudaGetLastError();
    }
