// azb_math.cuh — scalar helpers for the sm_100a force kernels: vector loads of HOOMD Scalar4 arrays through the
// read-only path, type-id decode, minimum image, and the float / double math flavours (the float flavour uses
// the single-instruction MUFU approximations the way HOOMD's hoomd::fast:: does on the GPU; the double flavour is
// IEEE).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace azb
    {
template<class S> struct vec4;
template<> struct vec4<float>
    {
    typedef float4 type;
    };
struct alignas(32) double4a
    {
    double x, y, z, w;
    };
template<> struct vec4<double>
    {
    typedef double4a type;
    };

// ---- Scalar4 gathers ------------------------------------------------------------------------------
__device__ __forceinline__ float4 ld4(const float4* p)
    {
    return __ldg(p);
    }
__device__ __forceinline__ vec4<double>::type ld4(const vec4<double>::type* p)
    {
    // one 256-bit read-only load (sm_100+): ld.global.nc.v4.f64
    vec4<double>::type r;
    asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
    return r;
    }
__device__ __forceinline__ void st4(float4* p, float x, float y, float z, float w)
    {
    *p = make_float4(x, y, z, w);
    }
__device__ __forceinline__ void st4(vec4<double>::type* p, double x, double y, double z, double w)
    {
    asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(x), "d"(y), "d"(z), "d"(w) : "memory");
    }

// request a 128-byte line into L2 without a destination register (nothing waits on it)
__device__ __forceinline__ void prefetch_l2(const void* p)
    {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    }

// ---- explicit shared-memory accesses at 32-bit shared addresses -----------------------------------------------------
// (the address stays a plain register add instead of a generic-pointer reconstruction per access)
__device__ __forceinline__ unsigned int smem_addr(const void* p)
    {
    // volatile: computed once where written, never rematerialised inside a loop
    unsigned int r;
    asm volatile("{ .reg .u64 t; cvta.to.shared.u64 t, %1; cvt.u32.u64 %0, t; }" : "=r"(r) : "l"(p));
    return r;
    }
__device__ __forceinline__ float4 lds4_at(unsigned int addr, float)
    {
    float4 r;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
    return r;
    }
__device__ __forceinline__ vec4<double>::type lds4_at(unsigned int addr, double)
    {
    vec4<double>::type r;
    asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "r"(addr));
    asm volatile("ld.shared.v2.f64 {%0,%1}, [%2+16];" : "=d"(r.z), "=d"(r.w) : "r"(addr));
    return r;
    }
__device__ __forceinline__ void sts4_at(unsigned int addr, const float4& v)
    {
    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    }
__device__ __forceinline__ void sts4_at(unsigned int addr, const vec4<double>::type& v)
    {
    asm volatile("st.shared.v2.f64 [%0], {%1,%2};" ::"r"(addr), "d"(v.x), "d"(v.y) : "memory");
    asm volatile("st.shared.v2.f64 [%0+16], {%1,%2};" ::"r"(addr), "d"(v.z), "d"(v.w) : "memory");
    }
// predicated stores (no branch: lanes with p == false issue nothing to memory)
__device__ __forceinline__ void sts4_if(bool p, unsigned int addr, const float4& v)
    {
    asm volatile("{ .reg .pred q; setp.ne.u32 q, %5, 0; @q st.shared.v4.f32 [%0], {%1,%2,%3,%4}; }" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w), "r"((unsigned int)p)
                 : "memory");
    }
__device__ __forceinline__ void sts4_if(bool p, unsigned int addr, const vec4<double>::type& v)
    {
    asm volatile("{ .reg .pred q; setp.ne.u32 q, %5, 0; @q st.shared.v2.f64 [%0], {%1,%2}; @q st.shared.v2.f64 [%0+16], {%3,%4}; }" ::"r"(addr),
                 "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w), "r"((unsigned int)p)
                 : "memory");
    }
__device__ __forceinline__ void sts32_if(bool p, unsigned int addr, unsigned int v)
    {
    asm volatile("{ .reg .pred q; setp.ne.u32 q, %2, 0; @q st.shared.u32 [%0], %1; }" ::"r"(addr), "r"(v), "r"((unsigned int)p) : "memory");
    }
__device__ __forceinline__ unsigned int lds32_at(unsigned int addr)
    {
    unsigned int r;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(addr));
    return r;
    }
__device__ __forceinline__ void sts32_at(unsigned int addr, unsigned int v)
    {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
    }
__device__ __forceinline__ float lds_scalar(unsigned int addr, float)
    {
    return __uint_as_float(lds32_at(addr));
    }
__device__ __forceinline__ double lds_scalar(unsigned int addr, double)
    {
    double r;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(r) : "r"(addr));
    return r;
    }
// Scalar4 element `slot` of a shared-memory array at shared address `base`
__device__ __forceinline__ float4 lds4(unsigned int base, unsigned int slot, float)
    {
    return lds4_at(base + slot * 16u, float());
    }
__device__ __forceinline__ vec4<double>::type lds4(unsigned int base, unsigned int slot, double)
    {
    return lds4_at(base + slot * 32u, double());
    }

__device__ __forceinline__ unsigned int type_of(float w)
    {
    return (unsigned int)__float_as_int(w);
    }
__device__ __forceinline__ unsigned int type_of(double w)
    {
    return (unsigned int)__double_as_longlong(w);
    }

// ---- math flavours ----------------------------------------------------------------------------------
// Single-instruction MUFU approximations with flush-to-zero (no denormal fix-up code around them): the float
// flavour of hoomd::fast::. Relative error <= 2 ulp each.
__device__ __forceinline__ float rcp(float x)
    {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
    }
__device__ __forceinline__ double rcp(double x)
    {
    return 1.0 / x;
    }
__device__ __forceinline__ float fdiv(float a, float b)
    {
    return __fdividef(a, b);
    }
__device__ __forceinline__ double fdiv(double a, double b)
    {
    return a / b;
    }
__device__ __forceinline__ float rsqrt_(float x)
    {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
    }
__device__ __forceinline__ double rsqrt_(double x)
    {
    return rsqrt(x);
    }
__device__ __forceinline__ float sqrt_(float x)
    {
    return __fsqrt_rn(x);
    }
__device__ __forceinline__ double sqrt_(double x)
    {
    return sqrt(x);
    }
__device__ __forceinline__ float exp_(float x)
    {
    // ex2.approx on a two-term product x*log2(e): __expf alone loses |x|*1e-7 relative to the rounding of x*log2(e)
    const float t = x * 1.4426950216293335f;
    const float r = __fmaf_rn(x, 1.4426950216293335f, -t) + x * 1.925963033500011e-8f;
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t));
    return e * __fmaf_rn(r, 0.6931471824645996f, 1.0f);
    }
__device__ __forceinline__ double exp_(double x)
    {
    return exp(x);
    }
__device__ __forceinline__ float log_(float x)
    {
    return logf(x);
    }
__device__ __forceinline__ double log_(double x)
    {
    return log(x);
    }
__device__ __forceinline__ float pow_(float x, float y)
    {
    return __powf(x, y);
    }
__device__ __forceinline__ double pow_(double x, double y)
    {
    return pow(x, y);
    }
__device__ __forceinline__ float rint_(float x)
    {
    return rintf(x);
    }
__device__ __forceinline__ double rint_(double x)
    {
    return rint(x);
    }
// Round-to-nearest-even without the XU pipe (FRND / F2I run at quarter rate): adding and subtracting 1.5 * 2^(p-1)
// leaves rint(x) for |x| < 2^(p-2). Image counts are tiny integers, far inside that range.
__device__ __forceinline__ float rint_small(float x)
    {
    return __fadd_rn(__fadd_rn(x, 12582912.0f), -12582912.0f);
    }
__device__ __forceinline__ double rint_small(double x)
    {
    return __dadd_rn(__dadd_rn(x, 6755399441055744.0), -6755399441055744.0);
    }

// ---- box ------------------------------------------------------------------------------------------------
template<class S> struct Box
    {
    S Lx, Ly, Lz, Lxinv, Lyinv, Lzinv, xy, xz, yz;
    unsigned int flags; // bit0..2 periodic x,y,z; bit3 = any tilt
    };

// Fully periodic orthorhombic box (flags == 7): straight-line code, no flag tests, no XU instructions.
template<class S> __device__ __forceinline__ void min_image_ortho(const Box<S>& b, S& x, S& y, S& z)
    {
    x -= b.Lx * rint_small(x * b.Lxinv);
    y -= b.Ly * rint_small(y * b.Lyinv);
    z -= b.Lz * rint_small(z * b.Lzinv);
    }

template<bool FASTBOX, class S> __device__ __forceinline__ void min_image_t(const Box<S>& b, S& x, S& y, S& z);

// BoxDim::minImage, GPU flavour: rint(w * Linv), z then y then x with tilt terms
template<class S> __device__ __forceinline__ void min_image(const Box<S>& b, S& x, S& y, S& z)
    {
    if (b.flags & 8u)
        {
        if (b.flags & 4u)
            {
            const S img = rint_(z * b.Lzinv);
            z -= b.Lz * img;
            y -= b.Lz * b.yz * img;
            x -= b.Lz * b.xz * img;
            }
        if (b.flags & 2u)
            {
            const S img = rint_(y * b.Lyinv);
            y -= b.Ly * img;
            x -= b.Ly * b.xy * img;
            }
        if (b.flags & 1u)
            x -= b.Lx * rint_(x * b.Lxinv);
        }
    else
        {
        if (b.flags & 4u)
            z -= b.Lz * rint_small(z * b.Lzinv);
        if (b.flags & 2u)
            y -= b.Ly * rint_small(y * b.Lyinv);
        if (b.flags & 1u)
            x -= b.Lx * rint_small(x * b.Lxinv);
        }
    }
template<bool FASTBOX, class S> __device__ __forceinline__ void min_image_t(const Box<S>& b, S& x, S& y, S& z)
    {
    if (FASTBOX)
        min_image_ortho(b, x, y, z);
    else
        min_image(b, x, y, z);
    }
    } // namespace azb
