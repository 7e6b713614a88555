// TEST INFRASTRUCTURE (oracle). Minimal stand-in for HOOMD-blue's hoomd/HOOMDMath.h so that the
// reference's evaluator headers (/root/reference/src/*Evaluator*.h, included at PairEvaluator.h:13)
// compile unmodified on the host. HOOMD-blue v5 itself is not vendored under /root/reference and is not
// installed here; this file restates only what those headers use: Scalar, Scalar3/4, make_scalar3/4,
// fast::{sqrt,rsqrt,exp,pow}, Scalar * Scalar3. Written from the documented HOOMD v5 semantics, not copied.
#ifndef AZB_ORACLE_SHIM_HOOMDMATH_H_
#define AZB_ORACLE_SHIM_HOOMDMATH_H_

#include <cmath>
#include <cstdint>

#ifndef HOOMD_LONGREAL_SIZE
#define HOOMD_LONGREAL_SIZE 64
#endif

namespace hoomd
    {
#if HOOMD_LONGREAL_SIZE == 32
typedef float Scalar;
#else
typedef double Scalar;
#endif

struct Scalar3
    {
    Scalar x, y, z;
    };
struct Scalar4
    {
    Scalar x, y, z, w;
    };

inline Scalar3 make_scalar3(Scalar x, Scalar y, Scalar z)
    {
    Scalar3 v = {x, y, z};
    return v;
    }
inline Scalar4 make_scalar4(Scalar x, Scalar y, Scalar z, Scalar w)
    {
    Scalar4 v = {x, y, z, w};
    return v;
    }
inline Scalar dot(const Scalar3& a, const Scalar3& b)
    {
    return a.x * b.x + a.y * b.y + a.z * b.z;
    }

// the float3/double3 operators of HOOMDMath.h that the barrier evaluators use (SphericalBarrierEvaluator.h:47)
inline Scalar3 operator*(const Scalar& a, const Scalar3& b)
    {
    return make_scalar3(a * b.x, a * b.y, a * b.z);
    }
inline Scalar3 operator*(const Scalar3& a, const Scalar& b)
    {
    return make_scalar3(a.x * b, a.y * b, a.z * b);
    }

// host flavour of hoomd::fast (no device intrinsics on the CPU path)
namespace fast
    {
inline float rsqrt(float x)
    {
    return 1.0f / ::sqrtf(x);
    }
inline double rsqrt(double x)
    {
    return 1.0 / ::sqrt(x);
    }
inline float sqrt(float x)
    {
    return ::sqrtf(x);
    }
inline double sqrt(double x)
    {
    return ::sqrt(x);
    }
inline float exp(float x)
    {
    return ::expf(x);
    }
inline double exp(double x)
    {
    return ::exp(x);
    }
inline float pow(float x, float y)
    {
    return ::powf(x, y);
    }
inline double pow(double x, double y)
    {
    return ::pow(x, y);
    }
    } // namespace fast
    } // namespace hoomd

// The reference headers use these unqualified inside namespace hoomd::azplugins::detail.
using std::log;
using std::pow;
using std::sqrt;

#endif
