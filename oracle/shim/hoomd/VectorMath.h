// TEST INFRASTRUCTURE (oracle). Minimal stand-in for hoomd/VectorMath.h as used by
// /root/reference/src/AnisoPairEvaluatorTwoPatchMorse.h:141-212: vec3, quat, rotate, cross, dot,
// unary minus, scalar*vec, vec+vec, vec_to_scalar3. Quaternion convention: Scalar4 (x,y,z,w) holds
// (s, v.x, v.y, v.z) — quat(Scalar4 a): s = a.x, v = (a.y, a.z, a.w).
#ifndef AZB_ORACLE_SHIM_VECTORMATH_H_
#define AZB_ORACLE_SHIM_VECTORMATH_H_
#include "HOOMDMath.h"

namespace hoomd
    {
template<class Real> struct vec3
    {
    Real x, y, z;
    vec3() : x(0), y(0), z(0) { }
    vec3(Real _x, Real _y, Real _z) : x(_x), y(_y), z(_z) { }
    explicit vec3(const Scalar3& a) : x(a.x), y(a.y), z(a.z) { }
    };
template<class Real> inline vec3<Real> operator+(const vec3<Real>& a, const vec3<Real>& b)
    {
    return vec3<Real>(a.x + b.x, a.y + b.y, a.z + b.z);
    }
template<class Real> inline vec3<Real> operator-(const vec3<Real>& a, const vec3<Real>& b)
    {
    return vec3<Real>(a.x - b.x, a.y - b.y, a.z - b.z);
    }
template<class Real> inline vec3<Real> operator-(const vec3<Real>& a)
    {
    return vec3<Real>(-a.x, -a.y, -a.z);
    }
template<class Real> inline vec3<Real> operator*(const vec3<Real>& a, const Real& b)
    {
    return vec3<Real>(a.x * b, a.y * b, a.z * b);
    }
template<class Real> inline vec3<Real> operator*(const Real& b, const vec3<Real>& a)
    {
    return vec3<Real>(a.x * b, a.y * b, a.z * b);
    }
template<class Real> inline Real dot(const vec3<Real>& a, const vec3<Real>& b)
    {
    return a.x * b.x + a.y * b.y + a.z * b.z;
    }
template<class Real> inline vec3<Real> cross(const vec3<Real>& a, const vec3<Real>& b)
    {
    return vec3<Real>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
    }
template<class Real> inline Scalar3 vec_to_scalar3(const vec3<Real>& a)
    {
    return make_scalar3(a.x, a.y, a.z);
    }

template<class Real> struct quat
    {
    Real s;
    vec3<Real> v;
    quat() : s(1), v() { }
    quat(Real _s, const vec3<Real>& _v) : s(_s), v(_v) { }
    explicit quat(const Scalar4& a) : s(a.x), v(a.y, a.z, a.w) { }
    };

// rotate(q, a) = (s^2 - v.v) a + 2 s (v x a) + 2 (v.a) v   (HOOMD VectorMath.h)
template<class Real> inline vec3<Real> rotate(const quat<Real>& a, const vec3<Real>& b)
    {
    return (a.s * a.s - dot(a.v, a.v)) * b + Real(2) * a.s * cross(a.v, b)
           + Real(2) * dot(a.v, b) * a.v;
    }
    } // namespace hoomd
#endif
