// azb_rng.cuh — counter-based random numbers for the DPD thermostat, device side.
//
// Spec being implemented (HOOMD-blue v5 hoomd/RandomNumbers.h over Random123's Philox4x32-10; HOOMD is not
// vendored in /root/reference — the call site is /root/reference/src/DPDPairEvaluatorGeneralWeight.h:227-233):
//   key     = { id<<24 | seed<<8 | (timestep>>32 & 0xff),  timestep & 0xffffffff }     hoomd::Seed(id, timestep, seed)
//   counter = { d<<16, c, b, a }                                                        hoomd::Counter(a, b, c=0, d=0)
//   draw    = philox4x32_10(counter, key)                                               first RandomGenerator call
//   u01     = x * 2^-W + 2^-(W+1)   (W = 32 using draw[0] for float; W = 64 using draw[0]<<32|draw[1] for double)
//   UniformDistribution<Scalar>(a, b) = a + (b - a) * u01
// For azplugins' DPD: id = 200 (RNGIdentifiers.h:23), timestep is truncated to 32 bits by the evaluator
// (DPDPairEvaluatorGeneralWeight.h:122,273) so key[0]'s low byte is always 0, a = min(tag_i, tag_j), b = max(...).
//
// PARITY UNPINNED: the Seed/Counter packing and the u01 mapping are restated from recollection of HOOMD/Random123;
// the reference's tests pin only the statistics (pytest/test_pair_dpd.py:46). The Philox core is pinned by
// Random123's known-answer vectors (tests/test_oracle_golden.py::test_philox_known_answers). Everything sits in this header (and, independently
// restated, in oracle/shim/hoomd/RandomNumbers.h) so a correction is a one-place change.
#pragma once
#include <stdint.h>

namespace azb
    {
#define AZB_PHILOX_M0 0xD2511F53u
#define AZB_PHILOX_M1 0xCD9E8D57u
#define AZB_PHILOX_W0 0x9E3779B9u
#define AZB_PHILOX_W1 0xBB67AE85u

// The ten round keys depend only on (seed, timestep): computed once on the host, passed by value to the kernel
// (they live in uniform registers / constant bank, not per-thread registers).
struct PhiloxRoundKeys
    {
    uint32_t k0[10];
    uint32_t k1[10];
    };

inline PhiloxRoundKeys make_round_keys(uint8_t id, uint64_t timestep, uint16_t seed)
    {
    PhiloxRoundKeys rk;
    uint32_t a = ((uint32_t)id << 24) | ((uint32_t)seed << 8) | (uint32_t)((timestep & 0x000000ff00000000ull) >> 32);
    uint32_t b = (uint32_t)(timestep & 0xffffffffull);
    for (int r = 0; r < 10; ++r)
        {
        rk.k0[r] = a;
        rk.k1[r] = b;
        a += AZB_PHILOX_W0;
        b += AZB_PHILOX_W1;
        }
    return rk;
    }

#ifdef __CUDACC__
// Philox4x32-10 on counter {c0,c1,c2,c3}; returns words 0 and 1 of the output block.
__device__ __forceinline__ void philox4x32_10_w01(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, const PhiloxRoundKeys& rk,
                                                   uint32_t& o0, uint32_t& o1)
    {
#pragma unroll
    for (int r = 0; r < 10; ++r)
        {
        const uint32_t hi0 = __umulhi(AZB_PHILOX_M0, c0);
        const uint32_t lo0 = AZB_PHILOX_M0 * c0;
        const uint32_t hi1 = __umulhi(AZB_PHILOX_M1, c2);
        const uint32_t lo1 = AZB_PHILOX_M1 * c2;
        c0 = hi1 ^ c1 ^ rk.k0[r];
        c1 = lo1;
        c2 = hi0 ^ c3 ^ rk.k1[r];
        c3 = lo0;
        }
    o0 = c0;
    o1 = c1;
    }

// hoomd::UniformDistribution<Scalar>(-1, 1) on the first draw of RandomGenerator(Seed, Counter(oi, oj))
template<class S> __device__ __forceinline__ S dpd_alpha(uint32_t tag_i, uint32_t tag_j, const PhiloxRoundKeys& rk);

template<> __device__ __forceinline__ float dpd_alpha<float>(uint32_t tag_i, uint32_t tag_j, const PhiloxRoundKeys& rk)
    {
    const uint32_t oi = min(tag_i, tag_j), oj = max(tag_i, tag_j);
    uint32_t o0, o1;
    philox4x32_10_w01(0u, 0u, oj, oi, rk, o0, o1);
    const float u = __fmaf_rn(__uint2float_rn(o0), 2.3283064365386963e-10f, 1.1641532182693481e-10f);
    return __fmaf_rn(2.0f, u, -1.0f);
    }
template<> __device__ __forceinline__ double dpd_alpha<double>(uint32_t tag_i, uint32_t tag_j, const PhiloxRoundKeys& rk)
    {
    const uint32_t oi = min(tag_i, tag_j), oj = max(tag_i, tag_j);
    uint32_t o0, o1;
    philox4x32_10_w01(0u, 0u, oj, oi, rk, o0, o1);
    const unsigned long long x = ((unsigned long long)o0 << 32) | o1;
    const double u = __fma_rn(__ull2double_rn(x), 5.421010862427522e-20, 2.710505431213761e-20);
    return __fma_rn(2.0, u, -1.0);
    }
#endif
    } // namespace azb
