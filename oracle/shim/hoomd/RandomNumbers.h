// TEST INFRASTRUCTURE (oracle). Stand-in for hoomd/RandomNumbers.h (HOOMD-blue v5.0.x, un-vendored; it
// wraps Random123's Philox4x32-10) as used by /root/reference/src/DPDPairEvaluatorGeneralWeight.h:227-233:
//   hoomd::RandomGenerator rng(hoomd::Seed(id, timestep, seed), hoomd::Counter(oi, oj));
//   Scalar alpha = hoomd::UniformDistribution<Scalar>(-1, 1)(rng);
//
// PARITY UNPINNED for the packing: the Philox4x32-10 core below is pinned by Random123's published
// known-answer vectors (tests/test_oracle_philox.py). The Seed/Counter word packing and the u01 mapping are
// restated from recollection of HOOMD v5 / Random123 uniform.hpp; no reference test pins them bit-wise
// (pytest/test_pair_dpd.py:46 is statistical only). Everything lives in this one header so that it can be
// corrected in one place if HOOMD's source becomes available. The CUDA side keeps an independent copy of
// the same spec (azplugins_b200/csrc/azb_rng.cuh); the two are cross-checked bit-for-bit.
#ifndef AZB_ORACLE_SHIM_RANDOMNUMBERS_H_
#define AZB_ORACLE_SHIM_RANDOMNUMBERS_H_
#include <cstdint>
#include "HOOMDMath.h"

namespace hoomd
    {
namespace detail
    {
struct philox_ctr
    {
    uint32_t v[4];
    };
struct philox_key
    {
    uint32_t v[2];
    };

// Philox4x32 with 10 rounds (Salmon et al., SC'11). Multipliers 0xD2511F53 / 0xCD9E8D57,
// Weyl key increments 0x9E3779B9 / 0xBB67AE85.
inline philox_ctr philox4x32_10(philox_ctr c, philox_key k)
    {
    for (int r = 0; r < 10; ++r)
        {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c.v[0];
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c.v[2];
        philox_ctr n;
        n.v[0] = (uint32_t)(p1 >> 32) ^ c.v[1] ^ k.v[0];
        n.v[1] = (uint32_t)p1;
        n.v[2] = (uint32_t)(p0 >> 32) ^ c.v[3] ^ k.v[1];
        n.v[3] = (uint32_t)p0;
        c = n;
        k.v[0] += 0x9E3779B9u;
        k.v[1] += 0xBB67AE85u;
        }
    return c;
    }
    } // namespace detail

//! Key: {id<<24 | seed<<8 | bits 32..39 of timestep, low 32 bits of timestep}
class Seed
    {
    public:
    Seed(uint8_t id, uint64_t timestep, uint16_t seed)
        {
        m_key.v[0] = ((uint32_t)id << 24) | ((uint32_t)seed << 8)
                     | (uint32_t)((timestep & 0x000000ff00000000ull) >> 32);
        m_key.v[1] = (uint32_t)(timestep & 0x00000000ffffffffull);
        }
    const detail::philox_key& getKey() const
        {
        return m_key;
        }

    private:
    detail::philox_key m_key;
    };

//! Counter: {d<<16 (low 16 bits reserved for the generator's draw count), c, b, a}
class Counter
    {
    public:
    Counter(uint32_t a = 0, uint32_t b = 0, uint32_t c = 0, uint16_t d = 0)
        {
        m_ctr.v[0] = (uint32_t)d << 16;
        m_ctr.v[1] = c;
        m_ctr.v[2] = b;
        m_ctr.v[3] = a;
        }
    const detail::philox_ctr& getCounter() const
        {
        return m_ctr;
        }

    private:
    detail::philox_ctr m_ctr;
    };

class RandomGenerator
    {
    public:
    RandomGenerator(const Seed& seed, const Counter& counter)
        : m_key(seed.getKey()), m_ctr(counter.getCounter())
        {
        }
    detail::philox_ctr operator()()
        {
        detail::philox_ctr u = detail::philox4x32_10(m_ctr, m_key);
        m_ctr.v[0] += 1;
        return u;
        }

    private:
    detail::philox_key m_key;
    detail::philox_ctr m_ctr;
    };

namespace detail
    {
inline uint32_t generate_u32(RandomGenerator& rng)
    {
    return rng().v[0];
    }
inline uint64_t generate_u64(RandomGenerator& rng)
    {
    philox_ctr u = rng();
    return ((uint64_t)u.v[0] << 32) | u.v[1];
    }
// Random123 uniform.hpp u01<F,I>: in * 2^-W + 2^-(W+1), a value in (0,1]
inline float generate_canonical(RandomGenerator& rng, float)
    {
    const float factor = 1.0f / 4294967296.0f;
    return (float)generate_u32(rng) * factor + 0.5f * factor;
    }
inline double generate_canonical(RandomGenerator& rng, double)
    {
    const double factor = 1.0 / 18446744073709551616.0;
    return (double)generate_u64(rng) * factor + 0.5 * factor;
    }
    } // namespace detail

template<class Real> class UniformDistribution
    {
    public:
    explicit UniformDistribution(Real a = Real(0), Real b = Real(1)) : m_a(a), m_width(b - a) { }
    Real operator()(RandomGenerator& rng)
        {
        return m_a + m_width * detail::generate_canonical(rng, Real());
        }

    private:
    Real m_a, m_width;
    };
    } // namespace hoomd
#endif
