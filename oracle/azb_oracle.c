/* TEST INFRASTRUCTURE — CPU oracle ("port") for the azplugins force path. NOT part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load the library built from
 * this file (oracle/_build/libazoracle.so); azplugins_b200/ never does.
 *
 * What it restates (plain C, float and double):
 *   - the reference's evaluators  /root/reference/src/{PairEvaluator*,DPDPairEvaluatorGeneralWeight,
 *     AnisoPairEvaluatorTwoPatchMorse,BondEvaluator*}.h — per-function citations in azb_oracle_impl.h;
 *   - the HOOMD-blue v5.0.x host loops that call them and hoomd/RandomNumbers.h (Philox4x32-10 from Random123).
 *     HOOMD is a third-party dependency that is absent from /root/reference (find_package(HOOMD 5.0.0),
 *     /root/reference/CMakeLists.txt:8; CI pins v5.0.1); the restatement follows SURVEY.md Appendix A.
 *   - a cell-list neighbour-list builder that emits HOOMD-layout n_neigh / nlist / head_list (full mode).
 *
 * Pinning: tests/test_oracle_golden.py replays every known-answer row of the reference's own tests
 * (pytest/test_pair.py, test_pair_aniso.py, test_bond.py — SURVEY.md Appendix C) through this oracle, and
 * tests/test_oracle_vs_ref.py compares it with oracle/_ref (the reference headers compiled unmodified) on random
 * inputs. PARITY UNPINNED for: the DPD random stream's Seed/Counter packing and u01 mapping (no reference test
 * pins them; the Philox core is pinned by Random123's published vectors), xplor smoothing and virials (no
 * reference test reads them).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "azb200.h"

/* Philox4x32-10 (Salmon, Moraes, Dror, Shaw, SC'11; Random123 v1.x as vendored by HOOMD under hoomd/extern) */
static void azo_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
    {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r)
        {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0;
        c1 = n1;
        c2 = n2;
        c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
        }
    out[0] = c0;
    out[1] = c1;
    out[2] = c2;
    out[3] = c3;
    }

#define AZO_CAT2(a, b) a##b
#define AZO_CAT(a, b) AZO_CAT2(a, b)

#define REAL float
#define SUF(x) AZO_CAT(x, _f32)
#define SQRT(x) sqrtf(x)
#define EXP(x) expf(x)
#define LOG(x) logf(x)
#define POW(x, y) powf(x, y)
#include "azb_oracle_impl.h"
#include "azb_oracle_nlist.h"
#undef REAL
#undef SUF
#undef SQRT
#undef EXP
#undef LOG
#undef POW

#define REAL double
#define SUF(x) AZO_CAT(x, _f64)
#define SQRT(x) sqrt(x)
#define EXP(x) exp(x)
#define LOG(x) log(x)
#define POW(x, y) pow(x, y)
#include "azb_oracle_impl.h"
#include "azb_oracle_nlist.h"
#undef REAL
#undef SUF

/* ---- exported C entry points (host pointers in the azb200.h structs) ---- */
void azo_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
    {
    azo_philox4x32_10(ctr, key, out);
    }

double azo_dpd_alpha(int scalar_bits, unsigned int tag_i, unsigned int tag_j, unsigned int timestep, unsigned int seed)
    {
    if (scalar_bits == 32)
        return (double)dpd_alpha_f32(tag_i, tag_j, timestep, (uint16_t)seed);
    return dpd_alpha_f64(tag_i, tag_j, timestep, (uint16_t)seed);
    }

int azo_pair_forces(int potential, int scalar_bits, const azb_pair_args* a, const void* params, int half_list, int nthreads)
    {
    if (scalar_bits == 32)
        return pairlike_loop_f32(AZO_KIND_PAIR, potential, a, NULL, NULL, params, half_list, nthreads);
    if (scalar_bits == 64)
        return pairlike_loop_f64(AZO_KIND_PAIR, potential, a, NULL, NULL, params, half_list, nthreads);
    return AZB_ERR_BAD_SCALAR;
    }

int azo_dpd_forces(int potential, int scalar_bits, const azb_dpd_args* a, const void* params, int half_list, int nthreads)
    {
    if (scalar_bits == 32)
        return pairlike_loop_f32(AZO_KIND_DPD, potential, &a->pair, a, NULL, params, half_list, nthreads);
    if (scalar_bits == 64)
        return pairlike_loop_f64(AZO_KIND_DPD, potential, &a->pair, a, NULL, params, half_list, nthreads);
    return AZB_ERR_BAD_SCALAR;
    }

int azo_aniso_forces(int potential, int scalar_bits, const azb_aniso_args* a, const void* params, int half_list, int nthreads)
    {
    if (scalar_bits == 32)
        return pairlike_loop_f32(AZO_KIND_ANISO, potential, &a->pair, NULL, a, params, half_list, nthreads);
    if (scalar_bits == 64)
        return pairlike_loop_f64(AZO_KIND_ANISO, potential, &a->pair, NULL, a, params, half_list, nthreads);
    return AZB_ERR_BAD_SCALAR;
    }

int azo_bond_forces(int potential, int scalar_bits, const azb_bond_args* a, const void* params, unsigned int* flags, int nthreads)
    {
    if (scalar_bits == 32)
        return bond_loop_f32(potential, a, params, flags, nthreads);
    if (scalar_bits == 64)
        return bond_loop_f64(potential, a, params, flags, nthreads);
    return AZB_ERR_BAD_SCALAR;
    }

int azo_barrier_forces(int geometry, int scalar_bits, const azb_barrier_args* a)
    {
    if (scalar_bits == 32)
        return barrier_loop_f32(geometry, a);
    if (scalar_bits == 64)
        return barrier_loop_f64(geometry, a);
    return AZB_ERR_BAD_SCALAR;
    }

/* neighbour list: see azb_oracle_nlist.h. *out_nlist is malloc'ed; release with azo_free. */
int azo_nlist_build(int scalar_bits, const void* pos, unsigned int N, unsigned int n_max, const azb_box* box,
                    const double* r_list, unsigned int ntypes, const size_t* excl_head, const unsigned int* excl_list,
                    unsigned int row_align, unsigned int* n_neigh, size_t* head_list, unsigned int** out_nlist,
                    size_t* out_size)
    {
    if (scalar_bits == 32)
        return nlist_build_f32(pos, N, n_max, box, r_list, ntypes, excl_head, excl_list, row_align, n_neigh, head_list,
                               out_nlist, out_size);
    if (scalar_bits == 64)
        return nlist_build_f64(pos, N, n_max, box, r_list, ntypes, excl_head, excl_list, row_align, n_neigh, head_list,
                               out_nlist, out_size);
    return AZB_ERR_BAD_SCALAR;
    }

void azo_free(void* p)
    {
    free(p);
    }
