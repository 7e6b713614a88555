// TEST INFRASTRUCTURE (oracle/_ref). Not part of the product; only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load the library built from this file.
//
// Drives the reference's evaluator headers — included UNMODIFIED from /root/reference/src at build time
// (never copied into this repo) — with a CPU restatement of the HOOMD-blue v5 host loops that call them
// (PotentialPair / PotentialPairDPDThermo / AnisoPotentialPair / PotentialBond ::computeForces; HOOMD is
// not vendored in /root/reference, the loops follow SURVEY.md Appendix A.1-A.4 and the call contract that
// is visible in the evaluator headers themselves: PairEvaluator.h:72,98;
// DPDPairEvaluatorGeneralWeight.h:121-155,198-255; AnisoPairEvaluator.h:103-109,166-173;
// BondEvaluator.h:62,84).
//
// Compiled twice (HOOMD_LONGREAL_SIZE = 32 and 64) with AZREF_PREFIX = azref32 / azref64 and linked into
// oracle/_ref/libazref32.so and libazref64.so (separately: the class names are identical in both builds). Arrays use the exact layouts of include/azb200.h but with HOST pointers.
#include <pybind11/pybind11.h>

#include <cstring>
#include <stdexcept>
#include <vector>

#include "hoomd/BoxDim.h"
#include "hoomd/HOOMDMath.h"
#include "hoomd/RandomNumbers.h"
#include "hoomd/VectorMath.h"

#include "AnisoPairEvaluatorTwoPatchMorse.h"
#include "BondEvaluatorDoubleWell.h"
#include "BondEvaluatorQuartic.h"
#include "DPDPairEvaluatorGeneralWeight.h"
#include "PairEvaluatorColloid.h"
#include "PairEvaluatorExpandedYukawa.h"
#include "PairEvaluatorHertz.h"
#include "PairEvaluatorPerturbedLennardJones.h"
#include "PlanarBarrierEvaluator.h"
#include "SphericalBarrierEvaluator.h"

#include "azb200.h"

#define AZREF_CAT2(a, b) a##_##b
#define AZREF_CAT(a, b) AZREF_CAT2(a, b)
#define AZREF(name) AZREF_CAT(AZREF_PREFIX, name)

namespace
    {
using namespace hoomd;
using namespace hoomd::azplugins::detail;

inline int type_of(const Scalar4& p)
    {
#if HOOMD_LONGREAL_SIZE == 32
    int32_t t;
    std::memcpy(&t, &p.w, 4);
    return (int)t;
#else
    int64_t t;
    std::memcpy(&t, &p.w, 8);
    return (int)t;
#endif
    }

// BoxDim::minImage, CPU flavour (general triclinic; wraps by at most one image per direction)
struct Box
    {
    Scalar Lx, Ly, Lz, xy, xz, yz;
    bool px, py, pz;
    explicit Box(const azb_box& b)
        : Lx((Scalar)b.Lx), Ly((Scalar)b.Ly), Lz((Scalar)b.Lz), xy((Scalar)b.xy), xz((Scalar)b.xz),
          yz((Scalar)b.yz), px(b.periodic[0]), py(b.periodic[1]), pz(b.periodic[2])
        {
        }
    inline Scalar3 minImage(Scalar3 w) const
        {
        const Scalar hx = Lx / Scalar(2), hy = Ly / Scalar(2), hz = Lz / Scalar(2);
        if (pz)
            {
            if (w.z >= hz)
                {
                w.z -= Lz;
                w.y -= Lz * yz;
                w.x -= Lz * xz;
                }
            else if (w.z < -hz)
                {
                w.z += Lz;
                w.y += Lz * yz;
                w.x += Lz * xz;
                }
            }
        if (py)
            {
            if (w.y >= hy)
                {
                w.y -= Ly;
                w.x -= Ly * xy;
                }
            else if (w.y < -hy)
                {
                w.y += Ly;
                w.x += Ly * xy;
                }
            }
        if (px)
            {
            if (w.x >= hx)
                w.x -= Lx;
            else if (w.x < -hx)
                w.x += Lx;
            }
        return w;
        }
    };

template<class E>
int pair_loop(const azb_pair_args& a, const typename E::param_type* params, int half_list, int nthreads)
    {
    const Scalar4* pos = (const Scalar4*)a.d_pos;
    Scalar4* force = (Scalar4*)a.d_force;
    Scalar* virial = (Scalar*)a.d_virial;
    const Scalar* rcutsq_t = (const Scalar*)a.d_rcutsq;
    const Scalar* ronsq_t = (const Scalar*)a.d_ronsq;
    const Box box(a.box);
    const unsigned int N = a.N, T = a.ntypes;
    const size_t pitch = a.virial_pitch;
    std::memset(force, 0, sizeof(Scalar4) * N);
    if (a.compute_virial)
        std::memset(virial, 0, sizeof(Scalar) * 6 * pitch);
    if (half_list)
        nthreads = 1;
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads) if (nthreads > 1)
    for (unsigned int i = 0; i < N; ++i)
        {
        const Scalar3 pi = make_scalar3(pos[i].x, pos[i].y, pos[i].z);
        const unsigned int ti = type_of(pos[i]);
        Scalar3 fi = make_scalar3(0, 0, 0);
        Scalar pei = 0;
        Scalar vir[6] = {0, 0, 0, 0, 0, 0};
        const size_t head = a.d_head_list[i];
        for (unsigned int k = 0; k < a.d_n_neigh[i]; ++k)
            {
            const unsigned int j = a.d_nlist[head + k];
            Scalar3 dx = make_scalar3(pi.x - pos[j].x, pi.y - pos[j].y, pi.z - pos[j].z);
            const unsigned int tj = type_of(pos[j]);
            dx = box.minImage(dx);
            const Scalar rsq = dot(dx, dx);
            const unsigned int tp = ti * T + tj;
            const Scalar rcutsq = rcutsq_t[tp];
            Scalar ronsq = Scalar(0);
            if (a.shift_mode == AZB_SHIFT_XPLOR)
                ronsq = ronsq_t[tp];
            bool energy_shift = false;
            if (a.shift_mode == AZB_SHIFT_SHIFT)
                energy_shift = true;
            else if (a.shift_mode == AZB_SHIFT_XPLOR && ronsq > rcutsq)
                energy_shift = true;
            Scalar force_divr = 0, pair_eng = 0;
            E eval(rsq, rcutsq, params[tp]);
            const bool evaluated = eval.evalForceAndEnergy(force_divr, pair_eng, energy_shift);
            if (!evaluated)
                continue;
            if (a.shift_mode == AZB_SHIFT_XPLOR && rsq >= ronsq && rsq < rcutsq)
                {
                const Scalar old_e = pair_eng, old_f = force_divr;
                const Scalar d = rcutsq - ronsq;
                const Scalar denom_inv = Scalar(1.0) / (d * d * d);
                const Scalar q = rsq - rcutsq;
                const Scalar s = q * q * (rcutsq + Scalar(2.0) * rsq - Scalar(3.0) * ronsq) * denom_inv;
                const Scalar ds = Scalar(12.0) * (rsq - ronsq) * q * denom_inv;
                pair_eng = old_e * s;
                force_divr = s * old_f - ds * old_e;
                }
            Scalar v[6] = {0, 0, 0, 0, 0, 0};
            if (a.compute_virial)
                {
                const Scalar h = Scalar(0.5) * force_divr;
                v[0] = h * dx.x * dx.x;
                v[1] = h * dx.x * dx.y;
                v[2] = h * dx.x * dx.z;
                v[3] = h * dx.y * dx.y;
                v[4] = h * dx.y * dx.z;
                v[5] = h * dx.z * dx.z;
                for (int r = 0; r < 6; ++r)
                    vir[r] += v[r];
                }
            fi.x += dx.x * force_divr;
            fi.y += dx.y * force_divr;
            fi.z += dx.z * force_divr;
            pei += pair_eng * Scalar(0.5);
            if (half_list && j < N)
                {
                force[j].x -= dx.x * force_divr;
                force[j].y -= dx.y * force_divr;
                force[j].z -= dx.z * force_divr;
                force[j].w += pair_eng * Scalar(0.5);
                if (a.compute_virial)
                    for (int r = 0; r < 6; ++r)
                        virial[r * pitch + j] += v[r];
                }
            }
        force[i].x += fi.x;
        force[i].y += fi.y;
        force[i].z += fi.z;
        force[i].w += pei;
        if (a.compute_virial)
            for (int r = 0; r < 6; ++r)
                virial[r * pitch + i] += vir[r];
        }
    return 0;
    }

int dpd_loop(const azb_dpd_args& d, const PairParametersDPDGeneralWeight* params, int half_list, int nthreads)
    {
    typedef DPDPairEvaluatorGeneralWeight E;
    const azb_pair_args& a = d.pair;
    const Scalar4* pos = (const Scalar4*)a.d_pos;
    const Scalar4* vel = (const Scalar4*)d.d_vel;
    Scalar4* force = (Scalar4*)a.d_force;
    Scalar* virial = (Scalar*)a.d_virial;
    const Scalar* rcutsq_t = (const Scalar*)a.d_rcutsq;
    const Box box(a.box);
    const unsigned int N = a.N, T = a.ntypes;
    const size_t pitch = a.virial_pitch;
    std::memset(force, 0, sizeof(Scalar4) * N);
    if (a.compute_virial)
        std::memset(virial, 0, sizeof(Scalar) * 6 * pitch);
    if (half_list)
        nthreads = 1;
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads) if (nthreads > 1)
    for (unsigned int i = 0; i < N; ++i)
        {
        const Scalar3 pi = make_scalar3(pos[i].x, pos[i].y, pos[i].z);
        const Scalar3 vi = make_scalar3(vel[i].x, vel[i].y, vel[i].z);
        const unsigned int ti = type_of(pos[i]);
        Scalar3 fi = make_scalar3(0, 0, 0);
        Scalar pei = 0;
        Scalar vir[6] = {0, 0, 0, 0, 0, 0};
        const size_t head = a.d_head_list[i];
        for (unsigned int k = 0; k < a.d_n_neigh[i]; ++k)
            {
            const unsigned int j = a.d_nlist[head + k];
            Scalar3 dx = make_scalar3(pi.x - pos[j].x, pi.y - pos[j].y, pi.z - pos[j].z);
            const unsigned int tj = type_of(pos[j]);
            dx = box.minImage(dx);
            const Scalar rsq = dot(dx, dx);
            const Scalar3 dv = make_scalar3(vi.x - vel[j].x, vi.y - vel[j].y, vi.z - vel[j].z);
            const Scalar rdotv = dot(dx, dv);
            const unsigned int tp = ti * T + tj;
            const Scalar rcutsq = rcutsq_t[tp];
            Scalar force_divr = 0, force_divr_cons = 0, pair_eng = 0;
            E eval(rsq, rcutsq, params[tp]);
            // HOOMD passes the 64-bit timestep; the evaluator's parameter is `unsigned int` (truncation)
            eval.set_seed_ij_timestep(d.seed, d.d_tag[i], d.d_tag[j], (unsigned int)d.timestep);
            eval.setDeltaT((Scalar)d.deltaT);
            eval.setRDotV(rdotv);
            eval.setT((Scalar)d.T);
            const bool evaluated
                = eval.evalForceEnergyThermo(force_divr, force_divr_cons, pair_eng, false);
            if (!evaluated)
                continue;
            Scalar v[6] = {0, 0, 0, 0, 0, 0};
            if (a.compute_virial)
                {
                const Scalar h = Scalar(0.5) * force_divr_cons;
                v[0] = h * dx.x * dx.x;
                v[1] = h * dx.x * dx.y;
                v[2] = h * dx.x * dx.z;
                v[3] = h * dx.y * dx.y;
                v[4] = h * dx.y * dx.z;
                v[5] = h * dx.z * dx.z;
                for (int r = 0; r < 6; ++r)
                    vir[r] += v[r];
                }
            fi.x += dx.x * force_divr;
            fi.y += dx.y * force_divr;
            fi.z += dx.z * force_divr;
            pei += pair_eng * Scalar(0.5);
            if (half_list && j < N)
                {
                force[j].x -= dx.x * force_divr;
                force[j].y -= dx.y * force_divr;
                force[j].z -= dx.z * force_divr;
                force[j].w += pair_eng * Scalar(0.5);
                if (a.compute_virial)
                    for (int r = 0; r < 6; ++r)
                        virial[r * pitch + j] += v[r];
                }
            }
        force[i].x += fi.x;
        force[i].y += fi.y;
        force[i].z += fi.z;
        force[i].w += pei;
        if (a.compute_virial)
            for (int r = 0; r < 6; ++r)
                virial[r * pitch + i] += vir[r];
        }
    return 0;
    }

int aniso_loop(const azb_aniso_args& d,
               const AnisoPairParametersTwoPatchMorse* params,
               int half_list,
               int nthreads)
    {
    typedef AnisoPairEvaluatorTwoPatchMorse E;
    const azb_pair_args& a = d.pair;
    const Scalar4* pos = (const Scalar4*)a.d_pos;
    const Scalar4* quat_arr = (const Scalar4*)d.d_orientation;
    Scalar4* force = (Scalar4*)a.d_force;
    Scalar4* torque = (Scalar4*)d.d_torque;
    Scalar* virial = (Scalar*)a.d_virial;
    const Scalar* rcutsq_t = (const Scalar*)a.d_rcutsq;
    const Box box(a.box);
    const unsigned int N = a.N, T = a.ntypes;
    const size_t pitch = a.virial_pitch;
    std::memset(force, 0, sizeof(Scalar4) * N);
    std::memset(torque, 0, sizeof(Scalar4) * N);
    if (a.compute_virial)
        std::memset(virial, 0, sizeof(Scalar) * 6 * pitch);
    if (half_list)
        nthreads = 1;
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads) if (nthreads > 1)
    for (unsigned int i = 0; i < N; ++i)
        {
        const Scalar3 pi = make_scalar3(pos[i].x, pos[i].y, pos[i].z);
        Scalar4 qi = quat_arr[i];
        const unsigned int ti = type_of(pos[i]);
        Scalar3 fi = make_scalar3(0, 0, 0), tqi = make_scalar3(0, 0, 0);
        Scalar pei = 0;
        Scalar vir[6] = {0, 0, 0, 0, 0, 0};
        const size_t head = a.d_head_list[i];
        for (unsigned int k = 0; k < a.d_n_neigh[i]; ++k)
            {
            const unsigned int j = a.d_nlist[head + k];
            Scalar3 dx = make_scalar3(pi.x - pos[j].x, pi.y - pos[j].y, pi.z - pos[j].z);
            const unsigned int tj = type_of(pos[j]);
            Scalar4 qj = quat_arr[j];
            dx = box.minImage(dx);
            const unsigned int tp = ti * T + tj;
            const Scalar rcutsq = rcutsq_t[tp];
            bool energy_shift = false;
            if (a.shift_mode == AZB_SHIFT_SHIFT && E::implementsEnergyShift())
                energy_shift = true;
            Scalar3 f = make_scalar3(0, 0, 0), t_i = make_scalar3(0, 0, 0), t_j = make_scalar3(0, 0, 0);
            Scalar pair_eng = 0;
            E eval(dx, qi, qj, rcutsq, params[tp]);
            const bool evaluated = eval.evaluate(f, pair_eng, energy_shift, t_i, t_j);
            if (!evaluated)
                continue;
            Scalar v[6] = {0, 0, 0, 0, 0, 0};
            if (a.compute_virial)
                {
                v[0] = Scalar(0.5) * dx.x * f.x;
                v[1] = Scalar(0.5) * dx.y * f.x;
                v[2] = Scalar(0.5) * dx.z * f.x;
                v[3] = Scalar(0.5) * dx.y * f.y;
                v[4] = Scalar(0.5) * dx.z * f.y;
                v[5] = Scalar(0.5) * dx.z * f.z;
                for (int r = 0; r < 6; ++r)
                    vir[r] += v[r];
                }
            fi.x += f.x;
            fi.y += f.y;
            fi.z += f.z;
            tqi.x += t_i.x;
            tqi.y += t_i.y;
            tqi.z += t_i.z;
            pei += pair_eng * Scalar(0.5);
            if (half_list && j < N)
                {
                force[j].x -= f.x;
                force[j].y -= f.y;
                force[j].z -= f.z;
                force[j].w += pair_eng * Scalar(0.5);
                torque[j].x += t_j.x;
                torque[j].y += t_j.y;
                torque[j].z += t_j.z;
                if (a.compute_virial)
                    for (int r = 0; r < 6; ++r)
                        virial[r * pitch + j] += v[r];
                }
            }
        force[i].x += fi.x;
        force[i].y += fi.y;
        force[i].z += fi.z;
        force[i].w += pei;
        torque[i].x += tqi.x;
        torque[i].y += tqi.y;
        torque[i].z += tqi.z;
        if (a.compute_virial)
            for (int r = 0; r < 6; ++r)
                virial[r * pitch + i] += vir[r];
        }
    return 0;
    }

// Per-particle bond-table walk (what HOOMD's PotentialBondGPU kernel does and what PotentialBond's CPU loop
// over global bonds sums to). Returns 0, sets *flags non-zero on evaluator failure.
template<class E>
int bond_loop(const azb_bond_args& a, const typename E::param_type* params, unsigned int* flags, int nthreads)
    {
    const Scalar4* pos = (const Scalar4*)a.d_pos;
    Scalar4* force = (Scalar4*)a.d_force;
    Scalar* virial = (Scalar*)a.d_virial;
    const Box box(a.box);
    const unsigned int N = a.N;
    const size_t pitch = a.virial_pitch;
    unsigned int bad = 0;
#pragma omp parallel for schedule(static) num_threads(nthreads) reduction(| : bad) if (nthreads > 1)
    for (unsigned int i = 0; i < N; ++i)
        {
        const Scalar3 pi = make_scalar3(pos[i].x, pos[i].y, pos[i].z);
        Scalar4 fi = make_scalar4(0, 0, 0, 0);
        Scalar vir[6] = {0, 0, 0, 0, 0, 0};
        for (unsigned int k = 0; k < a.d_gpu_n_bonds[i]; ++k)
            {
            const azb_bond_group g = a.d_gpu_bondlist[(size_t)k * a.pitch + i];
            Scalar3 dx = make_scalar3(pi.x - pos[g.idx].x, pi.y - pos[g.idx].y, pi.z - pos[g.idx].z);
            dx = box.minImage(dx);
            const Scalar rsq = dot(dx, dx);
            Scalar force_divr = 0, bond_eng = 0;
            E eval(rsq, params[g.type]);
            const bool evaluated = eval.evalForceAndEnergy(force_divr, bond_eng);
            if (!evaluated)
                {
                bad |= 1u;
                continue;
                }
            bond_eng *= Scalar(0.5);
            const Scalar h = Scalar(0.5) * force_divr;
            vir[0] += h * dx.x * dx.x;
            vir[1] += h * dx.x * dx.y;
            vir[2] += h * dx.x * dx.z;
            vir[3] += h * dx.y * dx.y;
            vir[4] += h * dx.y * dx.z;
            vir[5] += h * dx.z * dx.z;
            fi.x += dx.x * force_divr;
            fi.y += dx.y * force_divr;
            fi.z += dx.z * force_divr;
            fi.w += bond_eng;
            }
        force[i] = fi;
        if (a.compute_virial)
            for (int r = 0; r < 6; ++r)
                virial[r * pitch + i] = vir[r];
        }
    if (bad && flags)
        *flags = 1;
    return 0;
    }

template<class P> void make_param(PyObject* dict, void* out)
    {
    pybind11::dict d = pybind11::reinterpret_borrow<pybind11::dict>(dict);
    P p(d);
    std::memcpy(out, &p, sizeof(P));
    }
template<class P> PyObject* param_dict(const void* in)
    {
    P p;
    std::memcpy(&p, in, sizeof(P));
    return p.asDict().release().ptr();
    }
    } // namespace

#pragma GCC visibility push(default)
extern "C"
    {
    int AZREF(pair_forces)(int potential, const azb_pair_args* a, const void* params, int half_list, int nthreads)
        {
        switch (potential)
            {
        case AZB_PAIR_PERTURBED_LENNARD_JONES:
            return pair_loop<PairEvaluatorPerturbedLennardJones>(
                *a, (const PairParametersPerturbedLennardJones*)params, half_list, nthreads);
        case AZB_PAIR_COLLOID:
            return pair_loop<PairEvaluatorColloid>(*a, (const PairParametersColloid*)params, half_list, nthreads);
        case AZB_PAIR_EXPANDED_YUKAWA:
            return pair_loop<PairEvaluatorExpandedYukawa>(
                *a, (const PairParametersExpandedYukawa*)params, half_list, nthreads);
        case AZB_PAIR_HERTZ:
            return pair_loop<PairEvaluatorHertz>(*a, (const PairParametersHertz*)params, half_list, nthreads);
        case AZB_PAIR_DPD_CONSERVATIVE:
            return pair_loop<DPDPairEvaluatorGeneralWeight>(
                *a, (const PairParametersDPDGeneralWeight*)params, half_list, nthreads);
            }
        return AZB_ERR_BAD_POTENTIAL;
        }

    int AZREF(dpd_forces)(int potential, const azb_dpd_args* a, const void* params, int half_list, int nthreads)
        {
        if (potential != AZB_DPD_GENERAL_WEIGHT)
            return AZB_ERR_BAD_POTENTIAL;
        return dpd_loop(*a, (const PairParametersDPDGeneralWeight*)params, half_list, nthreads);
        }

    int AZREF(aniso_forces)(int potential, const azb_aniso_args* a, const void* params, int half_list, int nthreads)
        {
        if (potential != AZB_ANISO_TWO_PATCH_MORSE)
            return AZB_ERR_BAD_POTENTIAL;
        return aniso_loop(*a, (const AnisoPairParametersTwoPatchMorse*)params, half_list, nthreads);
        }

    int AZREF(bond_forces)(int potential, const azb_bond_args* a, const void* params, unsigned int* flags, int nthreads)
        {
        switch (potential)
            {
        case AZB_BOND_DOUBLE_WELL:
            return bond_loop<BondEvaluatorDoubleWell>(*a, (const BondParametersDoubleWell*)params, flags, nthreads);
        case AZB_BOND_QUARTIC:
            return bond_loop<BondEvaluatorQuartic>(*a, (const BondParametersQuartic*)params, flags, nthreads);
            }
        return AZB_ERR_BAD_POTENTIAL;
        }

    // one Philox draw exactly as DPDPairEvaluatorGeneralWeight.h:227-233 makes it
    double AZREF(dpd_alpha)(unsigned int tag_i, unsigned int tag_j, unsigned int timestep, unsigned int seed)
        {
        unsigned int oi = tag_i, oj = tag_j;
        if (tag_i > tag_j)
            {
            oi = tag_j;
            oj = tag_i;
            }
        hoomd::RandomGenerator rng(hoomd::Seed(RNGIdentifier::DPDEvaluatorGeneralWeight, timestep, (uint16_t)seed),
                                   hoomd::Counter(oi, oj));
        return (double)hoomd::UniformDistribution<Scalar>(-1, 1)(rng);
        }

    size_t AZREF(param_size)(int kind, int potential)
        {
        if (kind == AZB_KIND_PAIR)
            switch (potential)
                {
            case AZB_PAIR_PERTURBED_LENNARD_JONES:
                return sizeof(PairParametersPerturbedLennardJones);
            case AZB_PAIR_COLLOID:
                return sizeof(PairParametersColloid);
            case AZB_PAIR_EXPANDED_YUKAWA:
                return sizeof(PairParametersExpandedYukawa);
            case AZB_PAIR_HERTZ:
                return sizeof(PairParametersHertz);
            case AZB_PAIR_DPD_CONSERVATIVE:
                return sizeof(PairParametersDPDGeneralWeight);
                }
        if (kind == AZB_KIND_DPD && potential == 0)
            return sizeof(PairParametersDPDGeneralWeight);
        if (kind == AZB_KIND_ANISO && potential == 0)
            return sizeof(AnisoPairParametersTwoPatchMorse);
        if (kind == AZB_KIND_BOND)
            switch (potential)
                {
            case AZB_BOND_DOUBLE_WELL:
                return sizeof(BondParametersDoubleWell);
            case AZB_BOND_QUARTIC:
                return sizeof(BondParametersQuartic);
                }
        return 0;
        }

    // dict -> reference param struct through the reference's own pybind11 constructors. Call via ctypes.PyDLL.
    int AZREF(make_params)(int kind, int potential, PyObject* dict, void* out)
        {
        try
            {
            if (kind == AZB_KIND_PAIR || kind == AZB_KIND_DPD)
                {
                if (kind == AZB_KIND_DPD)
                    potential = AZB_PAIR_DPD_CONSERVATIVE;
                switch (potential)
                    {
                case AZB_PAIR_PERTURBED_LENNARD_JONES:
                    make_param<PairParametersPerturbedLennardJones>(dict, out);
                    return 0;
                case AZB_PAIR_COLLOID:
                    make_param<PairParametersColloid>(dict, out);
                    return 0;
                case AZB_PAIR_EXPANDED_YUKAWA:
                    make_param<PairParametersExpandedYukawa>(dict, out);
                    return 0;
                case AZB_PAIR_HERTZ:
                    make_param<PairParametersHertz>(dict, out);
                    return 0;
                case AZB_PAIR_DPD_CONSERVATIVE:
                    make_param<PairParametersDPDGeneralWeight>(dict, out);
                    return 0;
                    }
                }
            if (kind == AZB_KIND_ANISO && potential == 0)
                {
                make_param<AnisoPairParametersTwoPatchMorse>(dict, out);
                return 0;
                }
            if (kind == AZB_KIND_BOND && potential == AZB_BOND_DOUBLE_WELL)
                {
                make_param<BondParametersDoubleWell>(dict, out);
                return 0;
                }
            if (kind == AZB_KIND_BOND && potential == AZB_BOND_QUARTIC)
                {
                make_param<BondParametersQuartic>(dict, out);
                return 0;
                }
            }
        catch (...)
            {
            return AZB_ERR_BAD_ARGUMENT;
            }
        return AZB_ERR_BAD_POTENTIAL;
        }

    // HarmonicBarrier<BarrierEvaluatorT>::computeForces (HarmonicBarrier.h:148-178) around the reference's evaluators:
    // validity test (:125-133), wrap into the global box, evaluate, zero virial.
    int AZREF(barrier_forces)(int geometry, const azb_barrier_args* a)
        {
        const azb_box& b = a->global_box;
        const BoxDim box((Scalar)b.Lx, (Scalar)b.Ly, (Scalar)b.Lz, (Scalar)b.xy, (Scalar)b.xz, (Scalar)b.yz, b.periodic[0], b.periodic[1],
                         b.periodic[2]);
        const Scalar4* pos = (const Scalar4*)a->d_pos;
        const Scalar* params = (const Scalar*)a->d_params;
        Scalar4* force = (Scalar4*)a->d_force;
        auto run = [&](const auto& evaluator) -> int
        {
            if (!evaluator.valid(box))
                return AZB_ERR_BARRIER_INVALID;
            for (unsigned int idx = 0; idx < a->N; ++idx)
                {
                Scalar3 p = make_scalar3(pos[idx].x, pos[idx].y, pos[idx].z);
                const unsigned int type = type_of(pos[idx]);
                int3 img = make_int3(0, 0, 0);
                box.wrap(p, img);
                force[idx] = evaluator(p, params[2 * type], params[2 * type + 1]);
                }
            if (a->d_virial)
                std::memset(a->d_virial, 0, sizeof(Scalar) * 6 * a->virial_pitch);
            return 0;
        };
        if (geometry == AZB_BARRIER_PLANAR)
            return run(hoomd::azplugins::PlanarBarrierEvaluator((Scalar)a->location));
        if (geometry == AZB_BARRIER_SPHERICAL)
            return run(hoomd::azplugins::SphericalBarrierEvaluator((Scalar)a->location));
        return AZB_ERR_BAD_POTENTIAL;
        }

    // reference param struct -> dict (asDict / toPython). Returns a new reference or NULL.
    PyObject* AZREF(params_as_dict)(int kind, int potential, const void* in)
        {
        try
            {
            if (kind == AZB_KIND_PAIR || kind == AZB_KIND_DPD)
                {
                if (kind == AZB_KIND_DPD)
                    potential = AZB_PAIR_DPD_CONSERVATIVE;
                switch (potential)
                    {
                case AZB_PAIR_PERTURBED_LENNARD_JONES:
                    return param_dict<PairParametersPerturbedLennardJones>(in);
                case AZB_PAIR_COLLOID:
                    return param_dict<PairParametersColloid>(in);
                case AZB_PAIR_EXPANDED_YUKAWA:
                    return param_dict<PairParametersExpandedYukawa>(in);
                case AZB_PAIR_HERTZ:
                    return param_dict<PairParametersHertz>(in);
                case AZB_PAIR_DPD_CONSERVATIVE:
                    return param_dict<PairParametersDPDGeneralWeight>(in);
                    }
                }
            if (kind == AZB_KIND_ANISO && potential == 0)
                {
                AnisoPairParametersTwoPatchMorse p;
                std::memcpy(&p, in, sizeof(p));
                return p.toPython().release().ptr();
                }
            if (kind == AZB_KIND_BOND && potential == AZB_BOND_DOUBLE_WELL)
                return param_dict<BondParametersDoubleWell>(in);
            if (kind == AZB_KIND_BOND && potential == AZB_BOND_QUARTIC)
                return param_dict<BondParametersQuartic>(in);
            }
        catch (...)
            {
            }
        Py_RETURN_NONE;
        }
    }
#pragma GCC visibility pop
