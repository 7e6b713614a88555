/* TEST INFRASTRUCTURE (oracle) — included twice by azb_oracle.c with
 *   REAL = float,  SUF(x) = x##_f32   and   REAL = double, SUF(x) = x##_f64.
 *
 * Plain-C restatement of the reference's per-pair arithmetic (each function cites the file:line it follows)
 * and of the HOOMD-blue v5 host loops around it (HOOMD itself is not under
 * /root/reference). The operation order mirrors the reference so that, in the same precision, results agree
 * with oracle/_ref (the reference headers compiled unmodified) to the last bits.
 */

typedef struct
    {
    REAL x, y, z, w;
    } SUF(real4);

/* ---- parameter structs: field order of the reference's param_type ---- */
typedef struct
    {
    REAL sigma_6, epsilon_x_4, attraction_scale_factor, rwcasq;
    } SUF(plj_params); /* PairEvaluatorPerturbedLennardJones.h:57-60 */
typedef struct
    {
    REAL A, a_1, a_2, sigma_3;
    } SUF(colloid_params); /* PairEvaluatorColloid.h:48-51 */
typedef struct
    {
    REAL epsilon, kappa, delta, pad_;
    } SUF(yukawa_params); /* PairEvaluatorExpandedYukawa.h:45-47 (+ alignment padding to 4 Scalars) */
typedef struct
    {
    REAL epsilon;
    } SUF(hertz_params); /* PairEvaluatorHertz.h:41 */
typedef struct
    {
    REAL A, gamma, s, pad_;
    } SUF(dpd_params); /* DPDPairEvaluatorGeneralWeight.h:54-56 (+ padding) */
typedef struct
    {
    REAL M_d, M_rinv, r_eq, omega, alpha;
    unsigned char repulsion; /* bool; struct tail-padded to sizeof(REAL) multiple */
    } SUF(tpm_params); /* AnisoPairEvaluatorTwoPatchMorse.h:63-68 */
typedef struct
    {
    REAL r_1, r_diff, U_1, U_tilt;
    } SUF(dw_params); /* BondEvaluatorDoubleWell.h:51-54 */
typedef struct
    {
    REAL k, r_0, b_1, b_2, U_0, sigma_6, epsilon_x_4, delta;
    } SUF(quartic_params); /* BondEvaluatorQuartic.h:69-76 */

static inline int SUF(type_of)(REAL w)
    {
    if (sizeof(REAL) == 4)
        {
        int32_t t;
        memcpy(&t, &w, 4);
        return (int)t;
        }
    else
        {
        int64_t t;
        memcpy(&t, &w, 8);
        return (int)t;
        }
    }

/* BoxDim::minImage, CPU flavour: one wrap per direction, z then y then x with tilt terms */
static inline void SUF(min_image)(const azb_box* b, REAL* x, REAL* y, REAL* z)
    {
    const REAL Lx = (REAL)b->Lx, Ly = (REAL)b->Ly, Lz = (REAL)b->Lz;
    const REAL xy = (REAL)b->xy, xz = (REAL)b->xz, yz = (REAL)b->yz;
    const REAL hx = Lx / (REAL)2, hy = Ly / (REAL)2, hz = Lz / (REAL)2;
    if (b->periodic[2])
        {
        if (*z >= hz)
            {
            *z -= Lz;
            *y -= Lz * yz;
            *x -= Lz * xz;
            }
        else if (*z < -hz)
            {
            *z += Lz;
            *y += Lz * yz;
            *x += Lz * xz;
            }
        }
    if (b->periodic[1])
        {
        if (*y >= hy)
            {
            *y -= Ly;
            *x -= Ly * xy;
            }
        else if (*y < -hy)
            {
            *y += Ly;
            *x += Ly * xy;
            }
        }
    if (b->periodic[0])
        {
        if (*x >= hx)
            *x -= Lx;
        else if (*x < -hx)
            *x += Lx;
        }
    }

/* ------------------------------------------------------------------------------------------------- */
/* isotropic pair evaluators: return 1 when "evaluated"                                               */
/* ------------------------------------------------------------------------------------------------- */

/* PairEvaluatorPerturbedLennardJones.h:97-104 (ctor) and :117-155 (evalForceAndEnergy) */
static int SUF(eval_plj)(REAL rsq, REAL rcutsq, const SUF(plj_params) * p, int energy_shift, REAL* force_divr, REAL* pair_eng)
    {
    const REAL lj1 = p->epsilon_x_4 * p->sigma_6 * p->sigma_6;
    const REAL lj2 = p->epsilon_x_4 * p->sigma_6;
    const REAL lam = p->attraction_scale_factor;
    const REAL rwcasq = p->rwcasq;
    const REAL wca_shift = p->epsilon_x_4 * ((REAL)1.0 - lam) / (REAL)4.0;
    if (rsq < rcutsq && lj1 != 0)
        {
        const REAL r2inv = (REAL)1.0 / rsq;
        const REAL r6inv = r2inv * r2inv * r2inv;
        *force_divr = r2inv * r6inv * ((REAL)12.0 * lj1 * r6inv - (REAL)6.0 * lj2);
        *pair_eng = r6inv * (lj1 * r6inv - lj2);
        if (rsq < rwcasq)
            *pair_eng += wca_shift;
        else
            {
            *force_divr *= lam;
            *pair_eng *= lam;
            }
        if (energy_shift)
            {
            const REAL rcut2inv = (REAL)1.0 / rcutsq;
            const REAL rcut6inv = rcut2inv * rcut2inv * rcut2inv;
            REAL sh = rcut6inv * (lj1 * rcut6inv - lj2);
            if (rcutsq < rwcasq)
                sh += wca_shift;
            else
                sh *= lam;
            *pair_eng -= sh;
            }
        return 1;
        }
    return 0;
    }

/* PairEvaluatorColloid.h:101-113 */
static REAL SUF(colloid_ss)(REAL A, REAL sigma_6, int force, REAL* force_divr, REAL rsq)
    {
    const REAL r2inv = (REAL)1.0 / rsq;
    const REAL r6inv = r2inv * r2inv * r2inv;
    const REAL c1 = A * sigma_6 / (REAL)36.0;
    if (force)
        *force_divr = (REAL)6.0 * c1 * r2inv * r6inv * ((REAL)2.0 * sigma_6 * r6inv - (REAL)1.0);
    return c1 * r6inv * (sigma_6 * r6inv - (REAL)1.0);
    }

/* PairEvaluatorColloid.h:125-152. Constants 4/15, 2/9, 2.0, 5.0 (unsuffixed) and 4.2 follow the
 * reference's mixed Scalar()/double literals: `Scalar(4.0/15.0)`, `2.0 * (...)`, `5.0 * rsqsq`. */
static REAL SUF(colloid_cs)(REAL A, REAL sigma_3, REAL sigma_6, REAL ai, REAL aj, int force, REAL* force_divr, REAL rsq)
    {
    const REAL a = (ai > aj) ? ai : aj;
    const REAL asq = a * a;
    REAL asq_minus_rsq = asq - rsq;
    REAL rsqsq = rsq * rsq;
    REAL d3 = asq_minus_rsq * asq_minus_rsq * asq_minus_rsq;
    REAL d6 = d3 * d3;
    REAL fR = sigma_3 * A * a * asq / d3;
    if (force)
        {
        *force_divr = (REAL)((REAL)(4.0 / 15.0) * fR
                             * (2.0 * (asq + rsq) * (asq * ((REAL)5.0 * asq + (REAL)22.0 * rsq) + 5.0 * rsqsq)
                                    * sigma_6 / d6
                                - (REAL)5.0)
                             / asq_minus_rsq);
        }
    return (REAL)(2.0 / 9.0) * fR
           * ((REAL)1.0
              - (asq * (asq * (asq / (REAL)3.0 + (REAL)3.0 * rsq) + (REAL)4.2 * rsqsq) + rsq * rsqsq) * sigma_6
                    / d6);
    }

/* PairEvaluatorColloid.h:164-220 */
static REAL SUF(colloid_cc)(REAL A, REAL sigma_6, REAL ai, REAL aj, int force, REAL* force_divr, REAL rsq)
    {
    const REAL r = SQRT(rsq); /* unqualified sqrt() in the reference resolves per Scalar type */
    const REAL k0 = ai * aj, k1 = ai + aj, k2 = ai - aj;
    const REAL k3 = k1 + r, k4 = k1 - r, k5 = k2 + r, k6 = k2 - r;
    const REAL k7 = (REAL)1.0 / (k3 * k4);
    const REAL k8 = (REAL)1.0 / (k5 * k6);
    const REAL k3inv = (REAL)1.0 / k3;
    REAL g0 = k3inv * k3inv;
    g0 *= g0 * g0;
    g0 *= k3inv;
    const REAL k4inv = (REAL)1.0 / k4;
    REAL g1 = k4inv * k4inv;
    g1 *= g1 * g1;
    g1 *= k4inv;
    const REAL k5inv = (REAL)1.0 / k5;
    REAL g2 = k5inv * k5inv;
    g2 *= g2 * g2;
    g2 *= k5inv;
    const REAL k6inv = (REAL)1.0 / k6;
    REAL g3 = k6inv * k6inv;
    g3 *= g3 * g3;
    g3 *= k6inv;
    const REAL h0 = ((k3 + (REAL)5.0 * k1) * k3 + (REAL)30.0 * k0) * g0;
    const REAL h1 = ((k4 + (REAL)5.0 * k1) * k4 + (REAL)30.0 * k0) * g1;
    const REAL h2 = ((k5 + (REAL)5.0 * k2) * k5 - (REAL)30.0 * k0) * g2;
    const REAL h3 = ((k6 + (REAL)5.0 * k2) * k6 - (REAL)30.0 * k0) * g3;
    g0 *= (REAL)42.0 * k0 / k3 + (REAL)6.0 * k1 + k3;
    g1 *= (REAL)42.0 * k0 / k4 + (REAL)6.0 * k1 + k4;
    g2 *= (REAL)-42.0 * k0 / k5 + (REAL)6.0 * k2 + k5;
    g3 *= (REAL)-42.0 * k0 / k6 + (REAL)6.0 * k2 + k6;
    const REAL fR = A * sigma_6 / r / (REAL)37800.0;
    REAL pair_eng = fR * (h0 - h1 - h2 + h3);
    if (force)
        {
        const REAL dUR = pair_eng / r + (REAL)5.0 * fR * (g0 + g1 - g2 - g3);
        const REAL dUA = -A / (REAL)3.0 * r
                         * (((REAL)2.0 * k0 * k7 + (REAL)1.0) * k7 + ((REAL)2.0 * k0 * k8 - (REAL)1.0) * k8);
        *force_divr = (dUR + dUA) / r;
        }
    pair_eng += A / (REAL)6.0 * ((REAL)2.0 * k0 * (k7 + k8) - LOG(k8 / k7));
    return pair_eng;
    }

/* PairEvaluatorColloid.h:81-89 (ctor), :233-269 */
static int SUF(eval_colloid)(REAL rsq, REAL rcutsq, const SUF(colloid_params) * p, int energy_shift, REAL* force_divr, REAL* pair_eng)
    {
    const REAL A = p->A, sigma_3 = p->sigma_3, sigma_6 = sigma_3 * sigma_3, ai = p->a_1, aj = p->a_2;
    REAL dummy = 0;
    if (rsq < rcutsq && A != (REAL)0)
        {
        if (ai == (REAL)0 && aj == (REAL)0)
            {
            *pair_eng = SUF(colloid_ss)(A, sigma_6, 1, force_divr, rsq);
            if (energy_shift)
                *pair_eng -= SUF(colloid_ss)(A, sigma_6, 0, &dummy, rcutsq);
            }
        else if (ai != (REAL)0 && aj != (REAL)0)
            {
            *pair_eng = SUF(colloid_cc)(A, sigma_6, ai, aj, 1, force_divr, rsq);
            if (energy_shift)
                *pair_eng -= SUF(colloid_cc)(A, sigma_6, ai, aj, 0, &dummy, rcutsq);
            }
        else
            {
            *pair_eng = SUF(colloid_cs)(A, sigma_3, sigma_6, ai, aj, 1, force_divr, rsq);
            if (energy_shift)
                *pair_eng -= SUF(colloid_cs)(A, sigma_3, sigma_6, ai, aj, 0, &dummy, rcutsq);
            }
        return 1;
        }
    return 0;
    }

/* PairEvaluatorExpandedYukawa.h:73-79, :92-115 */
static int SUF(eval_yukawa)(REAL rsq, REAL rcutsq, const SUF(yukawa_params) * p, int energy_shift, REAL* force_divr, REAL* pair_eng)
    {
    const REAL epsilon = p->epsilon, kappa = p->kappa, delta = p->delta;
    if (rsq < rcutsq && epsilon != (REAL)0)
        {
        const REAL r = SQRT(rsq);
        const REAL r_delta = r - delta;
        const REAL r_delta_inv = (REAL)1.0 / r_delta;
        *pair_eng = epsilon * EXP(-kappa * r_delta) * r_delta_inv;
        *force_divr = *pair_eng * (kappa + r_delta_inv) / r;
        if (energy_shift)
            {
            const REAL rcut = SQRT(rcutsq);
            const REAL rcut_delta = rcut - delta;
            *pair_eng -= epsilon * EXP(-kappa * rcut_delta) / rcut_delta;
            }
        return 1;
        }
    return 0;
    }

/* PairEvaluatorHertz.h:77-80, :93-110 (energy_shift ignored) */
static int SUF(eval_hertz)(REAL rsq, REAL rcutsq, const SUF(hertz_params) * p, int energy_shift, REAL* force_divr, REAL* pair_eng)
    {
    const REAL epsilon = p->epsilon;
    (void)energy_shift;
    if (rsq < rcutsq && epsilon != (REAL)0)
        {
        const REAL r = SQRT(rsq);
        const REAL rcut = SQRT(rcutsq);
        const REAL x = (REAL)1.0 - (r / rcut);
        const REAL xsqrt = SQRT(x);
        const REAL ex3p2 = epsilon * x * xsqrt;
        *force_divr = (REAL)2.5 * ex3p2 / (r * rcut);
        *pair_eng = ex3p2 * x;
        return 1;
        }
    return 0;
    }

/* DPDPairEvaluatorGeneralWeight.h:165-183 (conservative only) */
static int SUF(eval_dpd_cons)(REAL rsq, REAL rcutsq, const SUF(dpd_params) * p, int energy_shift, REAL* force_divr, REAL* pair_eng)
    {
    const REAL A = p->A;
    (void)energy_shift;
    if (rsq < rcutsq)
        {
        REAL rinv = (REAL)1.0 / SQRT(rsq); /* fast::rsqrt, host flavour */
        REAL r = (REAL)1.0 / rinv;
        REAL rcutinv = (REAL)1.0 / SQRT(rcutsq);
        REAL rcut = (REAL)1.0 / rcutinv;
        *force_divr = A * (rinv - rcutinv);
        *pair_eng = A * (rcut - r) - (REAL)0.5 * A * rcutinv * (rcutsq - rsq);
        return 1;
        }
    return 0;
    }

/* ------------------------------------------------------------------------------------------------- */
/* DPD thermostat: DPDPairEvaluatorGeneralWeight.h:198-255                                            */
/* ------------------------------------------------------------------------------------------------- */

/* hoomd::UniformDistribution<Scalar>(-1,1)(rng) with rng = RandomGenerator(Seed(200, timestep, seed),
 * Counter(min(tag), max(tag))): exactly one Philox4x32-10 call (spec: oracle/shim/hoomd/RandomNumbers.h). */
static REAL SUF(dpd_alpha)(unsigned int tag_i, unsigned int tag_j, unsigned int timestep, uint16_t seed)
    {
    unsigned int oi = tag_i, oj = tag_j; /* :213-224 */
    uint32_t ctr[4], key[2], out[4];
    if (tag_i > tag_j)
        {
        oi = tag_j;
        oj = tag_i;
        }
    key[0] = ((uint32_t)200u << 24) | ((uint32_t)seed << 8) | 0u; /* RNGIdentifiers.h:23; timestep is u32 */
    key[1] = (uint32_t)timestep;
    ctr[0] = 0;
    ctr[1] = 0;
    ctr[2] = oj;
    ctr[3] = oi;
    azo_philox4x32_10(ctr, key, out);
    if (sizeof(REAL) == 4)
        {
        const float factor = 1.0f / 4294967296.0f;
        const float u = (float)out[0] * factor + 0.5f * factor;
        return (REAL)(-1.0f + 2.0f * u);
        }
    else
        {
        const double factor = 1.0 / 18446744073709551616.0;
        const uint64_t u64 = ((uint64_t)out[0] << 32) | out[1];
        const double u = (double)u64 * factor + 0.5 * factor;
        return (REAL)(-1.0 + 2.0 * u);
        }
    }

static int SUF(eval_dpd_thermo)(REAL rsq, REAL rcutsq, const SUF(dpd_params) * p, unsigned int tag_i, unsigned int tag_j,
                                unsigned int timestep, uint16_t seed, REAL deltaT, REAL rdotv, REAL T,
                                REAL* force_divr, REAL* force_divr_cons, REAL* pair_eng)
    {
    const REAL A = p->A, gamma = p->gamma, s = p->s;
    if (rsq < rcutsq)
        {
        REAL rinv = (REAL)1.0 / SQRT(rsq);
        REAL r = (REAL)1.0 / rinv;
        REAL rcutinv = (REAL)1.0 / SQRT(rcutsq);
        REAL rcut = (REAL)1.0 / rcutinv;
        const REAL alpha = SUF(dpd_alpha)(tag_i, tag_j, timestep, seed);
        *force_divr = A * (rinv - rcutinv);
        *force_divr_cons = *force_divr;
        const REAL wR = POW((REAL)1. - r * rcutinv, (REAL)0.5 * s) * rinv;
        *force_divr -= gamma * wR * wR * rdotv;
        *force_divr += ((REAL)1.0 / SQRT(deltaT / (T * gamma * (REAL)6.0))) * wR * alpha;
        *pair_eng = A * (rcut - r) - (REAL)0.5 * A * rcutinv * (rcutsq - rsq);
        return 1;
        }
    return 0;
    }

/* ------------------------------------------------------------------------------------------------- */
/* TwoPatchMorse: AnisoPairEvaluatorTwoPatchMorse.h:127-216                                           */
/* ------------------------------------------------------------------------------------------------- */
static inline void SUF(cross3)(const REAL* a, const REAL* b, REAL* o)
    {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
    }
/* rotate(quat, (1,0,0)) with the general formula (s^2 - v.v) b + 2 s (v x b) + 2 (v.b) v */
static inline void SUF(patch_axis)(const SUF(real4) * q, REAL* n)
    {
    const REAL s = q->x, v[3] = {q->y, q->z, q->w}, b[3] = {(REAL)1.0, 0, 0};
    REAL c[3];
    const REAL vv = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    const REAL vb = v[0] * b[0] + v[1] * b[1] + v[2] * b[2];
    SUF(cross3)(v, b, c);
    for (int d = 0; d < 3; ++d)
        n[d] = (s * s - vv) * b[d] + (REAL)2 * s * c[d] + (REAL)2 * vb * v[d];
    }

static int SUF(eval_tpm)(const REAL* dr, const SUF(real4) * qi, const SUF(real4) * qj, REAL rcutsq, const SUF(tpm_params) * p,
                         int energy_shift, REAL* force, REAL* pair_eng, REAL* torque_i, REAL* torque_j)
    {
    const REAL M_d = p->M_d, M_rinv = p->M_rinv, r_eq = p->r_eq, omega = p->omega, alpha = p->alpha;
    const int repulsion = p->repulsion != 0;
    REAL rsq = dr[0] * dr[0] + dr[1] * dr[1] + dr[2] * dr[2];
    if (rsq > rcutsq) /* strict '>' (:135) */
        return 0;
    REAL rinv = (REAL)1.0 / SQRT(rsq);
    REAL r = (REAL)1. / rinv;
    REAL unitr[3] = {dr[0] * rinv, dr[1] * rinv, dr[2] * rinv};
    REAL n_i[3], n_j[3];
    SUF(patch_axis)(qi, n_i);
    SUF(patch_axis)(qj, n_j);
    REAL e = (REAL)0.0;
    REAL UMorse = (REAL)-1.0 * M_d;
    REAL dUMorse_dr = (REAL)0.0;
    if (r > r_eq || repulsion)
        {
        REAL Morse_exp = EXP(-(r - r_eq) * M_rinv);
        REAL one_minus_exp = (REAL)1.0 - Morse_exp;
        UMorse = M_d * (one_minus_exp * one_minus_exp - (REAL)1.0);
        dUMorse_dr = (REAL)2.0 * M_d * M_rinv * Morse_exp * one_minus_exp;
        }
    REAL gamma_i = unitr[0] * n_i[0] + unitr[1] * n_i[1] + unitr[2] * n_i[2];
    REAL gamma_i_exp = EXP(-omega * (gamma_i * gamma_i - alpha));
    REAL Omega_i = (REAL)1.0 / ((REAL)1.0 + gamma_i_exp);
    REAL gamma_j = unitr[0] * n_j[0] + unitr[1] * n_j[1] + unitr[2] * n_j[2];
    REAL gamma_j_exp = EXP(-omega * (gamma_j * gamma_j - alpha));
    REAL Omega_j = (REAL)1.0 / ((REAL)1.0 + gamma_j_exp);
    e += UMorse * Omega_i * Omega_j;
    REAL dU_dr = dUMorse_dr * Omega_i * Omega_j;
    REAL dOmegai_dgi = (REAL)2.0 * omega * gamma_i * gamma_i_exp * Omega_i * Omega_i;
    REAL dOmegaj_dgj = (REAL)2.0 * omega * gamma_j * gamma_j_exp * Omega_j * Omega_j;
    REAL dU_dgi = dOmegai_dgi * UMorse * Omega_j;
    REAL dU_dgj = dOmegaj_dgj * UMorse * Omega_i;
    REAL neg_unitr[3] = {-unitr[0], -unitr[1], -unitr[2]};
    REAL ci[3], cj[3], n_i_perp[3], n_j_perp[3];
    SUF(cross3)(unitr, n_i, ci);
    SUF(cross3)(unitr, n_j, cj);
    SUF(cross3)(neg_unitr, ci, n_i_perp);
    SUF(cross3)(neg_unitr, cj, n_j_perp);
    for (int d = 0; d < 3; ++d)
        {
        force[d] = -dU_dr * unitr[d] - rinv * (dU_dgi * n_i_perp[d] + dU_dgj * n_j_perp[d]);
        torque_i[d] = dU_dgi * ci[d];
        torque_j[d] = dU_dgj * cj[d];
        }
    if (energy_shift)
        {
        REAL rcut = SQRT(rcutsq);
        REAL Morse_exp_shift = EXP(-(rcut - r_eq) * M_rinv);
        REAL one_minus_exp_shift = (REAL)1.0 - Morse_exp_shift;
        REAL UMorse_shift = M_d * (one_minus_exp_shift * one_minus_exp_shift - (REAL)1.0);
        e -= UMorse_shift * Omega_i * Omega_j;
        }
    *pair_eng = e;
    return 1;
    }

/* ------------------------------------------------------------------------------------------------- */
/* bonds                                                                                              */
/* ------------------------------------------------------------------------------------------------- */

/* BondEvaluatorDoubleWell.h:90-113 */
static int SUF(eval_dw)(REAL rsq, const SUF(dw_params) * p, REAL* force_divr, REAL* bond_eng)
    {
    const REAL r_1 = p->r_1, r_diff = p->r_diff, U_1 = p->U_1, U_tilt = p->U_tilt;
    *bond_eng = 0;
    *force_divr = 0;
    if (r_diff == (REAL)0.0)
        return 0;
    const REAL r = SQRT(rsq);
    const REAL x = (r_1 - r) / r_diff;
    const REAL x2 = x * x;
    const REAL y = (REAL)1.0 - x2;
    const REAL y2 = y * y;
    *bond_eng = U_1 * y2 + U_tilt * ((REAL)1.0 - x - y2);
    *force_divr = ((REAL)4.0 * x * y * (U_tilt - U_1) - U_tilt) / (r_diff * r);
    return 1;
    }

/* BondEvaluatorQuartic.h:116-127 (ctor), :129-200. Note the local `epsilon` that shadows the member in
 * the delta == 0 branch (:150) and the member epsilon = epsilon_x_4/4 used in the delta != 0 branch (:175). */
static int SUF(eval_quartic)(REAL rsq, const SUF(quartic_params) * p, REAL* force_divr, REAL* bond_eng)
    {
    const REAL k = p->k, r_0 = p->r_0, b_1 = p->b_1, b_2 = p->b_2, U_0 = p->U_0, delta = p->delta;
    const REAL lj1 = p->epsilon_x_4 * p->sigma_6 * p->sigma_6;
    const REAL lj2 = p->epsilon_x_4 * p->sigma_6;
    const REAL epsilon_m = p->epsilon_x_4 / (REAL)4.;
    *bond_eng = 0;
    *force_divr = 0;
    if (r_0 == (REAL)0.0)
        return 0;
    REAL r_red = (REAL)1.0;
    if (delta == (REAL)0.0)
        {
        const REAL r2inv = (REAL)1.0 / rsq;
        const REAL r6inv = r2inv * r2inv * r2inv;
        const REAL sigma6inv = lj2 / lj1;
        if (lj1 != (REAL)0 && r6inv > sigma6inv / (REAL)2.0)
            {
            REAL epsilon = lj2 * lj2 / (REAL)4.0 / lj1;
            *force_divr += r2inv * r6inv * ((REAL)12.0 * lj1 * r6inv - (REAL)6.0 * lj2);
            *bond_eng += r6inv * (lj1 * r6inv - lj2) + epsilon;
            }
        if (rsq < r_0 * r_0)
            r_red = SQRT(rsq) - r_0;
        }
    else
        {
        REAL r = SQRT(rsq) - delta;
        const REAL r2inv = (REAL)1.0 / r / r;
        const REAL r6inv = r2inv * r2inv * r2inv;
        const REAL sigma6inv = lj2 / lj1;
        if (lj1 != (REAL)0 && r6inv > sigma6inv / (REAL)2.0)
            {
            *force_divr += r6inv * ((REAL)12.0 * lj1 * r6inv - (REAL)6.0 * lj2) / r / (r + delta);
            *bond_eng += r6inv * (lj1 * r6inv - lj2) + epsilon_m;
            }
        if (r < r_0)
            r_red = r - r_0;
        }
    if (r_red < (REAL)0)
        {
        *force_divr += (REAL)-1 * k * r_red * (4 * r_red * r_red - 3 * (b_1 + b_2) * r_red + 2 * b_1 * b_2)
                       / (r_red + r_0 + delta);
        *bond_eng += k * (r_red - b_1) * (r_red - b_2) * r_red * r_red + U_0;
        }
    else
        *bond_eng += U_0;
    return 1;
    }

/* ------------------------------------------------------------------------------------------------- */
/* host loops. half_list != 0 applies Newton's third law (serial);      */
/* with a full list the loop over i is OpenMP-parallel (nthreads).                                    */
/* ------------------------------------------------------------------------------------------------- */
#define AZO_KIND_PAIR 0
#define AZO_KIND_DPD 1
#define AZO_KIND_ANISO 2

static int SUF(pairlike_loop)(int kind, int potential, const azb_pair_args* a, const azb_dpd_args* dpd, const azb_aniso_args* an,
                              const void* params, int half_list, int nthreads)
    {
    const SUF(real4)* pos = (const SUF(real4)*)a->d_pos;
    const SUF(real4)* vel = dpd ? (const SUF(real4)*)dpd->d_vel : NULL;
    const SUF(real4)* ori = an ? (const SUF(real4)*)an->d_orientation : NULL;
    SUF(real4)* force = (SUF(real4)*)a->d_force;
    SUF(real4)* torque = an ? (SUF(real4)*)an->d_torque : NULL;
    REAL* virial = (REAL*)a->d_virial;
    const REAL* rcutsq_t = (const REAL*)a->d_rcutsq;
    const REAL* ronsq_t = (const REAL*)a->d_ronsq;
    const unsigned int N = a->N, T = a->ntypes;
    const size_t pitch = a->virial_pitch;
    const int cv = a->compute_virial != 0;
    if (kind == AZO_KIND_PAIR && (potential < 0 || potential >= AZB_PAIR_COUNT_))
        return AZB_ERR_BAD_POTENTIAL;
    if (kind != AZO_KIND_PAIR && potential != 0)
        return AZB_ERR_BAD_POTENTIAL;
    memset(force, 0, sizeof(SUF(real4)) * N);
    if (torque)
        memset(torque, 0, sizeof(SUF(real4)) * N);
    if (cv)
        memset(virial, 0, sizeof(REAL) * 6 * pitch);
    if (half_list || nthreads < 1)
        nthreads = 1;
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads) if (nthreads > 1)
    for (unsigned int i = 0; i < N; ++i)
        {
        const REAL pix = pos[i].x, piy = pos[i].y, piz = pos[i].z;
        const unsigned int ti = (unsigned int)SUF(type_of)(pos[i].w);
        REAL fi[3] = {0, 0, 0}, tqi[3] = {0, 0, 0}, pei = 0, vir[6] = {0, 0, 0, 0, 0, 0};
        const size_t head = a->d_head_list[i];
        for (unsigned int k = 0; k < a->d_n_neigh[i]; ++k)
            {
            const unsigned int j = a->d_nlist[head + k];
            REAL dx[3] = {pix - pos[j].x, piy - pos[j].y, piz - pos[j].z};
            const unsigned int tj = (unsigned int)SUF(type_of)(pos[j].w);
            SUF(min_image)(&a->box, &dx[0], &dx[1], &dx[2]);
            const REAL rsq = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
            const unsigned int tp = ti * T + tj;
            const REAL rcutsq = rcutsq_t[tp];
            REAL f[3] = {0, 0, 0}, t_i[3] = {0, 0, 0}, t_j[3] = {0, 0, 0}, v[6] = {0, 0, 0, 0, 0, 0};
            REAL pair_eng = 0;
            int evaluated = 0;
            if (kind == AZO_KIND_ANISO)
                {
                /* AnisoPotentialPair: shift only when the evaluator implements it (TwoPatchMorse does, :122-125) */
                const int energy_shift = (a->shift_mode == AZB_SHIFT_SHIFT);
                evaluated = SUF(eval_tpm)(dx, &ori[i], &ori[j], rcutsq, &((const SUF(tpm_params)*)params)[tp], energy_shift,
                                          f, &pair_eng, t_i, t_j);
                if (!evaluated)
                    continue;
                if (cv)
                    {
                    v[0] = (REAL)0.5 * dx[0] * f[0];
                    v[1] = (REAL)0.5 * dx[1] * f[0];
                    v[2] = (REAL)0.5 * dx[2] * f[0];
                    v[3] = (REAL)0.5 * dx[1] * f[1];
                    v[4] = (REAL)0.5 * dx[2] * f[1];
                    v[5] = (REAL)0.5 * dx[2] * f[2];
                    }
                }
            else
                {
                REAL force_divr = 0, force_divr_vir = 0;
                if (kind == AZO_KIND_DPD)
                    {
                    const REAL dvx = vel[i].x - vel[j].x, dvy = vel[i].y - vel[j].y, dvz = vel[i].z - vel[j].z;
                    const REAL rdotv = dx[0] * dvx + dx[1] * dvy + dx[2] * dvz;
                    evaluated = SUF(eval_dpd_thermo)(rsq, rcutsq, &((const SUF(dpd_params)*)params)[tp], dpd->d_tag[i],
                                                     dpd->d_tag[j], (unsigned int)dpd->timestep, dpd->seed, (REAL)dpd->deltaT,
                                                     rdotv, (REAL)dpd->T, &force_divr, &force_divr_vir, &pair_eng);
                    if (!evaluated)
                        continue;
                    }
                else
                    {
                    REAL ronsq = (REAL)0;
                    int energy_shift = 0;
                    if (a->shift_mode == AZB_SHIFT_XPLOR)
                        ronsq = ronsq_t[tp];
                    if (a->shift_mode == AZB_SHIFT_SHIFT)
                        energy_shift = 1;
                    else if (a->shift_mode == AZB_SHIFT_XPLOR && ronsq > rcutsq)
                        energy_shift = 1;
                    switch (potential)
                        {
                    case AZB_PAIR_PERTURBED_LENNARD_JONES:
                        evaluated = SUF(eval_plj)(rsq, rcutsq, &((const SUF(plj_params)*)params)[tp], energy_shift, &force_divr, &pair_eng);
                        break;
                    case AZB_PAIR_COLLOID:
                        evaluated = SUF(eval_colloid)(rsq, rcutsq, &((const SUF(colloid_params)*)params)[tp], energy_shift, &force_divr, &pair_eng);
                        break;
                    case AZB_PAIR_EXPANDED_YUKAWA:
                        evaluated = SUF(eval_yukawa)(rsq, rcutsq, &((const SUF(yukawa_params)*)params)[tp], energy_shift, &force_divr, &pair_eng);
                        break;
                    case AZB_PAIR_HERTZ:
                        evaluated = SUF(eval_hertz)(rsq, rcutsq, &((const SUF(hertz_params)*)params)[tp], energy_shift, &force_divr, &pair_eng);
                        break;
                    default:
                        evaluated = SUF(eval_dpd_cons)(rsq, rcutsq, &((const SUF(dpd_params)*)params)[tp], energy_shift, &force_divr, &pair_eng);
                        break;
                        }
                    if (!evaluated)
                        continue;
                    if (a->shift_mode == AZB_SHIFT_XPLOR && rsq >= ronsq && rsq < rcutsq)
                        {
                        const REAL old_e = pair_eng, old_f = force_divr;
                        const REAL d = rcutsq - ronsq;
                        const REAL denom_inv = (REAL)1.0 / (d * d * d);
                        const REAL q = rsq - rcutsq;
                        const REAL s = q * q * (rcutsq + (REAL)2.0 * rsq - (REAL)3.0 * ronsq) * denom_inv;
                        const REAL ds = (REAL)12.0 * (rsq - ronsq) * q * denom_inv;
                        pair_eng = old_e * s;
                        force_divr = s * old_f - ds * old_e;
                        }
                    force_divr_vir = force_divr;
                    }
                if (cv)
                    {
                    const REAL h = (REAL)0.5 * force_divr_vir;
                    v[0] = h * dx[0] * dx[0];
                    v[1] = h * dx[0] * dx[1];
                    v[2] = h * dx[0] * dx[2];
                    v[3] = h * dx[1] * dx[1];
                    v[4] = h * dx[1] * dx[2];
                    v[5] = h * dx[2] * dx[2];
                    }
                f[0] = dx[0] * force_divr;
                f[1] = dx[1] * force_divr;
                f[2] = dx[2] * force_divr;
                }
            for (int d = 0; d < 3; ++d)
                {
                fi[d] += f[d];
                tqi[d] += t_i[d];
                }
            pei += pair_eng * (REAL)0.5;
            if (cv)
                for (int r = 0; r < 6; ++r)
                    vir[r] += v[r];
            if (half_list && j < N)
                {
                force[j].x -= f[0];
                force[j].y -= f[1];
                force[j].z -= f[2];
                force[j].w += pair_eng * (REAL)0.5;
                if (torque)
                    {
                    torque[j].x += t_j[0];
                    torque[j].y += t_j[1];
                    torque[j].z += t_j[2];
                    }
                if (cv)
                    for (int r = 0; r < 6; ++r)
                        virial[r * pitch + j] += v[r];
                }
            }
        force[i].x += fi[0];
        force[i].y += fi[1];
        force[i].z += fi[2];
        force[i].w += pei;
        if (torque)
            {
            torque[i].x += tqi[0];
            torque[i].y += tqi[1];
            torque[i].z += tqi[2];
            }
        if (cv)
            for (int r = 0; r < 6; ++r)
                virial[r * pitch + i] += vir[r];
        }
    return 0;
    }

/* per-particle bond table walk (Appendix A.4, GPU-table form) */
static int SUF(bond_loop)(int potential, const azb_bond_args* a, const void* params, unsigned int* flags, int nthreads)
    {
    const SUF(real4)* pos = (const SUF(real4)*)a->d_pos;
    SUF(real4)* force = (SUF(real4)*)a->d_force;
    REAL* virial = (REAL*)a->d_virial;
    const unsigned int N = a->N;
    const size_t pitch = a->virial_pitch;
    unsigned int bad = 0;
    if (potential < 0 || potential >= AZB_BOND_COUNT_)
        return AZB_ERR_BAD_POTENTIAL;
    if (nthreads < 1)
        nthreads = 1;
#pragma omp parallel for schedule(static) num_threads(nthreads) reduction(| : bad) if (nthreads > 1)
    for (unsigned int i = 0; i < N; ++i)
        {
        REAL fi[4] = {0, 0, 0, 0}, vir[6] = {0, 0, 0, 0, 0, 0};
        for (unsigned int k = 0; k < a->d_gpu_n_bonds[i]; ++k)
            {
            const azb_bond_group g = a->d_gpu_bondlist[(size_t)k * a->pitch + i];
            REAL dx[3] = {pos[i].x - pos[g.idx].x, pos[i].y - pos[g.idx].y, pos[i].z - pos[g.idx].z};
            SUF(min_image)(&a->box, &dx[0], &dx[1], &dx[2]);
            const REAL rsq = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2];
            REAL force_divr = 0, bond_eng = 0;
            int evaluated;
            if (potential == AZB_BOND_DOUBLE_WELL)
                evaluated = SUF(eval_dw)(rsq, &((const SUF(dw_params)*)params)[g.type], &force_divr, &bond_eng);
            else
                evaluated = SUF(eval_quartic)(rsq, &((const SUF(quartic_params)*)params)[g.type], &force_divr, &bond_eng);
            if (!evaluated)
                {
                bad |= 1u;
                continue;
                }
            bond_eng *= (REAL)0.5;
            const REAL h = (REAL)0.5 * force_divr;
            vir[0] += h * dx[0] * dx[0];
            vir[1] += h * dx[0] * dx[1];
            vir[2] += h * dx[0] * dx[2];
            vir[3] += h * dx[1] * dx[1];
            vir[4] += h * dx[1] * dx[2];
            vir[5] += h * dx[2] * dx[2];
            fi[0] += dx[0] * force_divr;
            fi[1] += dx[1] * force_divr;
            fi[2] += dx[2] * force_divr;
            fi[3] += bond_eng;
            }
        force[i].x = fi[0];
        force[i].y = fi[1];
        force[i].z = fi[2];
        force[i].w = fi[3];
        if (a->compute_virial)
            for (int r = 0; r < 6; ++r)
                virial[r * pitch + i] = vir[r];
        }
    if (bad && flags)
        *flags = 1;
    return 0;
    }

/* HarmonicBarrier<BarrierEvaluatorT>::computeForces (/root/reference/src/HarmonicBarrier.h:148-178): validity test of the
 * barrier location (:125-133; PlanarBarrierEvaluator.h:51-56, SphericalBarrierEvaluator.h:54-60), then per particle: wrap
 * the position into the global box by one image (z, y, x with tilt terms), evaluate PlanarBarrierEvaluator.h:36-48 or
 * SphericalBarrierEvaluator.h:36-51; the virial is zeroed (:176). */
static int SUF(barrier_loop)(int geometry, const azb_barrier_args* a)
    {
    const azb_box* b = &a->global_box;
    const REAL Lx = (REAL)b->Lx, Ly = (REAL)b->Ly, Lz = (REAL)b->Lz, xy = (REAL)b->xy, xz = (REAL)b->xz, yz = (REAL)b->yz;
    const REAL lox = -Lx / (REAL)2, loy = -Ly / (REAL)2, loz = -Lz / (REAL)2;
    const REAL hix = lox + Lx, hiy = loy + Ly, hiz = loz + Lz;
    const REAL loc = (REAL)a->location;
    const SUF(real4)* pos = (const SUF(real4)*)a->d_pos;
    const REAL* params = (const REAL*)a->d_params;
    SUF(real4)* force = (SUF(real4)*)a->d_force;
    if (geometry == AZB_BARRIER_PLANAR)
        {
        const REAL lo_y = loy + yz * loz, hi_y = hiy + yz * hiz; /* y of makeCoordinates(0,0,0) and (1,1,1) */
        if (!(loc >= lo_y && loc < hi_y))
            return AZB_ERR_BARRIER_INVALID;
        }
    else if (geometry == AZB_BARRIER_SPHERICAL)
        {
        const REAL term = xy * yz - xz;
        const REAL dx = Lx / SQRT((REAL)1 + xy * xy + term * term), dy = Ly / SQRT((REAL)1 + yz * yz), dz = Lz;
        const REAL two_R = (REAL)2 * loc;
        if (!(loc >= (REAL)0 && dx >= two_R && dy >= two_R && dz >= two_R))
            return AZB_ERR_BARRIER_INVALID;
        }
    else
        return AZB_ERR_BAD_POTENTIAL;
    for (unsigned int i = 0; i < a->N; ++i)
        {
        REAL x = pos[i].x, y = pos[i].y, z = pos[i].z;
        const unsigned int type = (unsigned int)SUF(type_of)(pos[i].w);
        const REAL k = params[2 * type], offset = params[2 * type + 1];
        if (b->periodic[2])
            {
            if (z >= hiz)
                {
                z -= Lz;
                y -= Lz * yz;
                x -= Lz * xz;
                }
            else if (z < loz)
                {
                z += Lz;
                y += Lz * yz;
                x += Lz * xz;
                }
            }
        if (b->periodic[1])
            {
            const REAL tilt_y = yz * z;
            if (y >= hiy + tilt_y)
                {
                y -= Ly;
                x -= Ly * xy;
                }
            else if (y < loy + tilt_y)
                {
                y += Ly;
                x += Ly * xy;
                }
            }
        if (b->periodic[0])
            {
            const REAL tilt_x = (xz - xy * yz) * z + xy * y;
            if (x >= hix + tilt_x)
                x -= Lx;
            else if (x < lox + tilt_x)
                x += Lx;
            }
        SUF(real4) f = {0, 0, 0, 0};
        if (geometry == AZB_BARRIER_PLANAR)
            {
            const REAL dy = y - (loc + offset);
            if (dy > (REAL)0)
                {
                f.y = -k * dy;
                f.w = (REAL)-0.5 * f.y * dy;
                }
            }
        else
            {
            const REAL r = SQRT(x * x + y * y + z * z);
            const REAL dr = r - (loc + offset);
            if (dr > (REAL)0)
                {
                const REAL k_dr = k * dr;
                const REAL c = -(k_dr / r);
                f.x = c * x;
                f.y = c * y;
                f.z = c * z;
                f.w = (REAL)0.5 * k_dr * dr;
                }
            }
        force[i] = f;
        }
    if (a->d_virial)
        memset(a->d_virial, 0, sizeof(REAL) * 6 * a->virial_pitch);
    return 0;
    }
