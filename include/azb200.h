/* azb200.h — C-ABI of the B200-native force-compute engine behind hoomd.azplugins.
 *
 * Every entry point replaces one explicit template instantiation that the reference stamps out of HOOMD's
 * kernel drivers (C++-mangled, namespace hoomd::md::kernel, returning hipError_t):
 *
 *   azb_pair_forces   <- gpu_compute_pair_forces<PairEvaluatorX>(const pair_args_t&, const param_type*)
 *                        /root/reference/src/PotentialPairGPUKernel.cu.inc:25-28
 *   azb_dpd_forces    <- gpu_compute_dpd_forces<DPDPairEvaluatorGeneralWeight>(const dpd_pair_args_t&, ...)
 *                        /root/reference/src/PotentialPairDPDThermoGPUKernel.cu.inc:21-24
 *   azb_aniso_forces  <- gpu_compute_pair_aniso_forces<AnisoPairEvaluatorTwoPatchMorse>(const a_pair_args_t&,
 *                        const param_type*, const shape_type*)
 *                        /root/reference/src/AnisoPotentialPairGPUKernel.cu.inc:21-25
 *   azb_bond_forces   <- gpu_compute_bond_forces<BondEvaluatorX, 2>(const bond_args_t<2>&, const param_type*,
 *                        unsigned int* d_flags)   /root/reference/src/PotentialBondGPUKernel.cu.inc:25-29
 *
 * Conventions (HOOMD-blue v5 array layouts):
 *  - "Scalar" is float (scalar_bits == 32) or double (scalar_bits == 64); Scalar4 is 4 packed Scalars.
 *  - d_pos[i]   = (x, y, z, type bit-cast into .w: __int_as_float / __longlong_as_double of the int type id)
 *  - d_vel[i]   = (vx, vy, vz, mass);  d_orientation[i] = quaternion (s, x, y, z)
 *  - d_force[i] = (fx, fy, fz, potential energy of particle i);  d_torque[i] = (tx, ty, tz, 0)
 *  - d_virial   = 6 rows (xx, xy, xz, yy, yz, zz) of virial_pitch Scalars
 *  - neighbour list: FULL storage mode. d_n_neigh[i] entries at d_nlist[d_head_list[i] + k]; entries are
 *    particle indices in [0, n_max) where indices >= N are ghosts (read, never written).
 *  - d_params: device array of the reference's own parameter structs (same field order, size and
 *    alignment as /root/reference/src/{Pair,AnisoPair,Bond,DPDPair}Evaluator<X>.h), indexed typei*ntypes+typej (pairs) or by bond type.
 *  - All pointers named d_* are DEVICE pointers owned by the caller for the duration of the call. Outputs
 *    are overwritten, not accumulated. Calls are asynchronous on `stream` (a cudaStream_t, NULL = default).
 *  - Return value: 0 on success, otherwise a cudaError_t value (or AZB_ERR_* below). Nothing throws.
 */
#ifndef AZB200_H_
#define AZB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AZB200_ABI_VERSION 3

#if defined(__GNUC__)
#define AZB_API __attribute__((visibility("default")))
#else
#define AZB_API
#endif

/* error codes outside the cudaError_t range */
#define AZB_ERR_BAD_POTENTIAL 10001
#define AZB_ERR_BAD_SCALAR 10002
#define AZB_ERR_BAD_ARGUMENT 10003

/* isotropic pair potentials (evaluator getName() strings in comments) */
enum azb_pair_potential
    {
    AZB_PAIR_PERTURBED_LENNARD_JONES = 0, /* "PerturbedLennardJones" */
    AZB_PAIR_COLLOID = 1,                 /* "colloid" */
    AZB_PAIR_EXPANDED_YUKAWA = 2,         /* "ExpandedYukawa" */
    AZB_PAIR_HERTZ = 3,                   /* "hertz" */
    AZB_PAIR_DPD_CONSERVATIVE = 4,        /* "dpd_gen" through evalForceAndEnergy (conservative only) */
    AZB_PAIR_COUNT_ = 5
    };
enum azb_dpd_potential
    {
    AZB_DPD_GENERAL_WEIGHT = 0, /* "dpd_gen" through evalForceEnergyThermo */
    AZB_DPD_COUNT_ = 1
    };
enum azb_aniso_potential
    {
    AZB_ANISO_TWO_PATCH_MORSE = 0, /* "TwoPatchMorse" */
    AZB_ANISO_COUNT_ = 1
    };
enum azb_bond_potential
    {
    AZB_BOND_DOUBLE_WELL = 0, /* "DoubleWell" */
    AZB_BOND_QUARTIC = 1,     /* "Quartic" */
    AZB_BOND_COUNT_ = 2
    };
enum azb_shift_mode
    {
    AZB_SHIFT_NONE = 0,
    AZB_SHIFT_SHIFT = 1,
    AZB_SHIFT_XPLOR = 2
    };
enum azb_kind
    {
    AZB_KIND_PAIR = 0,
    AZB_KIND_DPD = 1,
    AZB_KIND_ANISO = 2,
    AZB_KIND_BOND = 3
    };

/* Simulation box (HOOMD BoxDim): edge lengths, tilt factors, periodic flags. Always double here; the
 * kernels convert to Scalar (L, Linv = 1/L) exactly as a Scalar-typed BoxDim would hold them. */
typedef struct azb_box
    {
    double Lx, Ly, Lz;
    double xy, xz, yz;
    unsigned char periodic[3];
    unsigned char pad_[5];
    } azb_box;

/* Tile plan (optional, azb_tiles_build below): a derived view of one neighbour list that lets the force kernels stage
 * neighbour data in shared memory. Particles [t*tile_size, (t+1)*tile_size) form tile t; d_uniq holds, per tile, the
 * ascending list of distinct particle indices appearing in its rows; d_slots holds, entry for entry, the position of
 * each neighbour inside its tile's list. A slot row holds the same neighbours as the list row — in the same order, or
 * (group_width > 0) in bank-rotated order (csrc/azb_tiles.cu: dealt round-robin over the 8 shared-memory bank groups, so
 * that the lanes of a quarter-warp read 8 different bank groups at every step). Slot rows are laid out for the kernel that reads them: a group of W =
 * lane_width lanes walks a row in blocks of 4W entries, each lane fetching its four entries of a block with one 8-byte
 * load, and the lanes of a group touching consecutive list entries at the same time (consecutive entries are mostly
 * bank groups in sequence). So entry e = b*4W + u*W + l of particle i's (reordered) row is stored at
 * d_slots[AZB_SLOT_HEAD(head, i, W) + b*4W + 4*l + u], rows are padded with
 * slot 0 to a multiple of 4W entries and start on a multiple of max(4W, 8) entries, i.e. 16 bytes (W = 1 is the plain
 * order). This needs rows that
 * are disjoint and ordered by particle (head_list[i+1] >= head_list[i] + n_neigh[i], as HOOMD's prefix-sum head list
 * is), which azb_tiles_build verifies. The pair set is exactly the neighbour list's; only where pos[j] is read from
 * changes. The plan is valid for as long as the neighbour-list arrays it was built from are unchanged (HOOMD: until
 * the next rebuild). */
#define AZB_SLOT_ROW_ALIGN(W) ((size_t)((W) < 2 ? 8 : 4 * (W))) /* entries; rows are padded to a multiple of 4W, 16-byte aligned */
#define AZB_SLOT_HEAD(head, i, W) (((size_t)(head) + AZB_SLOT_ROW_ALIGN(W) * (size_t)(i)) & ~(AZB_SLOT_ROW_ALIGN(W) - 1))
#define AZB_SLOTS_SIZE(size_nlist, N, W) ((size_t)(size_nlist) + AZB_SLOT_ROW_ALIGN(W) * ((size_t)(N) + 1))
typedef struct azb_tile_plan
    {
    unsigned int tile_size;           /* particles per tile: 32, 64, 128 or 256 */
    unsigned int n_tiles;             /* ceil(N / tile_size) */
    unsigned int uniq_capacity;       /* stride of d_uniq per tile; power of two, <= 4096 */
    unsigned int max_uniq;            /* max over tiles of d_uniq_count (returned by azb_tiles_build) */
    unsigned int lane_width;          /* W of the slot-row layout: 1 (DPD kernels), or the pair kernel's lanes per particle */
    unsigned int max_slot_span;       /* max over tiles of the tile's slot-row range, in entries (returned by azb_tiles_build) */
    const unsigned int* d_uniq_count; /* [n_tiles] */
    const unsigned int* d_uniq;       /* [n_tiles * uniq_capacity] */
    const uint16_t* d_slots;          /* [AZB_SLOTS_SIZE(size_nlist, N, lane_width)] */
    const unsigned int* d_row_count;  /* [N] entries of each slot row, or NULL = the list's n_neigh. Shorter than n_neigh for a
                                         PRUNED plan (azb_tile_build_args.d_prune_pos): entries that cannot come inside this
                                         force's cutoff before the next list rebuild are left out of its slot rows */
    size_t n_entries;                 /* sum of the rows' lengths (0 = the list's size_nlist); launchers size their lane groups by it */
    unsigned int slot_row_stride;     /* 0: the slot row of particle i starts at AZB_SLOT_HEAD(head_list[i], i, W). > 0: at i * slot_row_stride
                                         (plans emitted by the list build write fixed-stride slot rows straight from the walk) */
    } azb_tile_plan;
#define AZB_VIEW_SLOT_STRIDE(row_stride, W) ((((size_t)(row_stride) + AZB_SLOT_ROW_ALIGN(W) - 1) / AZB_SLOT_ROW_ALIGN(W)) * AZB_SLOT_ROW_ALIGN(W))

/* mirrors HOOMD pair_args_t */
typedef struct azb_pair_args
    {
    void* d_force;                 /* Scalar4[N], overwritten */
    void* d_virial;                /* Scalar[6*virial_pitch], overwritten when compute_virial */
    size_t virial_pitch;           /* >= N */
    unsigned int N;                /* local particles */
    unsigned int n_max;            /* N + ghosts */
    const void* d_pos;             /* Scalar4[n_max]; as in HOOMD, the N local particles lie inside the box in its periodic directions
                                      (ghosts beyond it only where the box is not periodic): the packed ExpandedYukawa kernel skips
                                      the minimum-image test for particles farther than r_cut from every periodic face */
    const void* d_charge;          /* unused (needsCharge() == false for every azplugins evaluator) */
    azb_box box;
    const unsigned int* d_n_neigh; /* [N] */
    const unsigned int* d_nlist;   /* [size_nlist] */
    const size_t* d_head_list;     /* [N] */
    const void* d_rcutsq;          /* Scalar[ntypes*ntypes] */
    const void* d_ronsq;           /* Scalar[ntypes*ntypes] (xplor only; may be NULL otherwise) */
    size_t size_nlist;
    unsigned int ntypes;
    unsigned int block_size;           /* 0 = engine default */
    unsigned int shift_mode;           /* azb_shift_mode */
    unsigned int compute_virial;       /* 0/1 */
    unsigned int threads_per_particle; /* 0 = engine picks; else power of two <= 32 */
    const azb_tile_plan* tiles;        /* host pointer; NULL = gather neighbour data straight from global memory */
    unsigned int tile_begin;           /* with tiles: evaluate only tiles [tile_begin, tile_begin + tile_count) and leave the */
    unsigned int tile_count;           /* outputs of all other particles untouched; tile_count == 0 = every tile */
    const unsigned int* d_tile_order;  /* optional device permutation of the tiles: launch position p evaluates tile
                                          d_tile_order[p] (e.g. ghost-free tiles first, so they overlap the halo exchange) */
    } azb_pair_args;

/* mirrors HOOMD dpd_pair_args_t: pair args + velocities, tags, RNG key material, dt and kT */
typedef struct azb_dpd_args
    {
    azb_pair_args pair;
    const void* d_vel;         /* Scalar4[n_max] */
    const unsigned int* d_tag; /* [n_max] */
    uint16_t seed;             /* simulation seed */
    uint64_t timestep;         /* truncated to 32 bits inside, as DPDPairEvaluatorGeneralWeight.h:122 does */
    double deltaT;
    double T;                  /* kT(timestep) */
    } azb_dpd_args;

/* mirrors HOOMD a_pair_args_t */
typedef struct azb_aniso_args
    {
    azb_pair_args pair;
    const void* d_orientation; /* Scalar4[n_max] */
    void* d_torque;            /* Scalar4[N], overwritten */
    const void* d_shape_params; /* unused: AnisoPairShapeParameters is empty */
    } azb_aniso_args;

/* mirrors HOOMD bond_args_t<2>. Table element (k-th bond of particle i) lives at [k*pitch + i]. */
typedef struct azb_bond_group
    {
    unsigned int idx;  /* partner particle index (may be a ghost) */
    unsigned int type; /* bond type id */
    } azb_bond_group;

typedef struct azb_bond_args
    {
    void* d_force;
    void* d_virial;
    size_t virial_pitch;
    unsigned int N;
    unsigned int n_max;
    const void* d_pos;
    const void* d_charge; /* unused */
    azb_box box;
    const azb_bond_group* d_gpu_bondlist; /* [max_bonds*pitch] */
    const unsigned int* d_gpu_bond_pos;   /* [max_bonds*pitch] position of i inside the bond (unused by 2-body math) */
    const unsigned int* d_gpu_n_bonds;    /* [N] */
    unsigned int pitch;                   /* gpu_table_indexer width */
    unsigned int max_bonds;               /* gpu_table_indexer height */
    unsigned int n_bond_types;
    unsigned int block_size;
    unsigned int compute_virial; /* HOOMD's bond kernel always writes the virial; 0 skips the store */
    } azb_bond_args;

AZB_API int azb_abi_version(void);

/* sizeof / alignof of the reference's param struct for (kind, potential, scalar_bits); 0 if unknown */
AZB_API size_t azb_param_size(int kind, int potential, int scalar_bits);
AZB_API size_t azb_param_align(int kind, int potential, int scalar_bits);
/* evaluator getName() string, NULL if unknown */
AZB_API const char* azb_potential_name(int kind, int potential);

AZB_API int azb_pair_forces(int potential,
                    int scalar_bits,
                    const azb_pair_args* args,
                    const void* d_params,
                    void* stream);

AZB_API int azb_dpd_forces(int potential,
                   int scalar_bits,
                   const azb_dpd_args* args,
                   const void* d_params,
                   void* stream);

AZB_API int azb_aniso_forces(int potential,
                     int scalar_bits,
                     const azb_aniso_args* args,
                     const void* d_params,
                     void* stream);

/* *d_flags is set non-zero by the kernel when an evaluator rejects its parameters (DoubleWell r_0 == r_1,
 * Quartic r_0 == 0); the host class throws after the call, as HOOMD's PotentialBondGPU does. */
AZB_API int azb_bond_forces(int potential,
                    int scalar_bits,
                    const azb_bond_args* args,
                    const void* d_params,
                    unsigned int* d_flags,
                    void* stream);

/* ---- neighbour list: cell list -> HOOMD-layout n_neigh / head_list / nlist, FULL mode --------
 * Two calls, like HOOMD's NeighborListGPU (count, grow the buffer if needed, fill):
 *   azb_nlist_count: bins particles, counts neighbours within r_list(type_i,type_j), writes d_n_neigh and d_head_list
 *                    (rows padded to row_align entries) and returns the required nlist length in *h_total (this call
 *                    synchronises the stream);
 *   azb_nlist_fill : writes d_nlist[capacity >= *h_total]; must follow azb_nlist_count with the same arguments and the
 *                    same workspace contents. Row order is deterministic (cell traversal, then particle index). */
/* A tile-plan view emitted by the list build itself (azb_nlist_args.views): the slot rows of one consumer of the list. */
typedef struct azb_nlist_view
    {
    unsigned int lane_width;   /* W of the slot-row layout (1: DPD kernels; else the pair kernel's lanes per particle) */
    const void* d_keep_rsq;    /* NULL: every list entry. Else Scalar[ntypes*ntypes] of (r_cut + r_buff)^2 of THIS consumer: entries
                                  beyond it are left out of its rows (pruned view, see azb_tile_build_args.d_prune_pos) */
    uint16_t* d_slots;         /* out [N * AZB_VIEW_SLOT_STRIDE(row_stride_hint, lane_width)]: fixed-stride slot rows
                                  (azb_tile_plan.slot_row_stride), written by azb_nlist_count */
    unsigned int* d_row_count; /* out [N] row lengths of this view (required when d_keep_rsq is set, optional otherwise) */
    } azb_nlist_view;

typedef struct azb_nlist_args
    {
    const void* d_pos;               /* Scalar4[n_max] */
    unsigned int N;                  /* rows are built for the N local particles */
    unsigned int n_max;              /* candidates are all n_max particles (locals + ghosts) */
    azb_box box;                     /* orthorhombic only (tilt factors must be 0) */
    double ext_lo[3], ext_hi[3];     /* extent of the particle cloud along NON-periodic directions */
    double r_list_max;               /* max over type pairs of r_cut + r_buff: sets the cell width */
    const void* d_r_listsq;          /* Scalar[ntypes*ntypes], (r_cut + r_buff)^2; <= 0 disables the pair */
    unsigned int ntypes;
    unsigned int row_align;          /* 0/1 = tight rows */
    const size_t* d_excl_head;       /* [N+1] CSR of excluded partner indices per local particle, or NULL */
    const unsigned int* d_excl_list;
    unsigned int* d_n_neigh;         /* out [N] */
    size_t* d_head_list;             /* out [N] */
    unsigned int* d_nlist;           /* out [capacity] (fill only) */
    size_t capacity;
    void* d_workspace;               /* >= azb_nlist_workspace_size bytes, 256-byte aligned */
    size_t workspace_bytes;
    unsigned int row_stride_hint;    /* 0: two walks (count, then fill). > 0: an upper bound on the row length, e.g. 1.2 x the longest
                                        row of the previous build: azb_nlist_count then counts AND writes fixed-stride rows into the
                                        workspace in ONE walk and azb_nlist_fill only compacts them; a row that outgrows the bound
                                        makes azb_nlist_fill walk again (same result, deterministic row order either way). The
                                        workspace grows by 4 * N * row_stride_hint bytes. */
    /* Tile plans from the same walk (optional; needs row_stride_hint > 0). The staged set of a tile of tile_size consecutive
     * particles is taken from the cell structure instead of from a hash of its list entries: the cells of the tile's particles,
     * one cell further in every direction, as contiguous ranges of the cell-ordered particle array; the walk then knows the slot
     * of every entry it accepts (offset of its cell range in the tile's staging order + position in the range) at the cost of
     * one addition, and azb_nlist_fill writes the slot rows of every view while it compacts the list rows. The staged set is a
     * superset of the tile's distinct neighbours (about as large when particle order is spatially compact, e.g. HOOMD's SFC
     * order); d_uniq is NOT ascending (d_uniq_max holds each tile's largest index). A tile whose set exceeds uniq_capacity, or
     * whose cell box spans more than 64 (y, z) rows (a tile whose particles sit in two distant clumps), gets
     * d_uniq_count = 0xffffffff and is left to azb_tiles_build (d_tile_list); azb_nlist_plan_status counts them. */
    unsigned int n_views;            /* 0..4 */
    azb_nlist_view views[4];
    unsigned int tile_size;          /* 32, 64, 128 or 256 */
    unsigned int uniq_capacity;      /* power of two <= 4096 */
    unsigned int* d_uniq_count;      /* out [ceil(N / tile_size)] */
    unsigned int* d_uniq;            /* out [ceil(N / tile_size) * uniq_capacity] */
    unsigned int* d_uniq_max;        /* out [ceil(N / tile_size)] largest particle index the tile stages */
    unsigned int rows_in_place;      /* 1 (with row_stride_hint > 0, n_views == 0, and d_nlist / capacity >= N * row_stride_hint already
                                        set for azb_nlist_count): the single walk writes the rows straight into d_nlist at
                                        head_list[i] = i * row_stride_hint — a valid HOOMD-layout list (disjoint rows ordered by
                                        particle; the kernels read rows through head_list / n_neigh only) with gaps instead of tight
                                        packing: no temporary rows, no scan, no compaction pass, *h_total = N * row_stride_hint and
                                        azb_nlist_fill returns at once. A row that outgrows the bound makes azb_nlist_count fall
                                        back to the packed head_list (*h_total = packed size) and azb_nlist_fill walks again. */
    } azb_nlist_args;

typedef struct azb_nlist_plan_status
    {
    unsigned int ok;            /* 1: the views of the last azb_nlist_count / azb_nlist_fill pair were written (single walk succeeded) */
    unsigned int max_uniq;      /* largest staged set */
    unsigned int max_slot_span[4]; /* per view: azb_tile_plan.max_slot_span */
    unsigned int n_bad_tiles;   /* tiles whose cell box was too large (their uniq_count is 0xffffffff): give them to
                                   azb_tiles_build through azb_tile_build_args.d_tile_list and the views are complete */
    } azb_nlist_plan_status;

AZB_API int azb_nlist_workspace_size(const azb_nlist_args* args, size_t* bytes);
AZB_API int azb_nlist_count(int scalar_bits, const azb_nlist_args* args, size_t* h_total, void* stream);
AZB_API int azb_nlist_fill(int scalar_bits, const azb_nlist_args* args, void* stream);
/* The same list build (count + scan + fill, two walks) that allocates NOTHING: no cudaMalloc / cudaFree, no allocating library
 * routine, and no caller workspace — d_workspace / workspace_bytes / row_stride_hint / views of `args` are ignored; the scratch is a
 * file-scope `static __device__` array of AZB_STATIC_WORKSPACE_BYTES inside the library (16 MiB, held by every GPU that loads
 * the library until exit; about 330 000 particles incl. ghosts in float, 250 000 in double). The caller passes inputs and
 * outputs as for azb_nlist_fill, with d_nlist[capacity] allocated up front. Returns AZB_ERR_STATIC_SCRATCH_TOO_SMALL when the
 * system needs more scratch than the array holds, or when the list needs more than `capacity` entries (*h_total then says how many;
 * n_neigh / head_list are valid, d_nlist is untouched). Every call on a device shares the one array, which keeps the last call's
 * contents (the words that must start at zero are cleared by each call): calls that overlap on the GPU — two streams, two host
 * threads — race; serialise them. Synchronises `stream` like azb_nlist_count.
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid allocation). */
#define AZB_STATIC_WORKSPACE_BYTES (16u * 1024u * 1024u)
#define AZB_ERR_STATIC_SCRATCH_TOO_SMALL 10005
AZB_API int azb_nlist_build_static(int scalar_bits, const azb_nlist_args* args, size_t* h_total, void* stream);
/* after azb_nlist_fill (synchronises the stream): outcome of the tile-plan views requested in args */
AZB_API int azb_nlist_plan_result(int scalar_bits, const azb_nlist_args* args, azb_nlist_plan_status* status, void* stream);

/* ---- tile plan builder: one call per neighbour-list rebuild (the host class calls it when HOOMD's
 * NeighborList::hasBeenUpdated(timestep) reports a rebuild). Writes d_uniq_count / d_uniq / d_slots and returns the
 * largest per-tile unique count in *h_max_uniq and the largest per-tile slot-row range in *h_max_slot_span. *h_overflow is set non-zero when some tile has more than uniq_capacity
 * distinct neighbours (1) or the rows are not disjoint and ordered by particle (2): the plan is then unusable (retry
 * with a smaller tile_size or pass tiles = NULL). Synchronises
 * the stream. Rows may reference ghosts (indices in [N, n_max)). */
typedef struct azb_tile_build_args
    {
    const unsigned int* d_n_neigh;   /* [N] */
    const unsigned int* d_nlist;     /* [size_nlist] */
    const size_t* d_head_list;       /* [N] */
    size_t size_nlist;
    unsigned int N;
    unsigned int tile_size;
    unsigned int uniq_capacity;
    unsigned int lane_width;         /* 1, 2, 4, 8 or 16 */
    unsigned int group_width;        /* 0: slot rows keep the list's row order. G > 0: rows are re-ordered for a kernel that
                                        walks them with groups of G lanes (bank rotation, see above) */
    unsigned int* d_uniq_count;      /* out [ceil(N / tile_size)] */
    unsigned int* d_uniq;            /* out [ceil(N / tile_size) * uniq_capacity] */
    uint16_t* d_slots;               /* out [AZB_SLOTS_SIZE(size_nlist, N, lane_width)] */
    unsigned int* d_status;          /* scratch, 4 x u32 on the device */
    /* Optional pruning for a force whose cutoff is smaller than the list's (C5: DPD r_cut 1.0 on a list built for 2.0).
     * A neighbour list is rebuilt before any particle has moved r_buff / 2 (HOOMD: NeighborList::distanceCheck), so a listed pair
     * that was farther apart than r_cut(ti, tj) + r_buff WHEN THE LIST WAS BUILT cannot be inside r_cut before the next rebuild:
     * its entry is dropped from this force's slot rows (the rows keep their start and their order; d_row_count holds the new
     * lengths). The forces are bit-identical to the unpruned plan's; the walk just skips entries that could only fail the
     * cutoff test. d_prune_pos = NULL: no pruning. The plan must be built from the positions the list was built from. */
    const void* d_prune_pos;         /* Scalar4[n_max] positions at list-build time (x, y, z, type) */
    const void* d_prune_rsq;         /* Scalar[ntypes * ntypes]: (r_cut + r_buff)^2 per type pair; <= 0 drops the type pair */
    int prune_scalar_bits;           /* 32 or 64 */
    unsigned int prune_ntypes;
    azb_box prune_box;
    unsigned int* d_row_count;       /* out [N], required with d_prune_pos */
    unsigned int prune_full_set;     /* 1: prune the slot rows but stage the distinct neighbours of ALL list entries (views of one list
                                        that share d_uniq) */
    const unsigned int* d_tile_list; /* NULL: every tile. Else device array of the n_tile_list tiles to (re)build; the other tiles'
                                        rows of d_uniq_count / d_uniq / d_slots / d_row_count are left alone (the tiles the list
                                        build could not plan itself: azb_nlist_plan_status.n_bad_tiles) */
    unsigned int n_tile_list;
    unsigned int slot_row_stride;    /* 0: slot rows at AZB_SLOT_HEAD(head_list[i], i, lane_width); > 0: at i * slot_row_stride */
    const azb_tile_plan* base_plan;  /* optional, with d_prune_pos (host pointer): the UNPRUNED plan of the same list and tile_size
                                        (any lane_width, group_width 0, max_uniq <= uniq_capacity of this call). The pruned view is
                                        then derived from it — every entry is tested against its staged neighbour through the base
                                        plan's 16-bit slot, the surviving slots are ranked by a scan — instead of hashed and sorted
                                        from the list again: same plan, a fraction of the cost (several forces on one list) */
    unsigned int sort_rows;          /* 1 (group_width 0): every slot row holds its entries in ASCENDING slot order (= ascending particle
                                        index) instead of list order. The pair set is unchanged; the lanes of neighbouring particles
                                        then read neighbouring or identical staged positions at the same time, which removes most of
                                        the shared-memory bank conflicts of the tile kernels' gathers. A view derived through
                                        base_plan inherits the order of its base plan. */
    } azb_tile_build_args;

AZB_API int azb_tiles_build(const azb_tile_build_args* args, unsigned int* h_max_uniq, unsigned int* h_max_slot_span, unsigned int* h_overflow,
                            void* stream);

/* ---- library-owned tile plans: for callers whose argument struct cannot carry one -----------------------------------------------
 * HOOMD's pair_args_t / dpd_pair_args_t / a_pair_args_t have no field for a tile plan, and the kernel-driver stubs that replace
 * /root/reference/src/PotentialPairGPUKernel.cu.inc:25-28 (hoomd_glue/PotentialPairGPUKernel.cu.inc) see only that struct.
 * azb_tile_cache_get returns the plan of the list in `args` with the given slot layout (lane_width: 1 for azb_dpd_forces,
 * the lanes per particle — args->threads_per_particle — for azb_pair_forces / azb_aniso_forces), building it on first use
 * into buffers the library owns (tile_size 0 = 128; halved on overflow; *plan = NULL when no plan is possible: the direct
 * kernels then run). The plan is reused while the list arrays have the same addresses, N and size; the host class reports
 * a rebuild that rewrote the arrays in place with azb_tile_cache_invalidate(d_nlist) (HOOMD: NeighborList::hasBeenUpdated;
 * hoomd_glue/PotentialPairTiledGPU.h). d_nlist = NULL invalidates every entry. azb_tile_cache_clear frees all buffers. */
AZB_API int azb_tile_cache_get(const azb_pair_args* args, unsigned int lane_width, unsigned int tile_size, void* stream,
                               const azb_tile_plan** plan);
AZB_API void azb_tile_cache_invalidate(const void* d_nlist);
AZB_API void azb_tile_cache_clear(void);

/* ---- halo packing (SURVEY.md §8 e): d_out[e] = d_src[d_idx[e]] (+ d_shift[e].xyz when d_shift != NULL), Scalar4 rows.
 * Packs the boundary particles a neighbouring domain needs as ghosts into a contiguous send buffer; the periodic image
 * shift is applied on the sender so the receiver stores the message in place behind its local particles. */
AZB_API int azb_gather_scalar4(int scalar_bits, const void* d_src, const unsigned int* d_idx, const void* d_shift, void* d_out,
                               size_t n, void* stream);

/* ---- external.PlanarHarmonicBarrier / SphericalHarmonicBarrier -----------------------------------------
 * azb_barrier_forces <- azplugins::gpu::compute_harmonic_barrier<BarrierEvaluatorT>(d_force, d_virial, d_pos, d_params,
 *                       global_box, evaluator, N, ntypes, block_size)   (/root/reference/src/HarmonicBarrierGPU.cuh:106-133,
 *                       called from HarmonicBarrierGPU<T>::computeForces, HarmonicBarrierGPU.h:47-72)
 * One thread per particle: wrap the position into the GLOBAL box (HarmonicBarrierGPU.cuh:76-78), evaluate
 * PlanarBarrierEvaluator.h:36-48 (plane y = location + offset, pushes down) or SphericalBarrierEvaluator.h:36-51 (sphere
 * r = location + offset, pushes inward), overwrite force.xyz and energy (.w). The virial is NOT computed and is zeroed
 * (HarmonicBarrierGPU.cuh:130, HarmonicBarrier.h:176). d_params is the reference's per-type Scalar2 array {k, offset}
 * (HarmonicBarrier.h:33-34,95-99). The evaluator's own validity test (Planar: lo.y <= location < hi.y; Spherical:
 * location >= 0 and every nearest-plane distance >= 2 location; HarmonicBarrier.h:125-133) is made first and reported as
 * AZB_ERR_BARRIER_INVALID, for which the reference throws "Barrier position is invalid". */
#define AZB_ERR_BARRIER_INVALID 10004
enum azb_barrier_geometry
    {
    AZB_BARRIER_PLANAR = 0,    /* "PlanarHarmonicBarrier" */
    AZB_BARRIER_SPHERICAL = 1, /* "SphericalHarmonicBarrier" */
    AZB_BARRIER_COUNT_ = 2
    };
typedef struct azb_barrier_args
    {
    void* d_force;        /* Scalar4[N], overwritten */
    void* d_virial;       /* Scalar[6*virial_pitch], zeroed (may be NULL) */
    size_t virial_pitch;
    unsigned int N;
    unsigned int ntypes;
    const void* d_pos;    /* Scalar4[N]: x, y, z, type */
    const void* d_params; /* Scalar2[ntypes]: k, offset */
    azb_box global_box;
    double location;      /* the Variant evaluated at this timestep */
    unsigned int block_size;
    } azb_barrier_args;
AZB_API int azb_barrier_forces(int geometry, int scalar_bits, const azb_barrier_args* args, void* stream);

/* ---- the step around the force path: hoomd.md.methods.ConstantVolume without a thermostat (velocity Verlet,
 * what every reference test integrates with: test_pair.py:325-327, test_pair_dpd.py:31-32) and the neighbour-list rebuild
 * trigger. HOOMD-blue itself is not under /root/reference; the update order follows TwoStepConstantVolume (recalled):
 *   azb_nve_step_one   v += (dt/2) a;  x += dt v;  wrap x into the box by one image, counting images
 *   azb_nve_step_two   F = sum of the force arrays (Integrator::computeNetForce);  a = F / m;  v += (dt/2) a
 *                      optionally writes the net force / energy rows to d_net_force
 *   azb_displacement_check  *d_flag |= 1 if any local particle moved farther than sqrt(max_disp_sq) from d_last_pos
 *                      (NeighborList::distanceCheck: rebuild when a particle has moved r_buff / 2) */
typedef struct azb_nve_args
    {
    void* d_pos;          /* Scalar4[N] */
    void* d_vel;          /* Scalar4[N]: vx, vy, vz, mass */
    void* d_accel;        /* Scalar4[N] (xyz used; 16/32-byte rows keep the traffic vectorised) */
    int* d_image;         /* int[3N] x,y,z interleaved, or NULL */
    unsigned int N;
    azb_box box;          /* the box positions are wrapped into */
    double deltaT;
    const void* const* d_forces; /* HOST array of n_forces device pointers, Scalar4[N] each (step two) */
    unsigned int n_forces;       /* <= 8 */
    void* d_net_force;           /* Scalar4[N] or NULL (step two) */
    } azb_nve_args;
AZB_API int azb_nve_step_one(int scalar_bits, const azb_nve_args* args, void* stream);
AZB_API int azb_nve_step_two(int scalar_bits, const azb_nve_args* args, void* stream);
AZB_API int azb_displacement_check(int scalar_bits, const void* d_pos, const void* d_last_pos, unsigned int N, const azb_box* box,
                                   double max_disp_sq, unsigned int* d_flag, void* stream);

/* ---- ghost halo over peer memory (SURVEY.md §8 e; replaces HOOMD's CommunicatorGPU pack -> MPI -> unpack for ranks on one
 * NVLink/NVSwitch node). One kernel per step gathers every boundary row of every direction and field and stores it
 * straight into the neighbouring rank's particle arrays (behind its local particles, where HOOMD keeps ghosts), then
 * raises a flag in that rank's memory; the receiver's stream waits on its own flags. No staging buffer, no unpack.
 *
 *   azb_ipc_export   handle + byte offset of a device pointer inside its allocation, to hand to the other processes
 *   azb_ipc_open     map another process' allocation (cached per handle), returns the pointer at that offset
 *   azb_halo_push    rows of all routes -> destination memory; when every row is written, *d_peer_flags[i] = epoch
 *   azb_halo_wait    stream-ordered wait until d_flags[i] has reached epoch for all i (bounded: after timeout_ms it sets
 *                    *d_error = 1 and returns, it never hangs the device)
 *   azb_halo_signal  *d_peer_flags[i] = epoch, stream-ordered ("everything before this on the stream is done": the
 *                    receiver's acknowledgement that the previous ghosts have been read and may be overwritten)
 * Epochs are compared as signed differences, so the counter may wrap. */
#define AZB_IPC_HANDLE_BYTES 64
typedef struct azb_halo_route
    {
    const void* d_src;         /* Scalar4 array the rows are taken from (local) */
    void* d_dst;               /* first destination row (peer memory from azb_ipc_open, or local) */
    const unsigned int* d_idx; /* [n] rows of d_src */
    unsigned int n;
    unsigned int row_begin; /* sum of n over the routes before this one */
    double shift[3];        /* added to x, y, z: the periodic image the receiver sees */
    } azb_halo_route;

typedef struct azb_halo_push_args
    {
    const azb_halo_route* d_routes; /* device array */
    unsigned int n_routes;
    unsigned int total_rows;
    unsigned int* const* d_peer_flags; /* device array of n_flags pointers (peer memory) */
    unsigned int n_flags;
    unsigned int* d_counter; /* one u32 of scratch, zero before the first call */
    unsigned int epoch;
    } azb_halo_push_args;

AZB_API int azb_ipc_export(const void* d_ptr, unsigned char* handle, size_t* offset);
AZB_API int azb_ipc_open(const unsigned char* handle, size_t offset, void** d_ptr);
AZB_API int azb_halo_push(int scalar_bits, const azb_halo_push_args* args, void* stream);
AZB_API int azb_halo_wait(const unsigned int* d_flags, unsigned int n, unsigned int epoch, unsigned int timeout_ms, unsigned int* d_error,
                          void* stream);
AZB_API int azb_halo_signal(unsigned int* const* d_peer_flags, unsigned int n, unsigned int epoch, void* stream);

/* ---- one force evaluation from HOST buffers, pipelined inside the call ----------------------------------------------------
 * azb_host_step is what a host code that keeps its particle data in (pinned) host memory calls per timestep: "here are this
 * step's positions (+ velocities / orientations), give me the forces". HOOMD's own GPU path keeps the arrays resident
 * (GlobalArray, cf. /root/reference/src/HarmonicBarrierGPU.h:50-56); this entry is for callers on the CPU side of the
 * boundary (hoomd CPU integrators driving a GPU force, the bench's end-to-end leg). The three legs overlap:
 *   H2D   every input is copied in n_in_chunks ascending row ranges (in_edges) on stream_in;
 *   force launch c evaluates launch positions [tile_begin, tile_begin + tile_count) of the tile plan (positions map to tiles
 *         through force_args' d_tile_order, so a launch can hold exactly the tiles whose inputs are complete) as soon as
 *         input chunk launches[c].wait_chunk has arrived — the caller knows from the plan which input rows a tile reads;
 *   D2H   rows [row_lo, row_hi) of every output follow on stream_out once launch outs[k].after_launch has finished,
 *         while later launches compute.
 * force_args is the azb_pair_args / azb_dpd_args / azb_aniso_args of the whole evaluation (kind says which); its tiles must
 * be set; tile_begin / tile_count are overwritten per launch on a private copy. Several force objects over the same particles
 * and the same tiling (HOOMD: several pair forces subscribed to one neighbour list; C5 = DPD + ExpandedYukawa) are evaluated by
 * one call: `more` lists the further objects, every launch range is issued for each of them (their plans must have the same
 * tile_size and their d_tile_order the same launch positions), and `outputs` carries the force arrays of all of them.
 * Returns after every output row is in host memory; on an error every stream the call touched is synchronised first. */
typedef struct azb_host_buffer
    {
    void* d_ptr;      /* device array (a field of force_args points at it too) */
    void* h_ptr;      /* host array, same layout; pinned memory if the copies are to overlap */
    size_t row_bytes; /* bytes per particle row */
    } azb_host_buffer;

typedef struct azb_host_launch
    {
    unsigned int wait_chunk; /* index of the last input chunk this launch reads */
    unsigned int tile_begin, tile_count;
    } azb_host_launch;

typedef struct azb_host_out_range
    {
    unsigned int after_launch; /* index into launches: the last launch that writes rows of this range */
    size_t row_lo, row_hi;
    } azb_host_out_range;

typedef struct azb_host_force
    {
    int kind; /* AZB_KIND_PAIR, AZB_KIND_DPD or AZB_KIND_ANISO */
    int potential;
    const void* force_args;
    const void* d_params;
    } azb_host_force;

typedef struct azb_host_step_args
    {
    int kind; /* AZB_KIND_PAIR, AZB_KIND_DPD or AZB_KIND_ANISO */
    int potential;
    int scalar_bits;
    const void* force_args;
    const void* d_params;
    unsigned int n_inputs, n_outputs; /* <= 8 each */
    azb_host_buffer inputs[8];
    azb_host_buffer outputs[8];
    unsigned int n_in_chunks;
    const size_t* in_edges; /* host array [n_in_chunks + 1], ascending row indices */
    unsigned int n_launches;
    const azb_host_launch* launches; /* host array, in launch order */
    unsigned int n_outs;
    const azb_host_out_range* outs; /* host array, ascending after_launch */
    void* stream;     /* compute stream */
    void* stream_in;  /* H2D copies */
    void* stream_out; /* D2H copies */
    unsigned int n_more;       /* further force objects evaluated by the same launches (<= 7) */
    const azb_host_force* more; /* host array */
    unsigned int flags;        /* AZB_HOST_STEP_NO_SYNC: return once everything is queued; the caller orders later work behind
                                  `stream` / `stream_out` and synchronises them itself (a decomposed run queues its ghost exchange and
                                  the tiles that read ghosts behind this call: azplugins_b200/hostio.py) */
    } azb_host_step_args;
#define AZB_HOST_STEP_NO_SYNC 1u

/* Frees the per-thread event pool azb_host_step keeps between calls (optional; call from the thread that ran the steps). */
AZB_API void azb_host_step_release(void);

AZB_API int azb_host_step(const azb_host_step_args* args);

#ifdef __cplusplus
}
#endif
#endif /* AZB200_H_ */
