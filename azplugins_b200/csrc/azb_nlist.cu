// azb_nlist.cu — cell-list neighbour-list builder on the GPU, emitting the HOOMD-blue v5 arrays the force kernels
// consume through pair_args_t: n_neigh (u32), head_list (size_t), nlist (u32 particle indices, FULL storage mode).
// the step immediately before the force path each time the list is rebuilt; the reference uses
// hoomd.md.nlist.Cell(buffer=...) with exclusions=('bond',) (/root/reference/src/pytest/test_pair.py:337).
//
// Pipeline (all on `stream`):
//   1. cell index per particle (locals + ghosts)                      cell_index_kernel
//   2. stable radix sort of (cell, particle index)                    cub::DeviceRadixSort (plumbing)
//   3. cell offsets (empty cells included) + cell-ordered copy of the positions    cell_bounds_kernel
//   4. the walk: a group of 16 or 32 lanes per local particle, in cell order, over the 9 (y, z) rows of its 27 neighbouring
//      cells — the three x-cells of a row are one contiguous range of the cell-ordered copy, each range with its periodic
//      image shift — testing r_list^2(type_i, type_j) and the exclusion row. With a row-length bound from the previous build
//      (row_stride_hint) ONE walk counts and writes fixed-stride temporary rows       nlist_kernel<MODE_SINGLE>
//      otherwise (first build, or a row outgrew the bound) count ...                     nlist_kernel<MODE_COUNT>
//   5. exclusive scan of the (row_align-padded) counts -> head_list    cub::DeviceScan
//   6. ... and fill: same walk again, rows straight into nlist                           nlist_kernel<MODE_FILL>
//      or, after the single walk, a streaming compaction of the temporary rows            nlist_compact_kernel
//   Rows keep the deterministic cell-traversal order in every mode (ballot compaction, no atomics).
#include <cub/cub.cuh>
#include <thrust/iterator/transform_iterator.h>
#include <cuda_runtime.h>

#include <cstdlib>
#include <mutex>

#include "azb200.h"
#include "azb_launch.cuh"
#include "azb_math.cuh"

namespace azb
    {
struct CellGrid
    {
    double lo[3], inv_w[3];
    int nc[3];
    unsigned int periodic; // bits 0..2
    };

template<class S>
__global__ void cell_index_kernel(const typename vec4<S>::type* __restrict__ pos, unsigned int n_max, CellGrid g, unsigned int* __restrict__ cell,
                                  unsigned int* __restrict__ idx)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_max)
        return;
    const auto p = ld4(pos + i);
    const double c[3] = {(double)p.x, (double)p.y, (double)p.z};
    int ci[3];
#pragma unroll
    for (int d = 0; d < 3; ++d)
        {
        int v = (int)floor((c[d] - g.lo[d]) * g.inv_w[d]);
        v = v < 0 ? 0 : (v >= g.nc[d] ? g.nc[d] - 1 : v);
        ci[d] = v;
        }
    cell[i] = (unsigned int)((ci[2] * g.nc[1] + ci[1]) * g.nc[0] + ci[0]);
    idx[i] = i;
    }

// cell_begin[c] = first position in cell order whose cell is >= c (so cell c is [cell_begin[c], cell_begin[c+1]) and a run of
// cells that are consecutive in memory — the x-neighbours of a cell — is ONE contiguous range, empty cells included)
template<class S>
__global__ void cell_bounds_kernel(const unsigned int* __restrict__ cell_sorted, const unsigned int* __restrict__ idx_sorted, unsigned int n_max,
                                   const typename vec4<S>::type* __restrict__ pos, typename vec4<S>::type* __restrict__ pos_sorted,
                                   unsigned int* __restrict__ cell_begin, unsigned int ncell)
    {
    const unsigned int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_max)
        return;
    const unsigned int c = cell_sorted[s];
    const unsigned int prev = s ? cell_sorted[s - 1] + 1u : 0u;
    for (unsigned int k = prev; k <= c; ++k) // empty unless s starts a cell; then also covers the empty cells before it
        cell_begin[k] = s;
    if (s == n_max - 1)
        for (unsigned int k = c + 1u; k <= ncell; ++k)
            cell_begin[k] = n_max;
    const auto p = ld4(pos + idx_sorted[s]);
    st4(pos_sorted + s, p.x, p.y, p.z, p.w);
    }

// ---- tile plans from the cell structure (azb_nlist_args.views) ----------------------------------------------------------------
// The staged set of a tile = the cells of its particles, one cell further in every direction: (y, z) rows of cells, each row a
// contiguous range of the cell-ordered array, split in up to three segments where it crosses the periodic x boundary (cells
// reached through the low face, cells inside, cells reached through the high face). Slot of a particle = offset of its
// segment in the tile's staging order + its position in the segment.
constexpr unsigned int PLAN_ROWS_MAX = 64;
struct TileTable
    {
    int x0, y0, z0;       // first cell of the box in unwrapped coordinates (may be negative / beyond the grid along periodic axes)
    unsigned int ny, nz;  // rows of the box; 0 = unusable tile
    unsigned int total;   // staged particles
    unsigned int begin[PLAN_ROWS_MAX * 3];      // first position (cell order) of every segment
    unsigned int offset[PLAN_ROWS_MAX * 3 + 1]; // slot of the segment's first particle; [rows * 3] = total
    };

__device__ __forceinline__ int pos_mod(int a, int n)
    {
    const int r = a % n;
    return r < 0 ? r + n : r;
    }

// one warp per tile
__global__ void __launch_bounds__(128) tile_table_kernel(const unsigned int* __restrict__ cell_of, unsigned int N, unsigned int tile_size,
                                                         unsigned int n_tiles, CellGrid g, const unsigned int* __restrict__ cell_begin,
                                                         unsigned int ucap, TileTable* __restrict__ tables, unsigned int* __restrict__ uniq_count,
                                                         unsigned int* __restrict__ status)
    {
    constexpr unsigned int FULL = 0xffffffffu;
    const unsigned int T = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (T >= n_tiles)
        return;
    const unsigned int i0 = T * tile_size, i1 = min(i0 + tile_size, N);
    int cref[3];
        {
        unsigned int c = cell_of[i0];
        cref[0] = c % g.nc[0];
        c /= g.nc[0];
        cref[1] = c % g.nc[1];
        cref[2] = c / g.nc[1];
        }
    int lo[3] = {1 << 30, 1 << 30, 1 << 30}, hi[3] = {-(1 << 30), -(1 << 30), -(1 << 30)};
    for (unsigned int i = i0 + lane; i < i1; i += 32)
        {
        unsigned int c = cell_of[i];
        int cc[3];
        cc[0] = c % g.nc[0];
        c /= g.nc[0];
        cc[1] = c % g.nc[1];
        cc[2] = c / g.nc[1];
#pragma unroll
        for (int d = 0; d < 3; ++d)
            {
            int rel = cc[d] - cref[d];
            if ((g.periodic >> d) & 1u)
                {
                if (rel > g.nc[d] / 2)
                    rel -= g.nc[d];
                else if (rel < -(g.nc[d] / 2))
                    rel += g.nc[d];
                }
            lo[d] = min(lo[d], rel);
            hi[d] = max(hi[d], rel);
            }
        }
#pragma unroll
    for (int d = 0; d < 3; ++d)
        {
        lo[d] = __reduce_min_sync(FULL, lo[d]) - 1;
        hi[d] = __reduce_max_sync(FULL, hi[d]) + 1;
        if ((g.periodic >> d) & 1u)
            {
            if (hi[d] - lo[d] + 1 >= g.nc[d]) // the box wraps onto itself: the whole axis, every cell once
                {
                lo[d] = -cref[d];
                hi[d] = lo[d] + g.nc[d] - 1;
                }
            }
        else
            {
            lo[d] = max(lo[d], -cref[d]);
            hi[d] = min(hi[d], g.nc[d] - 1 - cref[d]);
            }
        }
    const unsigned int ny = (unsigned int)(hi[1] - lo[1] + 1), nz = (unsigned int)(hi[2] - lo[2] + 1);
    const unsigned int rows = ny * nz;
    bool ok = rows <= PLAN_ROWS_MAX;
    const int X0 = cref[0] + lo[0], X1 = cref[0] + hi[0];
    TileTable& tt = tables[T];
    // lane l prepares rows l and l + 32
    unsigned int b[2][3] = {{0, 0, 0}, {0, 0, 0}}, len[2][3] = {{0, 0, 0}, {0, 0, 0}};
#pragma unroll
    for (int h = 0; h < 2; ++h)
        {
        const unsigned int r = lane + 32u * h;
        if (ok && r < rows)
            {
            const int y = pos_mod(cref[1] + lo[1] + (int)(r % ny), g.nc[1]);
            const int z = pos_mod(cref[2] + lo[2] + (int)(r / ny), g.nc[2]);
            const unsigned int rb = (unsigned int)((z * g.nc[1] + y) * g.nc[0]);
            const int nx = g.nc[0];
            // low: unwrapped cells [X0, min(X1, -1)] live at +nx; main: [max(X0, 0), min(X1, nx - 1)]; high: [max(X0, nx), X1] at -nx
            const int sa[3] = {X0, max(X0, 0), max(X0, nx)}, sb[3] = {min(X1, -1), min(X1, nx - 1), X1}, sh[3] = {nx, 0, -nx};
#pragma unroll
            for (int k = 0; k < 3; ++k)
                if (sa[k] <= sb[k])
                    {
                    b[h][k] = cell_begin[rb + sa[k] + sh[k]];
                    len[h][k] = cell_begin[rb + sb[k] + sh[k] + 1] - b[h][k];
                    }
            }
        }
    // exclusive scan of the row totals: rows 0..31 over the lanes, then rows 32..63 behind them
    unsigned int base = 0, total = 0;
    unsigned int off_row[2];
#pragma unroll
    for (int h = 0; h < 2; ++h)
        {
        const unsigned int mine = len[h][0] + len[h][1] + len[h][2];
        unsigned int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
            {
            const unsigned int v = __shfl_up_sync(FULL, incl, o);
            if ((int)lane >= o)
                incl += v;
            }
        off_row[h] = base + incl - mine;
        base += __shfl_sync(FULL, incl, 31);
        }
    total = base;
    ok = ok && total <= ucap;
#pragma unroll
    for (int h = 0; h < 2; ++h)
        {
        const unsigned int r = lane + 32u * h;
        if (ok && r < rows)
            {
            unsigned int off = off_row[h];
#pragma unroll
            for (int k = 0; k < 3; ++k)
                {
                tt.begin[r * 3 + k] = b[h][k];
                tt.offset[r * 3 + k] = off;
                off += len[h][k];
                }
            }
        }
    if (lane == 0)
        {
        tt.x0 = X0;
        tt.y0 = cref[1] + lo[1];
        tt.z0 = cref[2] + lo[2];
        tt.ny = ok ? ny : 0u;
        tt.nz = ok ? nz : 0u;
        tt.total = ok ? total : 0u;
        if (ok)
            tt.offset[rows * 3] = total;
        uniq_count[T] = ok ? total : 0xffffffffu;
        if (ok)
            atomicMax(status + 0, total);
        else
            atomicAdd(status + 1, 1u); // tiles left to azb_tiles_build (azb_tile_build_args.d_tile_list)
        }
    }

// one CTA per tile: the staged set in staging order (a streaming copy of ranges of the cell-ordered index array)
__global__ void __launch_bounds__(128) tile_uniq_kernel(const TileTable* __restrict__ tables, const unsigned int* __restrict__ idx_sorted,
                                                        unsigned int ucap, unsigned int* __restrict__ uniq, unsigned int* __restrict__ uniq_max)
    {
    const TileTable& tt = tables[blockIdx.x];
    const unsigned int segs = tt.ny * tt.nz * 3u;
    const unsigned int warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    unsigned int* out = uniq + (size_t)blockIdx.x * ucap;
    unsigned int mx = 0;
    for (unsigned int f = warp; f < segs; f += 4)
        {
        const unsigned int off = tt.offset[f], len = tt.offset[f + 1] - off, b = tt.begin[f];
        for (unsigned int k = lane; k < len; k += 32)
            {
            const unsigned int j = idx_sorted[b + k];
            out[off + k] = j;
            mx = max(mx, j);
            }
        }
    __shared__ unsigned int s_max;
    if (threadIdx.x == 0)
        s_max = 0;
    __syncthreads();
    mx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0)
        atomicMax(&s_max, mx);
    __syncthreads();
    if (threadIdx.x == 0)
        uniq_max[blockIdx.x] = s_max;
    }

template<class S> struct NlistDev
    {
    const typename vec4<S>::type* pos_sorted;
    const unsigned int* idx_sorted;
    const unsigned int* cell_sorted;
    const unsigned int* cell_begin;
    const S* r_listsq;
    const size_t* excl_head;
    const unsigned int* excl_list;
    unsigned int* n_neigh;
    const size_t* head_list;
    unsigned int* nlist;
    size_t capacity;
    unsigned int* tmp_rows; // MODE_SINGLE: [N * row_stride] rows of fixed capacity, compacted afterwards
    unsigned int row_stride;
    unsigned int* overflow; // MODE_SINGLE: set when a row does not fit row_stride
    size_t* head_out;       // MODE_SINGLE with rows in place (tmp_rows = the caller's nlist): head_list[i] = i * row_stride
    // tile-plan views (MODE_SINGLE only): every view's slot rows are written straight from the walk, at a fixed stride
    unsigned int n_views, tile_size;
    const TileTable* tables;
    const S* keep_rsq[4];
    uint16_t* view_slots[4];
    unsigned int* view_rows[4];
    unsigned int view_stride[4], view_wlog[4]; // slot-row stride (entries); log2 of the lane width W of the row layout
    Box<S> box;
    CellGrid grid;
    unsigned int N, n_max, ntypes;
    };

enum
    {
    MODE_COUNT = 0, // n_neigh only (first build: row lengths unknown)
    MODE_FILL = 1,  // rows into nlist at head_list (after MODE_COUNT + scan)
    MODE_SINGLE = 2 // n_neigh AND rows, into fixed-stride temporary rows: one walk per rebuild instead of two
    };

// A group of G lanes (16 or 32) builds the row of one particle; groups run over the particles in CELL order, so neighbouring
// groups walk the same candidate ranges (L1 hits) and a group's own data is a coalesced read. The 27-cell neighbourhood is
// 9 (y, z) rows; the three x-cells of a row are ONE contiguous range of the cell-ordered copy (two where the row wraps around
// the periodic boundary). IMG: every range carries the periodic image shift of its cells, so a candidate costs three
// additions instead of a minimum-image computation — exact same arithmetic result as rint(dx / L) for every pair that can
// be inside r_list (which needs >= 3 cells along every periodic direction; otherwise the general minimum image is used).
// SINGLE_TYPE: one particle type (r_list^2 is one number); EXCL: an exclusion table is present.
template<class S, int MODE, bool IMG, bool SINGLE_TYPE, bool EXCL, int G>
__global__ void __launch_bounds__(128) nlist_kernel(const NlistDev<S> a)
    {
    constexpr unsigned int FULL = 0xffffffffu;
    const unsigned int gthread = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned int s_own = gthread / G; // position in cell order
    const unsigned int lane = threadIdx.x % G;
    const unsigned int gshift = (threadIdx.x & 31u) & ~(unsigned int)(G - 1);
    const unsigned int gmask = (G == 32 ? FULL : ((1u << G) - 1u)) << gshift;
    const unsigned int ltmask = ((1u << lane) - 1u) << gshift;
    unsigned int i = 0xffffffffu;
    if (s_own < a.n_max)
        i = a.idx_sorted[s_own];
    const bool active = i < a.N; // ghosts get no row
    typename vec4<S>::type pi = {};
    unsigned int c = 0;
    if (active)
        {
        pi = ld4(a.pos_sorted + s_own);
        c = a.cell_sorted[s_own];
        }
    const unsigned int trow = SINGLE_TYPE ? 0u : type_of(pi.w) * a.ntypes;
    const int cx = c % a.grid.nc[0];
    c /= a.grid.nc[0];
    const int cy = c % a.grid.nc[1];
    const int cz = c / a.grid.nc[1];
    const int ci[3] = {cx, cy, cz};
    int lo_o[3], hi_o[3];
#pragma unroll
    for (int d = 0; d < 3; ++d)
        {
        // a periodic dimension with fewer than 3 cells would visit the same cell twice: shrink the offset range
        const bool per = (a.grid.periodic >> d) & 1u;
        lo_o[d] = (per && a.grid.nc[d] < 3) ? 0 : -1;
        hi_o[d] = (per && a.grid.nc[d] < 2) ? 0 : 1;
        }
    size_t ex0 = 0, ex1 = 0;
    if (EXCL && active)
        {
        ex0 = a.excl_head[i];
        ex1 = a.excl_head[i + 1];
        }
    const S rl_single = SINGLE_TYPE ? a.r_listsq[0] : S(0);
    unsigned int count = 0;
    unsigned int* row = nullptr;
    size_t row_cap = 0;
    if (MODE == MODE_FILL && active)
        {
        row = a.nlist + a.head_list[i];
        row_cap = a.capacity - a.head_list[i];
        }
    if (MODE == MODE_SINGLE && active)
        {
        row = a.tmp_rows + (size_t)i * a.row_stride;
        row_cap = a.row_stride;
        }

    // Lane r < 9 prepares the (y, z) row r = 3 (oz + 1) + (oy + 1): ranges a (cells x-1.. in the list's row order) and b (the part
    // of the row on the other side of the periodic boundary), with the image shifts of the row (y, z) and of each range (x).
    unsigned int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
    S shy = S(0), shz = S(0), shxa = S(0), shxb = S(0);
    const bool plan = MODE == MODE_SINGLE && a.n_views > 0;
    unsigned int slot_a = 0, slot_b = 0; // slot of the first particle of range a / b in the staging order of my tile
    int xa_cell = 0, xb_cell = 0;        // first cell (actual x) of range a / b
    unsigned int vcount[4] = {0, 0, 0, 0}; // entries written to each view's slot row so far
    uint16_t* vrow[4] = {nullptr, nullptr, nullptr, nullptr};
    if (plan && active)
        {
#pragma unroll
        for (unsigned int v = 0; v < 4; ++v)
            if (v < a.n_views)
                vrow[v] = a.view_slots[v] + (size_t)i * a.view_stride[v];
        }
    if (lane < 9u && active)
        {
        const int oz = (int)lane / 3 - 1, oy = (int)lane % 3 - 1;
        bool ok = oz >= lo_o[2] && oz <= hi_o[2] && oy >= lo_o[1] && oy <= hi_o[1];
        int z = ci[2] + oz, y = ci[1] + oy;
        if ((a.grid.periodic >> 2) & 1u)
            {
            if (z < 0)
                {
                z += a.grid.nc[2];
                shz = a.box.Lz; // the cells sit one box length above their image next to me: d = (pi - pj) + L
                }
            else if (z >= a.grid.nc[2])
                {
                z -= a.grid.nc[2];
                shz = -a.box.Lz;
                }
            }
        else
            ok = ok && z >= 0 && z < a.grid.nc[2];
        if ((a.grid.periodic >> 1) & 1u)
            {
            if (y < 0)
                {
                y += a.grid.nc[1];
                shy = a.box.Ly;
                }
            else if (y >= a.grid.nc[1])
                {
                y -= a.grid.nc[1];
                shy = -a.box.Ly;
                }
            }
        else
            ok = ok && y >= 0 && y < a.grid.nc[1];
        if (ok)
            {
            const unsigned int rb = (unsigned int)((z * a.grid.nc[1] + y) * a.grid.nc[0]);
            int xlo = ci[0] + lo_o[0], xhi = ci[0] + hi_o[0];
            if (a.grid.periodic & 1u)
                {
                if (xlo < 0)
                    {
                    a0 = a.cell_begin[rb + xlo + a.grid.nc[0]];
                    a1 = a.cell_begin[rb + a.grid.nc[0]];
                    shxa = a.box.Lx;
                    xa_cell = xlo + a.grid.nc[0];
                    b0 = a.cell_begin[rb];
                    b1 = a.cell_begin[rb + xhi + 1];
                    xb_cell = 0;
                    }
                else if (xhi >= a.grid.nc[0])
                    {
                    a0 = a.cell_begin[rb + xlo];
                    a1 = a.cell_begin[rb + a.grid.nc[0]];
                    xa_cell = xlo;
                    b0 = a.cell_begin[rb];
                    b1 = a.cell_begin[rb + xhi - a.grid.nc[0] + 1];
                    shxb = -a.box.Lx;
                    xb_cell = 0;
                    }
                else
                    {
                    a0 = a.cell_begin[rb + xlo];
                    a1 = a.cell_begin[rb + xhi + 1];
                    xa_cell = xlo;
                    }
                }
            else
                {
                xlo = xlo < 0 ? 0 : xlo;
                xhi = xhi >= a.grid.nc[0] ? a.grid.nc[0] - 1 : xhi;
                a0 = a.cell_begin[rb + xlo];
                a1 = a.cell_begin[rb + xhi + 1];
                xa_cell = xlo;
                }
            if (plan)
                {
                // my tile's table: row of (y, z) in the tile's cell box, segment of the range's first cell
                const TileTable& tt = a.tables[i / a.tile_size];
                int ry = ci[1] + oy - tt.y0, rz = ci[2] + oz - tt.z0;
                if ((a.grid.periodic >> 1) & 1u)
                    ry = pos_mod(ry, a.grid.nc[1]);
                if ((a.grid.periodic >> 2) & 1u)
                    rz = pos_mod(rz, a.grid.nc[2]);
                const unsigned int trw = ((unsigned int)rz * tt.ny + (unsigned int)ry) * 3u;
                auto slot_of = [&](int x_cell, unsigned int first) -> unsigned int
                {
                    int xu = x_cell; // the cell's coordinate in the tile's unwrapped frame [x0, x0 + nc)
                    if (a.grid.periodic & 1u)
                        xu = tt.x0 + pos_mod(x_cell - tt.x0, a.grid.nc[0]);
                    const unsigned int sg = trw + (xu < 0 ? 0u : (xu >= a.grid.nc[0] ? 2u : 1u));
                    return tt.offset[sg] + (first - tt.begin[sg]);
                };
                if (tt.ny != 0u) // (a tile with too large a cell box gets its slots from azb_tiles_build afterwards)
                    {
                    if (a1 > a0)
                        slot_a = slot_of(xa_cell, a0);
                    if (b1 > b0)
                        slot_b = slot_of(xb_cell, b0);
                    }
                }
            }
        }
    auto visit = [&](unsigned int s0, unsigned int s1, const S sx, const S sy, const S sz, const unsigned int slot0)
    {
        // trip count uniform over the warp (two groups of 16 share it: the ballots below are warp-wide)
        unsigned int len = s1 - s0;
        if (G < 32)
            len = max(len, __shfl_xor_sync(FULL, len, 16));
        for (unsigned int off = 0; off < len; off += G)
            {
            const unsigned int s = s0 + off + lane;
            bool take = false;
            unsigned int packed = 0; // slot | view mask << 12
            if (s < s1)
                {
                const auto pj = ld4(a.pos_sorted + s);
                S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                if (IMG)
                    {
                    dx += sx;
                    dy += sy;
                    dz += sz;
                    }
                else
                    min_image(a.box, dx, dy, dz);
                const unsigned int tp = SINGLE_TYPE ? 0u : trow + type_of(pj.w);
                const S rl = SINGLE_TYPE ? rl_single : a.r_listsq[tp];
                const S rsq = dx * dx + dy * dy + dz * dz;
                take = (s != s_own) && (rl > S(0)) && (rsq < rl);
                if (plan)
                    {
                    packed = slot0 + (s - s0);
#pragma unroll
                    for (unsigned int v = 0; v < 4; ++v)
                        if (v < a.n_views && a.keep_rsq[v] != nullptr && rsq <= a.keep_rsq[v][tp] * S(1.00001))
                            packed |= 0x1000u << v; // inside this (pruned) view's radius
                    }
                }
            unsigned int j = 0;
            if ((MODE != MODE_COUNT || EXCL) && take)
                {
                j = a.idx_sorted[s];
                if (EXCL)
                    for (size_t e = ex0; e < ex1; ++e)
                        if (a.excl_list[e] == j)
                            take = false;
                }
            const unsigned int m = __ballot_sync(FULL, take);
            if (MODE != MODE_COUNT)
                {
                const unsigned int slot = count + __popc(m & ltmask);
                if (take && slot < row_cap)
                    row[slot] = j;
                }
            if (plan)
                {
#pragma unroll
                for (unsigned int v = 0; v < 4; ++v)
                    if (v < a.n_views)
                        {
                        // an unpruned view holds every accepted entry (same ballot as the list row); a pruned one its subset
                        const bool kv = take && (a.keep_rsq[v] == nullptr || ((packed >> (12u + v)) & 1u));
                        const unsigned int mv = a.keep_rsq[v] == nullptr ? m : __ballot_sync(FULL, kv);
                        const unsigned int o = vcount[v] + __popc(mv & ltmask);
                        if (kv && o < a.view_stride[v])
                            {
                            // lane-transposed block layout (include/azb200.h): entry o = b * 4W + u * W + l sits at b * 4W + 4 l + u
                            const unsigned int wl = a.view_wlog[v], r = o & ((4u << wl) - 1u);
                            vrow[v][(o - r) + 4u * (r & ((1u << wl) - 1u)) + (r >> wl)] = (uint16_t)(packed & 0xfffu);
                            }
                        vcount[v] += __popc(mv & gmask);
                        }
                }
            count += __popc(m & gmask);
            }
    };
#pragma unroll 1
    for (int r = 0; r < 9; ++r)
        {
        const int src = (int)gshift + r;
        const S sy = __shfl_sync(FULL, shy, src), sz = __shfl_sync(FULL, shz, src);
        visit(__shfl_sync(FULL, a0, src), __shfl_sync(FULL, a1, src), __shfl_sync(FULL, shxa, src), sy, sz, __shfl_sync(FULL, slot_a, src));
        visit(__shfl_sync(FULL, b0, src), __shfl_sync(FULL, b1, src), __shfl_sync(FULL, shxb, src), sy, sz, __shfl_sync(FULL, slot_b, src));
        }
    if (MODE != MODE_FILL && lane == 0 && active)
        {
        a.n_neigh[i] = count;
        if (MODE == MODE_SINGLE && count > a.row_stride)
            atomicExch(a.overflow, 1u);
        if (MODE == MODE_SINGLE && a.head_out)
            a.head_out[i] = (size_t)i * a.row_stride;
        }
    if (plan && active)
        {
#pragma unroll
        for (unsigned int v = 0; v < 4; ++v)
            if (v < a.n_views)
                {
                const unsigned int wl = a.view_wlog[v], M = 4u << wl;
                const unsigned int n = min(vcount[v], a.view_stride[v]), np = (n + M - 1u) & ~(M - 1u);
                for (unsigned int o = n + lane; o < np; o += G) // pad the row to a whole block of 4 W entries with slot 0
                    {
                    const unsigned int r = o & (M - 1u);
                    vrow[v][(o - r) + 4u * (r & ((1u << wl) - 1u)) + (r >> wl)] = 0;
                    }
                if (lane == 0 && a.view_rows[v])
                    a.view_rows[v][i] = n;
                }
        }
    }

// ---- the default walk: one LANE per particle, candidates broadcast ----------------------------------------------------------------
// nlist_kernel above spends one lane per CANDIDATE: every test costs its lane a 16-byte load, an index computation and a share of
// the ballot compaction (measured: 1030 warp instructions per particle on C5, issue-bound). Here a warp takes 32 consecutive
// particles of the cell order — almost always cells of one (y, z) row — and every lane owns ONE particle: the candidates of the
// row's nine neighbouring (y, z) rows, over the x cells [min cx - 1, max cx + 1] of the warp, are read by a broadcast load (one
// instruction per candidate and warp) and tested by all 32 lanes at once, ~18 instructions per candidate and warp. A lane only
// accepts candidates inside its own three-cell window of the range (so the accepted set and its order are exactly those of
// nlist_kernel: rows r = 0..8, low image, main range, high image, ascending position) and collects them in a 16-entry buffer in
// shared memory that it writes out 64 bytes at a time. Warps that straddle a row boundary process their rows one after the other.
// Needs the image-shift form (>= 3 cells along every periodic direction); tile-plan views stay with nlist_kernel.
constexpr unsigned int NLB_BUF = 16u; // entries buffered per lane

template<class S, int MODE, bool SINGLE_TYPE, bool EXCL, bool DIRECT>
__global__ void __launch_bounds__(128) nlist_bcast_kernel(const NlistDev<S> a)
    {
    typedef typename vec4<S>::type V4;
    constexpr unsigned int FULL = 0xffffffffu;
    extern __shared__ __align__(16) unsigned char nlb_smem[];
    unsigned int* s_buf = reinterpret_cast<unsigned int*>(nlb_smem);                  // [blockDim.x][NLB_BUF]
    S* s_rl = reinterpret_cast<S*>(s_buf + ((MODE == MODE_COUNT || DIRECT) ? 0u : blockDim.x * NLB_BUF)); // [ntypes^2] (multi-type)
    if (!SINGLE_TYPE)
        {
        for (unsigned int t = threadIdx.x; t < a.ntypes * a.ntypes; t += blockDim.x)
            s_rl[t] = a.r_listsq[t];
        __syncthreads();
        }
    const unsigned int s_own = blockIdx.x * blockDim.x + threadIdx.x; // position in cell order
    unsigned int i = 0xffffffffu;
    if (s_own < a.n_max)
        i = a.idx_sorted[s_own];
    const bool active = i < a.N; // ghosts get no row
    V4 pi = {};
    unsigned int c = 0;
    if (active)
        {
        pi = ld4(a.pos_sorted + s_own);
        c = a.cell_sorted[s_own];
        }
    const unsigned int trow = SINGLE_TYPE ? 0u : type_of(pi.w) * a.ntypes;
    const int nx = a.grid.nc[0], ny = a.grid.nc[1], nz = a.grid.nc[2];
    const int cx = (int)(c % (unsigned int)nx);
    const int cy = (int)((c / (unsigned int)nx) % (unsigned int)ny);
    const int cz = (int)(c / ((unsigned int)nx * (unsigned int)ny));
    const bool per_x = a.grid.periodic & 1u, per_y = (a.grid.periodic >> 1) & 1u, per_z = (a.grid.periodic >> 2) & 1u;
    const S rl_single = SINGLE_TYPE ? a.r_listsq[0] : S(0);

    // exclusions: the first four of my row in registers, the rest (rare) read from the table
    unsigned int ex[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
    size_t ex0 = 0, ex1 = 0;
    if (EXCL && active)
        {
        ex0 = a.excl_head[i];
        ex1 = a.excl_head[i + 1];
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (ex0 + k < ex1)
                ex[k] = a.excl_list[ex0 + k];
        }

    unsigned int* row = nullptr;
    size_t row_cap = 0;
    if (MODE == MODE_FILL && active)
        {
        row = a.nlist + a.head_list[i];
        row_cap = a.capacity - a.head_list[i];
        }
    if (MODE == MODE_SINGLE && active)
        {
        row = a.tmp_rows + (size_t)i * a.row_stride;
        row_cap = a.row_stride;
        }
    unsigned int* buf = s_buf + threadIdx.x * NLB_BUF;
    unsigned int count = 0;
    auto flush16 = [&](const unsigned int base) // buffered entries [base, base + 16) -> my row
    {
        if (base + NLB_BUF <= row_cap && ((reinterpret_cast<size_t>(row + base) & 15u) == 0u))
            {
#pragma unroll
            for (unsigned int q = 0; q < NLB_BUF / 4u; ++q)
                reinterpret_cast<uint4*>(row + base)[q] = reinterpret_cast<const uint4*>(buf)[q];
            }
        else
            for (unsigned int k = 0; k < NLB_BUF; ++k)
                if (base + k < row_cap)
                    row[base + k] = buf[k];
    };

    unsigned int todo = __ballot_sync(FULL, active);
    while (todo) // one pass per (y, z) row of cells among the warp's particles (one, except where the warp straddles a row end)
        {
        const int leader = __ffs(todo) - 1;
        const int ry = __shfl_sync(FULL, cy, leader), rz = __shfl_sync(FULL, cz, leader);
        const bool mine = active && cy == ry && cz == rz;
        todo &= ~__ballot_sync(FULL, mine);
        const int xmin = __reduce_min_sync(FULL, mine ? cx : 0x7fffffff);
        const int xmax = __reduce_max_sync(FULL, mine ? cx : -1);
#pragma unroll 1
        for (int r = 0; r < 9; ++r)
            {
            int z = rz + r / 3 - 1, y = ry + r % 3 - 1;
            S sy = S(0), sz = S(0);
            if (per_z)
                {
                if (z < 0)
                    {
                    z += nz;
                    sz = a.box.Lz; // the cells sit one box length above their image next to me: d = (pi - pj) + L
                    }
                else if (z >= nz)
                    {
                    z -= nz;
                    sz = -a.box.Lz;
                    }
                }
            else if (z < 0 || z >= nz)
                continue;
            if (per_y)
                {
                if (y < 0)
                    {
                    y += ny;
                    sy = a.box.Ly;
                    }
                else if (y >= ny)
                    {
                    y -= ny;
                    sy = -a.box.Ly;
                    }
                }
            else if (y < 0 || y >= ny)
                continue;
            const unsigned int rb = (unsigned int)((z * ny + y) * nx);
            // the warp's x range in unwrapped cell coordinates, in up to three segments: cells reached through the low face
            // (stored at + nx, shift + Lx), the cells inside the box, cells reached through the high face (- nx, - Lx)
            const int X0 = xmin - 1, X1 = xmax + 1;
#pragma unroll 1
            for (int seg = 0; seg < 3; ++seg)
                {
                int u0, u1, off;
                S sx = S(0);
                if (seg == 0)
                    {
                    u0 = X0, u1 = min(X1, -1), off = nx, sx = a.box.Lx;
                    if (!per_x)
                        continue;
                    }
                else if (seg == 1)
                    u0 = max(X0, 0), u1 = min(X1, nx - 1), off = 0;
                else
                    {
                    u0 = max(X0, nx), u1 = X1, off = -nx, sx = -a.box.Lx;
                    if (!per_x)
                        continue;
                    }
                if (u0 > u1)
                    continue;
                const unsigned int S0 = a.cell_begin[rb + u0 + off], S1 = a.cell_begin[rb + u1 + off + 1];
                if (S0 >= S1)
                    continue;
                // my own window of the range: cells [cx - 1, cx + 1] of this segment
                unsigned int w_lo = 0, w_len = 0;
                if (mine)
                    {
                    const int w0 = max(cx - 1, u0), w1 = min(cx + 1, u1);
                    if (w0 <= w1)
                        {
                        w_lo = a.cell_begin[rb + w0 + off];
                        w_len = a.cell_begin[rb + w1 + off + 1] - w_lo;
                        }
                    }
                if (!__any_sync(FULL, w_len != 0u))
                    continue;
                const S px = pi.x, py = pi.y, pz = pi.z;
                const bool shifted = (sx != S(0)) | (sy != S(0)) | (sz != S(0)); // warp-uniform
                // r^2 of one candidate: same operations, same order as nlist_kernel (identical lists); adding a zero shift is
                // the identity, so the unshifted body skips it
                auto in_list = [&](const V4& pj, const bool sh) -> bool
                {
                    S dx = px - pj.x, dy = py - pj.y, dz = pz - pj.z;
                    if (sh)
                        {
                        dx += sx;
                        dy += sy;
                        dz += sz;
                        }
                    const S rsq = dx * dx + dy * dy + dz * dz;
                    const S rl = SINGLE_TYPE ? rl_single : s_rl[trow + type_of(pj.w)];
                    return rsq < rl; // (implies rl > 0)
                };
                for (unsigned int s0 = S0; s0 < S1; s0 += 32u)
                    {
                    // 32 candidates at a time: every lane tests all of them (broadcast loads) and keeps a bit per hit; its own
                    // three-cell window and its own position are applied to the mask afterwards, not per candidate
                    const unsigned int n = min(32u, S1 - s0);
                    const unsigned int lo = max(w_lo, s0), hi = min(w_lo + w_len, s0 + n);
                    unsigned int wmask = 0;
                    if (hi > lo)
                        wmask = (hi - lo >= 32u ? FULL : ((1u << (hi - lo)) - 1u)) << (lo - s0);
                    if (s_own - s0 < 32u)
                        wmask &= ~(1u << (s_own - s0));
                    if (!__any_sync(FULL, wmask != 0u))
                        continue;
                    unsigned int mask = 0;
                    if (n == 32u)
                        {
                        if (shifted)
                            {
#pragma unroll
                            for (unsigned int u = 0; u < 32u; ++u)
                                if (in_list(ld4(a.pos_sorted + s0 + u), true))
                                    mask |= 1u << u;
                            }
                        else
                            {
#pragma unroll
                            for (unsigned int u = 0; u < 32u; ++u)
                                if (in_list(ld4(a.pos_sorted + s0 + u), false))
                                    mask |= 1u << u;
                            }
                        }
                    else
                        {
#pragma unroll 4
                        for (unsigned int u = 0; u < n; ++u)
                            if (in_list(ld4(a.pos_sorted + s0 + u), true))
                                mask |= 1u << u;
                        }
                    mask &= wmask;
                    if (MODE == MODE_COUNT && !EXCL)
                        count += __popc(mask);
                    else
                        while (mask)
                            {
                            const unsigned int b = __ffs(mask) - 1u;
                            mask &= mask - 1u;
                            const unsigned int j = __ldg(a.idx_sorted + s0 + b);
                            if (EXCL)
                                {
                                bool out = j == ex[0] || j == ex[1] || j == ex[2] || j == ex[3];
                                if (ex1 - ex0 > 4)
                                    for (size_t e = ex0 + 4; e < ex1; ++e)
                                        out = out || a.excl_list[e] == j;
                                if (out)
                                    continue;
                                }
                            if (MODE != MODE_COUNT)
                                {
                                if (DIRECT)
                                    {
                                    // straight into the row: consecutive entries of a row fill a 32-byte sector within a few
                                    // chunks, and L2 assembles the sectors
                                    if (count < row_cap)
                                        row[count] = j;
                                    }
                                else
                                    {
                                    buf[count & (NLB_BUF - 1u)] = j;
                                    if (((count + 1u) & (NLB_BUF - 1u)) == 0u)
                                        flush16(count + 1u - NLB_BUF);
                                    }
                                }
                            ++count;
                            }
                    }
                }
            }
        }
    if (!active)
        return;
    if (MODE != MODE_COUNT && !DIRECT)
        {
        const unsigned int base = count & ~(NLB_BUF - 1u);
        for (unsigned int k = 0; base + k < count; ++k)
            if (base + k < row_cap)
                row[base + k] = buf[k];
        }
    if (MODE != MODE_FILL)
        {
        a.n_neigh[i] = count;
        if (MODE == MODE_SINGLE && count > a.row_stride)
            atomicExch(a.overflow, 1u);
        if (MODE == MODE_SINGLE && a.head_out)
            a.head_out[i] = (size_t)i * a.row_stride;
        }
    }

// MODE_SINGLE, second half: fixed-stride temporary rows -> the packed HOOMD rows at head_list. A pure streaming copy; each warp
// moves four rows at a time so that their loads are in flight together (a row is only ~2 warp-wide loads long).
__global__ void __launch_bounds__(256) nlist_compact_kernel(const unsigned int* __restrict__ tmp_rows, unsigned int row_stride,
                                                            const unsigned int* __restrict__ n_neigh, const size_t* __restrict__ head_list,
                                                            unsigned int N, unsigned int* __restrict__ nlist, size_t capacity)
    {
    const unsigned int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    const unsigned int i0 = warp * 4u;
    if (i0 >= N)
        return;
    unsigned int nn[4];
    size_t head[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
        {
        const bool in = i0 + u < N;
        nn[u] = in ? n_neigh[i0 + u] : 0u;
        head[u] = in ? head_list[i0 + u] : 0;
        }
    const unsigned int longest = max(max(nn[0], nn[1]), max(nn[2], nn[3]));
    for (unsigned int k0 = 0; k0 < longest; k0 += 32)
        {
        const unsigned int k = k0 + lane;
        unsigned int v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (k < nn[u])
                v[u] = tmp_rows[(size_t)(i0 + u) * row_stride + k];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (k < nn[u] && head[u] + k < capacity)
                nlist[head[u] + k] = v[u];
        }
    }

template<class S, int MODE> static void launch_nlist_kernel(const NlistDev<S>& d, bool img, double mean_range, cudaStream_t stream)
    {
    const bool single = d.ntypes == 1, excl = d.excl_head != nullptr;
    // one lane per particle with broadcast candidates (nlist_bcast_kernel) whenever the image-shift form applies and no tile-plan
    // views ride on the walk; AZB_NLIST_LANEWISE=1 keeps the lane-per-candidate walk for A/B runs
    static const bool lanewise = getenv("AZB_NLIST_LANEWISE") != nullptr;
    const size_t rl_bytes = single ? 0 : (size_t)d.ntypes * d.ntypes * sizeof(S);
    if (img && !lanewise && d.n_views == 0 && rl_bytes <= 16 * 1024)
        {
        const unsigned int grid_b = (d.n_max + 127u) / 128u;
        // entries go straight into the rows (AZB_NLIST_BUFFERED=1: through a 16-entry buffer per lane in shared memory, A/B)
        static const bool buffered = getenv("AZB_NLIST_BUFFERED") != nullptr;
        const size_t smem_b = ((MODE == MODE_COUNT || !buffered) ? 0 : 128 * NLB_BUF * sizeof(unsigned int)) + rl_bytes;
#define AZB_NLB(T, E)                                                                      \
    do                                                                                     \
        {                                                                                  \
        if (buffered)                                                                      \
            nlist_bcast_kernel<S, MODE, T, E, false><<<grid_b, 128, smem_b, stream>>>(d);  \
        else                                                                               \
            nlist_bcast_kernel<S, MODE, T, E, true><<<grid_b, 128, smem_b, stream>>>(d);   \
        } while (0)
        if (single && excl)
            AZB_NLB(true, true);
        else if (single)
            AZB_NLB(true, false);
        else if (excl)
            AZB_NLB(false, true);
        else
            AZB_NLB(false, false);
#undef AZB_NLB
        return;
        }
    // lanes per particle: 16 when a candidate range (three cells) holds fewer than ~48 particles, so that the tail of a range
    // does not idle half a warp
    const bool narrow = mean_range < 48.0;
    const unsigned int g = narrow ? 16u : 32u;
    const unsigned int grid = (unsigned int)(((size_t)d.n_max * g + 127) / 128);
#define AZB_NL(I, T, E)                                                                   \
    do                                                                                    \
        {                                                                                 \
        if (narrow)                                                                       \
            nlist_kernel<S, MODE, I, T, E, 16><<<grid, 128, 0, stream>>>(d);              \
        else                                                                              \
            nlist_kernel<S, MODE, I, T, E, 32><<<grid, 128, 0, stream>>>(d);              \
        } while (0)
    if (excl)
        {
        if (img && single)
            AZB_NL(true, true, true);
        else if (img)
            AZB_NL(true, false, true);
        else if (single)
            AZB_NL(false, true, true);
        else
            AZB_NL(false, false, true);
        }
    else
        {
        if (img && single)
            AZB_NL(true, true, false);
        else if (img)
            AZB_NL(true, false, false);
        else if (single)
            AZB_NL(false, true, false);
        else
            AZB_NL(false, false, false);
        }
#undef AZB_NL
    }

struct PadTo
    {
    unsigned int align;
    __host__ __device__ size_t operator()(unsigned int n) const
        {
        return ((size_t)n + align - 1) / align * align;
        }
    };

__global__ void total_kernel(const unsigned int* n_neigh, const size_t* head, unsigned int N, unsigned int align, size_t* total)
    {
    if (threadIdx.x == 0 && blockIdx.x == 0)
        *total = N ? head[N - 1] + ((size_t)n_neigh[N - 1] + align - 1) / align * align : 0;
    }

static size_t align_up(size_t x)
    {
    return (x + 255) / 256 * 256;
    }

struct Workspace
    {
    unsigned int *cell, *idx, *cell_sorted, *idx_sorted, *cell_begin;
    void* pos_sorted;
    size_t* total;
    unsigned int* flags;    // [0] overflow of a fixed-stride row, [1] 1 = the single-walk rows are valid (fill only compacts)
    unsigned int* tmp_rows; // [N * row_stride] when row_stride_hint > 0
    unsigned int* plan_status; // 16 words: [0] max staged set, [1] tiles with too large a cell box, [2] views written, [4 + v] max slot span
    TileTable* tables;         // [n_tiles] when views are requested
    void* cub_tmp;
    size_t cub_bytes;
    size_t bytes;
    };

static int make_grid(const azb_nlist_args& a, CellGrid& g, size_t& ncell)
    {
    const double L[3] = {a.box.Lx, a.box.Ly, a.box.Lz};
    if (!(a.r_list_max > 0) || a.box.xy != 0 || a.box.xz != 0 || a.box.yz != 0)
        return AZB_ERR_BAD_ARGUMENT; // triclinic cell binning is not implemented
    g.periodic = 0;
    ncell = 1;
    for (int d = 0; d < 3; ++d)
        {
        double lo, ext;
        if (a.box.periodic[d])
            {
            g.periodic |= 1u << d;
            lo = -0.5 * L[d];
            ext = L[d];
            }
        else
            {
            lo = a.ext_lo[d];
            ext = a.ext_hi[d] - a.ext_lo[d];
            }
        if (!(ext > 0))
            return AZB_ERR_BAD_ARGUMENT;
        int nc = (int)floor(ext / a.r_list_max);
        nc = nc < 1 ? 1 : (nc > 1024 ? 1024 : nc);
        // keep the grid small enough that cells hold a few particles even in dilute systems
        g.nc[d] = nc;
        g.lo[d] = lo;
        g.inv_w[d] = nc / ext;
        ncell *= (size_t)nc;
        }
    return 0;
    }

// image shifts per candidate range are the minimum image only with >= 3 cells along every periodic direction
static bool image_shifts_ok(const CellGrid& g)
    {
    for (int d = 0; d < 3; ++d)
        if (((g.periodic >> d) & 1u) && g.nc[d] < 3)
            return false;
    return true;
    }

static Workspace carve(void* base, unsigned int n_max, unsigned int N, size_t ncell, size_t scalar4_bytes, size_t cub_bytes, unsigned int row_stride,
                       unsigned int n_tiles = 0)
    {
    Workspace w;
    char* p = (char*)base;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        char* r = p ? p + off : nullptr;
        off += align_up(bytes);
        return (void*)r;
    };
    const size_t n = n_max ? n_max : 1;
    w.cell = (unsigned int*)take(4 * n);
    w.idx = (unsigned int*)take(4 * n);
    w.cell_sorted = (unsigned int*)take(4 * n);
    w.idx_sorted = (unsigned int*)take(4 * n);
    w.cell_begin = (unsigned int*)take(4 * (ncell + 2));
    w.pos_sorted = take(scalar4_bytes * n);
    w.total = (size_t*)take(sizeof(size_t));
    w.flags = (unsigned int*)take(2 * sizeof(unsigned int));
    w.cub_tmp = take(cub_bytes);
    w.cub_bytes = cub_bytes;
    w.tmp_rows = row_stride ? (unsigned int*)take(4 * (size_t)row_stride * (N ? N : 1)) : nullptr;
    w.plan_status = (unsigned int*)take(16 * sizeof(unsigned int));
    w.tables = (row_stride && n_tiles) ? (TileTable*)take(sizeof(TileTable) * (size_t)n_tiles) : nullptr;
    w.bytes = off;
    return w;
    }

static size_t cub_temp_bytes(unsigned int n_max, unsigned int N)
    {
    size_t sort_bytes = 0, scan_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const unsigned int*)nullptr, (unsigned int*)nullptr, (const unsigned int*)nullptr,
                                    (unsigned int*)nullptr, (int)n_max);
    thrust::transform_iterator<PadTo, const unsigned int*, size_t> it((const unsigned int*)nullptr, PadTo {1});
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, it, (size_t*)nullptr, (int)N);
    return sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
    }

// rows written straight into the caller's nlist at a fixed stride (azb_nlist_args.rows_in_place): no temporary rows, no
// compaction, no scan. Not combined with tile-plan views.
static bool rows_in_place(const azb_nlist_args& a)
    {
    return a.rows_in_place != 0 && a.row_stride_hint > 0 && a.n_views == 0;
    }
// entries per row of the workspace's temporary rows
static unsigned int tmp_row_stride(const azb_nlist_args& a)
    {
    return rows_in_place(a) ? 0u : a.row_stride_hint;
    }

// tile-plan views ride on the single walk only
static unsigned int plan_tiles(const azb_nlist_args& a)
    {
    if (a.n_views == 0 || a.row_stride_hint == 0 || a.N == 0)
        return 0;
    return (a.N + a.tile_size - 1) / a.tile_size;
    }
static int check_views(const azb_nlist_args& a)
    {
    if (a.n_views == 0)
        return 0;
    const unsigned int ts = a.tile_size, ucap = a.uniq_capacity;
    if (a.n_views > 4 || !(ts == 32 || ts == 64 || ts == 128 || ts == 256) || ucap < 32 || ucap > 4096 || (ucap & (ucap - 1u)) || !a.d_uniq_count
        || !a.d_uniq || !a.d_uniq_max)
        return AZB_ERR_BAD_ARGUMENT;
    for (unsigned int v = 0; v < a.n_views; ++v)
        {
        const unsigned int W = a.views[v].lane_width;
        if (!(W == 1 || W == 2 || W == 4 || W == 8 || W == 16) || (a.views[v].d_keep_rsq && !a.views[v].d_row_count)
            || (a.row_stride_hint && !a.views[v].d_slots))
            return AZB_ERR_BAD_ARGUMENT;
        }
    return 0;
    }

template<class S> static NlistDev<S> make_dev(const azb_nlist_args& a, const Workspace& w, const CellGrid& g)
    {
    typedef typename vec4<S>::type s4;
    NlistDev<S> d;
    d.pos_sorted = (const s4*)w.pos_sorted;
    d.idx_sorted = w.idx_sorted;
    d.cell_sorted = w.cell_sorted;
    d.cell_begin = w.cell_begin;
    d.r_listsq = (const S*)a.d_r_listsq;
    d.excl_head = a.d_excl_head;
    d.excl_list = a.d_excl_list;
    d.n_neigh = a.d_n_neigh;
    d.head_list = a.d_head_list;
    d.nlist = a.d_nlist;
    d.capacity = a.capacity;
    d.tmp_rows = rows_in_place(a) ? a.d_nlist : w.tmp_rows;
    d.row_stride = a.row_stride_hint;
    d.overflow = w.flags;
    d.head_out = rows_in_place(a) ? a.d_head_list : nullptr;
    d.n_views = w.tables ? a.n_views : 0u;
    d.tile_size = a.tile_size ? a.tile_size : 128u;
    d.tables = w.tables;
    for (unsigned int v = 0; v < 4; ++v)
        {
        const bool on = v < d.n_views;
        d.keep_rsq[v] = on ? (const S*)a.views[v].d_keep_rsq : nullptr;
        d.view_slots[v] = on ? a.views[v].d_slots : nullptr;
        d.view_rows[v] = on ? a.views[v].d_row_count : nullptr;
        unsigned int wl = 0;
        while (on && (1u << wl) < a.views[v].lane_width)
            ++wl;
        d.view_wlog[v] = wl;
        d.view_stride[v] = on ? (unsigned int)AZB_VIEW_SLOT_STRIDE(a.row_stride_hint, a.views[v].lane_width) : 0u;
        }
    d.box = make_box<S>(a.box);
    d.grid = g;
    d.N = a.N;
    d.n_max = a.n_max;
    d.ntypes = a.ntypes;
    return d;
    }

static double mean_range_population(const azb_nlist_args& a, const CellGrid& g, size_t ncell)
    {
    const int span = g.nc[0] >= 3 ? 3 : g.nc[0];
    return (double)a.n_max / (double)ncell * span;
    }

template<class S> static int nlist_count(const azb_nlist_args& a, size_t* h_total, cudaStream_t stream)
    {
    CellGrid g;
    size_t ncell;
    int rc = make_grid(a, g, ncell);
    if (rc)
        return rc;
    const size_t cub_bytes = cub_temp_bytes(a.n_max, a.N);
    if ((rc = check_views(a)))
        return rc;
    const unsigned int n_tiles = plan_tiles(a);
    Workspace w = carve(a.d_workspace, a.n_max, a.N, ncell, sizeof(typename vec4<S>::type), cub_bytes, tmp_row_stride(a), n_tiles);
    if (w.bytes > a.workspace_bytes)
        return AZB_ERR_BAD_ARGUMENT;
    typedef typename vec4<S>::type s4;
    cudaError_t e = cudaMemsetAsync(w.flags, 0, 2 * sizeof(unsigned int), stream);
    if (e == cudaSuccess)
        e = cudaMemsetAsync(w.plan_status, 0, 16 * sizeof(unsigned int), stream);
    if (e != cudaSuccess)
        return (int)e;
    if (a.n_max)
        {
        const unsigned int blk = 256, grd = (a.n_max + blk - 1) / blk;
        cell_index_kernel<S><<<grd, blk, 0, stream>>>((const s4*)a.d_pos, a.n_max, g, w.cell, w.idx);
        int bits = 1;
        while (((size_t)1 << bits) < ncell)
            ++bits;
        size_t tmp = w.cub_bytes;
        cub::DeviceRadixSort::SortPairs(w.cub_tmp, tmp, w.cell, w.cell_sorted, w.idx, w.idx_sorted, (int)a.n_max, 0, bits, stream);
        cell_bounds_kernel<S><<<grd, blk, 0, stream>>>(w.cell_sorted, w.idx_sorted, a.n_max, (const s4*)a.d_pos, (s4*)w.pos_sorted, w.cell_begin,
                                                        (unsigned int)ncell);
        }
    if (n_tiles)
        {
        // the tiles' cell boxes and staging order first: the walk reads them to turn every accepted entry into a slot
        tile_table_kernel<<<(n_tiles * 32 + 127) / 128, 128, 0, stream>>>(w.cell, a.N, a.tile_size, n_tiles, g, w.cell_begin, a.uniq_capacity, w.tables,
                                                                          a.d_uniq_count, w.plan_status);
        tile_uniq_kernel<<<n_tiles, 128, 0, stream>>>(w.tables, w.idx_sorted, a.uniq_capacity, a.d_uniq, a.d_uniq_max);
        }
    const NlistDev<S> d = make_dev<S>(a, w, g);
    const bool img = image_shifts_ok(g);
    const double range = mean_range_population(a, g, ncell);
    const unsigned int align = a.row_align ? a.row_align : 1;
    const bool single_walk = a.row_stride_hint > 0;
    const bool in_place = rows_in_place(a);
    if (in_place && a.N)
        {
        if (!a.d_nlist || a.capacity < (size_t)a.N * a.row_stride_hint)
            return AZB_ERR_BAD_ARGUMENT;
        // one walk writes n_neigh, head_list and the rows themselves; nothing else to do unless a row outgrew the stride
        launch_nlist_kernel<S, MODE_SINGLE>(d, img, range, stream);
        unsigned int h_over = 0;
        e = cudaMemcpyAsync(&h_over, w.flags, sizeof(h_over), cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(stream);
        if (e == cudaSuccess)
            e = cudaGetLastError();
        if (e != cudaSuccess)
            return (int)e;
        if (!h_over)
            {
            *h_total = (size_t)a.N * a.row_stride_hint;
            const unsigned int done = 2u; // azb_nlist_fill has nothing left to do
            e = cudaMemcpyAsync(w.flags + 1, &done, sizeof(done), cudaMemcpyHostToDevice, stream);
            if (e == cudaSuccess)
                e = cudaStreamSynchronize(stream);
            return (int)e;
            }
        // overflow: n_neigh is right, the rows are not — packed head_list below, and azb_nlist_fill walks again
        }
    if (a.N)
        {
        if (in_place)
            ;
        else if (single_walk)
            launch_nlist_kernel<S, MODE_SINGLE>(d, img, range, stream);
        else
            launch_nlist_kernel<S, MODE_COUNT>(d, img, range, stream);
        thrust::transform_iterator<PadTo, const unsigned int*, size_t> it(a.d_n_neigh, PadTo {align});
        size_t tmp = w.cub_bytes;
        cub::DeviceScan::ExclusiveSum(w.cub_tmp, tmp, it, a.d_head_list, (int)a.N, stream);
        }
    total_kernel<<<1, 32, 0, stream>>>(a.d_n_neigh, a.d_head_list, a.N, align, w.total);
    unsigned int h_flags[2] = {0, 0};
    e = cudaMemcpyAsync(h_total, w.total, sizeof(size_t), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(h_flags, w.flags, sizeof(h_flags), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(stream);
    if (e == cudaSuccess)
        e = cudaGetLastError();
    if (e == cudaSuccess && single_walk && a.N)
        {
        // rows that all fitted their fixed stride are final: azb_nlist_fill only compacts them. Otherwise it walks again.
        const unsigned int valid = (h_flags[0] || in_place) ? 0u : 1u;
        e = cudaMemcpyAsync(w.flags + 1, &valid, sizeof(valid), cudaMemcpyHostToDevice, stream);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(stream);
        }
    return (int)e;
    }

template<class S> static int nlist_fill(const azb_nlist_args& a, cudaStream_t stream)
    {
    CellGrid g;
    size_t ncell;
    int rc = make_grid(a, g, ncell);
    if (rc)
        return rc;
    if ((rc = check_views(a)))
        return rc;
    const unsigned int n_tiles = plan_tiles(a);
    Workspace w = carve(a.d_workspace, a.n_max, a.N, ncell, sizeof(typename vec4<S>::type), cub_temp_bytes(a.n_max, a.N), tmp_row_stride(a), n_tiles);
    if (w.bytes > a.workspace_bytes || (a.N && !a.d_nlist))
        return AZB_ERR_BAD_ARGUMENT;
    if (!a.N)
        return 0;
    unsigned int valid = 0;
    if (a.row_stride_hint)
        {
        cudaError_t e = cudaMemcpyAsync(&valid, w.flags + 1, sizeof(valid), cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess)
            return (int)e;
        }
    if (valid == 2u)
        return 0; // rows in place: azb_nlist_count wrote the final list
    if (valid)
        {
        const unsigned int grd = (unsigned int)((((size_t)a.N + 3) / 4 * 32 + 255) / 256);
        nlist_compact_kernel<<<grd, 256, 0, stream>>>(w.tmp_rows, a.row_stride_hint, a.d_n_neigh, a.d_head_list, a.N, a.d_nlist, a.capacity);
        if (n_tiles)
            {
            const unsigned int one = 1u; // the views are complete (unless the table kernel flagged a tile: plan_status[1])
            cudaMemcpyAsync(w.plan_status + 2, &one, sizeof(one), cudaMemcpyHostToDevice, stream);
            }
        return (int)cudaGetLastError();
        }
    const NlistDev<S> d = make_dev<S>(a, w, g);
    launch_nlist_kernel<S, MODE_FILL>(d, image_shifts_ok(g), mean_range_population(a, g, ncell), stream);
    return (int)cudaGetLastError();
    }
// scratch of azb_nlist_build_static: held by every GPU that loads the library, until exit
static __device__ __align__(256) unsigned char g_static_workspace[AZB_STATIC_WORKSPACE_BYTES];
static void* g_static_ptr[AZB_MAX_DEVICES]; // device address of the array, per device (host code must not use the symbol itself)
static std::mutex g_static_mutex;
    } // namespace azb

extern "C"
    {
    int azb_nlist_workspace_size(const azb_nlist_args* a, size_t* bytes)
        {
        if (!a || !bytes)
            return AZB_ERR_BAD_ARGUMENT;
        azb::CellGrid g;
        size_t ncell;
        const int rc = azb::make_grid(*a, g, ncell);
        if (rc)
            return rc;
        *bytes = azb::carve(nullptr, a->n_max, a->N, ncell, 32, azb::cub_temp_bytes(a->n_max, a->N), azb::tmp_row_stride(*a), azb::plan_tiles(*a)).bytes;
        return 0;
        }

    int azb_nlist_plan_result(int scalar_bits, const azb_nlist_args* a, azb_nlist_plan_status* st, void* stream)
        {
        if (!a || !st || !a->d_workspace)
            return AZB_ERR_BAD_ARGUMENT;
        if (scalar_bits != 32 && scalar_bits != 64)
            return AZB_ERR_BAD_SCALAR;
        *st = azb_nlist_plan_status();
        const unsigned int n_tiles = azb::plan_tiles(*a);
        if (!n_tiles)
            return 0;
        azb::CellGrid g;
        size_t ncell;
        const int rc = azb::make_grid(*a, g, ncell);
        if (rc)
            return rc;
        // (the same carve as azb_nlist_count / azb_nlist_fill: the workspace layout depends on sizeof(Scalar4))
        const azb::Workspace w = azb::carve(a->d_workspace, a->n_max, a->N, ncell, scalar_bits == 32 ? 16 : 32, azb::cub_temp_bytes(a->n_max, a->N),
                                            azb::tmp_row_stride(*a), n_tiles);
        unsigned int h[16];
        cudaError_t e = cudaMemcpyAsync(h, w.plan_status, sizeof(h), cudaMemcpyDeviceToHost, (cudaStream_t)stream);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize((cudaStream_t)stream);
        if (e != cudaSuccess)
            return (int)e;
        st->ok = h[2] == 1u ? 1u : 0u;
        st->n_bad_tiles = h[1];
        st->max_uniq = h[0];
        for (unsigned int v = 0; v < 4; ++v)
            st->max_slot_span[v] = v < a->n_views ? (unsigned int)(a->tile_size * AZB_VIEW_SLOT_STRIDE(a->row_stride_hint, a->views[v].lane_width)) : 0u;
        return 0;
        }

    int azb_nlist_count(int scalar_bits, const azb_nlist_args* a, size_t* h_total, void* stream)
        {
        if (!a || !h_total || !a->d_workspace || (a->N && (!a->d_pos || !a->d_n_neigh || !a->d_head_list || !a->d_r_listsq)) || a->N > a->n_max)
            return AZB_ERR_BAD_ARGUMENT;
        if (scalar_bits == 32)
            return azb::nlist_count<float>(*a, h_total, (cudaStream_t)stream);
        if (scalar_bits == 64)
            return azb::nlist_count<double>(*a, h_total, (cudaStream_t)stream);
        return AZB_ERR_BAD_SCALAR;
        }

    int azb_nlist_fill(int scalar_bits, const azb_nlist_args* a, void* stream)
        {
        if (!a || !a->d_workspace)
            return AZB_ERR_BAD_ARGUMENT;
        if (scalar_bits == 32)
            return azb::nlist_fill<float>(*a, (cudaStream_t)stream);
        if (scalar_bits == 64)
            return azb::nlist_fill<double>(*a, (cudaStream_t)stream);
        return AZB_ERR_BAD_SCALAR;
        }

    // ---- azb_nlist_build_static: the whole list build with NO allocation anywhere (include/azb200.h). The workspace the count /
    // fill kernels use is the file-scope array below instead of caller memory. No counterpart in the reference: added
    // for callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses
    // that forbid it). In host code `g_static_workspace` is only a stand-in for the device symbol: its device address is looked
    // up once per device (cudaGetSymbolAddress, with that device current) and that pointer is what the kernels get.
    int azb_nlist_build_static(int scalar_bits, const azb_nlist_args* a, size_t* h_total, void* stream)
        {
        if (!a || !h_total)
            return AZB_ERR_BAD_ARGUMENT;
        *h_total = 0;
        if (scalar_bits != 32 && scalar_bits != 64)
            return AZB_ERR_BAD_SCALAR;
        if ((a->N && (!a->d_pos || !a->d_n_neigh || !a->d_head_list || !a->d_r_listsq || !a->d_nlist)) || a->N > a->n_max)
            return AZB_ERR_BAD_ARGUMENT;
        azb_nlist_args b = *a;
        b.row_stride_hint = 0; // the two-walk build: its scratch is O(n_max), independent of the row lengths
        b.n_views = 0;
        b.rows_in_place = 0;
        size_t need = 0;
        int rc = azb_nlist_workspace_size(&b, &need);
        if (rc)
            return rc;
        if (need > AZB_STATIC_WORKSPACE_BYTES)
            return AZB_ERR_STATIC_SCRATCH_TOO_SMALL; // n_max (or the cell grid) is beyond what the built-in scratch array holds
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess)
            return (int)e;
        if (dev < 0 || dev >= AZB_MAX_DEVICES)
            return AZB_ERR_BAD_ARGUMENT;
        void* ws = nullptr;
            {
            std::lock_guard<std::mutex> lock(azb::g_static_mutex);
            if (!azb::g_static_ptr[dev])
                {
                void* p = nullptr;
                e = cudaGetSymbolAddress(&p, azb::g_static_workspace); // device `dev` is current: its copy of the array
                if (e != cudaSuccess)
                    return (int)e;
                azb::g_static_ptr[dev] = p;
                }
            ws = azb::g_static_ptr[dev];
            }
        b.d_workspace = ws;
        b.workspace_bytes = AZB_STATIC_WORKSPACE_BYTES;
        // (azb_nlist_count clears the words of the workspace that must start at zero; everything else is overwritten before use)
        size_t total = 0;
        rc = azb_nlist_count(scalar_bits, &b, &total, stream);
        if (rc)
            return rc;
        *h_total = total;
        if (total > a->capacity)
            return AZB_ERR_STATIC_SCRATCH_TOO_SMALL; // d_nlist cannot hold the list: *h_total says what it takes
        return azb_nlist_fill(scalar_bits, &b, stream);
        }
    }
