// azb_nlist.cu — cell-list neighbour-list builder on the GPU, emitting the HOOMD-blue v5 arrays the force kernels
// consume through pair_args_t: n_neigh (u32), head_list (size_t), nlist (u32 particle indices, FULL storage mode).
// the step immediately before the force path each time the list is rebuilt; the reference uses
// hoomd.md.nlist.Cell(buffer=...) with exclusions=('bond',) (/root/reference/src/pytest/test_pair.py:337).
//
// Pipeline (all on `stream`):
//   1. cell index per particle (locals + ghosts)                      cell_index_kernel
//   2. stable radix sort of (cell, particle index)                    cub::DeviceRadixSort (plumbing)
//   3. cell offsets (empty cells included) + cell-ordered copy of the positions    cell_bounds_kernel
//   4. count pass: one warp per local particle walks the 9 (y, z) rows of its 27 neighbouring cells — the three x-cells of
//      a row are one contiguous range of the cell-ordered copy, so the 32 lanes stay busy even at 8 particles per cell —
//      testing r_list^2(type_i, type_j) and the exclusion row                                              nlist_kernel<COUNT>
//   5. exclusive scan of the (row_align-padded) counts -> head_list    cub::DeviceScan
//   6. fill pass: same walk, warp-ballot compaction keeps rows in deterministic cell-traversal order      nlist_kernel<FILL>
#include <cub/cub.cuh>
#include <thrust/iterator/transform_iterator.h>
#include <cuda_runtime.h>

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

template<class S> struct NlistDev
    {
    const typename vec4<S>::type* pos;
    const typename vec4<S>::type* pos_sorted;
    const unsigned int* idx_sorted;
    const unsigned int* cell_of;
    const unsigned int* cell_begin;
    const S* r_listsq;
    const size_t* excl_head;
    const unsigned int* excl_list;
    unsigned int* n_neigh;
    const size_t* head_list;
    unsigned int* nlist;
    size_t capacity;
    Box<S> box;
    CellGrid grid;
    unsigned int N, ntypes;
    };

// FAST: fully periodic orthorhombic box (straight-line minimum image); SINGLE: one particle type (r_list^2 is one number);
// EXCL: an exclusion table is present
template<class S, bool FILL, bool FAST, bool SINGLE, bool EXCL> __global__ void __launch_bounds__(128) nlist_kernel(const NlistDev<S> a)
    {
    const unsigned int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned int lane = threadIdx.x & 31u;
    if (warp >= a.N)
        return;
    const unsigned int i = warp;
    const auto pi = ld4(a.pos + i);
    const unsigned int trow = type_of(pi.w) * a.ntypes;
    unsigned int c = a.cell_of[i];
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
    if (EXCL)
        {
        ex0 = a.excl_head[i];
        ex1 = a.excl_head[i + 1];
        }
    const S rl_single = SINGLE ? a.r_listsq[0] : S(0);
    unsigned int count = 0;
    unsigned int* row = FILL ? a.nlist + a.head_list[i] : nullptr;
    const size_t row_cap = FILL ? a.capacity - a.head_list[i] : 0;

    // Lane r < 9 prepares the (y, z) row r = 3 (oz + 1) + (oy + 1) of the 27-cell neighbourhood: the cells x-1, x, x+1 of a row
    // are consecutive in cell order, so a row is ONE range of the cell-ordered copy, or two where it wraps around the periodic
    // boundary (visited x-1, x, x+1: the list's row order). All cell_begin loads of a particle are in flight together.
    unsigned int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
    if (lane < 9u)
        {
        const int oz = (int)lane / 3 - 1, oy = (int)lane % 3 - 1;
        bool ok = oz >= lo_o[2] && oz <= hi_o[2] && oy >= lo_o[1] && oy <= hi_o[1];
        int z = ci[2] + oz, y = ci[1] + oy;
        if ((a.grid.periodic >> 2) & 1u)
            z += z < 0 ? a.grid.nc[2] : (z >= a.grid.nc[2] ? -a.grid.nc[2] : 0);
        else
            ok = ok && z >= 0 && z < a.grid.nc[2];
        if ((a.grid.periodic >> 1) & 1u)
            y += y < 0 ? a.grid.nc[1] : (y >= a.grid.nc[1] ? -a.grid.nc[1] : 0);
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
                    b0 = a.cell_begin[rb];
                    b1 = a.cell_begin[rb + xhi + 1];
                    }
                else if (xhi >= a.grid.nc[0])
                    {
                    a0 = a.cell_begin[rb + xlo];
                    a1 = a.cell_begin[rb + a.grid.nc[0]];
                    b0 = a.cell_begin[rb];
                    b1 = a.cell_begin[rb + xhi - a.grid.nc[0] + 1];
                    }
                else
                    {
                    a0 = a.cell_begin[rb + xlo];
                    a1 = a.cell_begin[rb + xhi + 1];
                    }
                }
            else
                {
                xlo = xlo < 0 ? 0 : xlo;
                xhi = xhi >= a.grid.nc[0] ? a.grid.nc[0] - 1 : xhi;
                a0 = a.cell_begin[rb + xlo];
                a1 = a.cell_begin[rb + xhi + 1];
                }
            }
        }
    auto visit = [&](unsigned int s0, unsigned int s1)
    {
        for (unsigned int base = s0; base < s1; base += 32)
            {
            const unsigned int s = base + lane;
            bool take = false;
            unsigned int j = 0;
            if (s < s1)
                {
                j = a.idx_sorted[s];
                const auto pj = ld4(a.pos_sorted + s);
                S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                if (FAST)
                    min_image_ortho(a.box, dx, dy, dz);
                else
                    min_image(a.box, dx, dy, dz);
                const S rl = SINGLE ? rl_single : a.r_listsq[trow + type_of(pj.w)];
                take = (j != i) && (rl > S(0)) && (dx * dx + dy * dy + dz * dz < rl);
                if (EXCL && take)
                    for (size_t e = ex0; e < ex1; ++e)
                        if (a.excl_list[e] == j)
                            take = false;
                }
            const unsigned int m = __ballot_sync(0xffffffffu, take);
            if constexpr (FILL)
                {
                const unsigned int slot = count + __popc(m & ((1u << lane) - 1u));
                if (take && slot < row_cap)
                    row[slot] = j;
                }
            count += __popc(m);
            }
    };
#pragma unroll 1
    for (int r = 0; r < 9; ++r)
        {
        visit(__shfl_sync(0xffffffffu, a0, r), __shfl_sync(0xffffffffu, a1, r));
        visit(__shfl_sync(0xffffffffu, b0, r), __shfl_sync(0xffffffffu, b1, r));
        }
    if (!FILL && lane == 0)
        a.n_neigh[i] = count;
    }

template<class S, bool FILL> static void launch_nlist_kernel(const NlistDev<S>& d, unsigned int grid, cudaStream_t stream)
    {
    const bool fast = d.box.flags == 7u, single = d.ntypes == 1, excl = d.excl_head != nullptr;
#define AZB_NL(F, T, E) nlist_kernel<S, FILL, F, T, E><<<grid, 128, 0, stream>>>(d)
    if (excl)
        {
        if (fast)
            single ? AZB_NL(true, true, true) : AZB_NL(true, false, true);
        else
            single ? AZB_NL(false, true, true) : AZB_NL(false, false, true);
        }
    else
        {
        if (fast)
            single ? AZB_NL(true, true, false) : AZB_NL(true, false, false);
        else
            single ? AZB_NL(false, true, false) : AZB_NL(false, false, false);
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

static Workspace carve(void* base, unsigned int n_max, size_t ncell, size_t scalar4_bytes, size_t cub_bytes)
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
    w.cub_tmp = take(cub_bytes);
    w.cub_bytes = cub_bytes;
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

template<class S> static int nlist_count(const azb_nlist_args& a, size_t* h_total, cudaStream_t stream)
    {
    CellGrid g;
    size_t ncell;
    int rc = make_grid(a, g, ncell);
    if (rc)
        return rc;
    const size_t cub_bytes = cub_temp_bytes(a.n_max, a.N);
    Workspace w = carve(a.d_workspace, a.n_max, ncell, sizeof(typename vec4<S>::type), cub_bytes);
    if (w.bytes > a.workspace_bytes)
        return AZB_ERR_BAD_ARGUMENT;
    typedef typename vec4<S>::type s4;
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
    NlistDev<S> d;
    d.pos = (const s4*)a.d_pos;
    d.pos_sorted = (const s4*)w.pos_sorted;
    d.idx_sorted = w.idx_sorted;
    d.cell_of = w.cell;
    d.cell_begin = w.cell_begin;
    d.r_listsq = (const S*)a.d_r_listsq;
    d.excl_head = a.d_excl_head;
    d.excl_list = a.d_excl_list;
    d.n_neigh = a.d_n_neigh;
    d.head_list = a.d_head_list;
    d.nlist = a.d_nlist;
    d.capacity = a.capacity;
    d.box = make_box<S>(a.box);
    d.grid = g;
    d.N = a.N;
    d.ntypes = a.ntypes;
    const unsigned int align = a.row_align ? a.row_align : 1;
    if (a.N)
        {
        const unsigned int grd = (a.N + 3) / 4;
        launch_nlist_kernel<S, false>(d, grd, stream);
        thrust::transform_iterator<PadTo, const unsigned int*, size_t> it(a.d_n_neigh, PadTo {align});
        size_t tmp = w.cub_bytes;
        cub::DeviceScan::ExclusiveSum(w.cub_tmp, tmp, it, a.d_head_list, (int)a.N, stream);
        }
    total_kernel<<<1, 32, 0, stream>>>(a.d_n_neigh, a.d_head_list, a.N, align, w.total);
    cudaError_t e = cudaMemcpyAsync(h_total, w.total, sizeof(size_t), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(stream);
    if (e == cudaSuccess)
        e = cudaGetLastError();
    return (int)e;
    }

template<class S> static int nlist_fill(const azb_nlist_args& a, cudaStream_t stream)
    {
    CellGrid g;
    size_t ncell;
    int rc = make_grid(a, g, ncell);
    if (rc)
        return rc;
    Workspace w = carve(a.d_workspace, a.n_max, ncell, sizeof(typename vec4<S>::type), cub_temp_bytes(a.n_max, a.N));
    if (w.bytes > a.workspace_bytes || (a.N && !a.d_nlist))
        return AZB_ERR_BAD_ARGUMENT;
    typedef typename vec4<S>::type s4;
    NlistDev<S> d;
    d.pos = (const s4*)a.d_pos;
    d.pos_sorted = (const s4*)w.pos_sorted;
    d.idx_sorted = w.idx_sorted;
    d.cell_of = w.cell;
    d.cell_begin = w.cell_begin;
    d.r_listsq = (const S*)a.d_r_listsq;
    d.excl_head = a.d_excl_head;
    d.excl_list = a.d_excl_list;
    d.n_neigh = a.d_n_neigh;
    d.head_list = a.d_head_list;
    d.nlist = a.d_nlist;
    d.capacity = a.capacity;
    d.box = make_box<S>(a.box);
    d.grid = g;
    d.N = a.N;
    d.ntypes = a.ntypes;
    if (a.N)
        launch_nlist_kernel<S, true>(d, (a.N + 3) / 4, stream);
    return (int)cudaGetLastError();
    }
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
        *bytes = azb::carve(nullptr, a->n_max, ncell, 32, azb::cub_temp_bytes(a->n_max, a->N)).bytes;
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
    }
