// azb_tiles.cu — tile plan builder (include/azb200.h, azb_tiles_build).
//
// The force kernels are bound by the L1 data pipe when every list entry costs a scattered 16 B gather from global
// memory (one 128 B line per lane in the worst case). A tile of consecutive particles shares most of its neighbours,
// so the kernels stage each tile's distinct neighbours in shared memory once (coalesced runs) and then address them
// with 16-bit slots. This file derives that view from the HOOMD-layout list: one CTA per tile,
//   1. insert every list entry of the tile into a shared-memory hash set (open addressing, atomicCAS),
//   2. compact the set, bitonic-sort it ascending (runs of consecutive indices -> coalesced staging loads),
//   3. record each key's rank in the table, and 4. rewrite every list entry as its rank (u16).
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdlib>

#include "azb200.h"
#include "azb_launch.cuh"
#include "azb_math.cuh"

namespace azb
    {
constexpr unsigned int EMPTY_KEY = 0xffffffffu;

// Pruning test of one list entry (azb_tile_build_args.d_prune_pos): was the pair within r_cut(ti, tj) + r_buff when the list was
// built? A hair of slack keeps the test conservative against the rounding of the cutoff test the force kernel will make.
template<class S> struct PruneDev
    {
    const typename vec4<S>::type* pos;
    const S* keep_rsq;
    unsigned int ntypes;
    Box<S> box;
    };
struct NoPrune
    {
    };
template<class S> __device__ __forceinline__ bool keep_entry(const PruneDev<S>& p, const typename vec4<S>::type& pi, unsigned int trow, unsigned int j)
    {
    const auto pj = ld4(p.pos + j);
    S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
    min_image(p.box, dx, dy, dz);
    const S lim = p.keep_rsq[trow + type_of(pj.w)];
    return dx * dx + dy * dy + dz * dz <= lim * S(1.00001);
    }

__device__ __forceinline__ unsigned int tile_hash(unsigned int j, unsigned int shift)
    {
    return (j * 2654435761u) >> shift;
    }

struct NoOwn
    {
    };
__device__ __forceinline__ NoOwn load_own(const NoPrune&, unsigned int)
    {
    return NoOwn();
    }
__device__ __forceinline__ bool kept(const NoPrune&, const NoOwn&, unsigned int)
    {
    return true;
    }
template<class S> struct Own
    {
    typename vec4<S>::type pi;
    unsigned int trow;
    };
template<class S> __device__ __forceinline__ Own<S> load_own(const PruneDev<S>& p, unsigned int i)
    {
    Own<S> o;
    o.pi = ld4(p.pos + i);
    o.trow = type_of(o.pi.w) * p.ntypes;
    return o;
    }
template<class S> __device__ __forceinline__ bool kept(const PruneDev<S>& p, const Own<S>& o, unsigned int j)
    {
    return keep_entry(p, o.pi, o.trow, j);
    }

// dynamic shared memory: keys[hcap] u32 | list[ucap] u32 | vals[hcap] u16
template<class Prune>
__global__ void __launch_bounds__(256) tiles_build_kernel(const Prune prune, const bool full_set, const unsigned int* __restrict__ tile_list,
                                                          const unsigned int slot_stride, unsigned int* __restrict__ row_count,
                                                          const unsigned int* __restrict__ n_neigh,
                                                          const unsigned int* __restrict__ nlist,
                                                          const size_t* __restrict__ head_list,
                                                          const unsigned int N,
                                                          const unsigned int tile_size,
                                                          const unsigned int ucap,
                                                          const unsigned int hcap_log2,
                                                          const unsigned int W,
                                                          const unsigned int G,
                                                          unsigned int* __restrict__ uniq_count,
                                                          unsigned int* __restrict__ uniq,
                                                          uint16_t* __restrict__ slots,
                                                          unsigned int* __restrict__ status)
    {
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int hcap = 1u << hcap_log2;
    const unsigned int hmask = hcap - 1u;
    const unsigned int hshift = 32u - hcap_log2;
    unsigned int* keys = reinterpret_cast<unsigned int*>(smem_raw);
    unsigned int* list = keys + hcap;
    uint16_t* vals = reinterpret_cast<uint16_t*>(list + ucap);
    __shared__ unsigned int s_count, s_fill;

    const unsigned int tile = tile_list ? tile_list[blockIdx.x] : blockIdx.x;
    const unsigned int first = tile * tile_size;
    const unsigned int last = min(first + tile_size, N);
    const unsigned int warp = threadIdx.x >> 5, lane = threadIdx.x & 31u, nwarps = blockDim.x >> 5;

    for (unsigned int h = threadIdx.x; h < hcap; h += blockDim.x)
        keys[h] = EMPTY_KEY;
    if (threadIdx.x == 0)
        {
        s_count = 0;
        s_fill = 0;
        }
    __syncthreads();

    // 1. hash-set insert (a warp walks one row at a time: coalesced index reads)
    for (unsigned int i = first + warp; i < last; i += nwarps)
        {
        const unsigned int nn = n_neigh[i];
        const unsigned int* row = nlist + head_list[i];
        decltype(load_own(prune, i)) own = load_own(prune, i);
        for (unsigned int k = lane; k < nn; k += 32)
            {
            const unsigned int j = row[k];
            if (!full_set && !kept(prune, own, j))
                continue;
            unsigned int h = tile_hash(j, hshift);
            while (*(volatile unsigned int*)&s_count <= ucap)
                {
                // most entries are repeats (every neighbour is shared by 6-12 rows of the tile): a plain read finds them
                // without an atomic; only an empty cell is claimed with a CAS
                unsigned int old = *(volatile unsigned int*)&keys[h];
                if (old == EMPTY_KEY)
                    {
                    old = atomicCAS(&keys[h], EMPTY_KEY, j);
                    if (old == EMPTY_KEY)
                        {
                        atomicAdd(&s_count, 1u);
                        break;
                        }
                    }
                if (old == j)
                    break;
                h = (h + 1u) & hmask;
                }
            }
        }
    __syncthreads();
    const unsigned int U = s_count;
    if (U > ucap)
        {
        if (threadIdx.x == 0)
            {
            uniq_count[tile] = EMPTY_KEY;
            atomicMax(status + 0, U);
            atomicExch(status + 1, 1u);
            }
        return;
        }

    // 2. compact + sort ascending
    for (unsigned int h = threadIdx.x; h < hcap; h += blockDim.x)
        {
        const unsigned int key = keys[h];
        if (key != EMPTY_KEY)
            list[atomicAdd(&s_fill, 1u)] = key;
        }
    unsigned int P = 1;
    while (P < U)
        P <<= 1;
    __syncthreads();
    for (unsigned int s = U + threadIdx.x; s < P; s += blockDim.x)
        list[s] = EMPTY_KEY; // P <= ucap because ucap is a power of two
    __syncthreads();
    for (unsigned int size = 2; size <= P; size <<= 1)
        for (unsigned int stride = size >> 1; stride > 0; stride >>= 1)
            {
            for (unsigned int t = threadIdx.x; t < (P >> 1); t += blockDim.x)
                {
                const unsigned int lo = 2 * t - (t & (stride - 1u));
                const unsigned int hi = lo + stride;
                const bool up = (lo & size) == 0;
                const unsigned int a = list[lo], b = list[hi];
                if ((a > b) == up)
                    {
                    list[lo] = b;
                    list[hi] = a;
                    }
                }
            __syncthreads();
            }

    // 3. publish the list; remember each key's rank next to its table position
    unsigned int* out = uniq + (size_t)tile * ucap;
    for (unsigned int s = threadIdx.x; s < U; s += blockDim.x)
        {
        const unsigned int key = list[s];
        out[s] = key;
        unsigned int h = tile_hash(key, hshift);
        while (keys[h] != key)
            h = (h + 1u) & hmask;
        vals[h] = (uint16_t)s;
        }
    if (threadIdx.x == 0)
        {
        uniq_count[tile] = U;
        atomicMax(status + 0, U);
        }
    __syncthreads();

    // 4. list entry -> slot, stored in the lane-transposed, padded row layout (include/azb200.h, AZB_SLOT_HEAD).
    // With group_width = 0 a slot row keeps the order of the list row. Otherwise (opt-in: it halves the shared-memory bank
    // conflicts of the force kernels' neighbour reads, which round 1's kernels are not bound by, and doubles this step)
    // the entries of a row are emitted in BANK-ROTATED order: bucket the row's slots by shared-memory bank group
    // (slot mod 8: the 16-byte Scalar4 of slot s lives in bank group s mod 8) and deal the buckets round-robin, starting
    // at bucket (particle * G) mod 8 for a kernel that walks rows with groups of G lanes. Emitted position e then holds
    // bank group (start + e) mod 8 while no bucket has run dry, so the 8 lanes of a quarter-warp — 8/G particles x G
    // lanes, lane (q, l) reading position t*G + l at step t — touch 8 different bank groups: the neighbour reads of the
    // force kernels become conflict-free except in the tail of a row. Deterministic (ranks come from ballots, not
    // atomics), and only the order inside a row changes: the set of pairs is the list's.
    const unsigned int M = 4u * W;
    for (unsigned int i = first + warp; i < last; i += nwarps)
        {
        const unsigned int nn = n_neigh[i];
        const size_t head = head_list[i];
        const size_t shead = slot_stride ? (size_t)i * slot_stride : AZB_SLOT_HEAD(head, i, W);
        const unsigned int nnp = (nn + M - 1u) / M * M;
        if (G == 0u)
            {
            // plain order: slot row = list row (the kept entries of it, in list order, when pruning)
            decltype(load_own(prune, i)) own = load_own(prune, i);
            unsigned int out_n = 0;
            for (unsigned int e0 = 0; e0 < nn; e0 += 32)
                {
                const unsigned int e = e0 + lane;
                unsigned int j = 0;
                bool keep = false;
                if (e < nn)
                    {
                    j = nlist[head + e];
                    keep = kept(prune, own, j);
                    }
                const unsigned int m = __ballot_sync(0xffffffffu, keep);
                if (keep)
                    {
                    unsigned int h = tile_hash(j, hshift);
                    while (keys[h] != j)
                        h = (h + 1u) & hmask;
                    const unsigned int o = out_n + __popc(m & ((1u << lane) - 1u));
                    const unsigned int r = o % M;
                    slots[shead + (o - r) + 4u * (r % W) + r / W] = vals[h];
                    }
                out_n += __popc(m);
                }
            const unsigned int out_p = (out_n + M - 1u) / M * M;
            for (unsigned int o = out_n + lane; o < out_p; o += 32) // padding
                {
                const unsigned int r = o % M;
                slots[shead + (o - r) + 4u * (r % W) + r / W] = 0;
                }
            if (row_count && lane == 0)
                row_count[i] = out_n;
            }
        else
            {
        // pass 1: bucket sizes
        unsigned int cnt[8];
#pragma unroll
        for (int g = 0; g < 8; ++g)
            cnt[g] = 0;
        for (unsigned int e0 = 0; e0 < nn; e0 += 32)
            {
            const unsigned int e = e0 + lane;
            unsigned int v = 0xffffffffu;
            if (e < nn)
                {
                const unsigned int j = nlist[head + e];
                unsigned int h = tile_hash(j, hshift);
                while (keys[h] != j)
                    h = (h + 1u) & hmask;
                v = vals[h];
                }
#pragma unroll
            for (int g = 0; g < 8; ++g)
                cnt[g] += __popc(__ballot_sync(0xffffffffu, v != 0xffffffffu && (v & 7u) == (unsigned int)g));
            }
        // pass 2: rank inside the bucket -> position in the round-robin deal -> position in the lane-transposed row
        const unsigned int start = (i * G) & 7u;
        unsigned int base[8];
#pragma unroll
        for (int g = 0; g < 8; ++g)
            base[g] = 0;
        for (unsigned int e0 = 0; e0 < nn; e0 += 32)
            {
            const unsigned int e = e0 + lane;
            unsigned int v = 0xffffffffu;
            if (e < nn)
                {
                const unsigned int j = nlist[head + e];
                unsigned int h = tile_hash(j, hshift);
                while (keys[h] != j)
                    h = (h + 1u) & hmask;
                v = vals[h];
                }
            unsigned int rank = 0;
#pragma unroll
            for (int g = 0; g < 8; ++g)
                {
                const unsigned int mask = __ballot_sync(0xffffffffu, v != 0xffffffffu && (v & 7u) == (unsigned int)g);
                if (v != 0xffffffffu && (v & 7u) == (unsigned int)g)
                    rank = base[g] + __popc(mask & ((1u << lane) - 1u));
                base[g] += __popc(mask);
                }
            if (v != 0xffffffffu)
                {
                // round r = rank deals one entry from every bucket that still has one, in rotated bucket order
                const unsigned int mine = ((v & 7u) - start) & 7u; // my bucket's place in the rotation
                unsigned int pos = 0;
#pragma unroll
                for (int g = 0; g < 8; ++g)
                    {
                    const unsigned int place = ((unsigned int)g - start) & 7u;
                    pos += min(cnt[g], rank) + ((place < mine && cnt[g] > rank) ? 1u : 0u);
                    }
                const unsigned int r = pos % M;
                slots[shead + (pos - r) + 4u * (r % W) + r / W] = (uint16_t)v;
                }
            }
        for (unsigned int e = nn + lane; e < nnp; e += 32) // padding
            {
            const unsigned int r = e % M;
            slots[shead + (e - r) + 4u * (r % W) + r / W] = 0;
            }
            }
        // rows must be disjoint and ordered by particle for the padded layout to be collision-free
        if (lane == 0 && i + 1 < N && head_list[i + 1] < head + nn)
            atomicExch(status + 1, 2u);
        if (lane == 0 && i + 1 == last) // extent of the tile's slot rows (kernels that stage them size their buffer by it)
            atomicMax(status + 2, (unsigned int)(shead + nnp - (slot_stride ? (size_t)first * slot_stride : AZB_SLOT_HEAD(head_list[first], first, W))));
        }
    }

// ---- the default builder (group_width == 0): same four steps, organised for memory-level parallelism -------------------------
// tiles_build_kernel above walks a tile row by row, one warp per row: n_neigh -> head_list -> row entries (-> pos[j] when
// pruning) is a chain of dependent global loads per row, twice per build, and a warp has one such chain in flight — measured
// 6.9 ms for the 4 M LJ list (545 M entries), ~15x what streaming the list costs. Here the row headers of the tile go to
// shared memory first (one coalesced read), an exclusive scan turns the tile's rows into ONE flat entry range [0, E), and
// every warp streams a contiguous chunk of it with four independent 32-entry loads in flight (eight when pruning: the
// pos[j] gathers of four blocks are issued together). A pruning build keeps its per-entry decision in a shared bitmask
// (E <= 65536 entries per tile), so the slot pass neither gathers pos[j] again nor recomputes a distance; the rank of a
// kept entry in its slot row is a popcount over the row's bit range. Output is identical to tiles_build_kernel's.
constexpr unsigned int KEEP_WORDS = 2048u;

template<class P> struct PruneTraits
    {
    static constexpr bool on = false;
    typedef float4 V4;
    };
template<class S> struct PruneTraits<PruneDev<S>>
    {
    static constexpr bool on = true;
    typedef typename vec4<S>::type V4;
    };
__device__ __forceinline__ bool kept_staged(const NoPrune&, const float4&, const float4&)
    {
    return true;
    }
template<class S>
__device__ __forceinline__ bool kept_staged(const PruneDev<S>& p, const typename vec4<S>::type& pi, const typename vec4<S>::type& pj)
    {
    S dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
    min_image(p.box, dx, dy, dz);
    const S lim = p.keep_rsq[type_of(pi.w) * p.ntypes + type_of(pj.w)];
    return dx * dx + dy * dy + dz * dz <= lim * S(1.00001); // same expression as keep_entry()
    }
template<class S> __device__ __forceinline__ typename vec4<S>::type load_pos(const PruneDev<S>& p, unsigned int j)
    {
    return ld4(p.pos + j);
    }
__device__ __forceinline__ float4 load_pos(const NoPrune&, unsigned int)
    {
    return make_float4(0.f, 0.f, 0.f, 0.f);
    }
// number of set bits in [b0, b1) of a bitmask
__device__ __forceinline__ unsigned int count_bits(const unsigned int* m, unsigned int b0, unsigned int b1)
    {
    if (b1 <= b0)
        return 0u;
    const unsigned int w0 = b0 >> 5, w1 = (b1 - 1u) >> 5;
    const unsigned int lo = 0xffffffffu << (b0 & 31u), hi = 0xffffffffu >> (31u - ((b1 - 1u) & 31u));
    if (w0 == w1)
        return __popc(m[w0] & lo & hi);
    unsigned int c = __popc(m[w0] & lo);
    for (unsigned int w = w0 + 1u; w < w1; ++w)
        c += __popc(m[w]);
    return c + __popc(m[w1] & hi);
    }

// block-wide exclusive scan of one value per thread (blockDim.x <= 256); s_wtot: 8 words of scratch
__device__ __forceinline__ unsigned int block_scan_exclusive(const unsigned int v, unsigned int* s_wtot, unsigned int& total)
    {
    const unsigned int warp = threadIdx.x >> 5, lane = threadIdx.x & 31u, nwarps = blockDim.x >> 5;
    unsigned int incl = v;
#pragma unroll
    for (unsigned int d = 1; d < 32; d <<= 1)
        {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d)
            incl += t;
        }
    __syncthreads(); // (s_wtot may still be read from a previous scan)
    if (lane == 31u)
        s_wtot[warp] = incl;
    __syncthreads();
    unsigned int wbase = 0;
    total = 0;
    for (unsigned int w = 0; w < nwarps; ++w)
        {
        const unsigned int t = s_wtot[w];
        wbase += w < warp ? t : 0u;
        total += t;
        }
    return wbase + incl - v;
    }

// Ascending sort of the U distinct keys in list[0 .. U) into sorted[0 .. U). list must hold roundup(U, 128) entries.
// Runs of 128 keys are sorted by one warp each in registers (4 keys per lane, bitonic network over shuffles), then every key
// finds its final rank as its position in its own run + the number of smaller keys in every other run (binary searches).
__device__ __forceinline__ void tile_sort(unsigned int* list, unsigned int* sorted, const unsigned int U)
    {
    const unsigned int warp = threadIdx.x >> 5, lane = threadIdx.x & 31u, nwarps = blockDim.x >> 5;
    const unsigned int runs = (U + 127u) >> 7;
    for (unsigned int s = U + threadIdx.x; s < (runs << 7); s += blockDim.x)
        list[s] = EMPTY_KEY;
    __syncthreads();
    for (unsigned int r = warp; r < runs; r += nwarps)
        {
        unsigned int* run = list + (r << 7);
        unsigned int v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
            v[q] = run[lane * 4u + q];
#pragma unroll
        for (unsigned int size = 2; size <= 128u; size <<= 1)
            {
#pragma unroll
            for (unsigned int j = size >> 1; j > 0; j >>= 1)
                {
                if (j >= 4u)
                    {
                    const bool lower = (lane & (j >> 2)) == 0u;
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        {
                        const unsigned int o = __shfl_xor_sync(0xffffffffu, v[q], j >> 2);
                        const bool up = ((lane * 4u + q) & size) == 0u;
                        v[q] = (lower == up) ? min(v[q], o) : max(v[q], o);
                        }
                    }
                else
                    {
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if ((q & j) == 0u)
                            {
                            const bool up = ((lane * 4u + q) & size) == 0u;
                            const unsigned int a = v[q], b = v[q | j];
                            if ((a > b) == up)
                                {
                                v[q] = b;
                                v[q | j] = a;
                                }
                            }
                    }
                }
            }
#pragma unroll
        for (int q = 0; q < 4; ++q)
            run[lane * 4u + q] = v[q];
        }
    __syncthreads();
    for (unsigned int s = threadIdx.x; s < (runs << 7); s += blockDim.x)
        {
        const unsigned int key = list[s];
        if (key == EMPTY_KEY)
            continue;
        unsigned int rank = s & 127u;
        const unsigned int mine = s >> 7;
        for (unsigned int r = 0; r < runs; ++r)
            if (r != mine)
                {
                const unsigned int* run = list + (r << 7);
                unsigned int lo = 0; // number of keys of this run below `key`
#pragma unroll
                for (unsigned int step = 64; step > 0; step >>= 1)
                    if (run[lo + step - 1u] < key)
                        lo += step;
                if (run[lo] < key) // (lo <= 127 here: the seven steps add up to at most 127)
                    ++lo;
                rank += lo;
                }
        sorted[rank] = key;
        }
    __syncthreads();
    }

// dynamic shared memory: own[ts] V4 (pruning) | head[ts] u64 | keys[hcap] u32 | list[max(ucap, 128)] u32 | sorted[ucap] u32 | off[ts + 1] u32 |
//                        keep[KEEP_WORDS] u32 (pruning) | vals[hcap] u16
template<class Prune>
__global__ void __launch_bounds__(256) tiles_build_flat_kernel(const Prune prune, const bool full_set, const unsigned int* __restrict__ tile_list,
                                                               const unsigned int slot_stride, unsigned int* __restrict__ row_count,
                                                               const unsigned int* __restrict__ n_neigh,
                                                               const unsigned int* __restrict__ nlist,
                                                               const size_t* __restrict__ head_list,
                                                               const unsigned int N,
                                                               const unsigned int tile_size,
                                                               const unsigned int ucap,
                                                               const unsigned int hcap_log2,
                                                               const unsigned int wlog,
                                                               const bool sort_rows,
                                                               unsigned int* __restrict__ uniq_count,
                                                               unsigned int* __restrict__ uniq,
                                                               uint16_t* __restrict__ slots,
                                                               unsigned int* __restrict__ status)
    {
    typedef PruneTraits<Prune> PT;
    typedef typename PT::V4 V4;
    constexpr bool PRUNE = PT::on;
    constexpr int UN = 4;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    const unsigned int hcap = 1u << hcap_log2;
    const unsigned int hmask = hcap - 1u;
    const unsigned int hshift = 32u - hcap_log2;
    V4* s_own = reinterpret_cast<V4*>(smem_raw);
    unsigned long long* s_head = reinterpret_cast<unsigned long long*>(s_own + (PRUNE ? tile_size : 0u));
    unsigned int* keys = reinterpret_cast<unsigned int*>(s_head + tile_size);
    unsigned int* list = keys + hcap;
    unsigned int* sorted = list + max(ucap, 128u);
    unsigned int* s_off = sorted + ucap;
    unsigned int* s_keep = s_off + tile_size + 1u;
    uint16_t* vals = reinterpret_cast<uint16_t*>(s_keep + (PRUNE ? KEEP_WORDS : 0u));
    unsigned int* s_bitmap = reinterpret_cast<unsigned int*>(vals + hcap); // [8 warps][128 words] (sort_rows)
    __shared__ unsigned int s_count, s_fill, s_wtot[8];

    const unsigned int tile = tile_list ? tile_list[blockIdx.x] : blockIdx.x;
    const unsigned int first = tile * tile_size;
    const unsigned int last = min(first + tile_size, N);
    const unsigned int warp = threadIdx.x >> 5, lane = threadIdx.x & 31u, nwarps = blockDim.x >> 5;
    const unsigned int W = 1u << wlog, M = 4u << wlog;
    const unsigned int A = W < 2u ? 8u : M; // AZB_SLOT_ROW_ALIGN(W)
    auto slot_head = [&](const unsigned int p) -> size_t
    {
        const unsigned int i = first + p;
        return slot_stride ? (size_t)i * slot_stride : (((size_t)s_head[p] + (size_t)A * i) & ~((size_t)A - 1u));
    };
    auto slot_at = [&](const size_t shead, const unsigned int o) -> size_t
    {
        const unsigned int r = o & (M - 1u); // entry o = b 4W + u W + l sits at b 4W + 4 l + u
        return shead + (o - r) + 4u * (r & (W - 1u)) + (r >> wlog);
    };

    // 0. row headers (and own positions) -> shared; exclusive scan of the row lengths
    for (unsigned int h = threadIdx.x; h < hcap; h += blockDim.x)
        keys[h] = EMPTY_KEY;
    if (threadIdx.x == 0)
        {
        s_count = 0;
        s_fill = 0;
        }
    const unsigned int p_own = threadIdx.x; // blockDim.x = 256 >= tile_size
    unsigned int nn_own = 0;
    if (p_own < tile_size && first + p_own < last)
        {
        nn_own = n_neigh[first + p_own];
        s_head[p_own] = head_list[first + p_own];
        if (PRUNE)
            s_own[p_own] = load_pos(prune, first + p_own);
        }
    unsigned int incl = nn_own;
#pragma unroll
    for (unsigned int d = 1; d < 32; d <<= 1)
        {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d)
            incl += t;
        }
    if (lane == 31u)
        s_wtot[warp] = incl;
    __syncthreads();
    unsigned int wbase = 0, total = 0;
    for (unsigned int w = 0; w < nwarps; ++w)
        {
        const unsigned int t = s_wtot[w];
        wbase += w < warp ? t : 0u;
        total += t;
        }
    if (p_own < tile_size)
        s_off[p_own] = wbase + incl - nn_own;
    if (threadIdx.x == 0)
        s_off[tile_size] = total;
    __syncthreads();
    const unsigned int E = total;
    if (PRUNE && E > KEEP_WORDS * 32u) // the keep mask cannot hold this tile: reported like an overflowing set (the host halves the tile)
        {
        if (threadIdx.x == 0)
            {
            uniq_count[tile] = EMPTY_KEY;
            atomicMax(status + 0, ucap + 1u);
            atomicExch(status + 1, 1u);
            }
        return;
        }

    // my warp's chunk of the flat entry range, and the row my first entry falls into
    const unsigned int chunk = ((E + nwarps - 1u) / nwarps + 31u) & ~31u;
    const unsigned int e0 = min(E, warp * chunk), e1 = min(E, e0 + chunk);
    unsigned int r_first = 0;
    if (e0 < e1)
        {
        const unsigned int e = min(e0 + lane, e1 - 1u);
        unsigned int lo = 0, hi = tile_size; // s_off[lo] <= e < s_off[hi]
        while (hi - lo > 1u)
            {
            const unsigned int mid = (lo + hi) >> 1;
            if (s_off[mid] <= e)
                lo = mid;
            else
                hi = mid;
            }
        r_first = lo;
        }

    // 1. hash-set insert
    unsigned int r = r_first;
    for (unsigned int base = e0; base < e1; base += 32u * UN)
        {
        unsigned int j[UN], rr[UN];
        bool valid[UN], keep[UN];
#pragma unroll
        for (int u = 0; u < UN; ++u)
            {
            const unsigned int e = base + 32u * u + lane;
            valid[u] = e < e1;
            j[u] = 0;
            if (valid[u])
                {
                while (e >= s_off[r + 1u])
                    ++r;
                j[u] = nlist[s_head[r] + (e - s_off[r])];
                }
            rr[u] = r;
            }
        if (PRUNE)
            {
            V4 pj[UN];
#pragma unroll
            for (int u = 0; u < UN; ++u)
                if (valid[u])
                    pj[u] = load_pos(prune, j[u]);
#pragma unroll
            for (int u = 0; u < UN; ++u)
                {
                keep[u] = valid[u] && kept_staged(prune, s_own[rr[u]], pj[u]);
                const unsigned int m = __ballot_sync(0xffffffffu, keep[u]);
                if (lane == 0 && base + 32u * u < e1)
                    s_keep[(base + 32u * u) >> 5] = m;
                }
            }
#pragma unroll
        for (int u = 0; u < UN; ++u)
            {
            if (!(PRUNE && !full_set ? keep[u] : valid[u]))
                continue;
            unsigned int h = tile_hash(j[u], hshift);
            for (;;)
                {
                // most entries are repeats (every neighbour is shared by 6-12 rows of the tile): a plain read finds them
                // without an atomic; only an empty cell is claimed with a CAS
                unsigned int old = *(volatile unsigned int*)&keys[h];
                if (old == EMPTY_KEY)
                    {
                    old = atomicCAS(&keys[h], EMPTY_KEY, j[u]);
                    if (old == EMPTY_KEY)
                        {
                        atomicAdd(&s_count, 1u);
                        break;
                        }
                    }
                if (old == j[u])
                    break;
                h = (h + 1u) & hmask;
                if (*(volatile unsigned int*)&s_count > ucap) // overflowing set: give up (the table could fill)
                    break;
                }
            }
        }
    __syncthreads();
    const unsigned int U = s_count;
    if (U > ucap)
        {
        if (threadIdx.x == 0)
            {
            uniq_count[tile] = EMPTY_KEY;
            atomicMax(status + 0, U);
            atomicExch(status + 1, 1u);
            }
        return;
        }

    // 2. compact + sort ascending
    for (unsigned int h = threadIdx.x; h < hcap; h += blockDim.x)
        {
        const unsigned int key = keys[h];
        if (key != EMPTY_KEY)
            list[atomicAdd(&s_fill, 1u)] = key;
        }
    __syncthreads();
    tile_sort(list, sorted, U);

    // 3. publish the list; remember each key's rank next to its table position
    unsigned int* out = uniq + (size_t)tile * ucap;
    for (unsigned int s = threadIdx.x; s < U; s += blockDim.x)
        {
        const unsigned int key = sorted[s];
        out[s] = key;
        unsigned int h = tile_hash(key, hshift);
        while (keys[h] != key)
            h = (h + 1u) & hmask;
        vals[h] = (uint16_t)s;
        }
    if (threadIdx.x == 0)
        {
        uniq_count[tile] = U;
        atomicMax(status + 0, U);
        }
    __syncthreads();

    // 4'. (sort_rows) slot rows in ASCENDING slot order instead of list order: the lanes of neighbouring particles then read
    // neighbouring or identical staged positions at the same time (fewer shared-memory bank conflicts in the pair kernels). A warp
    // takes a row: every (kept) entry sets the bit of its slot in a U-bit map; the row is the enumeration of the set bits.
    if (sort_rows)
        {
        unsigned int* bm = s_bitmap + warp * 128u;
        const unsigned int nwords = (U + 31u) >> 5;
        const unsigned int wpl = (nwords + 31u) >> 5; // words per lane (<= 4)
        for (unsigned int p = warp; p < last - first; p += nwarps)
            {
            for (unsigned int w = lane; w < nwords; w += 32u)
                bm[w] = 0u;
            __syncwarp();
            const unsigned int b0 = s_off[p], nn = s_off[p + 1u] - b0;
            const unsigned int* row = nlist + s_head[p];
            for (unsigned int kk = lane; kk < nn; kk += 32u)
                {
                if (PRUNE && !((s_keep[(b0 + kk) >> 5] >> ((b0 + kk) & 31u)) & 1u))
                    continue;
                const unsigned int j = row[kk];
                unsigned int h = tile_hash(j, hshift);
                while (keys[h] != j)
                    h = (h + 1u) & hmask;
                const unsigned int v = vals[h];
                atomicOr(&bm[v >> 5], 1u << (v & 31u));
                }
            __syncwarp();
            unsigned int mine = 0;
            for (unsigned int q = 0; q < wpl; ++q)
                {
                const unsigned int w = lane * wpl + q;
                mine += w < nwords ? __popc(bm[w]) : 0u;
                }
            unsigned int incl = mine;
#pragma unroll
            for (unsigned int d = 1; d < 32; d <<= 1)
                {
                const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d)
                    incl += t;
                }
            unsigned int o = incl - mine;
            uint16_t* srow = slots + slot_head(p);
            for (unsigned int q = 0; q < wpl; ++q)
                {
                const unsigned int w = lane * wpl + q;
                unsigned int m = w < nwords ? bm[w] : 0u;
                while (m)
                    {
                    const unsigned int b = __ffs(m) - 1u;
                    m &= m - 1u;
                    srow[slot_at(0, o) ] = (uint16_t)(w * 32u + b);
                    ++o;
                    }
                }
            __syncwarp();
            }
        }
    else
        {
    // 4. list entry -> slot, in the lane-transposed, padded row layout (include/azb200.h, AZB_SLOT_HEAD); a pruned row holds
    // its kept entries in list order
    r = r_first;
    for (unsigned int base = e0; base < e1; base += 32u * UN)
        {
        unsigned int j[UN], rr[UN], ee[UN];
        bool valid[UN];
#pragma unroll
        for (int u = 0; u < UN; ++u)
            {
            const unsigned int e = base + 32u * u + lane;
            ee[u] = e;
            valid[u] = e < e1;
            j[u] = 0;
            if (valid[u])
                {
                while (e >= s_off[r + 1u])
                    ++r;
                if (PRUNE)
                    valid[u] = (s_keep[e >> 5] >> (e & 31u)) & 1u;
                if (valid[u])
                    j[u] = nlist[s_head[r] + (e - s_off[r])];
                }
            rr[u] = r;
            }
#pragma unroll
        for (int u = 0; u < UN; ++u)
            if (valid[u])
                {
                unsigned int h = tile_hash(j[u], hshift);
                while (keys[h] != j[u])
                    h = (h + 1u) & hmask;
                const unsigned int b0 = s_off[rr[u]];
                const unsigned int o = PRUNE ? count_bits(s_keep, b0, ee[u]) : ee[u] - b0;
                slots[slot_at(slot_head(rr[u]), o)] = vals[h];
                }
        }
        }
    // per row: padding, kept count, layout checks
    if (p_own < tile_size && first + p_own < last)
        {
        const unsigned int i = first + p_own;
        const unsigned int b0 = s_off[p_own], b1 = s_off[p_own + 1u];
        const unsigned int out_n = PRUNE ? count_bits(s_keep, b0, b1) : nn_own;
        const unsigned int out_p = (out_n + M - 1u) & ~(M - 1u);
        const size_t shead = slot_head(p_own);
        for (unsigned int o = out_n; o < out_p; ++o)
            slots[slot_at(shead, o)] = 0;
        if (row_count)
            row_count[i] = out_n;
        // rows must be disjoint and ordered by particle for the padded layout to be collision-free
        if (i + 1u < N)
            {
            const size_t next_head = (i + 1u < last) ? (size_t)s_head[p_own + 1u] : head_list[i + 1u];
            if (next_head < (size_t)s_head[p_own] + nn_own)
                atomicExch(status + 1, 2u);
            }
        if (i + 1u == last) // extent of the tile's slot rows (kernels that stage them size their buffer by it)
            atomicMax(status + 2, (unsigned int)(shead + ((nn_own + M - 1u) & ~(M - 1u)) - slot_head(0u)));
        }
    }

// ---- a pruned view derived from the unpruned plan of the same list (azb_tile_build_args.base_plan) ----------------------------
// Several forces on one list (C5: ExpandedYukawa walks every entry, DPD a pruned view) used to pay the hash build twice. The
// unpruned plan already names every entry's neighbour by a 16-bit slot into an ASCENDING distinct list, so the pruned view
// needs no hash and no sort: stage the positions of the tile's distinct neighbours once (U x 16 B), test every entry against
// its staged neighbour (2-byte slot read, shared-memory position), flag the slots that survive, scan the flags — the ranks of
// the surviving slots ARE the pruned plan's slots, and its distinct list is the flagged subsequence, still ascending.
// Output identical to tiles_build_flat_kernel<PruneDev> with full_set = false.
// dynamic shared memory: pos[ucap_b] V4 | own[ts] V4 | head[ts] u64 | off[ts + 1] u32 | keep[KEEP_WORDS] u32 | rank[ucap_b] u32
template<class S>
__global__ void __launch_bounds__(256) tiles_prune_kernel(const PruneDev<S> prune, const unsigned int* __restrict__ b_uniq_count,
                                                          const unsigned int* __restrict__ b_uniq, const uint16_t* __restrict__ b_slots,
                                                          const unsigned int b_ucap, const unsigned int b_smem_uniq, const unsigned int b_wlog,
                                                          const unsigned int b_slot_stride, const unsigned int slot_stride,
                                                          unsigned int* __restrict__ row_count, const unsigned int* __restrict__ n_neigh,
                                                          const size_t* __restrict__ head_list, const unsigned int N, const unsigned int tile_size,
                                                          const unsigned int ucap, const unsigned int wlog, unsigned int* __restrict__ uniq_count,
                                                          unsigned int* __restrict__ uniq, uint16_t* __restrict__ slots,
                                                          unsigned int* __restrict__ status)
    {
    typedef typename vec4<S>::type V4;
    constexpr int UN = 4;
    extern __shared__ __align__(32) unsigned char smem_raw[];
    V4* s_pos = reinterpret_cast<V4*>(smem_raw);
    V4* s_own = s_pos + b_smem_uniq;
    unsigned long long* s_head = reinterpret_cast<unsigned long long*>(s_own + tile_size);
    unsigned int* s_off = reinterpret_cast<unsigned int*>(s_head + tile_size);
    unsigned int* s_keep = s_off + tile_size + 1u;
    unsigned int* s_rank = s_keep + KEEP_WORDS;
    __shared__ unsigned int s_wtot[8];

    const unsigned int tile = blockIdx.x;
    const unsigned int first = tile * tile_size;
    const unsigned int last = min(first + tile_size, N);
    const unsigned int warp = threadIdx.x >> 5, lane = threadIdx.x & 31u, nwarps = blockDim.x >> 5;
    auto layout = [](const unsigned int o, const unsigned int wl) -> unsigned int
    {
        const unsigned int W = 1u << wl, M = 4u << wl, r = o & (M - 1u); // entry o = b 4W + u W + l sits at b 4W + 4 l + u
        return (o - r) + 4u * (r & (W - 1u)) + (r >> wl);
    };
    auto slot_head = [&](const unsigned int p, const unsigned int wl, const unsigned int stride) -> size_t
    {
        const unsigned int i = first + p;
        const size_t A = wl < 1u ? 8u : (4u << wl); // AZB_SLOT_ROW_ALIGN(W)
        return stride ? (size_t)i * stride : (((size_t)s_head[p] + A * i) & ~(A - 1u));
    };

    const unsigned int Ub = b_uniq_count[tile];
    if (Ub == EMPTY_KEY) // the base plan has no set for this tile
        {
        if (threadIdx.x == 0)
            {
            uniq_count[tile] = EMPTY_KEY;
            atomicExch(status + 1, 1u);
            }
        return;
        }
    // 0. the tile's distinct neighbours (positions), row headers, own positions; exclusive scan of the row lengths
    const unsigned int* buq = b_uniq + (size_t)tile * b_ucap;
    for (unsigned int s0 = threadIdx.x; s0 < Ub; s0 += 4u * blockDim.x)
        {
        unsigned int g[4];
        V4 pv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            g[u] = (s0 + u * blockDim.x < Ub) ? buq[s0 + u * blockDim.x] : EMPTY_KEY;
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (g[u] != EMPTY_KEY)
                pv[u] = ld4(prune.pos + g[u]);
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (g[u] != EMPTY_KEY)
                {
                s_pos[s0 + u * blockDim.x] = pv[u];
                s_rank[s0 + u * blockDim.x] = 0u;
                }
        }
    const unsigned int p_own = threadIdx.x; // blockDim.x = 256 >= tile_size
    unsigned int nn_own = 0;
    if (p_own < tile_size && first + p_own < last)
        {
        nn_own = n_neigh[first + p_own];
        s_head[p_own] = head_list[first + p_own];
        s_own[p_own] = ld4(prune.pos + first + p_own);
        }
    unsigned int incl = nn_own;
#pragma unroll
    for (unsigned int d = 1; d < 32; d <<= 1)
        {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d)
            incl += t;
        }
    if (lane == 31u)
        s_wtot[warp] = incl;
    __syncthreads();
    unsigned int wbase = 0, total = 0;
    for (unsigned int w = 0; w < nwarps; ++w)
        {
        const unsigned int t = s_wtot[w];
        wbase += w < warp ? t : 0u;
        total += t;
        }
    if (p_own < tile_size)
        s_off[p_own] = wbase + incl - nn_own;
    if (threadIdx.x == 0)
        s_off[tile_size] = total;
    __syncthreads();
    const unsigned int E = total;
    if (E > KEEP_WORDS * 32u)
        {
        if (threadIdx.x == 0)
            {
            uniq_count[tile] = EMPTY_KEY;
            atomicMax(status + 0, ucap + 1u);
            atomicExch(status + 1, 1u);
            }
        return;
        }
    const unsigned int chunk = ((E + nwarps - 1u) / nwarps + 31u) & ~31u;
    const unsigned int e0 = min(E, warp * chunk), e1 = min(E, e0 + chunk);
    unsigned int r_first = 0;
    if (e0 < e1)
        {
        const unsigned int e = min(e0 + lane, e1 - 1u);
        unsigned int lo = 0, hi = tile_size; // s_off[lo] <= e < s_off[hi]
        while (hi - lo > 1u)
            {
            const unsigned int mid = (lo + hi) >> 1;
            if (s_off[mid] <= e)
                lo = mid;
            else
                hi = mid;
            }
        r_first = lo;
        }

    // 1. keep test of every entry against its staged neighbour; surviving slots are flagged
    unsigned int r = r_first;
    for (unsigned int base = e0; base < e1; base += 32u * UN)
        {
        unsigned int sl[UN], rr[UN];
        bool valid[UN];
#pragma unroll
        for (int u = 0; u < UN; ++u)
            {
            const unsigned int e = base + 32u * u + lane;
            valid[u] = e < e1;
            sl[u] = 0;
            if (valid[u])
                {
                while (e >= s_off[r + 1u])
                    ++r;
                sl[u] = b_slots[slot_head(r, b_wlog, b_slot_stride) + layout(e - s_off[r], b_wlog)];
                }
            rr[u] = r;
            }
#pragma unroll
        for (int u = 0; u < UN; ++u)
            {
            const bool keep = valid[u] && kept_staged(prune, s_own[rr[u]], s_pos[sl[u]]);
            const unsigned int m = __ballot_sync(0xffffffffu, keep);
            if (lane == 0 && base + 32u * u < e1)
                s_keep[(base + 32u * u) >> 5] = m;
            if (keep)
                s_rank[sl[u]] = 1u; // (benign race: every writer stores the same value)
            }
        }
    __syncthreads();

    // 2. ranks of the surviving slots (exclusive scan of the flags, 8 slots per thread) = the pruned plan's slots; its distinct
    // list is the surviving subsequence of the base list, still ascending
    unsigned int f[8], mine = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q)
        {
        const unsigned int sidx = threadIdx.x * 8u + q;
        f[q] = sidx < Ub ? s_rank[sidx] : 0u;
        mine += f[q];
        }
    incl = mine;
#pragma unroll
    for (unsigned int d = 1; d < 32; d <<= 1)
        {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d)
            incl += t;
        }
    __syncthreads(); // (s_wtot is reused)
    if (lane == 31u)
        s_wtot[warp] = incl;
    __syncthreads();
    wbase = 0, total = 0;
    for (unsigned int w = 0; w < nwarps; ++w)
        {
        const unsigned int t = s_wtot[w];
        wbase += w < warp ? t : 0u;
        total += t;
        }
    const unsigned int U = total;
    unsigned int* out = uniq + (size_t)tile * ucap;
    unsigned int run = wbase + incl - mine;
    if (U <= ucap)
        {
#pragma unroll
        for (int q = 0; q < 8; ++q)
            {
            const unsigned int sidx = threadIdx.x * 8u + q;
            if (sidx < Ub)
                {
                s_rank[sidx] = run;
                if (f[q])
                    out[run] = buq[sidx];
                run += f[q];
                }
            }
        }
    if (threadIdx.x == 0)
        {
        uniq_count[tile] = U <= ucap ? U : EMPTY_KEY;
        atomicMax(status + 0, U);
        if (U > ucap)
            atomicExch(status + 1, 1u);
        }
    if (U > ucap)
        return;
    __syncthreads();

    // 3. surviving entries -> pruned slot rows (list order), in the view's own lane-transposed layout
    r = r_first;
    for (unsigned int base = e0; base < e1; base += 32u * UN)
        {
#pragma unroll
        for (int u = 0; u < UN; ++u)
            {
            const unsigned int e = base + 32u * u + lane;
            if (e < e1)
                {
                while (e >= s_off[r + 1u])
                    ++r;
                if ((s_keep[e >> 5] >> (e & 31u)) & 1u)
                    {
                    const unsigned int b0 = s_off[r];
                    const unsigned int s_old = b_slots[slot_head(r, b_wlog, b_slot_stride) + layout(e - b0, b_wlog)];
                    slots[slot_head(r, wlog, slot_stride) + layout(count_bits(s_keep, b0, e), wlog)] = (uint16_t)s_rank[s_old];
                    }
                }
            }
        }
    if (p_own < tile_size && first + p_own < last)
        {
        const unsigned int i = first + p_own;
        const unsigned int M = 4u << wlog;
        const unsigned int out_n = count_bits(s_keep, s_off[p_own], s_off[p_own + 1u]);
        const unsigned int out_p = (out_n + M - 1u) & ~(M - 1u);
        const size_t shead = slot_head(p_own, wlog, slot_stride);
        for (unsigned int o = out_n; o < out_p; ++o)
            slots[shead + layout(o, wlog)] = 0;
        row_count[i] = out_n;
        if (i + 1u == last) // extent of the tile's slot rows, as tiles_build_kernel reports it
            atomicMax(status + 2, (unsigned int)(shead + ((nn_own + M - 1u) & ~(M - 1u)) - slot_head(0u, wlog, slot_stride)));
        }
    }
    } // namespace azb

extern "C" int azb_tiles_build(const azb_tile_build_args* a, unsigned int* h_max_uniq, unsigned int* h_max_slot_span, unsigned int* h_overflow,
                               void* stream)
    {
    if (!a || !h_max_uniq || !h_max_slot_span || !h_overflow)
        return AZB_ERR_BAD_ARGUMENT;
    const unsigned int ts = a->tile_size, ucap = a->uniq_capacity;
    const unsigned int W = a->lane_width;
    const unsigned int G = a->group_width; // 0: rows keep the list's order
    if (!(ts == 32 || ts == 64 || ts == 128 || ts == 256) || ucap < 32 || ucap > 4096 || (ucap & (ucap - 1u))
        || !(W == 1 || W == 2 || W == 4 || W == 8 || W == 16) || G > 32 || (G & (G - 1u)))
        return AZB_ERR_BAD_ARGUMENT;
    *h_max_uniq = 0;
    *h_max_slot_span = 0;
    *h_overflow = 0;
    if (a->N == 0)
        return 0;
    if (!a->d_n_neigh || !a->d_nlist || !a->d_head_list || !a->d_uniq_count || !a->d_uniq || !a->d_slots || !a->d_status)
        return AZB_ERR_BAD_ARGUMENT;
    cudaStream_t s = (cudaStream_t)stream;
    unsigned int hcap_log2 = 0;
    while ((1u << hcap_log2) < 2u * ucap)
        ++hcap_log2;
    const size_t hcap = (size_t)1 << hcap_log2;
    const size_t smem = hcap * sizeof(unsigned int) + ucap * sizeof(unsigned int) + hcap * sizeof(uint16_t);
    cudaError_t err = cudaMemsetAsync(a->d_status, 0, 4 * sizeof(unsigned int), s);
    if (err != cudaSuccess)
        return (int)err;
    const unsigned int n_tiles = a->d_tile_list ? a->n_tile_list : (a->N + ts - 1) / ts;
    if (n_tiles == 0)
        return 0;
    const bool prune = a->d_prune_pos != nullptr;
    if (a->base_plan)
        {
        // a pruned view derived from the unpruned plan of the same list (tiles_prune_kernel)
        const azb_tile_plan& b = *a->base_plan;
        const unsigned int Wb = b.lane_width;
        if (!prune || !a->d_prune_rsq || !a->d_row_count || a->prune_ntypes == 0 || G != 0u || a->d_tile_list || a->prune_full_set
            || (a->prune_scalar_bits != 32 && a->prune_scalar_bits != 64) || b.tile_size != ts || b.n_tiles != n_tiles || b.d_row_count
            || !b.d_uniq_count || !b.d_uniq || !b.d_slots || b.max_uniq > b.uniq_capacity || b.max_uniq > ucap
            || !(Wb == 1 || Wb == 2 || Wb == 4 || Wb == 8 || Wb == 16))
            return AZB_ERR_BAD_ARGUMENT;
        unsigned int wlog_b = 0, wlog_v = 0;
        while ((1u << wlog_b) < Wb)
            ++wlog_b;
        while ((1u << wlog_v) < W)
            ++wlog_v;
        const unsigned int su = ((b.max_uniq > 0 ? b.max_uniq : 1u) + 7u) & ~7u;
        if (su > 2048u)
            return AZB_ERR_BAD_ARGUMENT; // (the rank scan handles 256 x 8 slots)
        const size_t v4 = a->prune_scalar_bits == 32 ? 16 : 32;
        const size_t psmem = (size_t)su * v4 + (size_t)ts * v4 + (size_t)ts * 8 + ((size_t)ts + 1) * 4 + azb::KEEP_WORDS * 4 + (size_t)su * 4;
        if (a->prune_scalar_bits == 32)
            {
            azb::PruneDev<float> p = {(const float4*)a->d_prune_pos, (const float*)a->d_prune_rsq, a->prune_ntypes, azb::make_box<float>(a->prune_box)};
            if (psmem > 48 * 1024 && (err = cudaFuncSetAttribute(azb::tiles_prune_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)psmem)) != cudaSuccess)
                return (int)err;
            azb::tiles_prune_kernel<float><<<n_tiles, 256, psmem, s>>>(p, b.d_uniq_count, b.d_uniq, b.d_slots, b.uniq_capacity, su, wlog_b, b.slot_row_stride,
                                                                      a->slot_row_stride, a->d_row_count, a->d_n_neigh, a->d_head_list, a->N, ts, ucap, wlog_v,
                                                                      a->d_uniq_count, a->d_uniq, a->d_slots, a->d_status);
            }
        else
            {
            azb::PruneDev<double> p = {(const azb::double4a*)a->d_prune_pos, (const double*)a->d_prune_rsq, a->prune_ntypes,
                                       azb::make_box<double>(a->prune_box)};
            if (psmem > 48 * 1024 && (err = cudaFuncSetAttribute(azb::tiles_prune_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)psmem)) != cudaSuccess)
                return (int)err;
            azb::tiles_prune_kernel<double><<<n_tiles, 256, psmem, s>>>(p, b.d_uniq_count, b.d_uniq, b.d_slots, b.uniq_capacity, su, wlog_b, b.slot_row_stride,
                                                                       a->slot_row_stride, a->d_row_count, a->d_n_neigh, a->d_head_list, a->N, ts, ucap, wlog_v,
                                                                       a->d_uniq_count, a->d_uniq, a->d_slots, a->d_status);
            }
        err = cudaGetLastError();
        if (err != cudaSuccess)
            return (int)err;
        unsigned int hs[4] = {0, 0, 0, 0};
        err = cudaMemcpyAsync(hs, a->d_status, sizeof(hs), cudaMemcpyDeviceToHost, s);
        if (err == cudaSuccess)
            err = cudaStreamSynchronize(s);
        if (err != cudaSuccess)
            return (int)err;
        *h_max_uniq = hs[0];
        *h_overflow = hs[1];
        *h_max_slot_span = hs[2];
        return 0;
        }
    // group_width == 0 (rows keep the list's order; the default): the flat-range builder. AZB_TILES_ROWWISE=1 keeps the row-wise
    // kernel for A/B runs; bank-rotated rows (group_width > 0, opt-in) always take it.
    static const bool rowwise = getenv("AZB_TILES_ROWWISE") != nullptr;
    const bool flat = G == 0u && !rowwise;
    unsigned int wlog = 0;
    while ((1u << wlog) < W)
        ++wlog;
    if (prune && (!a->d_prune_rsq || !a->d_row_count || a->prune_ntypes == 0 || G != 0u || (a->prune_scalar_bits != 32 && a->prune_scalar_bits != 64)))
        return AZB_ERR_BAD_ARGUMENT;
#define AZB_TILES_LAUNCH(PRUNE_T, PRUNE_V)                                                                                              \
    do                                                                                                                                  \
        {                                                                                                                               \
        if (flat)                                                                                                                       \
            {                                                                                                                           \
            const size_t fsmem = smem + (size_t)(ucap < 128 ? 128 - ucap : 0) * 4 + (size_t)ucap * 4 + (size_t)ts * 8 + ((size_t)ts + 1) * 4                                                           \
                                 + (azb::PruneTraits<PRUNE_T>::on ? (size_t)ts * sizeof(azb::PruneTraits<PRUNE_T>::V4) + azb::KEEP_WORDS * 4 : 0) \
                                 + (a->sort_rows ? 8 * 128 * 4 : 0);                                                                     \
            if (fsmem > 48 * 1024)                                                                                                      \
                {                                                                                                                       \
                err = cudaFuncSetAttribute(azb::tiles_build_flat_kernel<PRUNE_T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsmem); \
                if (err != cudaSuccess)                                                                                                 \
                    return (int)err;                                                                                                    \
                }                                                                                                                       \
            azb::tiles_build_flat_kernel<PRUNE_T><<<n_tiles, 256, fsmem, s>>>(PRUNE_V, a->prune_full_set != 0, a->d_tile_list, a->slot_row_stride, a->d_row_count, a->d_n_neigh, a->d_nlist, \
                                                                            a->d_head_list, a->N, ts, ucap, hcap_log2, wlog, a->sort_rows != 0, a->d_uniq_count, a->d_uniq, a->d_slots, a->d_status); \
            break;                                                                                                                      \
            }                                                                                                                           \
        if (smem > 48 * 1024)                                                                                                           \
            {                                                                                                                           \
            err = cudaFuncSetAttribute(azb::tiles_build_kernel<PRUNE_T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);       \
            if (err != cudaSuccess)                                                                                                     \
                return (int)err;                                                                                                        \
            }                                                                                                                           \
        azb::tiles_build_kernel<PRUNE_T><<<n_tiles, 256, smem, s>>>(PRUNE_V, a->prune_full_set != 0, a->d_tile_list, a->slot_row_stride, a->d_row_count, a->d_n_neigh, a->d_nlist, a->d_head_list, a->N, ts, \
                                                                   ucap, hcap_log2, W, G, a->d_uniq_count, a->d_uniq, a->d_slots, a->d_status); \
        } while (0)
    if (!prune)
        AZB_TILES_LAUNCH(azb::NoPrune, azb::NoPrune());
    else if (a->prune_scalar_bits == 32)
        {
        azb::PruneDev<float> p = {(const float4*)a->d_prune_pos, (const float*)a->d_prune_rsq, a->prune_ntypes, azb::make_box<float>(a->prune_box)};
        AZB_TILES_LAUNCH(azb::PruneDev<float>, p);
        }
    else
        {
        azb::PruneDev<double> p = {(const azb::double4a*)a->d_prune_pos, (const double*)a->d_prune_rsq, a->prune_ntypes,
                                   azb::make_box<double>(a->prune_box)};
        AZB_TILES_LAUNCH(azb::PruneDev<double>, p);
        }
#undef AZB_TILES_LAUNCH
    err = cudaGetLastError();
    if (err != cudaSuccess)
        return (int)err;
    unsigned int h[4] = {0, 0, 0, 0};
    err = cudaMemcpyAsync(h, a->d_status, sizeof(h), cudaMemcpyDeviceToHost, s);
    if (err != cudaSuccess)
        return (int)err;
    err = cudaStreamSynchronize(s);
    if (err != cudaSuccess)
        return (int)err;
    *h_max_uniq = h[0];
    *h_overflow = h[1];
    *h_max_slot_span = h[2];
    return 0;
    }
