// azb_tile_cache.cu — library-owned tile plans for callers that cannot carry one (include/azb200.h, azb_tile_cache_*).
//
// HOOMD's pair_args_t has no room for a tile plan, and the kernel-driver stubs that replace
// /root/reference/src/PotentialPairGPUKernel.cu.inc:25-28 (hoomd_glue/*.cu.inc) see nothing but that struct. They ask this
// cache for the plan of the list they were handed: it is built on first use (azb_tiles_build into buffers the library owns)
// and reused until the host class reports a list rebuild (azb_tile_cache_invalidate, from NeighborList::hasBeenUpdated) or
// the list's shape changes. One entry per (d_nlist, lane_width): a DPD force and a pair force on the same list need
// different slot layouts.
#include <cuda_runtime.h>

#include <map>
#include <mutex>
#include <utility>

#include "azb200.h"

namespace
    {
struct Entry
    {
    azb_tile_plan plan {};
    const void* d_head_list = nullptr;
    const void* d_n_neigh = nullptr;
    size_t size_nlist = 0;
    unsigned int N = 0;
    bool valid = false;  // arrays match the list's current content
    bool usable = false; // a plan exists (false: every tile size overflowed -> direct kernels)
    unsigned int* d_uniq_count = nullptr;
    unsigned int* d_uniq = nullptr;
    uint16_t* d_slots = nullptr;
    unsigned int* d_status = nullptr;
    size_t cap_tiles = 0, cap_uniq = 0, cap_slots = 0, cap_status = 0;
    void release()
        {
        cudaFree(d_uniq_count);
        cudaFree(d_uniq);
        cudaFree(d_slots);
        cudaFree(d_status);
        d_uniq_count = d_uniq = d_status = nullptr;
        d_slots = nullptr;
        cap_tiles = cap_uniq = cap_slots = cap_status = 0;
        valid = usable = false;
        }
    };
std::mutex g_mutex;
std::map<std::pair<const void*, unsigned int>, Entry> g_cache;

template<class T> int grow(T*& p, size_t& cap, size_t need)
    {
    if (need <= cap)
        return 0;
    cudaFree(p);
    p = nullptr;
    cap = 0;
    const size_t want = need + need / 8;
    const cudaError_t err = cudaMalloc((void**)&p, want * sizeof(T));
    if (err != cudaSuccess)
        return (int)err;
    cap = want;
    return 0;
    }
    } // namespace

extern "C" AZB_API int azb_tile_cache_get(const azb_pair_args* a, unsigned int lane_width, unsigned int tile_size, void* stream,
                                          const azb_tile_plan** plan)
    {
    if (!a || !plan || !a->d_nlist || !a->d_head_list || !a->d_n_neigh)
        return AZB_ERR_BAD_ARGUMENT;
    *plan = nullptr;
    if (a->N == 0)
        return 0;
    std::lock_guard<std::mutex> lock(g_mutex);
    Entry& e = g_cache[std::make_pair((const void*)a->d_nlist, lane_width)];
    size_t size_nlist = a->size_nlist;
    if (size_nlist == 0 && !(e.valid && e.N == a->N && e.d_head_list == a->d_head_list && e.d_n_neigh == a->d_n_neigh))
        {
        // a caller whose struct does not carry the list size (HOOMD's a_pair_args_t): end of the last row, read once per rebuild
        size_t head = 0;
        unsigned int nn = 0;
        cudaStream_t s = (cudaStream_t)stream;
        cudaError_t err = cudaMemcpyAsync(&head, (const size_t*)a->d_head_list + (a->N - 1), sizeof(head), cudaMemcpyDeviceToHost, s);
        if (err == cudaSuccess)
            err = cudaMemcpyAsync(&nn, (const unsigned int*)a->d_n_neigh + (a->N - 1), sizeof(nn), cudaMemcpyDeviceToHost, s);
        if (err == cudaSuccess)
            err = cudaStreamSynchronize(s);
        if (err != cudaSuccess)
            return (int)err;
        size_nlist = head + nn;
        }
    else if (size_nlist == 0)
        size_nlist = e.size_nlist;
    if (e.valid && e.N == a->N && e.size_nlist == size_nlist && e.d_head_list == a->d_head_list && e.d_n_neigh == a->d_n_neigh)
        {
        *plan = e.usable ? &e.plan : nullptr;
        return 0;
        }
    const unsigned int ucap = 2048;
    int rc = grow(e.d_status, e.cap_status, 4); // 4 words of scratch
    if (rc)
        return rc;
    e.usable = false;
    for (unsigned int ts = tile_size ? tile_size : 128u; ts >= 32u; ts /= 2u)
        {
        const size_t n_tiles = ((size_t)a->N + ts - 1) / ts;
        if ((rc = grow(e.d_uniq_count, e.cap_tiles, n_tiles)) || (rc = grow(e.d_uniq, e.cap_uniq, n_tiles * ucap))
            || (rc = grow(e.d_slots, e.cap_slots, AZB_SLOTS_SIZE(size_nlist, a->N, lane_width))))
            return rc;
        azb_tile_build_args b {};
        b.d_n_neigh = a->d_n_neigh;
        b.d_nlist = a->d_nlist;
        b.d_head_list = a->d_head_list;
        b.size_nlist = size_nlist;
        b.N = a->N;
        b.tile_size = ts;
        b.uniq_capacity = ucap;
        b.lane_width = lane_width;
        b.group_width = 0;
        b.d_uniq_count = e.d_uniq_count;
        b.d_uniq = e.d_uniq;
        b.d_slots = e.d_slots;
        b.d_status = e.d_status;
        unsigned int max_uniq = 0, max_span = 0, overflow = 0;
        rc = azb_tiles_build(&b, &max_uniq, &max_span, &overflow, stream);
        if (rc)
            return rc;
        if (overflow == 2u)
            break; // rows are not disjoint / ordered: no plan for this list layout
        if (!overflow)
            {
            e.plan.tile_size = ts;
            e.plan.n_tiles = (unsigned int)n_tiles;
            e.plan.uniq_capacity = ucap;
            e.plan.max_uniq = max_uniq;
            e.plan.lane_width = lane_width;
            e.plan.max_slot_span = max_span;
            e.plan.d_uniq_count = e.d_uniq_count;
            e.plan.d_uniq = e.d_uniq;
            e.plan.d_slots = e.d_slots;
            e.usable = true;
            break;
            }
        }
    e.N = a->N;
    e.size_nlist = size_nlist;
    e.d_head_list = a->d_head_list;
    e.d_n_neigh = a->d_n_neigh;
    e.valid = true;
    *plan = e.usable ? &e.plan : nullptr;
    return 0;
    }

extern "C" AZB_API void azb_tile_cache_invalidate(const void* d_nlist)
    {
    std::lock_guard<std::mutex> lock(g_mutex);
    for (auto& kv : g_cache)
        if (d_nlist == nullptr || kv.first.first == d_nlist)
            kv.second.valid = false;
    }

extern "C" AZB_API void azb_tile_cache_clear(void)
    {
    std::lock_guard<std::mutex> lock(g_mutex);
    for (auto& kv : g_cache)
        kv.second.release();
    g_cache.clear();
    }
