// azb_hostio.cu — azb_host_step: one force evaluation driven from host buffers, with the H2D copies, the tile-range
// launches and the D2H copies overlapped inside the call (include/azb200.h). Everything here is host code: it issues
// cudaMemcpyAsync / event waits / the ordinary azb_*_forces launches; the kernels are the same ones a resident caller runs.
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "azb200.h"

namespace
    {
// events are created once per thread and reused (creation costs more than a launch)
struct EventPool
    {
    std::vector<cudaEvent_t> ev;
    cudaEvent_t get(size_t i)
        {
        while (ev.size() <= i)
            {
            cudaEvent_t e = nullptr;
            if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess)
                return nullptr;
            ev.push_back(e);
            }
        return ev[i];
        }
    };
thread_local EventPool g_pool;

#define AZB_TRY(expr)                 \
    do                                \
        {                             \
        const int _e = (int)(expr);   \
        if (_e != 0)                  \
            return _e;                \
        } while (0)
    } // namespace

extern "C" AZB_API int azb_host_step(const azb_host_step_args* a)
    {
    if (!a || !a->force_args || !a->in_edges || !a->launches || (a->n_outs && !a->outs) || a->n_inputs > 4 || a->n_outputs > 4 || a->n_in_chunks == 0)
        return AZB_ERR_BAD_ARGUMENT;
    if (a->kind != AZB_KIND_PAIR && a->kind != AZB_KIND_DPD && a->kind != AZB_KIND_ANISO)
        return AZB_ERR_BAD_ARGUMENT;
    cudaStream_t cur = (cudaStream_t)a->stream, s_in = (cudaStream_t)a->stream_in, s_out = (cudaStream_t)a->stream_out;
    // private copies of the argument struct: tile_begin / tile_count change per chunk
    azb_pair_args pa;
    azb_dpd_args da;
    azb_aniso_args aa;
    azb_pair_args* range = nullptr;
    if (a->kind == AZB_KIND_PAIR)
        {
        pa = *(const azb_pair_args*)a->force_args;
        range = &pa;
        }
    else if (a->kind == AZB_KIND_DPD)
        {
        da = *(const azb_dpd_args*)a->force_args;
        range = &da.pair;
        }
    else
        {
        aa = *(const azb_aniso_args*)a->force_args;
        range = &aa.pair;
        }
    if (!range->tiles)
        return AZB_ERR_BAD_ARGUMENT;

    static const bool trace = std::getenv("AZB_TRACE_HOST_STEP") != nullptr; // prints how long the host spent issuing
    const auto t_begin = std::chrono::steady_clock::now();
    const size_t n_ev = (size_t)a->n_in_chunks + a->n_launches + 1;
    if (!g_pool.get(n_ev))
        return (int)cudaGetLastError();
    // the copy streams start after whatever the compute stream has queued (the previous step's readers)
    cudaEvent_t start = g_pool.get(n_ev - 1);
    AZB_TRY(cudaEventRecord(start, cur));
    AZB_TRY(cudaStreamWaitEvent(s_in, start, 0));
    AZB_TRY(cudaStreamWaitEvent(s_out, start, 0));
    for (unsigned int c = 0; c < a->n_in_chunks; ++c)
        {
        const size_t lo = a->in_edges[c], up = a->in_edges[c + 1];
        if (up < lo)
            return AZB_ERR_BAD_ARGUMENT;
        for (unsigned int k = 0; k < a->n_inputs && up > lo; ++k)
            {
            const azb_host_buffer& b = a->inputs[k];
            AZB_TRY(cudaMemcpyAsync((char*)b.d_ptr + lo * b.row_bytes, (const char*)b.h_ptr + lo * b.row_bytes, (up - lo) * b.row_bytes,
                                    cudaMemcpyHostToDevice, s_in));
            }
        AZB_TRY(cudaEventRecord(g_pool.get(c), s_in));
        }
    unsigned int o = 0;
    for (unsigned int c = 0; c < a->n_launches; ++c)
        {
        const azb_host_launch& ln = a->launches[c];
        if (ln.wait_chunk >= a->n_in_chunks)
            return AZB_ERR_BAD_ARGUMENT;
        if (ln.tile_count)
            {
            AZB_TRY(cudaStreamWaitEvent(cur, g_pool.get(ln.wait_chunk), 0));
            range->tile_begin = ln.tile_begin;
            range->tile_count = ln.tile_count;
            if (a->kind == AZB_KIND_PAIR)
                AZB_TRY(azb_pair_forces(a->potential, a->scalar_bits, &pa, a->d_params, cur));
            else if (a->kind == AZB_KIND_DPD)
                AZB_TRY(azb_dpd_forces(a->potential, a->scalar_bits, &da, a->d_params, cur));
            else
                AZB_TRY(azb_aniso_forces(a->potential, a->scalar_bits, &aa, a->d_params, cur));
            }
        if (o < a->n_outs && a->outs[o].after_launch == c)
            {
            cudaEvent_t done = g_pool.get(a->n_in_chunks + c);
            AZB_TRY(cudaEventRecord(done, cur));
            AZB_TRY(cudaStreamWaitEvent(s_out, done, 0));
            for (; o < a->n_outs && a->outs[o].after_launch == c; ++o)
                {
                const azb_host_out_range& r = a->outs[o];
                if (r.row_hi < r.row_lo)
                    return AZB_ERR_BAD_ARGUMENT;
                for (unsigned int k = 0; k < a->n_outputs && r.row_hi > r.row_lo; ++k)
                    {
                    const azb_host_buffer& buf = a->outputs[k];
                    AZB_TRY(cudaMemcpyAsync((char*)buf.h_ptr + r.row_lo * buf.row_bytes, (const char*)buf.d_ptr + r.row_lo * buf.row_bytes,
                                            (r.row_hi - r.row_lo) * buf.row_bytes, cudaMemcpyDeviceToHost, s_out));
                    }
                }
            }
        }
    if (o != a->n_outs)
        return AZB_ERR_BAD_ARGUMENT; // outs must be sorted by after_launch, each naming an existing launch
    // the compute stream also owns the tail of the input copies (a chunk nobody waited for), then the host waits for the
    // last output row
    AZB_TRY(cudaStreamWaitEvent(cur, g_pool.get(a->n_in_chunks - 1), 0));
    const auto t_issued = std::chrono::steady_clock::now();
    AZB_TRY(cudaStreamSynchronize(s_out));
    AZB_TRY(cudaStreamSynchronize(cur));
    if (trace)
        std::fprintf(stderr, "azb_host_step: issue %.1f us, total %.1f us\n", std::chrono::duration<double, std::micro>(t_issued - t_begin).count(),
                     std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_begin).count());
    return 0;
    }
