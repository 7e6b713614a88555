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
    void release()
        {
        for (cudaEvent_t e : ev)
            cudaEventDestroy(e);
        ev.clear();
        }
    ~EventPool()
        {
        // thread exit: the events go with the thread (a context torn down earlier makes the destroy a harmless error)
        release();
        }
    };
thread_local EventPool g_pool;

// One force object of the step: a private copy of its argument struct (tile_begin / tile_count change per launch).
struct ForceCall
    {
    int kind = 0, potential = 0;
    const void* d_params = nullptr;
    azb_pair_args pa;
    azb_dpd_args da;
    azb_aniso_args aa;
    azb_pair_args* range = nullptr;
    int init(int kind_, int potential_, const void* force_args, const void* d_params_)
        {
        kind = kind_;
        potential = potential_;
        d_params = d_params_;
        if (!force_args)
            return AZB_ERR_BAD_ARGUMENT;
        if (kind == AZB_KIND_PAIR)
            {
            pa = *(const azb_pair_args*)force_args;
            range = &pa;
            }
        else if (kind == AZB_KIND_DPD)
            {
            da = *(const azb_dpd_args*)force_args;
            range = &da.pair;
            }
        else if (kind == AZB_KIND_ANISO)
            {
            aa = *(const azb_aniso_args*)force_args;
            range = &aa.pair;
            }
        else
            return AZB_ERR_BAD_ARGUMENT;
        return range->tiles ? 0 : AZB_ERR_BAD_ARGUMENT;
        }
    int launch(int scalar_bits, unsigned int tile_begin, unsigned int tile_count, cudaStream_t s)
        {
        range->tile_begin = tile_begin;
        range->tile_count = tile_count;
        if (kind == AZB_KIND_PAIR)
            return azb_pair_forces(potential, scalar_bits, &pa, d_params, s);
        if (kind == AZB_KIND_DPD)
            return azb_dpd_forces(potential, scalar_bits, &da, d_params, s);
        return azb_aniso_forces(potential, scalar_bits, &aa, d_params, s);
        }
    };

int issue(const azb_host_step_args* a, std::vector<ForceCall>& calls, cudaStream_t cur, cudaStream_t s_in, cudaStream_t s_out)
    {
#define AZB_TRY(expr)                 \
    do                                \
        {                             \
        const int _e = (int)(expr);   \
        if (_e != 0)                  \
            return _e;                \
        } while (0)
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
            for (ForceCall& f : calls)
                AZB_TRY(f.launch(a->scalar_bits, ln.tile_begin, ln.tile_count, cur));
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
    // the compute stream also owns the tail of the input copies (a chunk nobody waited for)
    AZB_TRY(cudaStreamWaitEvent(cur, g_pool.get(a->n_in_chunks - 1), 0));
    return 0;
#undef AZB_TRY
    }
    } // namespace

extern "C" AZB_API void azb_host_step_release(void)
    {
    g_pool.release();
    }

extern "C" AZB_API int azb_host_step(const azb_host_step_args* a)
    {
    if (!a || !a->force_args || !a->in_edges || !a->launches || (a->n_outs && !a->outs) || a->n_inputs > 8 || a->n_outputs > 8 || a->n_in_chunks == 0
        || a->n_more > 7 || (a->n_more && !a->more))
        return AZB_ERR_BAD_ARGUMENT;
    cudaStream_t cur = (cudaStream_t)a->stream, s_in = (cudaStream_t)a->stream_in, s_out = (cudaStream_t)a->stream_out;
    std::vector<ForceCall> calls(1 + a->n_more);
    int rc = calls[0].init(a->kind, a->potential, a->force_args, a->d_params);
    for (unsigned int k = 0; k < a->n_more && rc == 0; ++k)
        rc = calls[1 + k].init(a->more[k].kind, a->more[k].potential, a->more[k].force_args, a->more[k].d_params);
    if (rc != 0)
        return rc;
    static const bool trace = std::getenv("AZB_TRACE_HOST_STEP") != nullptr; // prints how long the host spent issuing
    const auto t_begin = std::chrono::steady_clock::now();
    rc = issue(a, calls, cur, s_in, s_out);
    const auto t_issued = std::chrono::steady_clock::now();
    if (rc == 0 && (a->flags & AZB_HOST_STEP_NO_SYNC))
        return 0;
    // the host waits for the last output row — and, after an error in the middle of the sequence, for whatever copies and
    // launches were already queued: the caller's buffers must not be written behind its back
    const int e_in = (int)cudaStreamSynchronize(s_in), e_out = (int)cudaStreamSynchronize(s_out), e_cur = (int)cudaStreamSynchronize(cur);
    if (rc == 0)
        rc = e_in ? e_in : (e_out ? e_out : e_cur);
    if (trace)
        std::fprintf(stderr, "azb_host_step: issue %.1f us, total %.1f us\n", std::chrono::duration<double, std::micro>(t_issued - t_begin).count(),
                     std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_begin).count());
    return rc;
    }
