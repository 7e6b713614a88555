// azb_halo.cu — packing of boundary particles for the ghost halo exchange (SURVEY.md §8 e). HOOMD's CommunicatorGPU does
// this step over MPI; here either the packed buffer is handed to NCCL send/recv by the host layer (azplugins_b200/domain.py),
// or — ranks of one NVLink/NVSwitch node — a single kernel stores the rows straight into the neighbours' particle arrays
// and raises a flag there (azb_halo_push / azb_halo_wait / azb_halo_signal).
#include <cuda_runtime.h>

#include <cstring>
#include <map>
#include <mutex>
#include <string>

#include "azb200.h"
#include "azb_math.cuh"

namespace azb
    {
template<class S>
__global__ void gather_scalar4_kernel(const typename vec4<S>::type* __restrict__ src, const unsigned int* __restrict__ idx,
                                      const typename vec4<S>::type* __restrict__ shift, typename vec4<S>::type* __restrict__ out, size_t n)
    {
    const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n)
        return;
    auto v = ld4(src + __ldg(idx + e));
    if (shift)
        {
        const auto s = ld4(shift + e);
        v.x += s.x;
        v.y += s.y;
        v.z += s.z;
        }
    st4(out + e, v.x, v.y, v.z, v.w);
    }
    } // namespace azb

extern "C" int azb_gather_scalar4(int scalar_bits, const void* d_src, const unsigned int* d_idx, const void* d_shift, void* d_out, size_t n,
                                  void* stream)
    {
    if (n == 0)
        return 0;
    if (!d_src || !d_idx || !d_out)
        return AZB_ERR_BAD_ARGUMENT;
    const unsigned int block = 256;
    const unsigned int grid = (unsigned int)((n + block - 1) / block);
    cudaStream_t s = (cudaStream_t)stream;
    if (scalar_bits == 32)
        azb::gather_scalar4_kernel<float><<<grid, block, 0, s>>>((const float4*)d_src, d_idx, (const float4*)d_shift, (float4*)d_out, n);
    else if (scalar_bits == 64)
        azb::gather_scalar4_kernel<double><<<grid, block, 0, s>>>((const azb::double4a*)d_src, d_idx, (const azb::double4a*)d_shift,
                                                                  (azb::double4a*)d_out, n);
    else
        return AZB_ERR_BAD_SCALAR;
    return (int)cudaGetLastError();
    }

// ---- halo over peer memory -------------------------------------------------------------------------------------------------
namespace azb
    {
__device__ __forceinline__ void store_release_sys(unsigned int* p, unsigned int v)
    {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
    }
__device__ __forceinline__ unsigned int load_acquire_sys(const unsigned int* p)
    {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
    }
__device__ __forceinline__ unsigned long long global_timer_ns()
    {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
    }

constexpr unsigned int HALO_MAX_ROUTES = 128; // 26 directions x (pos, vel | orientation) = 52 .. 78

// One thread per row: routes are found by a search over the staged prefix table, rows go out as 16- or 32-byte stores that
// are consecutive within a route (coalesced on the NVLink side). The last CTA to finish raises the receivers' flags.
template<class S> __global__ void __launch_bounds__(256) halo_push_kernel(const azb_halo_push_args a)
    {
    typedef typename vec4<S>::type S4;
    __shared__ unsigned int s_begin[HALO_MAX_ROUTES + 1];
    __shared__ bool s_last;
    for (unsigned int r = threadIdx.x; r <= a.n_routes; r += blockDim.x)
        s_begin[r] = r < a.n_routes ? a.d_routes[r].row_begin : a.total_rows;
    __syncthreads();
    const unsigned int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < a.total_rows)
        {
        unsigned int lo = 0, hi = a.n_routes; // last route with row_begin <= e (empty routes share a row_begin: skip them)
        while (hi - lo > 1)
            {
            const unsigned int mid = (lo + hi) >> 1;
            if (s_begin[mid] <= e)
                lo = mid;
            else
                hi = mid;
            }
        const azb_halo_route& rt = a.d_routes[lo];
        const unsigned int k = e - s_begin[lo];
        auto v = ld4((const S4*)rt.d_src + __ldg(rt.d_idx + k));
        st4((S4*)rt.d_dst + k, v.x + (S)rt.shift[0], v.y + (S)rt.shift[1], v.z + (S)rt.shift[2], v.w);
        }
    // all rows of this CTA are out before its arrival is counted; the last arrival publishes the epoch to every receiver
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0)
        {
        const unsigned int arrived = atomicAdd(a.d_counter, 1u);
        s_last = (arrived == gridDim.x - 1u);
        }
    __syncthreads();
    if (s_last)
        {
        __threadfence_system();
        for (unsigned int i = threadIdx.x; i < a.n_flags; i += blockDim.x)
            store_release_sys(a.d_peer_flags[i], a.epoch);
        if (threadIdx.x == 0)
            *a.d_counter = 0u;
        }
    }

__global__ void halo_wait_kernel(const unsigned int* flags, unsigned int n, unsigned int epoch, unsigned long long timeout_ns, unsigned int* error)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const unsigned long long t0 = global_timer_ns();
    while ((int)(load_acquire_sys(flags + i) - epoch) < 0)
        {
        if (global_timer_ns() - t0 > timeout_ns)
            {
            atomicExch(error, 1u);
            break;
            }
        __nanosleep(200);
        }
    }

__global__ void halo_signal_kernel(unsigned int* const* peer_flags, unsigned int n, unsigned int epoch)
    {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
        store_release_sys(peer_flags[i], epoch);
    }

typedef int (*mem_get_address_range_t)(unsigned long long*, size_t*, unsigned long long);
static std::mutex g_ipc_mutex;
static std::map<std::string, void*> g_ipc_open; // handle bytes -> mapped base
    } // namespace azb

extern "C" int azb_ipc_export(const void* d_ptr, unsigned char* handle, size_t* offset)
    {
    if (!d_ptr || !handle || !offset)
        return AZB_ERR_BAD_ARGUMENT;
    // the IPC handle names the whole allocation: find its base through the driver (no link-time libcuda dependency)
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult found;
    cudaError_t err = cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &found);
    if (err != cudaSuccess)
        return (int)err;
    if (!fn || found != cudaDriverEntryPointSuccess)
        return (int)cudaErrorNotSupported;
    unsigned long long base = 0;
    size_t size = 0;
    if (((azb::mem_get_address_range_t)fn)(&base, &size, (unsigned long long)(uintptr_t)d_ptr) != 0)
        return (int)cudaErrorInvalidValue;
    cudaIpcMemHandle_t h;
    err = cudaIpcGetMemHandle(&h, (void*)(uintptr_t)base);
    if (err != cudaSuccess)
        return (int)err;
    static_assert(sizeof(h) == AZB_IPC_HANDLE_BYTES, "IPC handle size");
    std::memcpy(handle, &h, sizeof(h));
    *offset = (size_t)((unsigned long long)(uintptr_t)d_ptr - base);
    return 0;
    }

extern "C" int azb_ipc_open(const unsigned char* handle, size_t offset, void** d_ptr)
    {
    if (!handle || !d_ptr)
        return AZB_ERR_BAD_ARGUMENT;
    std::lock_guard<std::mutex> lock(azb::g_ipc_mutex);
    const std::string key((const char*)handle, AZB_IPC_HANDLE_BYTES);
    auto it = azb::g_ipc_open.find(key);
    void* base = nullptr;
    if (it == azb::g_ipc_open.end())
        {
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handle, sizeof(h));
        const cudaError_t err = cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess);
        if (err != cudaSuccess)
            return (int)err;
        azb::g_ipc_open[key] = base;
        }
    else
        base = it->second;
    *d_ptr = (char*)base + offset;
    return 0;
    }

extern "C" int azb_halo_push(int scalar_bits, const azb_halo_push_args* a, void* stream)
    {
    if (!a || !a->d_counter || (a->n_routes && !a->d_routes) || (a->n_flags && !a->d_peer_flags) || a->n_routes > azb::HALO_MAX_ROUTES)
        return AZB_ERR_BAD_ARGUMENT;
    const unsigned int block = 256;
    const unsigned int grid = a->total_rows ? (a->total_rows + block - 1) / block : 1u;
    cudaStream_t s = (cudaStream_t)stream;
    if (scalar_bits == 32)
        azb::halo_push_kernel<float><<<grid, block, 0, s>>>(*a);
    else if (scalar_bits == 64)
        azb::halo_push_kernel<double><<<grid, block, 0, s>>>(*a);
    else
        return AZB_ERR_BAD_SCALAR;
    return (int)cudaGetLastError();
    }

extern "C" int azb_halo_wait(const unsigned int* d_flags, unsigned int n, unsigned int epoch, unsigned int timeout_ms, unsigned int* d_error,
                             void* stream)
    {
    if (n == 0)
        return 0;
    if (!d_flags || !d_error)
        return AZB_ERR_BAD_ARGUMENT;
    azb::halo_wait_kernel<<<(n + 31) / 32, 32, 0, (cudaStream_t)stream>>>(d_flags, n, epoch, (unsigned long long)timeout_ms * 1000000ull, d_error);
    return (int)cudaGetLastError();
    }

extern "C" int azb_halo_signal(unsigned int* const* d_peer_flags, unsigned int n, unsigned int epoch, void* stream)
    {
    if (n == 0)
        return 0;
    if (!d_peer_flags)
        return AZB_ERR_BAD_ARGUMENT;
    azb::halo_signal_kernel<<<(n + 31) / 32, 32, 0, (cudaStream_t)stream>>>(d_peer_flags, n, epoch);
    return (int)cudaGetLastError();
    }
