// Micro-benchmark: throughput of the XU pipe (MUFU.EX2 / MUFU.RSQ / MUFU.RCP) on sm_100a, alone and mixed with packed FFMA2 at
// the ratio of the ExpandedYukawa tile kernel's inner loop (6 MUFU in ~60 issue slots per pair of list entries).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu mufu.cu ; prints warp-instructions per clock and SM
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float ex2(float x)
    {
    float r;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
    }
__device__ __forceinline__ float rsq(float x)
    {
    float r;
    asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
    }
__device__ __forceinline__ float rcpf(float x)
    {
    float r;
    asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
    }
// MODE 0: ex2 only, 1: rsqrt only, 2: rcp only, 3: 6 MUFU + NF packed FMAs per iteration (independent chains)
template<int MODE, int NF> __global__ void k(float* out, int iters, float a, float b)
    {
    float m[6];
    float2 x[8];
#pragma unroll
    for (int i = 0; i < 6; ++i)
        m[i] = 0.5f + threadIdx.x * 1e-3f + i;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        x[i] = make_float2(threadIdx.x * 1e-3f + i, threadIdx.x * 2e-3f - i);
    const float2 aa = make_float2(a, a), bb = make_float2(b, b);
    for (int it = 0; it < iters; ++it)
        {
#pragma unroll
        for (int i = 0; i < 6; ++i)
            m[i] = MODE == 0 ? ex2(m[i]) : MODE == 1 ? rsq(m[i]) : MODE == 2 ? rcpf(m[i]) : (i < 2 ? rsq(m[i]) : i < 4 ? rcpf(m[i]) : ex2(m[i]));
        if (MODE == 3)
            {
#pragma unroll
            for (int i = 0; i < NF; ++i)
                x[i % 8] = __ffma2_rn(x[i % 8], aa, bb);
            }
        }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i)
        s += m[i];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        s += x[i].x + x[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    }
template<int MODE, int NF> void run(const char* name, float* out)
    {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 20000;
    for (int rep = 0; rep < 2; ++rep)
        {
        cudaEventRecord(e0);
        k<MODE, NF><<<148 * 8, 256>>>(out, iters, 1.0001f, 0.5f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        const double warps = 148.0 * 8 * 8;
        const double mufu = warps * iters * 6, all = warps * iters * (6 + (MODE == 3 ? NF : 0));
        const double clk = ms * 1e-3 * 1.965e9 * 148;
        if (rep)
            printf("%-28s %.3f ms: %.3f MUFU warp-instr/clk/SM (%.1f lanes/clk/SM), %.3f warp-instr/clk/SM in all\n", name, ms, mufu / clk,
                   32 * mufu / clk, all / clk);
        }
    }
int main()
    {
    float* out;
    cudaMalloc(&out, 148 * 8 * 256 * sizeof(float));
    run<0, 0>("ex2 only", out);
    run<1, 0>("rsqrt only", out);
    run<2, 0>("rcp only", out);
    run<3, 24>("6 MUFU + 24 FFMA2", out);
    run<3, 40>("6 MUFU + 40 FFMA2", out);
    run<3, 54>("6 MUFU + 54 FFMA2", out);
    run<3, 66>("6 MUFU + 66 FFMA2", out);
    return 0;
    }
