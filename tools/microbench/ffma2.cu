// Micro-benchmark: throughput of scalar FFMA vs packed FFMA2 (fma.rn.f32x2) on sm_100a.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2 ffma2.cu ; prints Gflop-pairs/s
#include <cstdio>
#include <cuda_runtime.h>
template<int MODE> __global__ void k(float* out, int iters, float a, float b)
    {
    float2 x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        x[i] = make_float2(threadIdx.x * 1e-3f + i, threadIdx.x * 2e-3f - i);
    const float2 aa = make_float2(a, a), bb = make_float2(b, b);
    for (int it = 0; it < iters; ++it)
        {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            {
            if (MODE == 0)
                {
                x[i].x = fmaf(x[i].x, a, b);
                x[i].y = fmaf(x[i].y, a, b);
                }
            else
                x[i] = __ffma2_rn(x[i], aa, bb);
            }
        }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        s += x[i].x + x[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    }
int main()
    {
    float* out;
    cudaMalloc(&out, 148 * 8 * 256 * sizeof(float));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 20000;
    for (int mode = 0; mode < 2; ++mode)
        for (int rep = 0; rep < 2; ++rep)
            {
            cudaEventRecord(e0);
            if (mode == 0)
                k<0><<<148 * 8, 256>>>(out, iters, 1.0001f, 0.5f);
            else
                k<1><<<148 * 8, 256>>>(out, iters, 1.0001f, 0.5f);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            const double fmas = 148.0 * 8 * 256 * (double)iters * 16;
            printf("%s: %.3f ms, %.1f G fma/s (%.1f fma/clk/SM at 1.965 GHz)\n", mode ? "FFMA2" : "FFMA ", ms, fmas / ms * 1e-6,
                   fmas / (ms * 1e-3) / 148 / 1.965e9);
            }
    return 0;
    }
