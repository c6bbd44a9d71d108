// FP64 FMA (non-tensor) issue rate on one GPU: independent DFMA chains per thread, all SMs busy.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dfma_rate dfma_rate.cu && ./dfma_rate
#include <cstdio>
#include <cuda_runtime.h>
template <int CH>
__global__ void k(double *out, int iters, double a, double b)
{
    double x[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) x[c] = threadIdx.x + c;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < CH; ++c) x[c] = fma(x[c], a, b);
    }
    double s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main()
{
    cudaDeviceProp pr;
    cudaGetDeviceProperties(&pr, 0);
    double *out;
    cudaMalloc(&out, 8ull * 148 * 8 * 1024);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 20000;
    for (int warps_per_sm : {4, 8, 16, 32, 64}) {
        const int threads = 256, blocks = pr.multiProcessorCount * warps_per_sm * 32 / threads;
        if (blocks == 0) continue;
        k<8><<<blocks, threads>>>(out, 100, 1.0000001, 1e-9);
        cudaEventRecord(e0);
        k<8><<<blocks, threads>>>(out, iters, 1.0000001, 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        const double fma = (double)blocks * threads * 8.0 * iters;
        printf("%s: %2d warps/SM  %8.3f ms  %.2f TFLOP/s FP64 FMA  (%.1f DFMA lanes/clk/SM at 1.965 GHz)\n", pr.name,
               warps_per_sm, ms, 2 * fma / ms / 1e9, fma / (ms * 1e-3) / pr.multiProcessorCount / 1.965e9);
    }
    return 0;
}
