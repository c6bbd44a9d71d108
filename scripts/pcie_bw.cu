// scripts/pcie_bw.cu -- development aid: D2H / H2D bandwidth with pinned memory, 1-D vs 2-D copies
#include <cstdio>
#include <cuda_runtime.h>
int main() {
    const size_t n = 24310, bytes = n * n * 8;
    double *d, *h; cudaMalloc(&d, bytes); cudaMallocHost(&h, bytes);
    cudaMemset(d, 0, bytes);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1); float ms;
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0); cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, 0); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1); printf("D2H 1D  %.1f GB/s\n", bytes / ms / 1e6);
        cudaEventRecord(e0); cudaMemcpy2DAsync(h, n * 8, d, n * 8, n * 8, n, cudaMemcpyDeviceToHost, 0); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1); printf("D2H 2D  %.1f GB/s\n", bytes / ms / 1e6);
        cudaEventRecord(e0); cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, 0); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1); printf("H2D 1D  %.1f GB/s\n", bytes / ms / 1e6);
    }
    return 0;
}
