// scripts/gram_variants.cu -- development aid: sweeps tile shapes of the K3 kernel on a B200.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I quemb_b200/csrc -I scripts scripts/gram_variants.cu quemb_b200/csrc/k3_gram.cu -lcuda -o gpurun_out/gram_variants
#include <cstdio>
#include <vector>
#include "gram_impl.cuh"
using namespace qemb;

__global__ void fillk(double *p, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = 1e-3 * (double)((i * 2654435761u) % 1000) - 0.5;
}

template <class C>
void run(const char *name, int64_t npair, int64_t naux, double *X, int64_t ldx, double *G, int group)
{
    std::vector<int2> h(gram_build_tiles<C>(npair, nullptr, group));
    gram_build_tiles<C>(npair, h.data(), group);
    int2 *dt; cudaMalloc(&dt, h.size() * sizeof(int2));
    cudaMemcpy(dt, h.data(), h.size() * sizeof(int2), cudaMemcpyHostToDevice);
    GramParams p; p.X = X; p.ldx = ldx; p.ktiles = (int)((naux + 15) / 16); p.npair = npair; p.G = G; p.ldg = npair;
    p.tiles = dt; p.ntiles = (int)h.size();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaError_t e = gram_launch<C>(p, 0);
    if (e != cudaSuccess) { printf("%-28s launch failed: %s\n", name, cudaGetErrorString(e)); return; }
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        cudaEventRecord(e0); gram_launch<C>(p, 0); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    e = cudaGetLastError();
    const double alg = (double)naux * npair * (npair + 1);
    const double exe = 2.0 * C::TM * C::TN * (double)(p.ktiles * 16) * h.size();
    printf("%-28s tiles %6zu smem %3d KB  %8.3f ms  algorithmic %6.2f TF/s  executed %6.2f TF/s  %s\n", name, h.size(),
           C::SMEM_BYTES / 1024, best, alg / best / 1e9, exe / best / 1e9, e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaFree(dt);
}

void run_tma(const char *name, int64_t npair, int64_t naux, double *X, int64_t ldx, double *G, int group, int sms)
{
    using C = GramCfg<128, 128, 64, 32, 16, 4, 1>;
    std::vector<int2> h(gram_build_tiles<C>(npair, nullptr, group));
    gram_build_tiles<C>(npair, h.data(), group);
    int2 *dt; cudaMalloc(&dt, h.size() * sizeof(int2));
    cudaMemcpy(dt, h.data(), h.size() * sizeof(int2), cudaMemcpyHostToDevice);
    GramParams p; p.X = X; p.ldx = ldx; p.ktiles = (int)((naux + 15) / 16); p.npair = npair; p.G = G; p.ldg = npair;
    p.tiles = dt; p.ntiles = (int)h.size();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaError_t e = launch_gram(p, sms, 0);
    if (e != cudaSuccess) { printf("%-28s launch failed: %s\n", name, cudaGetErrorString(e)); return; }
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%-28s run failed: %s\n", name, cudaGetErrorString(e)); return; }
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        cudaEventRecord(e0); launch_gram(p, sms, 0); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double alg = (double)naux * npair * (npair + 1);
    printf("%-28s tiles %6zu             %8.3f ms  algorithmic %6.2f TF/s\n", name, h.size(), best, alg / best / 1e9);
    // compare against the cp.async kernel
    double *G2; cudaMalloc(&G2, npair * npair * 8);
    p.G = G2; gram_launch<C>(p, 0); cudaDeviceSynchronize();
    std::vector<double> a(npair * 64), b(npair * 64);
    cudaMemcpy(a.data(), G + (npair / 2) * npair, a.size() * 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(b.data(), G2 + (npair / 2) * npair, b.size() * 8, cudaMemcpyDeviceToHost);
    double md = 0, mx = 0; for (size_t i = 0; i < a.size(); ++i) { md = fmax(md, fabs(a[i] - b[i])); mx = fmax(mx, fabs(b[i])); }
    printf("    max |G_tma - G_cpasync| over 64 columns = %.3e (max |G| %.3e)\n", md, mx);
    cudaFree(G2); cudaFree(dt);
}

int main(int argc, char **argv)
{
    const int64_t npair = argc > 1 ? atoll(argv[1]) : 24310, naux = argc > 2 ? atoll(argv[2]) : 2716;
    const int64_t ldx = (naux + 127) / 128 * 128;
    double *X, *G;
    cudaMalloc(&X, ldx * npair * 8); cudaMalloc(&G, npair * npair * 8);
    fillk<<<148 * 8, 256>>>(X, ldx * npair);
    cudaDeviceSynchronize();
    printf("npair %lld naux %lld\n", (long long)npair, (long long)naux);
    run_tma("TMA 128x128 persistent g12", npair, naux, X, ldx, G, 12, 148);
    run_tma("TMA 128x128 persistent g16", npair, naux, X, ldx, G, 16, 148);
    //            TM   TN   WM  WN  BK  ST MINB
    run<GramCfg<128, 128, 64, 32, 16, 4, 1>>("128x128 w64x32 k16 s4", npair, naux, X, ldx, G, 12);
    run<GramCfg<128, 128, 64, 32, 16, 3, 1>>("128x128 w64x32 k16 s3", npair, naux, X, ldx, G, 12);
    run<GramCfg<128, 128, 64, 32, 32, 3, 1>>("128x128 w64x32 k32 s3", npair, naux, X, ldx, G, 12);
    run<GramCfg<128, 128, 64, 32, 32, 2, 1>>("128x128 w64x32 k32 s2", npair, naux, X, ldx, G, 12);
    run<GramCfg<128, 128, 32, 32, 16, 4, 1>>("128x128 w32x32 k16 s4 (16w)", npair, naux, X, ldx, G, 12);
    run<GramCfg<128, 128, 32, 32, 32, 3, 1>>("128x128 w32x32 k32 s3 (16w)", npair, naux, X, ldx, G, 12);
    run<GramCfg<128, 128, 32, 64, 16, 4, 1>>("128x128 w32x64 k16 s4", npair, naux, X, ldx, G, 12);
    run<GramCfg<128, 64, 32, 32, 16, 3, 2>>("128x64 w32x32 k16 s3 x2", npair, naux, X, ldx, G, 12);
    run<GramCfg<128, 64, 32, 32, 32, 2, 2>>("128x64 w32x32 k32 s2 x2", npair, naux, X, ldx, G, 12);
    run<GramCfg<128, 64, 32, 32, 16, 4, 2>>("128x64 w32x32 k16 s4 x2", npair, naux, X, ldx, G, 12);
    run<GramCfg<64, 128, 32, 32, 16, 3, 2>>("64x128 w32x32 k16 s3 x2", npair, naux, X, ldx, G, 24);
    run<GramCfg<128, 128, 64, 32, 16, 4, 1>>("128x128 w64x32 k16 s4 g8", npair, naux, X, ldx, G, 8);
    run<GramCfg<128, 128, 64, 32, 16, 4, 1>>("128x128 w64x32 k16 s4 g16", npair, naux, X, ldx, G, 16);
    run<GramCfg<256, 128, 64, 64, 16, 3, 1>>("256x128 w64x64 k16 s3", npair, naux, X, ldx, G, 6);
    // independent CTAs per SM instead of one big one (barriers of different CTAs do not line up)
    run<GramCfg<64, 128, 32, 64, 16, 3, 2>>("64x128 w32x64 k16 s3 x2 (4w)", npair, naux, X, ldx, G, 24);
    run<GramCfg<128, 64, 64, 32, 16, 3, 2>>("128x64 w64x32 k16 s3 x2 (4w)", npair, naux, X, ldx, G, 12);
    run<GramCfg<128, 64, 64, 32, 16, 4, 2>>("128x64 w64x32 k16 s4 x2 (4w)", npair, naux, X, ldx, G, 12);
    run<GramCfg<128, 64, 64, 32, 32, 2, 2>>("128x64 w64x32 k32 s2 x2 (4w)", npair, naux, X, ldx, G, 12);
    run<GramCfg<64, 64, 32, 32, 16, 4, 3>>("64x64 w32x32 k16 s4 x3 (4w)", npair, naux, X, ldx, G, 24);
    run<GramCfg<64, 128, 32, 64, 16, 3, 2>>("64x128 w32x64 k16 s3 x2 g12", npair, naux, X, ldx, G, 12);
    return 0;
}
