// scripts/host_bw.cpp -- development aid: host memory bandwidth of the library's thread pool (memcpy) and the
// lower->upper mirror rate, per thread count.  g++ -O3 -pthread -Iquemb_b200/csrc scripts/host_bw.cpp quemb_b200/csrc/host_pool.cpp
#include "host_pool.h"
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
using namespace qemb;
#ifndef RBV
#define RBV 128
#define CBV 64
#endif
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main()
{
    const size_t n = 2ull << 30;
    char *a = (char *)aligned_alloc(4096, n), *b = (char *)aligned_alloc(4096, n);
    memset(a, 1, n); memset(b, 2, n);
    for (int i = 0; i < 3; i++) { double t = now(); parallel_memcpy(b, a, n); double dt = now() - t;
        printf("memcpy %.1f GB/s copied, pool threads %d\n", n / dt / 1e9, host_pool().size()); }
    free(a); free(b);
    const int64_t N = 16384;
    double *G = (double *)aligned_alloc(4096, N * N * 8);
    memset(G, 0, N * N * 8);
    { double t = now(); mirror_block(G, N, 0, 256, 256, N); double dt = now() - t; printf("mirror 1 thread: %.2f GB/s mirrored\n", 256.0 * (N - 256) * 8 / dt / 1e9); }
    for (int rep = 0; rep < 3; rep++) {
        std::mutex m; std::condition_variable cv; int64_t left = 0;
        const int64_t W = 256, R = 1024;
        for (int64_t c0 = 0; c0 < N; c0 += W) left += (N - c0 + R - 1) / R;
        double t = now();
        for (int64_t c0 = 0; c0 < N; c0 += W)
            for (int64_t r0 = c0; r0 < N; r0 += R)
                host_pool().submit([=, &m, &cv, &left] { mirror_block(G, N, c0, c0 + W, r0, r0 + R < N ? r0 + R : N);
                    std::lock_guard<std::mutex> lk(m); if (--left == 0) cv.notify_all(); });
        { std::unique_lock<std::mutex> lk(m); cv.wait(lk, [&] { return left == 0; }); }
        double dt = now() - t;
        printf("mirror %d threads (RB %d CB %d): %.1f GB/s mirrored\n", host_pool().size(), RBV, CBV, N * N * 4.0 / dt / 1e9);
    }
}
