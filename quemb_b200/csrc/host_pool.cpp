// host_pool.cpp -- see host_pool.h
#include "host_pool.h"

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

#include <immintrin.h>
#include <sched.h>

namespace qemb {

// ---------------------------------------------------------------------------------------------
// pool
// ---------------------------------------------------------------------------------------------
struct HostPool::Impl {
    std::mutex m;
    std::condition_variable cv;
    std::deque<std::function<void()>> q;
    std::vector<std::thread> th;
    bool stop = false;
    void run()
    {
        for (;;) {
            std::function<void()> f;
            {
                std::unique_lock<std::mutex> lk(m);
                cv.wait(lk, [&] { return stop || !q.empty(); });
                if (q.empty()) return; // stop requested and nothing left
                f = std::move(q.front());
                q.pop_front();
            }
            f();
        }
    }
};

HostPool::HostPool(int nthreads) : impl_(new Impl), n_(std::max(1, nthreads))
{
    for (int i = 0; i < n_; ++i) impl_->th.emplace_back([this] { impl_->run(); });
}

HostPool::~HostPool()
{
    {
        std::lock_guard<std::mutex> lk(impl_->m);
        impl_->stop = true;
    }
    impl_->cv.notify_all();
    for (auto &t : impl_->th) t.join();
    delete impl_;
}

void HostPool::submit(std::function<void()> f)
{
    {
        std::lock_guard<std::mutex> lk(impl_->m);
        impl_->q.push_back(std::move(f));
    }
    impl_->cv.notify_one();
}

static int default_threads()
{
    if (const char *e = std::getenv("QEMB_HOST_THREADS")) {
        const int v = std::atoi(e);
        if (v > 0) return std::min(v, 256);
    }
    int n = 0;
    cpu_set_t set;
    if (sched_getaffinity(0, sizeof(set), &set) == 0) n = CPU_COUNT(&set);
    if (n <= 0) n = (int)std::thread::hardware_concurrency();
    // one process per GPU (torchrun sets LOCAL_WORLD_SIZE): the ranks of a node share its cores and memory bandwidth
    int ranks = 1;
    if (const char *e = std::getenv("LOCAL_WORLD_SIZE")) ranks = std::max(1, std::atoi(e));
    // the callers' own threads sleep while outputs drain (blocking event waits), so with several ranks per node every
    // core is handed to a pool; a single rank leaves one core to its caller
    return std::min(std::max(ranks > 1 ? n / ranks : n - 1, 2), 32);
}

HostPool &host_pool()
{
    // leaked on purpose: worker threads must not be joined from a static destructor at exit
    static HostPool *pool = new HostPool(default_threads());
    return *pool;
}

// ---------------------------------------------------------------------------------------------
// mirror
// ---------------------------------------------------------------------------------------------
namespace {

#ifndef QEMB_MIRROR_RB
#define QEMB_MIRROR_RB 128
#endif
#ifndef QEMB_MIRROR_CB
#define QEMB_MIRROR_CB 64
#endif
constexpr int RB = QEMB_MIRROR_RB;    // rows (= destination columns) per pass: 1 KB contiguous read per panel column
constexpr int CBMAX = QEMB_MIRROR_CB; // panel columns (= destination run length) per pass

void mirror_scalar(double *G, int64_t ld, int64_t c0, int64_t c1, int64_t r0, int64_t r1)
{
    for (int64_t r = r0; r < r1; ++r) {
        double *dst = G + ld * r;
        const int64_t ce = std::min(c1, r); // r > c only
        for (int64_t c = c0; c < ce; ++c) dst[c] = G[r + ld * c];
    }
}

// contiguous run -> destination, non-temporal where 32-byte aligned
__attribute__((target("avx2"))) inline void store_run(double *dst, const double *src, int n)
{
    int i = 0;
    while (i < n && (reinterpret_cast<uintptr_t>(dst + i) & 31)) {
        dst[i] = src[i];
        ++i;
    }
    for (; i + 4 <= n; i += 4) _mm256_stream_pd(dst + i, _mm256_loadu_pd(src + i));
    for (; i < n; ++i) dst[i] = src[i];
}

// rows [r, r + RB) x columns [c0, c0 + nc), all strictly below the diagonal (r >= c0 + nc), nc % 4 == 0
__attribute__((target("avx2"))) void mirror_pass_avx2(double *G, int64_t ld, int64_t c0, int nc, int64_t r)
{
    alignas(64) double t[RB][CBMAX];
    for (int cc = 0; cc < nc; cc += 4) {
        const double *s0 = G + r + ld * (c0 + cc);
        const double *s1 = s0 + ld, *s2 = s1 + ld, *s3 = s2 + ld;
#pragma GCC unroll 4
        for (int rr = 0; rr < RB; rr += 4) {
            const __m256d a = _mm256_loadu_pd(s0 + rr), b = _mm256_loadu_pd(s1 + rr), c = _mm256_loadu_pd(s2 + rr),
                          d = _mm256_loadu_pd(s3 + rr);
            const __m256d ab_lo = _mm256_unpacklo_pd(a, b), ab_hi = _mm256_unpackhi_pd(a, b);
            const __m256d cd_lo = _mm256_unpacklo_pd(c, d), cd_hi = _mm256_unpackhi_pd(c, d);
            _mm256_store_pd(&t[rr + 0][cc], _mm256_permute2f128_pd(ab_lo, cd_lo, 0x20));
            _mm256_store_pd(&t[rr + 1][cc], _mm256_permute2f128_pd(ab_hi, cd_hi, 0x20));
            _mm256_store_pd(&t[rr + 2][cc], _mm256_permute2f128_pd(ab_lo, cd_lo, 0x31));
            _mm256_store_pd(&t[rr + 3][cc], _mm256_permute2f128_pd(ab_hi, cd_hi, 0x31));
        }
    }
    for (int rr = 0; rr < RB; ++rr) store_run(G + c0 + ld * (r + rr), t[rr], nc);
}

__attribute__((target("avx2"))) void mirror_avx2(double *G, int64_t ld, int64_t c0, int64_t c1, int64_t r0, int64_t r1)
{
    // rows that are not strictly below the panel go through the scalar path
    const int64_t rlow = std::max(r0, c1);
    if (r0 < rlow) mirror_scalar(G, ld, c0, c1, r0, std::min(rlow, r1));
    if (rlow >= r1) return;
    const int64_t rfull = rlow + (r1 - rlow) / RB * RB;
    for (int64_t cb = c0; cb < c1; cb += CBMAX) {
        const int64_t ce = std::min(c1, cb + CBMAX);
        const int nc4 = (int)((ce - cb) / 4 * 4);
        if (nc4 > 0)
            for (int64_t r = rlow; r < rfull; r += RB) mirror_pass_avx2(G, ld, cb, nc4, r);
        if (cb + nc4 < ce) mirror_scalar(G, ld, cb + nc4, ce, rlow, rfull);
    }
    if (rfull < r1) mirror_scalar(G, ld, c0, c1, rfull, r1);
    _mm_sfence();
}

// ---- AVX-512: 8 x 8 register transposes into an L1-resident tile, full-line non-temporal stores ---------------
// Measured on a Sapphire-Rapids-class Xeon (7 threads, n = 25425): the AVX2 version above moves 2.0 GB/s per core
// (a plain copy: ~7).  What limits a transposing copy is the access pattern, not the shuffles: reads want FEW LONG
// streams (8 columns x 2 KB at a time: 9 GB/s read-only, against 3.6 for 64 concurrent column streams of 512 B), the
// non-temporal writes want long contiguous runs (2 KB rows).  Both at once need a tile much larger than L1: 256 x 256
// doubles = 512 KB, resident in the core's 2 MB L2, filled 8 columns at a time -> 3.7-4.1 GB/s per core.
#ifndef QEMB_RB5
#define QEMB_RB5 256
#endif
#ifndef QEMB_CB5
#define QEMB_CB5 256
#endif
constexpr int RB5 = QEMB_RB5, CB5 = QEMB_CB5; // tile: 64 rows x 64 columns of doubles = 32 KB (L1d is 48 KB)

#define QEMB_AVX512 __attribute__((target("avx512f")))
#ifdef QEMB_MIRROR_CACHED
#define QEMB_ST512 _mm512_store_pd
#else
#define QEMB_ST512 _mm512_stream_pd
#endif

QEMB_AVX512 inline void store_run512(double *dst, const double *src, int n)
{
    int i = 0;
    while (i < n && (reinterpret_cast<uintptr_t>(dst + i) & 63)) {
        dst[i] = src[i];
        ++i;
    }
    for (; i + 8 <= n; i += 8) QEMB_ST512(dst + i, _mm512_loadu_pd(src + i));
    for (; i < n; ++i) dst[i] = src[i];
}

// rows [r, r + nr) x columns [c0, c0 + nc) (nr <= RB5, nr % 8 == 0, nc <= CB5, nc % 8 == 0), all strictly below the diagonal
QEMB_AVX512 void mirror_pass_avx512(double *G, int64_t ld, int64_t c0, int nc, int64_t r, int nr)
{
    alignas(64) static thread_local double t[RB5][CB5];
    for (int cc = 0; cc < nc; cc += 8) {
        const double *s = G + r + ld * (c0 + cc);
        for (int rr = 0; rr < nr; rr += 8) {
            const __m512d r0 = _mm512_loadu_pd(s + rr), r1 = _mm512_loadu_pd(s + rr + ld), r2 = _mm512_loadu_pd(s + rr + 2 * ld),
                          r3 = _mm512_loadu_pd(s + rr + 3 * ld), r4 = _mm512_loadu_pd(s + rr + 4 * ld),
                          r5 = _mm512_loadu_pd(s + rr + 5 * ld), r6 = _mm512_loadu_pd(s + rr + 6 * ld),
                          r7 = _mm512_loadu_pd(s + rr + 7 * ld);
            const __m512d t0 = _mm512_unpacklo_pd(r0, r1), t1 = _mm512_unpackhi_pd(r0, r1), t2 = _mm512_unpacklo_pd(r2, r3),
                          t3 = _mm512_unpackhi_pd(r2, r3), t4 = _mm512_unpacklo_pd(r4, r5), t5 = _mm512_unpackhi_pd(r4, r5),
                          t6 = _mm512_unpacklo_pd(r6, r7), t7 = _mm512_unpackhi_pd(r6, r7);
            const __m512d e0 = _mm512_shuffle_f64x2(t0, t2, 0x44), e1 = _mm512_shuffle_f64x2(t0, t2, 0xEE),
                          e2 = _mm512_shuffle_f64x2(t4, t6, 0x44), e3 = _mm512_shuffle_f64x2(t4, t6, 0xEE),
                          o0 = _mm512_shuffle_f64x2(t1, t3, 0x44), o1 = _mm512_shuffle_f64x2(t1, t3, 0xEE),
                          o2 = _mm512_shuffle_f64x2(t5, t7, 0x44), o3 = _mm512_shuffle_f64x2(t5, t7, 0xEE);
            _mm512_store_pd(&t[rr + 0][cc], _mm512_shuffle_f64x2(e0, e2, 0x88));
            _mm512_store_pd(&t[rr + 1][cc], _mm512_shuffle_f64x2(o0, o2, 0x88));
            _mm512_store_pd(&t[rr + 2][cc], _mm512_shuffle_f64x2(e0, e2, 0xDD));
            _mm512_store_pd(&t[rr + 3][cc], _mm512_shuffle_f64x2(o0, o2, 0xDD));
            _mm512_store_pd(&t[rr + 4][cc], _mm512_shuffle_f64x2(e1, e3, 0x88));
            _mm512_store_pd(&t[rr + 5][cc], _mm512_shuffle_f64x2(o1, o3, 0x88));
            _mm512_store_pd(&t[rr + 6][cc], _mm512_shuffle_f64x2(e1, e3, 0xDD));
            _mm512_store_pd(&t[rr + 7][cc], _mm512_shuffle_f64x2(o1, o3, 0xDD));
        }
    }
    for (int rr = 0; rr < nr; ++rr) store_run512(G + c0 + ld * (r + rr), t[rr], nc);
}

QEMB_AVX512 void mirror_avx512(double *G, int64_t ld, int64_t c0, int64_t c1, int64_t r0, int64_t r1)
{
    const int64_t rlow = std::max(r0, c1);
    if (r0 < rlow) mirror_scalar(G, ld, c0, c1, r0, std::min(rlow, r1));
    if (rlow >= r1) return;
    const int64_t rfull = rlow + (r1 - rlow) / 8 * 8; // whole 8-row groups; passes of up to RB5 rows
    for (int64_t cb = c0; cb < c1; cb += CB5) {
        const int64_t ce = std::min(c1, cb + CB5);
        const int nc8 = (int)((ce - cb) / 8 * 8);
        if (nc8 > 0)
            for (int64_t r = rlow; r < rfull; r += RB5)
                mirror_pass_avx512(G, ld, cb, nc8, r, (int)std::min<int64_t>(RB5, rfull - r));
        if (cb + nc8 < ce) mirror_scalar(G, ld, cb + nc8, ce, rlow, rfull);
    }
    if (rfull < r1) mirror_scalar(G, ld, c0, c1, rfull, r1);
    _mm_sfence();
}

} // namespace

void mirror_block(double *G, int64_t ld, int64_t c0, int64_t c1, int64_t r0, int64_t r1)
{
    if (c0 >= c1 || r0 >= r1) return;
    static const bool have_avx512 = __builtin_cpu_supports("avx512f") && !std::getenv("QEMB_NO_AVX512");
    static const bool have_avx2 = __builtin_cpu_supports("avx2");
    if (have_avx512)
        mirror_avx512(G, ld, c0, c1, r0, r1);
    else if (have_avx2)
        mirror_avx2(G, ld, c0, c1, r0, r1);
    else
        mirror_scalar(G, ld, c0, c1, r0, r1);
}

// ---------------------------------------------------------------------------------------------
// parallel for / memcpy
// ---------------------------------------------------------------------------------------------
void parallel_for(int64_t n, int64_t grain, const std::function<void(int64_t, int64_t)> &fn)
{
    if (n <= 0) return;
    HostPool &pool = host_pool();
    grain = std::max<int64_t>(grain, 1);
    const int64_t nchunks = (n + grain - 1) / grain;
    if (nchunks == 1 || pool.size() == 1) {
        fn(0, n);
        return;
    }
    std::mutex m;
    std::condition_variable cv;
    int64_t left = nchunks;
    for (int64_t i = 0; i < nchunks; ++i) {
        const int64_t i0 = i * grain, i1 = std::min(n, i0 + grain);
        pool.submit([=, &fn, &m, &cv, &left] {
            fn(i0, i1);
            std::lock_guard<std::mutex> lk(m);
            if (--left == 0) cv.notify_all();
        });
    }
    std::unique_lock<std::mutex> lk(m);
    cv.wait(lk, [&] { return left == 0; });
}

__attribute__((target("avx2"))) static void stream_fill_avx2(char *c, size_t bytes)
{
    size_t i = 0;
    while (i < bytes && (reinterpret_cast<uintptr_t>(c + i) & 31)) c[i++] = 0;
    const __m256i z = _mm256_setzero_si256();
    for (; i + 32 <= bytes; i += 32) _mm256_stream_si256(reinterpret_cast<__m256i *>(c + i), z);
    for (; i < bytes; ++i) c[i] = 0;
    _mm_sfence();
}

void stream_fill(void *p, size_t bytes)
{
    static const bool have_avx2 = __builtin_cpu_supports("avx2");
    if (have_avx2)
        stream_fill_avx2(static_cast<char *>(p), bytes);
    else
        std::memset(p, 0, bytes);
}

void parallel_memcpy(void *dst, const void *src, size_t bytes)
{
    const int64_t chunk = 4 << 20;
    parallel_for((int64_t)bytes, chunk, [=](int64_t b0, int64_t b1) {
        std::memcpy(static_cast<char *>(dst) + b0, static_cast<const char *>(src) + b0, (size_t)(b1 - b0));
    });
}

} // namespace qemb
