// host_pool.h -- host-side helpers of libqemb_b200: a small persistent thread pool and the
// lower -> upper mirror of a symmetric matrix that arrived from the device as its lower triangle.
//
// Why: (ij|kl) is symmetric, and the PCIe link (53 GB/s measured) is the slowest stage of the
// all-fragment build once G has to reach host memory (4.4 GB per C32H66 fragment).  Only the lower
// triangle travels; the host cores rebuild the upper one while the next fragment computes.  The
// reference's own GPU variant does the same mirror pass single-threaded on the host after a full
// D2H (src/quemb/molbe/_cpp/eri_sparse_DF.cpp:684-708).
#pragma once
#include <cstdint>
#include <functional>

namespace qemb {

class HostPool {
public:
    explicit HostPool(int nthreads);
    ~HostPool();
    HostPool(const HostPool &) = delete;
    HostPool &operator=(const HostPool &) = delete;
    void submit(std::function<void()> f);
    int size() const { return n_; }

private:
    struct Impl;
    Impl *impl_;
    int n_;
};

// process-wide pool, created on first use; size = QEMB_HOST_THREADS or (cores in the affinity mask - 1), <= 32
HostPool &host_pool();

// G is n x n column-major (ld).  For every c in [c0, c1) and r in [r0, r1) with r > c:
//     G[c + ld * r] = G[r + ld * c]
// Reads run down the columns of the panel, writes are contiguous runs of (c1 - c0) doubles.
void mirror_block(double *G, int64_t ld, int64_t c0, int64_t c1, int64_t r0, int64_t r1);

// fn(i0, i1) over [0, n) in chunks of `grain` on the pool; blocks until done
void parallel_for(int64_t n, int64_t grain, const std::function<void(int64_t, int64_t)> &fn);

// fills `bytes` at p with zeros using non-temporal stores where aligned (host write-bandwidth probe)
void stream_fill(void *p, size_t bytes);

// multi-threaded memcpy (used to stage pageable host data through pinned memory); blocks until done
void parallel_memcpy(void *dst, const void *src, size_t bytes);

} // namespace qemb
