// api.cu -- host side of libqemb_b200: handles, index maps, per-fragment plans and the C ABI declared in
// include/qemb_b200.h.  Reference citations: "cpp:" = src/quemb/molbe/_cpp/eri_sparse_DF.cpp,
// "hpp:" = src/quemb/molbe/_cpp/indexers.hpp of troyvvgroup/quemb.
//
// Streams:
//   ctx compute  tensor / metric level work (uploads, block staging, metric application), in order
//   ctx copy     device -> host copy of a finished G while the next fragment computes
//   ws compute   K1, K2a, K2b, K3 (and the J/K consumer) of the fragments of ONE workspace, in order
//   ws prep      TA upload, screening kernel, plan upload of that workspace's NEXT fragment
// Every workspace owns its two streams, so concurrent callers (one workspace each; the reference's
// ThreadPoolExecutor(n_threads), eri_sparse_DF.py:649-653) overlap on the device and fragments too small to fill
// 148 SMs share them.  A fragment forks from the ctx compute stream (it sees every tensor update enqueued before
// it); tensor updates, marks and stream hand-offs join the workspace streams first (fork_from_ctx / join_workspaces).
// Fragment-local inputs (TA, mask, plan tables) are double buffered ("slots"), and so is G when it has
// to travel to the host, so that back-to-back qemb_fragment_transform calls pipeline without any
// host-side wait on the compute stream.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <condition_variable>
#include <cstring>
#include <limits>
#include <map>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/qemb_b200.h"
#include "common.cuh"
#include "host_pool.h"
#include "kernels.cuh"

#include <nvtx3/nvToolsExt.h> // header-only NVTX v3: ranges show up under nsys / ncu --nvtx, no-ops otherwise

using namespace qemb;

// ------------------------------------------------------------------------------------------------
// log level (hpp:24-46: Python logging levels, process global) and NVTX ranges per step
// ------------------------------------------------------------------------------------------------
static std::atomic<int> g_log_level{40}; // LogLevel::Error, the reference's default (hpp:34)

extern "C" int qemb_get_log_level(void) { return g_log_level.load(std::memory_order_relaxed); }
extern "C" int qemb_set_log_level(int level)
{
    g_log_level.store(level, std::memory_order_relaxed);
    return QEMB_OK;
}
static inline bool log_info() { return g_log_level.load(std::memory_order_relaxed) <= 20; }

namespace {
struct NvtxRange {
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};
} // namespace

// ------------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_err;

static int fail(int code, const std::string &msg)
{
    g_err = msg;
    return code;
}
#define CU(call)                                                                                        \
    do {                                                                                                \
        cudaError_t e__ = (call);                                                                       \
        if (e__ != cudaSuccess) {                                                                       \
            const int code__ = (e__ == cudaErrorMemoryAllocation) ? QEMB_ENOMEM                         \
                               : (e__ == cudaErrorNoDevice || e__ == cudaErrorInsufficientDriver)       \
                                   ? QEMB_ENODEV                                                        \
                                   : QEMB_ECUDA;                                                        \
            return fail(code__, std::string(#call) + ": " + cudaGetErrorString(e__));                   \
        }                                                                                               \
    } while (0)
#define REQ(cond, msg)                                                                                  \
    do {                                                                                                \
        if (!(cond)) return fail(QEMB_EINVAL, std::string(msg) + " [" #cond "]");                       \
    } while (0)
#define TRY(expr)                                                                                       \
    do {                                                                                                \
        int rc__ = (expr);                                                                              \
        if (rc__ != QEMB_OK) return rc__;                                                               \
    } while (0)

// for the other translation units of the library (gram_noalloc.cu): same thread-local message, same codes
namespace qemb {
int api_fail(int code, const char *msg) { return fail(code, msg ? msg : ""); }
} // namespace qemb

extern "C" const char *qemb_last_error(void) { return g_err.c_str(); }
extern "C" const char *qemb_version(void) { return "quemb-b200 0.1 (sm_100a, FP64 DMMA)"; }

// ------------------------------------------------------------------------------------------------
// handles
// ------------------------------------------------------------------------------------------------
struct qemb_ctx {
    int device = 0;
    cudaStream_t compute = nullptr, prep = nullptr, copy = nullptr;
    bool timing = false;
    std::atomic<int64_t> launches{0};
    cudaEvent_t marks[16] = {};
    // how a host-bound G travels: 1 = lower triangle over PCIe + mirror on the host cores (default),
    // 0 = the full square over PCIe
    int out_mode = 1;
    // mode 1 only: the last `full_frac` of the columns cross PCIe whole (rows 0..n) and are not mirrored on the host:
    // balances the copy engine (slack) against the host threads (the bottleneck of a pure triangle transfer)
    double full_frac = 0.0;
    // bytes per copy group of a triangle transfer: the host starts mirroring a group as soon as it has landed
    int64_t out_group_bytes = 256ll << 20; // measured at N = 1 (profiles/e2e_group_size_r2_n1.txt): 8 MB 2.78 s ... 64 MB 1.95 s, 256 MB 1.85 s per build
    // 0: K3 on the FP64 tensor path (default, the parity reference); 5 .. 8: opt-in int8-slice K3 (k3_int8.cu)
    int k3_slices = 0;
    // > 0: guaranteed-tolerance mode of the int8-slice K3: per fragment the smallest slice count whose worst-case bound
    // meets this ABSOLUTE tolerance on every (ij|kl), the FP64 kernel when none does (overrides k3_slices)
    double k3_tol = 0.0;
    double tune_p0 = 0.0; // device -> host rate with idle host threads, last autotune
    double tune_secs[3] = {-1.0, -1.0, -1.0}; // trial deliveries of the last autotune: full squares, triangle, model split
    int num_sms = 148;
    std::atomic<int64_t> h2d_bytes{0}, d2h_bytes{0}; // bytes handed to the copy engines (uploads / G transfers)
    // every workspace runs its fragments on streams of its own (concurrent callers overlap on the device, small
    // fragments that do not fill 148 SMs share them); operations on the context's compute stream that must see or
    // must not overtake fragment work join these streams first (join_workspaces)
    std::mutex ws_mutex;
    std::vector<qemb_workspace *> workspaces;
    // host-bound G transfers whose host mirror has not finished yet (qemb_ctx_synchronize waits for them)
    std::mutex out_mutex;
    std::condition_variable out_cv;
    int64_t out_pending = 0;
    // pinned bounce buffers for uploads out of pageable host memory
    void *bounce[2] = {nullptr, nullptr};
    cudaEvent_t bounce_free[2] = {nullptr, nullptr};
    std::mutex bounce_mutex;
};

namespace {

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes)
    {
        if (bytes <= cap) return QEMB_OK;
        if (p) {
            CU(cudaFree(p));
            p = nullptr;
            cap = 0;
        }
        const size_t want = bytes + bytes / 16;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) {
            (void)cudaGetLastError();
            e = cudaMalloc(&p, bytes);
            if (e != cudaSuccess) {
                (void)cudaGetLastError();
                p = nullptr;
                return fail(QEMB_ENOMEM, "cudaMalloc of " + std::to_string(bytes) + " bytes failed");
            }
            cap = bytes;
        } else {
            cap = want;
        }
        return QEMB_OK;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T>
    T *as() const
    {
        return static_cast<T *>(p);
    }
};

struct HostBuf { // pinned
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes)
    {
        if (bytes <= cap) return QEMB_OK;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        const size_t want = bytes + bytes / 4 + 4096;
        CU(cudaHostAlloc(&p, want, cudaHostAllocMapped | cudaHostAllocPortable));
        cap = want;
        return QEMB_OK;
    }
    void release()
    {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
};

inline int64_t gauss_sum(int64_t n) { return n * (n + 1) / 2; }

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev)
    {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
    }
    ~DeviceGuard()
    {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

} // namespace

namespace {

constexpr size_t BOUNCE_BYTES = 64u << 20;

// host (pitch spitch) -> device (pitch dpitch), `height` rows of `width` bytes, on stream s.  Pinned sources go
// straight to the copy engine; pageable ones are packed into pinned bounce buffers by the host pool (60 GB/s
// measured on the box vs ~10 GB/s for the driver's own pageable path) while the previous chunk is in flight.
int upload_2d(qemb_ctx *c, void *dst, size_t dpitch, const void *src, size_t spitch, size_t width, size_t height,
              cudaStream_t s)
{
    if (width == 0 || height == 0) return QEMB_OK;
    c->h2d_bytes += (int64_t)(width * height);
    cudaPointerAttributes at;
    cudaError_t e = cudaPointerGetAttributes(&at, src);
    if (e != cudaSuccess) (void)cudaGetLastError();
    const bool pinned = (e == cudaSuccess && at.type == cudaMemoryTypeHost);
    if (pinned || width * height < (8u << 20) || width > BOUNCE_BYTES) {
        CU(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, height, cudaMemcpyHostToDevice, s));
        return QEMB_OK;
    }
    std::lock_guard<std::mutex> lk(c->bounce_mutex);
    for (int i = 0; i < 2; ++i)
        if (!c->bounce[i]) {
            CU(cudaHostAlloc(&c->bounce[i], BOUNCE_BYTES, cudaHostAllocPortable));
            CU(cudaEventCreateWithFlags(&c->bounce_free[i], cudaEventDisableTiming));
        }
    const size_t rows_per = std::max<size_t>(1, BOUNCE_BYTES / width);
    int k = 0;
    for (size_t r0 = 0; r0 < height; r0 += rows_per, k ^= 1) {
        const size_t nr = std::min(rows_per, height - r0);
        CU(cudaEventSynchronize(c->bounce_free[k])); // a never-recorded event counts as complete
        char *b = static_cast<char *>(c->bounce[k]);
        const char *sp = static_cast<const char *>(src) + r0 * spitch;
        parallel_for((int64_t)nr, std::max<int64_t>(1, (int64_t)((2u << 20) / width)), [=](int64_t i0, int64_t i1) {
            for (int64_t i = i0; i < i1; ++i) std::memcpy(b + (size_t)i * width, sp + (size_t)i * spitch, width);
        });
        CU(cudaMemcpy2DAsync(static_cast<char *>(dst) + r0 * dpitch, dpitch, b, width, width, nr, cudaMemcpyHostToDevice, s));
        CU(cudaEventRecord(c->bounce_free[k], s));
    }
    return QEMB_OK;
}

} // namespace

struct qemb_tensor {
    qemb_ctx *ctx = nullptr;
    int64_t naux = 0, nao = 0, n_unique = 0;
    int64_t ldp = 0;    // leading dimension on the device: naux rounded up to 128, pad rows zero
    int64_t naux16 = 0; // naux rounded up to 16: rows the kernels touch
    std::vector<int64_t> rptr, ridx, rcol; // exch_reachable lists + column of every entry (cpp:260-279)
    std::vector<int64_t> uptr, uidx;       // exch_reachable_unique (hpp:149-162)
    DevBuf data, d_rcol, d_ridx;           // FP64 ldp x n_unique ; int32 copies of rcol / ridx
    DevBuf scratch;                        // column chunk being whitened (qemb_tensor_upload_whitened_rows)
    // (mu nu|P) block staging (qemb_tensor_scatter_block): two pinned host slots + two device slots, so that
    // the caller evaluates libcint block b+1 while block b is copied and scattered
    HostBuf stage_h[2];
    DevBuf stage_d[2];
    cudaEvent_t stage_done[2] = {nullptr, nullptr};
    int stage_k = 0;
    DevBuf nan_count;
    bool whitened = false;
};

struct qemb_metric {
    qemb_ctx *ctx = nullptr;
    int64_t naux = 0, naux_pad = 0;
    int nblocks = 0;
    DevBuf L, Dinv;
};

struct qemb_screen {
    qemb_ctx *ctx = nullptr;
    int64_t nao = 0;
    DevBuf S;
};

namespace {

struct PlanDev { // device pointers into a slot's plan buffer
    const int *bcol1 = nullptr, *ccol1 = nullptr;
    const GatherGroup *groups1 = nullptr;
    const int4 *tiles1 = nullptr;
    int ntiles1 = 0;
    const int *acol2 = nullptr, *brow2 = nullptr, *bcol2 = nullptr, *ccol2 = nullptr;
    const GatherGroup *groups2 = nullptr;
    const int4 *tiles2 = nullptr;
    int ntiles2 = 0;
};

struct Slot {
    int64_t nao = 0, nmo = 0, npair = 0, nA = 0;
    int64_t n_borderline = 0; // screened entries within 1e-12 of the threshold (last qemb_fragment_screen)
    std::vector<int64_t> aptr, aidx;
    DevBuf TA, mask, plan1, plan2;
    HostBuf stage1, stage2, hmask;
    PlanDev pd;
    cudaEvent_t uploaded1 = nullptr, uploaded2 = nullptr, consumed = nullptr, ta_ready = nullptr;
    bool consumed_pending = false;
    bool have_plan1 = false, have_plan2 = false;
    const qemb_tensor *plan1_tensor = nullptr;
    double work[8] = {0, 0, 0, 0, 0, 0, 0, 0};
};

// completion state of one host-bound G (lower-triangle copy groups + host mirror tasks)
struct OutJob {
    std::mutex m;
    std::condition_variable cv;
    int64_t remaining = 0; // outstanding mirror tasks + 1 while copy groups are still in flight
    bool active = false;   // this ticket travels as a triangle (wait on the job, not only on the event)
    bool counted = false;  // ctx->out_pending holds one count for this job (dropped exactly once)
    qemb_ctx *ctx = nullptr;
};

// called with job.m held, when job.remaining has just dropped to 0
inline void job_finished(OutJob &job)
{
    job.cv.notify_all();
    if (job.ctx && job.counted) {
        job.counted = false;
        std::lock_guard<std::mutex> lk(job.ctx->out_mutex);
        if (--job.ctx->out_pending == 0) job.ctx->out_cv.notify_all();
    }
}
constexpr int OUT_RING = 4;

inline void wait_job(OutJob &job)
{
    std::unique_lock<std::mutex> lk(job.m);
    job.cv.wait(lk, [&] { return job.remaining == 0; });
    job.active = false;
}

} // namespace

struct qemb_workspace {
    qemb_ctx *ctx = nullptr;
    cudaStream_t compute = nullptr, prep = nullptr; // this workspace's own streams
    cudaEvent_t fork_ev = nullptr, join_ev = nullptr;
    Slot slot[2];
    int cur = 0;
    bool started = false;
    // intermediates of the current fragment (compute stream order protects them)
    DevBuf Bint, sym;
    int64_t ldp = 0, mrows = 0; // leading dimension / valid rows of Bint and sym for the current fragment
    int64_t naux = 0;
    bool have_B = false, have_sym = false, solved = false;
    // G
    DevBuf G[2];
    cudaEvent_t g_copied[2] = {nullptr, nullptr};
    bool g_copy_pending[2] = {false, false};
    int gi = 0;
    int64_t ticket = 0;        // number of host-output transforms issued
    int ticket_gi[2] = {0, 0};
    OutJob jobs[OUT_RING];     // by ticket % OUT_RING
    cudaEvent_t out_done[OUT_RING] = {}; // copy-stream event after the last copy of ticket % OUT_RING
    double *G_last = nullptr; // where the last K3 wrote
    int64_t G_last_ld = 0;
    bool have_G = false;
    DevBuf gram_partial; // split-K partial tiles of K3 for small fragments
    // J/K consumer (jk.cu)
    DevBuf jk_in, jk_scratch, jk_out;
    DevBuf gram_i8; // scratch of the opt-in int8-slice K3
    DevBuf k3_expo;       // device int: largest column exponent of X (guaranteed-tolerance mode)
    HostBuf k3_expo_host; // its pinned landing place
    int k3_last_slices = 0, k3_last_expo = 0; // what the last K3 of this workspace ran with (0 = FP64 kernel)
    cudaEvent_t jk_ev[2] = {nullptr, nullptr};
    float jk_ms = 0.f;
    HostBuf jk_stage[2];
    cudaEvent_t jk_stage_free[2] = {nullptr, nullptr};
    int jk_stage_k = 0;
    // gram tile lists by number of row tiles
    std::map<int, std::pair<DevBuf, int>> gram_tiles;
    // timing
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t evs[2] = {nullptr, nullptr};
    cudaEvent_t ev2s = nullptr; // start of K2a (its launch follows host-side plan building, unlike the others)
    bool ev2s_valid = false;
    bool ev_valid[5] = {false, false, false, false, false};
    bool evs_valid = false;
};

namespace {

// the workspace's streams continue behind everything enqueued on the context's compute stream so far (tensor
// uploads, metric application, collectives handed over with qemb_ctx_stream_handoff)
int fork_from_ctx(qemb_workspace *w)
{
    CU(cudaEventRecord(w->fork_ev, w->ctx->compute));
    CU(cudaStreamWaitEvent(w->compute, w->fork_ev, 0));
    CU(cudaStreamWaitEvent(w->prep, w->fork_ev, 0));
    return QEMB_OK;
}

// the context's compute stream continues behind the fragment work of every workspace enqueued so far
int join_workspaces(qemb_ctx *c)
{
    std::lock_guard<std::mutex> lk(c->ws_mutex);
    for (qemb_workspace *w : c->workspaces) {
        CU(cudaEventRecord(w->join_ev, w->compute));
        CU(cudaStreamWaitEvent(c->compute, w->join_ev, 0));
    }
    return QEMB_OK;
}

int sync_workspaces(qemb_ctx *c)
{
    std::lock_guard<std::mutex> lk(c->ws_mutex);
    for (qemb_workspace *w : c->workspaces) {
        CU(cudaStreamSynchronize(w->prep));
        CU(cudaStreamSynchronize(w->compute));
    }
    return QEMB_OK;
}

} // namespace

// ------------------------------------------------------------------------------------------------
// device / ctx
// ------------------------------------------------------------------------------------------------
extern "C" int qemb_device_count(int *n)
{
    REQ(n != nullptr, "null output");
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        *n = 0;
        return fail(QEMB_ENODEV, std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e));
    }
    *n = c;
    return QEMB_OK;
}

extern "C" int qemb_ctx_create(int device, qemb_ctx **out)
{
    REQ(out != nullptr, "null output");
    *out = nullptr;
    int n = 0;
    TRY(qemb_device_count(&n));
    if (n <= 0) return fail(QEMB_ENODEV, "no CUDA device visible: libqemb_b200 has no CPU fallback");
    REQ(device >= 0 && device < n, "device index out of range");
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(QEMB_ENODEV, std::string("device '") + prop.name + "' is sm_" + std::to_string(prop.major) +
                                     std::to_string(prop.minor) + "; this library is built for sm_100a only");
    qemb_ctx *c = new (std::nothrow) qemb_ctx;
    if (!c) return fail(QEMB_ENOMEM, "host allocation failed");
    c->device = device;
    c->num_sms = prop.multiProcessorCount;
    if (const char *e = std::getenv("QEMB_FULL_COLUMN_FRACTION")) c->full_frac = std::min(1.0, std::max(0.0, std::atof(e)));
    if (const char *e = std::getenv("QEMB_K3_INT8")) {
        const int n = std::atoi(e);
        c->k3_slices = (n >= 5 && n <= 8) ? n : 0;
    }
    // Experiment knob, default untouched: an L2 set-aside for persisting lines, which the evict_last loads
    // of K3's row panels then compete for.  The device limit is process-wide for this GPU.
    if (const char *e = std::getenv("QEMB_L2_PERSIST_MB")) {
        size_t want = (size_t)std::max(0, std::atoi(e)) << 20;
        if (want > (size_t)prop.persistingL2CacheMaxSize) want = (size_t)prop.persistingL2CacheMaxSize;
        const cudaError_t r = cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want);
        if (r != cudaSuccess) (void)cudaGetLastError();
        std::fprintf(stderr, "[qemb] QEMB_L2_PERSIST_MB: persisting L2 set-aside %zu MB of at most %d MB (%s), L2 %d MB\n",
                     want >> 20, prop.persistingL2CacheMaxSize >> 20, r == cudaSuccess ? "set" : cudaGetErrorString(r),
                     prop.l2CacheSize >> 20);
    }
    CU(cudaStreamCreateWithFlags(&c->compute, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&c->prep, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&c->copy, cudaStreamNonBlocking));
    *out = c;
    return QEMB_OK;
}

extern "C" void qemb_ctx_destroy(qemb_ctx *c)
{
    if (!c) return;
    DeviceGuard g(c->device);
    cudaStreamSynchronize(c->compute);
    cudaStreamSynchronize(c->prep);
    cudaStreamSynchronize(c->copy);
    cudaStreamDestroy(c->compute);
    cudaStreamDestroy(c->prep);
    cudaStreamDestroy(c->copy);
    for (auto &m : c->marks)
        if (m) cudaEventDestroy(m);
    for (int i = 0; i < 2; ++i) {
        if (c->bounce[i]) cudaFreeHost(c->bounce[i]);
        if (c->bounce_free[i]) cudaEventDestroy(c->bounce_free[i]);
    }
    delete c;
}

extern "C" int qemb_ctx_synchronize(qemb_ctx *c)
{
    REQ(c != nullptr, "null ctx");
    DeviceGuard g(c->device);
    TRY(sync_workspaces(c));
    CU(cudaStreamSynchronize(c->prep));
    CU(cudaStreamSynchronize(c->compute));
    CU(cudaStreamSynchronize(c->copy));
    {   // host threads still mirroring a G that has landed
        std::unique_lock<std::mutex> lk(c->out_mutex);
        c->out_cv.wait(lk, [&] { return c->out_pending == 0; });
    }
    return QEMB_OK;
}

extern "C" int qemb_ctx_device(const qemb_ctx *c, int *device)
{
    REQ(c != nullptr && device != nullptr, "null argument");
    *device = c->device;
    return QEMB_OK;
}

extern "C" int qemb_ctx_set_timing(qemb_ctx *c, int enabled)
{
    REQ(c != nullptr, "null ctx");
    c->timing = enabled != 0;
    return QEMB_OK;
}

extern "C" int qemb_ctx_set_output_mode(qemb_ctx *c, int mode)
{
    REQ(c != nullptr, "null ctx");
    REQ(mode == 0 || mode == 1, "output mode must be 0 (full square over PCIe) or 1 (lower triangle + host mirror)");
    c->out_mode = mode;
    return QEMB_OK;
}

extern "C" int qemb_ctx_set_output_group_bytes(qemb_ctx *c, int64_t bytes)
{
    REQ(c && bytes >= (1 << 20) && bytes <= (1ll << 32), "group size must be between 1 MB and 4 GB");
    c->out_group_bytes = bytes;
    return QEMB_OK;
}

extern "C" int qemb_ctx_set_full_column_fraction(qemb_ctx *c, double frac)
{
    REQ(c && frac >= 0.0 && frac <= 1.0, "fraction must be in [0, 1]");
    c->full_frac = frac;
    return QEMB_OK;
}

extern "C" int qemb_ctx_set_k3_int8_slices(qemb_ctx *c, int slices)
{
    REQ(c && (slices == 0 || (slices >= 5 && slices <= 8)), "slices must be 0 (FP64 K3) or 5 .. 8");
    c->k3_slices = slices;
    return QEMB_OK;
}

extern "C" int qemb_ctx_set_k3_int8_tolerance(qemb_ctx *c, double abs_tol)
{
    REQ(c && abs_tol >= 0.0 && abs_tol == abs_tol, "tolerance must be >= 0 (0 switches the mode off)");
    c->k3_tol = abs_tol;
    return QEMB_OK;
}

extern "C" int qemb_k3_int8_bound(int64_t naux, int slices, double *bound)
{
    REQ(bound && naux >= 0 && slices >= 5 && slices <= 8, "bad argument");
    *bound = gram_i8_bound(naux, slices);
    return QEMB_OK;
}

extern "C" int qemb_fragment_k3_choice(const qemb_workspace *w, int *slices, int *max_expo)
{
    REQ(w && slices, "null argument");
    *slices = w->k3_last_slices;
    if (max_expo) *max_expo = w->k3_last_expo;
    return QEMB_OK;
}

extern "C" int qemb_host_threads(int *n)
{
    REQ(n != nullptr, "null argument");
    *n = host_pool().size();
    return QEMB_OK;
}

// G[c, r] = G[r, c] for all r > c of an n x n column-major host matrix, on the library's host threads
extern "C" int qemb_host_mirror_lower(double *G, int64_t ld, int64_t n)
{
    REQ(G != nullptr && n >= 0 && ld >= n, "bad argument");
    OutJob job;
    const int64_t W = 256, R = 1024;
    int64_t ntasks = 0;
    for (int64_t c0 = 0; c0 < n; c0 += W) ntasks += (n - c0 + R - 1) / R;
    job.remaining = ntasks;
    for (int64_t c0 = 0; c0 < n; c0 += W) {
        const int64_t c1 = std::min(n, c0 + W);
        for (int64_t r0 = c0; r0 < n; r0 += R) {
            const int64_t r1 = std::min(n, r0 + R);
            OutJob *jp = &job;
            host_pool().submit([=] {
                mirror_block(G, ld, c0, c1, r0, r1);
                std::lock_guard<std::mutex> lk(jp->m);
                if (--jp->remaining == 0) jp->cv.notify_all();
            });
        }
    }
    wait_job(job);
    return QEMB_OK;
}

extern "C" int qemb_ctx_mark(qemb_ctx *c, int id, int stream)
{
    REQ(c != nullptr && id >= 0 && id < 16 && stream >= 0 && stream < 3, "bad argument");
    DeviceGuard g(c->device);
    if (!c->marks[id]) CU(cudaEventCreate(&c->marks[id]));
    if (stream == 0) TRY(join_workspaces(c)); // a mark on the compute stream brackets the fragment work as well
    CU(cudaEventRecord(c->marks[id], stream == 0 ? c->compute : stream == 1 ? c->prep : c->copy));
    return QEMB_OK;
}

extern "C" int qemb_ctx_mark_elapsed(qemb_ctx *c, int id0, int id1, float *ms)
{
    REQ(c != nullptr && ms != nullptr && id0 >= 0 && id0 < 16 && id1 >= 0 && id1 < 16, "bad argument");
    REQ(c->marks[id0] && c->marks[id1], "mark was never recorded");
    DeviceGuard g(c->device);
    CU(cudaEventSynchronize(c->marks[id0]));
    CU(cudaEventSynchronize(c->marks[id1]));
    CU(cudaEventElapsedTime(ms, c->marks[id0], c->marks[id1]));
    return QEMB_OK;
}

extern "C" int qemb_ctx_launch_count(const qemb_ctx *c, int64_t *n)
{
    REQ(c != nullptr && n != nullptr, "null argument");
    *n = c->launches;
    return QEMB_OK;
}

extern "C" int qemb_ctx_transfer_bytes(const qemb_ctx *c, int64_t *h2d, int64_t *d2h)
{
    REQ(c != nullptr && h2d != nullptr && d2h != nullptr, "null argument");
    *h2d = c->h2d_bytes;
    *d2h = c->d2h_bytes;
    return QEMB_OK;
}

extern "C" int qemb_host_alloc(void **p, int64_t bytes)
{
    REQ(p != nullptr && bytes >= 0, "bad argument");
    *p = nullptr;
    if (bytes == 0) return QEMB_OK;
    cudaError_t e = cudaMallocHost(p, (size_t)bytes);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return fail(QEMB_ENOMEM, std::string("cudaMallocHost: ") + cudaGetErrorString(e));
    }
    return QEMB_OK;
}
extern "C" int qemb_host_free(void *p)
{
    if (p) CU(cudaFreeHost(p));
    return QEMB_OK;
}

// Host DRAM bandwidth as the library's own host threads see it: non-temporal fill of `bytes` (write) and a
// parallel memcpy of bytes/2 -> bytes/2 (copy: read + write), best of 3.  The denominator of e2e.host_roofline_frac
// in bench.py: every (ij|kl) matrix has to be WRITTEN into host DRAM once, by the copy engines or by the mirror pass.
extern "C" int qemb_host_bandwidth(int64_t bytes, double *write_gbs, double *copy_gbs)
{
    REQ(bytes >= (1 << 20) && write_gbs && copy_gbs, "bad argument");
    void *buf = nullptr;
    if (posix_memalign(&buf, 4096, (size_t)bytes) != 0 || !buf) return fail(QEMB_ENOMEM, "host allocation failed");
    char *b = static_cast<char *>(buf);
    parallel_for(bytes, 4 << 20, [=](int64_t i0, int64_t i1) { std::memset(b + i0, 1, (size_t)(i1 - i0)); }); // touch
    double wbest = 0, cbest = 0;
    for (int rep = 0; rep < 3; ++rep) {
        auto t0 = std::chrono::steady_clock::now();
        parallel_for(bytes, 4 << 20, [=](int64_t i0, int64_t i1) { stream_fill(b + i0, (size_t)(i1 - i0)); });
        auto t1 = std::chrono::steady_clock::now();
        parallel_memcpy(b + bytes / 2, b, (size_t)(bytes / 2));
        auto t2 = std::chrono::steady_clock::now();
        wbest = std::max(wbest, (double)bytes / std::chrono::duration<double>(t1 - t0).count() / 1e9);
        cbest = std::max(cbest, (double)(bytes / 2) / std::chrono::duration<double>(t2 - t1).count() / 1e9);
    }
    std::free(buf);
    *write_gbs = wbest;
    *copy_gbs = cbest;
    return QEMB_OK;
}

// Self-tuning of how a host-bound (ij|kl) travels (output mode + full-column fraction).  Three rates are measured on
// this context, all ranks of a node calling at the same moment so that the contention is the real one:
//   P0  device -> host copy into pinned memory with the host threads idle      (what a full-square transfer gets)
//   P   the same copy WHILE the library's host threads mirror a test matrix    (what a triangle transfer gets)
//   M   the mirror rate during that time (bytes of upper triangle produced per second)
// With L ~ U bytes of lower / upper triangle:  full squares take 2U / P0;  sending the fraction f of the upper triangle
// over PCIe and mirroring the rest takes max((1 + f) U / P, (1 - f) U / M), least at f* = max(0, (P - M) / (P + M)).
// The cheaper of the two is chosen (the copy engine loses up to half its rate to the mirror's DRAM traffic on some
// hosts, so neither extreme wins everywhere: N = 1 boxes want the pure triangle, the 2-GPU box full squares, four
// GPUs behind one PCIe switch a split).
static int autotune_model(qemb_ctx *c, double *d2h_gbs, double *mirror_gbs, double *upper_fraction_pcie)
{
    REQ(c != nullptr, "null ctx");
    DeviceGuard g(c->device);
    const int64_t n = 6144;                       // mirror test matrix: 6144^2 doubles = 302 MB
    const size_t mbytes = (size_t)n * n * 8, cbytes = 256u << 20;
    const int ncopies = 8;
    void *hm = nullptr, *hc = nullptr, *dv = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    auto cleanup = [&] {
        for (auto &e : ev)
            if (e) cudaEventDestroy(e);
        if (hm) cudaFreeHost(hm);
        if (hc) cudaFreeHost(hc);
        if (dv) cudaFree(dv);
    };
    cudaError_t e = cudaMallocHost(&hm, mbytes);
    if (e == cudaSuccess) e = cudaMallocHost(&hc, cbytes);
    if (e == cudaSuccess) e = cudaMalloc(&dv, cbytes);
    for (int i = 0; i < 4 && e == cudaSuccess; ++i) e = cudaEventCreate(&ev[i]);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        cleanup();
        return fail(QEMB_ENOMEM, std::string("autotune buffers: ") + cudaGetErrorString(e));
    }
    std::memset(hm, 0, mbytes);
    int reps = 0;
    double tm = 0.0;
    float ms0 = 0.f, ms1 = 0.f;
    auto run = [&]() -> cudaError_t {
        cudaError_t r = cudaMemsetAsync(dv, 0, cbytes, c->copy);
        if (r == cudaSuccess) r = cudaMemcpyAsync(hc, dv, cbytes, cudaMemcpyDeviceToHost, c->copy); // warm-up
        if (r == cudaSuccess) r = cudaStreamSynchronize(c->copy);
        if (r != cudaSuccess) return r;
        qemb_host_mirror_lower(static_cast<double *>(hm), n, 1024); // warm-up (threads, page faults)
        // phase 1: the copy engine alone
        r = cudaEventRecord(ev[0], c->copy);
        for (int i = 0; i < ncopies && r == cudaSuccess; ++i) r = cudaMemcpyAsync(hc, dv, cbytes, cudaMemcpyDeviceToHost, c->copy);
        if (r == cudaSuccess) r = cudaEventRecord(ev[1], c->copy);
        if (r == cudaSuccess) r = cudaEventSynchronize(ev[1]);
        // phase 2: the same copies while the host threads mirror
        if (r == cudaSuccess) r = cudaEventRecord(ev[2], c->copy);
        for (int i = 0; i < ncopies && r == cudaSuccess; ++i) r = cudaMemcpyAsync(hc, dv, cbytes, cudaMemcpyDeviceToHost, c->copy);
        if (r == cudaSuccess) r = cudaEventRecord(ev[3], c->copy);
        if (r != cudaSuccess) return r;
        const auto t0 = std::chrono::steady_clock::now();
        do {
            qemb_host_mirror_lower(static_cast<double *>(hm), n, n);
            ++reps;
        } while (cudaEventQuery(ev[3]) == cudaErrorNotReady && reps < 64);
        tm = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        (void)cudaGetLastError();
        r = cudaEventSynchronize(ev[3]);
        if (r == cudaSuccess) r = cudaEventElapsedTime(&ms0, ev[0], ev[1]);
        if (r == cudaSuccess) r = cudaEventElapsedTime(&ms1, ev[2], ev[3]);
        return r;
    };
    e = run();
    cleanup();
    if (e != cudaSuccess) return fail(QEMB_ECUDA, std::string("autotune: ") + cudaGetErrorString(e));
    const double bytes = (double)ncopies * (double)cbytes;
    const double P0 = bytes / (ms0 * 1e-3) / 1e9, P = bytes / (ms1 * 1e-3) / 1e9;
    const double M = (double)reps * (double)n * (double)(n - 1) / 2.0 * 8.0 / tm / 1e9;
    double f = (P - M) / (P + M);
    f = f < 0.0 ? 0.0 : f;
    const double t_split = std::max((1.0 + f) / P, (1.0 - f) / M), t_full = 2.0 / P0; // per byte of U
    if (f >= 0.9 || t_full <= t_split) {
        c->out_mode = 0;
        c->full_frac = 0.0;
        f = 1.0;
    } else {
        c->out_mode = 1;
        c->full_frac = 1.0 - std::sqrt(1.0 - f); // whole columns [cs, n) carry 1 - (cs / n)^2 of the upper triangle
    }
    c->tune_p0 = P0;
    if (d2h_gbs) *d2h_gbs = P;
    if (mirror_gbs) *mirror_gbs = M;
    if (upper_fraction_pcie) *upper_fraction_pcie = f;
    return QEMB_OK;
}

// copy rate with the host threads idle, from the last qemb_ctx_autotune_output (GB/s; 0 before the first call)
extern "C" int qemb_ctx_autotune_copy_alone(const qemb_ctx *c, double *d2h_alone_gbs)
{
    REQ(c && d2h_alone_gbs, "null argument");
    *d2h_alone_gbs = c->tune_p0;
    return QEMB_OK;
}

// Orders an EXTERNAL stream (e.g. torch's current stream, which NCCL collectives are enqueued on) against the
// library's compute stream without a host synchronisation.  direction 0: work enqueued on `external_stream` after
// this call waits for everything enqueued on the compute stream so far; direction 1: the reverse.
extern "C" int qemb_ctx_stream_handoff(qemb_ctx *c, void *external_stream, int direction)
{
    REQ(c && (direction == 0 || direction == 1), "bad argument");
    DeviceGuard g(c->device);
    cudaStream_t ext = static_cast<cudaStream_t>(external_stream);
    if (direction == 0) TRY(join_workspaces(c));
    cudaEvent_t e;
    CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    cudaError_t rc = cudaEventRecord(e, direction == 0 ? c->compute : ext);
    if (rc == cudaSuccess) rc = cudaStreamWaitEvent(direction == 0 ? ext : c->compute, e, 0);
    cudaEventDestroy(e); // released once the wait has been satisfied
    CU(rc);
    return QEMB_OK;
}

// ------------------------------------------------------------------------------------------------
// index maps
// ------------------------------------------------------------------------------------------------
extern "C" int64_t qemb_ravel_symmetric(int64_t a, int64_t b)
{
    return (a > b) ? gauss_sum(a) + b : gauss_sum(b) + a;
}

extern "C" void qemb_unravel_symmetric(int64_t I, int64_t *lo, int64_t *hi)
{
    // largest a with a(a+1)/2 <= I, exact: start from the floating estimate and correct
    int64_t a = (int64_t)((std::sqrt(8.0 * (double)I + 1.0) - 1.0) / 2.0);
    while (gauss_sum(a) > I) --a;
    while (gauss_sum(a + 1) <= I) ++a;
    const int64_t b = I - gauss_sum(a);
    if (a <= b) {
        *lo = a;
        *hi = b;
    } else {
        *lo = b;
        *hi = a;
    }
}

extern "C" int64_t qemb_extract_unique(int64_t nao, const int64_t *ptr, const int64_t *idx, int64_t *uptr,
                                       int64_t *uidx)
{
    if (nao < 0 || !ptr || !uptr || (!idx && ptr[nao] > 0)) {
        fail(QEMB_EINVAL, "bad CSR input");
        return QEMB_EINVAL;
    }
    int64_t n = 0;
    uptr[0] = 0;
    for (int64_t mu = 0; mu < nao; ++mu) {
        for (int64_t k = ptr[mu]; k < ptr[mu + 1]; ++k) {
            if (idx[k] > mu) break; // sorted lists: the reference stops at the first nu > mu (hpp:155-157)
            if (uidx) uidx[n] = idx[k];
            ++n;
        }
        uptr[mu + 1] = n;
    }
    return n;
}

namespace {
// column of the pair (mu,nu): counter over mu, then over the unique list of mu (cpp:248-258).
// returns -1 when the key was never assigned
int64_t find_column(const std::vector<int64_t> &uptr, const std::vector<int64_t> &uidx, int64_t mu, int64_t nu)
{
    const int64_t hi = std::max(mu, nu), lo = std::min(mu, nu);
    const int64_t *b = uidx.data() + uptr[hi], *e = uidx.data() + uptr[hi + 1];
    const int64_t *it = std::lower_bound(b, e, lo);
    if (it != e && *it == lo) return it - uidx.data();
    return -1;
}
} // namespace

// ------------------------------------------------------------------------------------------------
// tensor
// ------------------------------------------------------------------------------------------------
extern "C" int qemb_tensor_create(qemb_ctx *ctx, int64_t naux, int64_t nao, const int64_t *exch_ptr,
                                  const int64_t *exch_idx, qemb_tensor **out)
{
    REQ(ctx && out && exch_ptr, "null argument");
    REQ(naux > 0 && nao > 0, "naux and nao must be positive");
    REQ(exch_ptr[0] == 0, "exch_ptr[0] must be 0");
    *out = nullptr;
    const int64_t nnz = exch_ptr[nao];
    REQ(nnz >= 0 && nnz < std::numeric_limits<int32_t>::max(), "exch_reachable too large");
    REQ(nnz == 0 || exch_idx, "null exch_idx");
    for (int64_t mu = 0; mu < nao; ++mu) {
        REQ(exch_ptr[mu + 1] >= exch_ptr[mu], "exch_ptr must be non-decreasing");
        for (int64_t k = exch_ptr[mu]; k < exch_ptr[mu + 1]; ++k) {
            REQ(exch_idx[k] >= 0 && exch_idx[k] < nao, "exch_reachable entry out of range");
            REQ(k == exch_ptr[mu] || exch_idx[k] > exch_idx[k - 1], "exch_reachable lists must be sorted ascending");
        }
    }
    DeviceGuard g(ctx->device);
    qemb_tensor *t = new (std::nothrow) qemb_tensor;
    if (!t) return fail(QEMB_ENOMEM, "host allocation failed");
    t->ctx = ctx;
    t->naux = naux;
    t->nao = nao;
    t->ldp = round_up(naux, 128);
    t->naux16 = round_up(naux, 16);
    t->rptr.assign(exch_ptr, exch_ptr + nao + 1);
    t->ridx.assign(exch_idx, exch_idx + nnz);
    t->uptr.resize(nao + 1);
    const int64_t nu = qemb_extract_unique(nao, exch_ptr, exch_idx, t->uptr.data(), nullptr);
    t->uidx.resize(std::max<int64_t>(nu, 1));
    qemb_extract_unique(nao, exch_ptr, exch_idx, t->uptr.data(), t->uidx.data());
    t->uidx.resize(nu);
    t->n_unique = nu;
    REQ(nu < std::numeric_limits<int32_t>::max(), "too many unique pairs");
    t->rcol.resize(nnz);
    std::vector<int32_t> rcol32(nnz), ridx32(nnz);
    for (int64_t mu = 0; mu < nao; ++mu)
        for (int64_t k = exch_ptr[mu]; k < exch_ptr[mu + 1]; ++k) {
            const int64_t c = find_column(t->uptr, t->uidx, mu, exch_idx[k]);
            t->rcol[k] = c < 0 ? 0 : c; // operator[] default-insert of the reference (cpp:271)
            rcol32[k] = (int32_t)t->rcol[k];
            ridx32[k] = (int32_t)exch_idx[k];
        }
    int rc = t->data.reserve((size_t)t->ldp * (size_t)std::max<int64_t>(nu, 1) * 8);
    if (rc == QEMB_OK) rc = t->d_rcol.reserve((size_t)std::max<int64_t>(nnz, 1) * 4);
    if (rc == QEMB_OK) rc = t->d_ridx.reserve((size_t)std::max<int64_t>(nnz, 1) * 4);
    if (rc != QEMB_OK) {
        qemb_tensor_destroy(t);
        return rc;
    }
    cudaStream_t s = ctx->compute;
    // NaN-fill like cpp:160 so that a column nobody uploaded is detected, pad rows zero
    CU(launch_fill(t->data.as<double>(), t->ldp * nu, std::numeric_limits<double>::quiet_NaN(), s));
    CU(launch_zero_rows(t->data.as<double>(), t->ldp, naux, t->ldp, nu, s));
    ctx->launches += 2;
    if (nnz) {
        CU(cudaMemcpyAsync(t->d_rcol.p, rcol32.data(), (size_t)nnz * 4, cudaMemcpyHostToDevice, s));
        CU(cudaMemcpyAsync(t->d_ridx.p, ridx32.data(), (size_t)nnz * 4, cudaMemcpyHostToDevice, s));
    }
    CU(cudaStreamSynchronize(s));
    *out = t;
    return QEMB_OK;
}

extern "C" void qemb_tensor_destroy(qemb_tensor *t)
{
    if (!t) return;
    DeviceGuard g(t->ctx->device);
    sync_workspaces(t->ctx);
    cudaStreamSynchronize(t->ctx->compute);
    t->data.release();
    t->d_rcol.release();
    t->d_ridx.release();
    t->scratch.release();
    t->nan_count.release();
    for (int i = 0; i < 2; ++i) {
        t->stage_h[i].release();
        t->stage_d[i].release();
        if (t->stage_done[i]) cudaEventDestroy(t->stage_done[i]);
    }
    delete t;
}

extern "C" int qemb_tensor_n_unique(const qemb_tensor *t, int64_t *n)
{
    REQ(t && n, "null argument");
    *n = t->n_unique;
    return QEMB_OK;
}

extern "C" int qemb_tensor_shape(const qemb_tensor *t, int64_t *naux, int64_t *nao, int64_t *n_unique)
{
    REQ(t, "null tensor");
    if (naux) *naux = t->naux;
    if (nao) *nao = t->nao;
    if (n_unique) *n_unique = t->n_unique;
    return QEMB_OK;
}

extern "C" int qemb_tensor_column(const qemb_tensor *t, int64_t mu, int64_t nu, int64_t *col)
{
    REQ(t && col, "null argument");
    if (mu < 0 || nu < 0 || mu >= t->nao || nu >= t->nao) return fail(QEMB_EINDEX, "AO index out of range");
    const int64_t c = find_column(t->uptr, t->uidx, mu, nu);
    if (c < 0) return fail(QEMB_EINDEX, "pair (" + std::to_string(mu) + "," + std::to_string(nu) + ") is not in the screened set");
    *col = c;
    return QEMB_OK;
}

extern "C" int qemb_tensor_reachable_columns(const qemb_tensor *t, int64_t *col)
{
    REQ(t && (col || t->rcol.empty()), "null argument");
    std::copy(t->rcol.begin(), t->rcol.end(), col);
    return QEMB_OK;
}

extern "C" int qemb_tensor_upload(qemb_tensor *t, const double *host, int64_t ld_host, int64_t col0, int64_t ncols)
{
    REQ(t && (host || ncols == 0), "null argument");
    REQ(col0 >= 0 && ncols >= 0 && col0 + ncols <= t->n_unique, "column range outside the tensor");
    REQ(ld_host >= t->naux, "ld_host < naux");
    if (ncols == 0) return QEMB_OK;
    // a partial refresh of a whitened tensor would leave raw and L^-1-scaled columns side by side under one flag:
    // replace everything, or clear the flag first (qemb_tensor_set_whitened(t, 0)) and whiten again afterwards
    REQ(!t->whitened || (col0 == 0 && ncols == t->n_unique),
        "partial upload into a whitened tensor (call qemb_tensor_set_whitened(t, 0) first)");
    DeviceGuard g(t->ctx->device);
    NvtxRange nv("qemb:tensor_upload");
    TRY(join_workspaces(t->ctx)); // fragments still reading the old content finish first
    cudaStream_t s = t->ctx->compute;
    TRY(upload_2d(t->ctx, t->data.as<double>() + col0 * t->ldp, (size_t)t->ldp * 8, host, (size_t)ld_host * 8,
                  (size_t)t->naux * 8, (size_t)ncols, s));
    CU(cudaStreamSynchronize(s));
    t->whitened = false;
    return QEMB_OK;
}

// (mu nu|P) staging without the Python triple loop (py:459-488): one libcint block `integrals[i, j, P]` (C order,
// P contiguous) for the AOs mu = p0 + i, nu = q0 + j is uploaded once and scattered on the device into the columns
// of the screened pairs with nu <= mu (pairs outside the screened set are skipped, like the key_to_offset test).
extern "C" int qemb_tensor_scatter_block(qemb_tensor *t, const double *block, int64_t n_i, int64_t n_j, int64_t p0,
                                         int64_t q0)
{
    REQ(t && (block || n_i * n_j == 0), "null argument");
    REQ(n_i >= 0 && n_j >= 0 && p0 >= 0 && q0 >= 0 && p0 + n_i <= t->nao && q0 + n_j <= t->nao, "AO block outside the tensor");
    const int64_t npairs = n_i * n_j;
    if (npairs == 0) return QEMB_OK;
    REQ(npairs < std::numeric_limits<int32_t>::max(), "block too large");
    std::vector<int32_t> col((size_t)npairs, -1);
    int64_t live = 0;
    for (int64_t i = 0; i < n_i; ++i)
        for (int64_t j = 0; j < n_j; ++j) {
            const int64_t mu = p0 + i, nu = q0 + j;
            if (nu > mu) break; // "still ensure p <= q" of the reference loop (py:476-480)
            const int64_t c = find_column(t->uptr, t->uidx, mu, nu);
            if (c >= 0) {
                col[(size_t)(i * n_j + j)] = (int32_t)c;
                ++live;
            }
        }
    if (live == 0) return QEMB_OK;
    REQ(!t->whitened, "block staging into a whitened tensor (call qemb_tensor_set_whitened(t, 0) first)");
    DeviceGuard g(t->ctx->device);
    NvtxRange nv("qemb:scatter_block");
    TRY(join_workspaces(t->ctx));
    cudaStream_t s = t->ctx->compute;
    const size_t bbytes = (size_t)npairs * (size_t)t->naux * 8, cbytes = (size_t)npairs * 4;
    const size_t coff = (bbytes + 15) & ~size_t(15);
    // two staging slots: the host copy of this block into pinned memory overlaps the H2D copy + scatter kernel of the
    // previous one, and the call returns as soon as the block is in the slot (the caller's buffer is free again)
    const int k = t->stage_k;
    t->stage_k ^= 1;
    if (!t->stage_done[k]) CU(cudaEventCreateWithFlags(&t->stage_done[k], cudaEventDisableTiming));
    CU(cudaEventSynchronize(t->stage_done[k])); // never recorded = complete
    TRY(t->stage_h[k].reserve(coff + cbytes));
    TRY(t->stage_d[k].reserve(coff + cbytes));
    char *h = static_cast<char *>(t->stage_h[k].p);
    if (bbytes >= (8u << 20))
        parallel_memcpy(h, block, bbytes);
    else
        std::memcpy(h, block, bbytes);
    std::memcpy(h + coff, col.data(), cbytes);
    char *d = static_cast<char *>(t->stage_d[k].p);
    CU(cudaMemcpyAsync(d, h, coff + cbytes, cudaMemcpyHostToDevice, s));
    t->ctx->h2d_bytes += (int64_t)(coff + cbytes);
    CU(launch_scatter_block(reinterpret_cast<const double *>(d), reinterpret_cast<const int *>(d + coff), (int)npairs,
                            (int)t->naux, t->data.as<double>(), t->ldp, s));
    t->ctx->launches += 1;
    CU(cudaEventRecord(t->stage_done[k], s));
    return QEMB_OK;
}

// NaN entries left in the tensor (a column nobody filled keeps the NaN of qemb_tensor_create): the reference asserts
// `not np.isnan(result.unique_dense_data).any()` after staging (py:492)
extern "C" int qemb_tensor_count_nan(qemb_tensor *t, int64_t *n)
{
    REQ(t && n, "null argument");
    DeviceGuard g(t->ctx->device);
    cudaStream_t s = t->ctx->compute;
    TRY(t->nan_count.reserve(8));
    CU(cudaMemsetAsync(t->nan_count.p, 0, 8, s));
    CU(launch_count_nan(t->data.as<double>(), t->ldp, t->naux, t->n_unique, t->nan_count.as<unsigned long long>(), s));
    t->ctx->launches += 1;
    unsigned long long v = 0;
    CU(cudaMemcpyAsync(&v, t->nan_count.p, 8, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    *n = (int64_t)v;
    return QEMB_OK;
}

extern "C" int qemb_tensor_download(const qemb_tensor *t, double *host, int64_t ld_host, int64_t col0, int64_t ncols)
{
    REQ(t && (host || ncols == 0), "null argument");
    REQ(col0 >= 0 && ncols >= 0 && col0 + ncols <= t->n_unique, "column range outside the tensor");
    REQ(ld_host >= t->naux, "ld_host < naux");
    if (ncols == 0) return QEMB_OK;
    DeviceGuard g(t->ctx->device);
    cudaStream_t s = t->ctx->compute;
    CU(cudaMemcpy2DAsync(host, (size_t)ld_host * 8, t->data.as<double>() + col0 * t->ldp, (size_t)t->ldp * 8,
                         (size_t)t->naux * 8, (size_t)ncols, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    return QEMB_OK;
}

// ------------------------------------------------------------------------------------------------
// metric
// ------------------------------------------------------------------------------------------------
extern "C" int qemb_metric_create(qemb_ctx *ctx, const double *L_host, int64_t ld, int64_t naux, qemb_metric **out)
{
    REQ(ctx && L_host && out, "null argument");
    REQ(naux > 0 && ld >= naux, "bad shape");
    *out = nullptr;
    for (int64_t i = 0; i < naux; ++i) {
        const double d = L_host[i + i * ld];
        REQ(d == d && d != 0.0, "zero or NaN on the diagonal of L");
    }
    DeviceGuard g(ctx->device);
    qemb_metric *m = new (std::nothrow) qemb_metric;
    if (!m) return fail(QEMB_ENOMEM, "host allocation failed");
    m->ctx = ctx;
    m->naux = naux;
    m->naux_pad = round_up(naux, TRSM_BLOCK);
    m->nblocks = (int)(m->naux_pad / TRSM_BLOCK);
    int rc = m->L.reserve((size_t)m->naux_pad * m->naux_pad * 8);
    if (rc == QEMB_OK) rc = m->Dinv.reserve((size_t)m->nblocks * TRSM_BLOCK * TRSM_BLOCK * 8);
    if (rc != QEMB_OK) {
        qemb_metric_destroy(m);
        return rc;
    }
    cudaStream_t s = ctx->compute;
    CU(cudaMemsetAsync(m->L.p, 0, (size_t)m->naux_pad * m->naux_pad * 8, s));
    TRY(upload_2d(ctx, m->L.p, (size_t)m->naux_pad * 8, L_host, (size_t)ld * 8, (size_t)naux * 8, (size_t)naux, s));
    CU(launch_set_diag(m->L.as<double>(), m->naux_pad, naux, m->naux_pad, 1.0, s));
    CU(launch_invert_diag_blocks(m->L.as<double>(), m->naux_pad, m->nblocks, m->Dinv.as<double>(), s));
    ctx->launches += 2;
    CU(cudaStreamSynchronize(s));
    *out = m;
    return QEMB_OK;
}

extern "C" void qemb_metric_destroy(qemb_metric *m)
{
    if (!m) return;
    DeviceGuard g(m->ctx->device);
    sync_workspaces(m->ctx);
    cudaStreamSynchronize(m->ctx->compute);
    m->L.release();
    m->Dinv.release();
    delete m;
}

extern "C" int qemb_tensor_whiten(qemb_tensor *t, const qemb_metric *m)
{
    REQ(t && m, "null argument");
    REQ(t->ctx == m->ctx, "tensor and metric live on different contexts");
    REQ(t->naux == m->naux, "naux mismatch between tensor and metric");
    REQ(!t->whitened, "tensor is already whitened");
    DeviceGuard g(t->ctx->device);
    NvtxRange nv("qemb:tensor_whiten");
    TRY(join_workspaces(t->ctx));
    TrsmParams p;
    p.L = m->L.as<double>();
    p.ldl = m->naux_pad;
    p.Dinv = m->Dinv.as<double>();
    p.nblocks = m->nblocks;
    p.X = t->data.as<double>();
    p.ldx = t->ldp;
    p.ncol = t->n_unique;
    CU(launch_trsm_rows(p, (int)t->naux16, t->ctx->compute));
    t->ctx->launches += 1;
    CU(cudaStreamSynchronize(t->ctx->compute));
    t->whitened = true;
    return QEMB_OK;
}

// One column chunk of an AUXILIARY ROW SLAB of the whitened tensor, without ever holding the full
// (P|mu nu) on this device: the chunk (all naux rows, `ncols` pair columns) is uploaded into scratch,
// solved against L there (rows >= p_end are not needed by forward substitution and never travel) and
// only rows [p_begin, p_begin + slab->naux) are kept.
extern "C" int qemb_tensor_upload_whitened_rows(qemb_tensor *t, const qemb_metric *m, const double *host,
                                                int64_t ld_host, int64_t col0, int64_t ncols, int64_t p_begin)
{
    REQ(t && m && (host || ncols == 0), "null argument");
    REQ(t->ctx == m->ctx, "tensor and metric live on different contexts");
    REQ(col0 >= 0 && ncols >= 0 && col0 + ncols <= t->n_unique, "column range outside the tensor");
    REQ(p_begin >= 0 && p_begin % 16 == 0 && p_begin + t->naux <= m->naux, "bad auxiliary slab");
    REQ(ld_host >= m->naux, "ld_host < naux of the metric");
    if (ncols == 0) return QEMB_OK;
    DeviceGuard g(t->ctx->device);
    qemb_ctx *c = t->ctx;
    TRY(join_workspaces(c));
    cudaStream_t s = c->compute;
    const int64_t p_end = p_begin + t->naux;
    const int64_t mrows = round_up(p_end, 16);
    const int64_t lds = m->naux_pad;
    TRY(t->scratch.reserve((size_t)lds * (size_t)ncols * 8));
    double *S = t->scratch.as<double>();
    CU(cudaMemsetAsync(S, 0, (size_t)lds * (size_t)ncols * 8, s));
    TRY(upload_2d(c, S, (size_t)lds * 8, host, (size_t)ld_host * 8, (size_t)p_end * 8, (size_t)ncols, s));
    TrsmParams p;
    p.L = m->L.as<double>();
    p.ldl = m->naux_pad;
    p.Dinv = m->Dinv.as<double>();
    p.nblocks = m->nblocks;
    p.X = S;
    p.ldx = lds;
    p.ncol = ncols;
    CU(launch_trsm_rows(p, (int)mrows, s));
    c->launches += 1;
    CU(cudaMemcpy2DAsync(t->data.as<double>() + col0 * t->ldp, (size_t)t->ldp * 8, S + p_begin, (size_t)lds * 8,
                         (size_t)t->naux * 8, (size_t)ncols, cudaMemcpyDeviceToDevice, s));
    CU(cudaStreamSynchronize(s));
    t->whitened = true;
    return QEMB_OK;
}

extern "C" int qemb_tensor_release_scratch(qemb_tensor *t)
{
    REQ(t, "null tensor");
    DeviceGuard g(t->ctx->device);
    CU(cudaStreamSynchronize(t->ctx->compute));
    t->scratch.release();
    return QEMB_OK;
}

// Column range of the one-off metric application (design option O1), for ranks that split the columns of (P|mu nu)
// among themselves and exchange the results (bench.py --gpus N: every rank needs the whole whitened tensor, but
// only 1/N of the naux^2 n_unique flops).  The tensor is marked whitened by qemb_tensor_set_whitened once the
// caller has gathered all ranges.
extern "C" int qemb_tensor_whiten_range(qemb_tensor *t, const qemb_metric *m, int64_t col0, int64_t ncols)
{
    REQ(t && m, "null argument");
    REQ(t->ctx == m->ctx, "tensor and metric live on different contexts");
    REQ(t->naux == m->naux, "naux mismatch between tensor and metric");
    REQ(!t->whitened, "tensor is already whitened");
    REQ(col0 >= 0 && ncols >= 0 && col0 + ncols <= t->n_unique, "column range outside the tensor");
    if (ncols == 0) return QEMB_OK;
    DeviceGuard g(t->ctx->device);
    TRY(join_workspaces(t->ctx));
    TrsmParams p;
    p.L = m->L.as<double>();
    p.ldl = m->naux_pad;
    p.Dinv = m->Dinv.as<double>();
    p.nblocks = m->nblocks;
    p.X = t->data.as<double>() + col0 * t->ldp;
    p.ldx = t->ldp;
    p.ncol = ncols;
    CU(launch_trsm_rows(p, (int)t->naux16, t->ctx->compute));
    t->ctx->launches += 1;
    return QEMB_OK;
}

extern "C" int qemb_tensor_set_whitened(qemb_tensor *t, int flag)
{
    REQ(t, "null tensor");
    t->whitened = flag != 0;
    return QEMB_OK;
}

// device address and leading dimension (in doubles) of the ldp x n_unique data, for collectives on it
extern "C" int qemb_tensor_data_dev(const qemb_tensor *t, void **dev, int64_t *ld)
{
    REQ(t && dev, "null argument");
    *dev = t->data.p;
    if (ld) *ld = t->ldp;
    return QEMB_OK;
}

extern "C" int qemb_tensor_copy_data(qemb_tensor *dst, const qemb_tensor *src)
{
    REQ(dst && src, "null argument");
    REQ(dst->ctx == src->ctx, "tensors live on different contexts");
    REQ(dst->naux == src->naux && dst->nao == src->nao && dst->n_unique == src->n_unique && dst->rcol == src->rcol,
        "tensors have different shapes or lists");
    DeviceGuard g(dst->ctx->device);
    TRY(join_workspaces(dst->ctx));
    CU(cudaMemcpyAsync(dst->data.p, src->data.p, (size_t)src->ldp * (size_t)src->n_unique * 8, cudaMemcpyDeviceToDevice,
                       dst->ctx->compute));
    dst->whitened = src->whitened;
    return QEMB_OK;
}

extern "C" int qemb_tensor_is_whitened(const qemb_tensor *t, int *flag)
{
    REQ(t && flag, "null argument");
    *flag = t->whitened ? 1 : 0;
    return QEMB_OK;
}

// ------------------------------------------------------------------------------------------------
// screen
// ------------------------------------------------------------------------------------------------
extern "C" int qemb_screen_create(qemb_ctx *ctx, const double *S, int64_t ld, int64_t nao, qemb_screen **out)
{
    REQ(ctx && S && out, "null argument");
    REQ(nao > 0 && ld >= nao, "bad shape");
    *out = nullptr;
    DeviceGuard g(ctx->device);
    qemb_screen *sc = new (std::nothrow) qemb_screen;
    if (!sc) return fail(QEMB_ENOMEM, "host allocation failed");
    sc->ctx = ctx;
    sc->nao = nao;
    int rc = sc->S.reserve((size_t)nao * nao * 8);
    if (rc != QEMB_OK) {
        delete sc;
        return rc;
    }
    CU(cudaMemcpy2DAsync(sc->S.p, (size_t)nao * 8, S, (size_t)ld * 8, (size_t)nao * 8, (size_t)nao,
                         cudaMemcpyHostToDevice, ctx->compute));
    CU(cudaStreamSynchronize(ctx->compute));
    *out = sc;
    return QEMB_OK;
}

extern "C" void qemb_screen_destroy(qemb_screen *s)
{
    if (!s) return;
    DeviceGuard g(s->ctx->device);
    sync_workspaces(s->ctx);
    s->S.release();
    delete s;
}

// ------------------------------------------------------------------------------------------------
// workspace
// ------------------------------------------------------------------------------------------------
extern "C" int qemb_workspace_create(qemb_ctx *ctx, qemb_workspace **out)
{
    REQ(ctx && out, "null argument");
    *out = nullptr;
    DeviceGuard g(ctx->device);
    qemb_workspace *w = new (std::nothrow) qemb_workspace;
    if (!w) return fail(QEMB_ENOMEM, "host allocation failed");
    w->ctx = ctx;
    CU(cudaStreamCreateWithFlags(&w->compute, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&w->prep, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&w->fork_ev, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&w->join_ev, cudaEventDisableTiming));
    {
        std::lock_guard<std::mutex> lk(ctx->ws_mutex);
        ctx->workspaces.push_back(w);
    }
    for (int i = 0; i < 2; ++i) {
        CU(cudaEventCreateWithFlags(&w->slot[i].uploaded1, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&w->slot[i].uploaded2, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&w->slot[i].consumed, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&w->slot[i].ta_ready, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&w->g_copied[i], cudaEventDisableTiming));
        // blocking sync: the caller's thread sleeps while a multi-GB (ij|kl) drains instead of spinning on a host core that
        // the mirror threads (and, with several ranks per node, the other ranks) need
        for (int k = i; k < OUT_RING; k += 2)
            CU(cudaEventCreateWithFlags(&w->out_done[k], cudaEventDisableTiming | cudaEventBlockingSync));
    }
    for (int i = 0; i < 5; ++i) CU(cudaEventCreate(&w->ev[i]));
    for (int i = 0; i < 2; ++i) CU(cudaEventCreate(&w->evs[i]));
    CU(cudaEventCreate(&w->ev2s));
    *out = w;
    return QEMB_OK;
}

extern "C" void qemb_workspace_destroy(qemb_workspace *w)
{
    if (!w) return;
    DeviceGuard g(w->ctx->device);
    qemb_ctx_synchronize(w->ctx);
    for (int i = 0; i < OUT_RING; ++i) wait_job(w->jobs[i]);
    {
        std::lock_guard<std::mutex> lk(w->ctx->ws_mutex);
        auto &v = w->ctx->workspaces;
        v.erase(std::remove(v.begin(), v.end(), w), v.end());
    }
    cudaStreamDestroy(w->compute);
    cudaStreamDestroy(w->prep);
    cudaEventDestroy(w->fork_ev);
    cudaEventDestroy(w->join_ev);
    for (int i = 0; i < 2; ++i) {
        Slot &s = w->slot[i];
        s.TA.release();
        s.mask.release();
        s.plan1.release();
        s.plan2.release();
        s.stage1.release();
        s.stage2.release();
        s.hmask.release();
        cudaEventDestroy(s.uploaded1);
        cudaEventDestroy(s.uploaded2);
        cudaEventDestroy(s.consumed);
        cudaEventDestroy(s.ta_ready);
        cudaEventDestroy(w->g_copied[i]);
        for (int k = i; k < OUT_RING; k += 2) cudaEventDestroy(w->out_done[k]);
        w->G[i].release();
    }
    for (int i = 0; i < 5; ++i) cudaEventDestroy(w->ev[i]);
    for (int i = 0; i < 2; ++i) cudaEventDestroy(w->evs[i]);
    cudaEventDestroy(w->ev2s);
    for (auto &e : w->jk_ev)
        if (e) cudaEventDestroy(e);
    w->gram_i8.release();
    w->k3_expo.release();
    w->k3_expo_host.release();
    for (int i = 0; i < 2; ++i) {
        w->jk_stage[i].release();
        if (w->jk_stage_free[i]) cudaEventDestroy(w->jk_stage_free[i]);
    }
    w->jk_in.release();
    w->gram_partial.release();
    w->jk_scratch.release();
    w->jk_out.release();
    w->Bint.release();
    w->sym.release();
    for (auto &kv : w->gram_tiles) kv.second.first.release();
    delete w;
}

extern "C" int qemb_workspace_bytes(const qemb_workspace *w, int64_t *bytes)
{
    REQ(w && bytes, "null argument");
    size_t b = w->Bint.cap + w->sym.cap + w->G[0].cap + w->G[1].cap;
    for (int i = 0; i < 2; ++i) b += w->slot[i].TA.cap + w->slot[i].mask.cap + w->slot[i].plan1.cap + w->slot[i].plan2.cap;
    for (auto &kv : w->gram_tiles) b += kv.second.first.cap;
    *bytes = (int64_t)b;
    return QEMB_OK;
}

// ------------------------------------------------------------------------------------------------
// fragment: lists and plans
// ------------------------------------------------------------------------------------------------
namespace {

// next slot; waits (host side) until the kernels of the fragment that used it two calls ago are done
// with its TA / plan tables
int begin_fragment(qemb_workspace *w, int64_t nao, int64_t nmo)
{
    TRY(fork_from_ctx(w));
    if (w->started) w->cur ^= 1;
    w->started = true;
    Slot &s = w->slot[w->cur];
    if (s.consumed_pending) {
        CU(cudaEventSynchronize(s.consumed));
        s.consumed_pending = false;
    }
    s.nao = nao;
    s.nmo = nmo;
    s.npair = gauss_sum(nmo);
    s.nA = 0;
    s.have_plan1 = s.have_plan2 = false;
    s.plan1_tensor = nullptr;
    std::fill(s.work, s.work + 8, 0.0);
    w->have_B = w->have_sym = w->solved = w->have_G = false;
    return QEMB_OK;
}

int upload_TA(qemb_workspace *w, const double *TA, int64_t ld)
{
    Slot &s = w->slot[w->cur];
    TRY(s.TA.reserve((size_t)s.nao * s.nmo * 8));
    CU(cudaMemcpy2DAsync(s.TA.p, (size_t)s.nao * 8, TA, (size_t)ld * 8, (size_t)s.nao * 8, (size_t)s.nmo,
                         cudaMemcpyHostToDevice, w->prep));
    w->ctx->h2d_bytes += s.nao * s.nmo * 8;
    CU(cudaEventRecord(s.ta_ready, w->prep));
    return QEMB_OK;
}

inline size_t align16(size_t x) { return (x + 15) & ~size_t(15); }

// Reorders a gather-GEMM tile list for the persistent kernel (CTA c runs tiles c, c + nsm, c + 2 nsm, ...):
// tiles are dealt to the CTAs in serpentine order of decreasing cost, so every CTA gets the same amount of
// tensor-pipe work, and each CTA alternates its expensive (many live columns: tensor bound) and cheap (few live
// columns: bound by the A-operand traffic) tiles, odd CTAs starting with a cheap one, so the two kinds overlap
// in time instead of leaving an HBM-bound tail of small tiles.
void balance_tiles(int4 *tiles, const std::vector<float> &cost, int64_t n, int nsm)
{
    if (n <= nsm || nsm <= 1) return;
    std::vector<int32_t> order((size_t)n);
    for (int64_t i = 0; i < n; ++i) order[(size_t)i] = (int32_t)i;
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return cost[(size_t)a] > cost[(size_t)b]; });
    const int64_t rounds = (n + nsm - 1) / nsm;
    std::vector<int4> out((size_t)n);
    std::vector<int32_t> mine;
    for (int c = 0; c < nsm; ++c) {
        mine.clear();
        for (int64_t r = 0; r < rounds; ++r) {
            const int64_t full = (r + 1) * nsm <= n ? nsm : n - r * nsm; // tiles in this round
            const int64_t slot = (r & 1) && full == nsm ? nsm - 1 - c : c;
            if (slot < full) mine.push_back(order[(size_t)(r * nsm + slot)]);
        }
        // alternate large / small
        const int64_t k = (int64_t)mine.size();
        int64_t lo = 0, hi = k - 1;
        bool take_big = (c & 1) == 0;
        for (int64_t q = 0; q < k; ++q) {
            const int32_t t = take_big ? mine[(size_t)lo++] : mine[(size_t)hi--];
            take_big = !take_big;
            out[(size_t)(q * nsm + c)] = tiles[t];
        }
    }
    std::copy(out.begin(), out.end(), tiles);
}

// K2a tables: depend only on the lists
int build_plan2(qemb_workspace *w)
{
    Slot &s = w->slot[w->cur];
    const int64_t nmo = s.nmo, nA = s.nA, npair = s.npair;
    const int mtiles = (int)((w->mrows + GATHER_TILE_M - 1) / GATHER_TILE_M);
    int64_t ntiles = 0;
    for (int64_t a = 0; a < nmo; ++a) ntiles += (int64_t)mtiles * ((nmo - a + GATHER_TILE_N - 1) / GATHER_TILE_N);
    REQ(ntiles < std::numeric_limits<int32_t>::max(), "too many K2a tiles");
    size_t o_acol = 0, o_brow = align16(o_acol + nA * 4), o_bcol = align16(o_brow + nA * 4),
           o_ccol = align16(o_bcol + npair * 4), o_grp = align16(o_ccol + npair * 4),
           o_tile = align16(o_grp + nmo * sizeof(GatherGroup)), total = align16(o_tile + ntiles * sizeof(int4));
    TRY(s.stage2.reserve(total));
    TRY(s.plan2.reserve(total));
    char *h = static_cast<char *>(s.stage2.p);
    int *acol = (int *)(h + o_acol), *brow = (int *)(h + o_brow), *bcol = (int *)(h + o_bcol), *ccol = (int *)(h + o_ccol);
    GatherGroup *grp = (GatherGroup *)(h + o_grp);
    int4 *tiles = (int4 *)(h + o_tile);
    for (int64_t k = 0; k < nA; ++k) {
        acol[k] = (int)k;
        brow[k] = (int)s.aidx[k];
    }
    int64_t bo = 0, nt = 0;
    double f2 = 0;
    std::vector<float> cost((size_t)ntiles);
    for (int64_t a = 0; a < nmo; ++a) {
        const int K = (int)(s.aptr[a + 1] - s.aptr[a]), N = (int)(nmo - a);
        grp[a] = GatherGroup{K, N, (int)s.aptr[a], (int)bo};
        for (int64_t n = 0; n < N; ++n) {
            bcol[bo + n] = (int)(a + n);
            ccol[bo + n] = (int)(gauss_sum(a + n) + a); // T(max) + min, hpp:75-79
        }
        bo += N;
        const int ntn = (N + GATHER_TILE_N - 1) / GATHER_TILE_N;
        for (int m = 0; m < mtiles; ++m)
            for (int n = 0; n < ntn; ++n) {
                const int nv = std::min(GATHER_TILE_N, N - n * GATHER_TILE_N);
                cost[(size_t)nt] = (float)((K + 15) / 16) * (2.f + (float)((nv + 7) / 8));
                tiles[nt++] = make_int4((int)a, m, n, 0);
            }
        f2 += 2.0 * (double)K * (double)N;
    }
    balance_tiles(tiles, cost, nt, w->ctx->num_sms);
    s.work[1] = f2 * (double)w->naux;                               // F_K2a
    s.work[5] = 8.0 * (double)w->naux * ((double)nA + (double)npair); // Bytes_K2a
    w->ctx->h2d_bytes += (int64_t)total;
    CU(cudaMemcpyAsync(s.plan2.p, h, total, cudaMemcpyHostToDevice, w->prep));
    CU(cudaEventRecord(s.uploaded2, w->prep));
    char *d = static_cast<char *>(s.plan2.p);
    s.pd.acol2 = (const int *)(d + o_acol);
    s.pd.brow2 = (const int *)(d + o_brow);
    s.pd.bcol2 = (const int *)(d + o_bcol);
    s.pd.ccol2 = (const int *)(d + o_ccol);
    s.pd.groups2 = (const GatherGroup *)(d + o_grp);
    s.pd.tiles2 = (const int4 *)(d + o_tile);
    s.pd.ntiles2 = (int)nt;
    s.have_plan2 = true;
    return QEMB_OK;
}

// K1 tables: transpose of the lists (per AO mu: the MOs that reach it) + the tensor's R(mu) lists
int build_plan1(qemb_workspace *w, const qemb_tensor *t)
{
    Slot &s = w->slot[w->cur];
    const int64_t nao = s.nao, nmo = s.nmo, nA = s.nA;
    std::vector<int64_t> tptr(nao + 1, 0);
    for (int64_t k = 0; k < nA; ++k) tptr[s.aidx[k] + 1]++;
    int64_t ngroups = 0;
    for (int64_t mu = 0; mu < nao; ++mu) {
        if (tptr[mu + 1] > 0) ++ngroups;
        tptr[mu + 1] += tptr[mu];
    }
    const int mtiles = (int)((w->mrows + GATHER_TILE_M - 1) / GATHER_TILE_M);
    int64_t ntiles = 0;
    for (int64_t mu = 0; mu < nao; ++mu) {
        const int64_t N = tptr[mu + 1] - tptr[mu];
        ntiles += (int64_t)mtiles * ((N + GATHER_TILE_N - 1) / GATHER_TILE_N);
    }
    REQ(ntiles < std::numeric_limits<int32_t>::max(), "too many K1 tiles");
    size_t o_bcol = 0, o_ccol = align16(o_bcol + nA * 4), o_grp = align16(o_ccol + nA * 4),
           o_tile = align16(o_grp + ngroups * sizeof(GatherGroup)), total = align16(o_tile + ntiles * sizeof(int4));
    TRY(s.stage1.reserve(total));
    TRY(s.plan1.reserve(total));
    char *h = static_cast<char *>(s.stage1.p);
    int *bcol = (int *)(h + o_bcol), *ccol = (int *)(h + o_ccol);
    GatherGroup *grp = (GatherGroup *)(h + o_grp);
    int4 *tiles = (int4 *)(h + o_tile);
    {
        std::vector<int64_t> fillp(tptr.begin(), tptr.end() - 1);
        for (int64_t i = 0; i < nmo; ++i)
            for (int64_t k = s.aptr[i]; k < s.aptr[i + 1]; ++k) {
                const int64_t q = fillp[s.aidx[k]]++;
                bcol[q] = (int)i;
                ccol[q] = (int)k; // offset of (i, mu): running counter over i then mu (cpp:467-482)
            }
    }
    int64_t gi = 0, nt = 0;
    double f1 = 0;
    std::vector<float> cost((size_t)ntiles);
    std::vector<uint8_t> col_touched; // U_f of SURVEY 8(d): distinct unique columns read
    col_touched.assign((size_t)std::max<int64_t>(t->n_unique, 1), 0);
    int64_t U = 0;
    for (int64_t mu = 0; mu < nao; ++mu) {
        const int N = (int)(tptr[mu + 1] - tptr[mu]);
        if (N == 0) continue;
        const int K = (int)(t->rptr[mu + 1] - t->rptr[mu]);
        grp[gi] = GatherGroup{K, N, (int)t->rptr[mu], (int)tptr[mu]};
        const int ntn = (N + GATHER_TILE_N - 1) / GATHER_TILE_N;
        for (int m = 0; m < mtiles; ++m)
            for (int n = 0; n < ntn; ++n) {
                const int nv = std::min(GATHER_TILE_N, N - n * GATHER_TILE_N);
                cost[(size_t)nt] = (float)((K + 15) / 16) * (2.f + (float)((nv + 7) / 8));
                tiles[nt++] = make_int4((int)gi, m, n, 0);
            }
        ++gi;
        f1 += 2.0 * (double)K * (double)N;
        for (int64_t k = t->rptr[mu]; k < t->rptr[mu + 1]; ++k)
            if (!col_touched[t->rcol[k]]) {
                col_touched[t->rcol[k]] = 1;
                ++U;
            }
    }
    balance_tiles(tiles, cost, nt, w->ctx->num_sms);
    s.work[0] = f1 * (double)w->naux;                                                          // F_K1
    s.work[4] = 8.0 * (double)w->naux * ((double)U + (double)nA) + 8.0 * (double)nao * (double)nmo; // Bytes_K1
    w->ctx->h2d_bytes += (int64_t)total;
    CU(cudaMemcpyAsync(s.plan1.p, h, total, cudaMemcpyHostToDevice, w->prep));
    CU(cudaEventRecord(s.uploaded1, w->prep));
    char *d = static_cast<char *>(s.plan1.p);
    s.pd.bcol1 = (const int *)(d + o_bcol);
    s.pd.ccol1 = (const int *)(d + o_ccol);
    s.pd.groups1 = (const GatherGroup *)(d + o_grp);
    s.pd.tiles1 = (const int4 *)(d + o_tile);
    s.pd.ntiles1 = (int)nt;
    s.have_plan1 = true;
    s.plan1_tensor = t;
    return QEMB_OK;
}

int set_lists(qemb_workspace *w, const int64_t *A_ptr, const int64_t *A_idx)
{
    Slot &s = w->slot[w->cur];
    REQ(A_ptr[0] == 0, "A_ptr[0] must be 0");
    const int64_t nA = A_ptr[s.nmo];
    REQ(nA >= 0 && nA < std::numeric_limits<int32_t>::max(), "AO_by_MO too large");
    for (int64_t i = 0; i < s.nmo; ++i) {
        REQ(A_ptr[i + 1] >= A_ptr[i], "A_ptr must be non-decreasing");
        for (int64_t k = A_ptr[i]; k < A_ptr[i + 1]; ++k) REQ(A_idx[k] >= 0 && A_idx[k] < s.nao, "AO index out of range");
    }
    s.aptr.assign(A_ptr, A_ptr + s.nmo + 1);
    s.aidx.assign(A_idx, A_idx + nA);
    s.nA = nA;
    return QEMB_OK;
}

int get_gram_tiles(qemb_workspace *w, int nt, const int2 **tiles, int *count)
{
    auto it = w->gram_tiles.find(nt);
    if (it == w->gram_tiles.end()) {
        std::vector<int2> h((size_t)nt * (nt + 1) / 2);
        const int64_t n = build_gram_tiles(nt, h.data());
        DevBuf b;
        TRY(b.reserve(std::max<size_t>(h.size(), 1) * sizeof(int2)));
        // blocking copy on the prep stream: happens once per distinct tile count
        CU(cudaMemcpyAsync(b.p, h.data(), h.size() * sizeof(int2), cudaMemcpyHostToDevice, w->prep));
        CU(cudaStreamSynchronize(w->prep));
        it = w->gram_tiles.emplace(nt, std::make_pair(b, (int)n)).first;
    }
    *tiles = it->second.first.as<int2>();
    *count = it->second.second;
    return QEMB_OK;
}

} // namespace

extern "C" int qemb_fragment_screen(qemb_workspace *w, const qemb_screen *sc, const double *TA, int64_t ld,
                                    int64_t nao, int64_t nmo, double eps, int64_t *n_mu_i)
{
    REQ(w && sc && TA, "null argument");
    REQ(nao == sc->nao, "TA rows != nao of S_abs");
    REQ(nmo > 0 && ld >= nao, "bad TA shape");
    REQ(w->ctx == sc->ctx, "workspace and S_abs live on different contexts");
    DeviceGuard g(w->ctx->device);
    TRY(begin_fragment(w, nao, nmo));
    Slot &s = w->slot[w->cur];
    TRY(upload_TA(w, TA, ld));
    // the mask is written straight into mapped pinned host memory by the kernel's own stores (UVA): a
    // device -> host memcpy here would queue behind the multi-GB G transfer of the previous fragment on the
    // D2H copy engine and stall the host (measured: 110-137 ms per fragment instead of 85)
    TRY(s.hmask.reserve((size_t)nao * nmo));
    cudaStream_t ps = w->prep;
    if (w->ctx->timing) CU(cudaEventRecord(w->evs[0], ps));
    CU(launch_screen(sc->S.as<double>(), nao, s.TA.as<double>(), nao, (int)nao, (int)nmo, eps, static_cast<uint8_t *>(s.hmask.p), ps));
    w->ctx->launches += 1;
    if (w->ctx->timing) {
        CU(cudaEventRecord(w->evs[1], ps));
        w->evs_valid = true;
    }
    CU(cudaStreamSynchronize(ps));
    const uint8_t *m = static_cast<const uint8_t *>(s.hmask.p);
    s.aptr.assign(nmo + 1, 0);
    s.aidx.clear();
    s.n_borderline = 0;
    for (int64_t i = 0; i < nmo; ++i) {
        const uint8_t *col = m + i * nao;
        for (int64_t mu = 0; mu < nao; ++mu) {
            if (col[mu] & 1) s.aidx.push_back(mu);
            if (col[mu] & 2) ++s.n_borderline;
        }
        s.aptr[i + 1] = (int64_t)s.aidx.size();
    }
    s.nA = (int64_t)s.aidx.size();
    REQ(s.nA < std::numeric_limits<int32_t>::max(), "AO_by_MO too large");
    if (n_mu_i) *n_mu_i = s.nA;
    return QEMB_OK;
}

extern "C" int qemb_fragment_borderline(const qemb_workspace *w, int64_t *n)
{
    REQ(w && n && w->started, "no fragment planned");
    *n = w->slot[w->cur].n_borderline;
    return QEMB_OK;
}

extern "C" int qemb_fragment_set_lists(qemb_workspace *w, const double *TA, int64_t ld, int64_t nao, int64_t nmo,
                                       const int64_t *A_ptr, const int64_t *A_idx)
{
    REQ(w && TA && A_ptr, "null argument");
    REQ(nao > 0 && nmo > 0 && ld >= nao, "bad TA shape");
    REQ(A_idx || A_ptr[nmo] == 0, "null A_idx");
    DeviceGuard g(w->ctx->device);
    TRY(begin_fragment(w, nao, nmo));
    TRY(upload_TA(w, TA, ld));
    TRY(set_lists(w, A_ptr, A_idx));
    return QEMB_OK;
}

extern "C" int qemb_fragment_get_lists(const qemb_workspace *w, int64_t *A_ptr, int64_t *A_idx)
{
    REQ(w && w->started && A_ptr, "no fragment planned");
    const Slot &s = w->slot[w->cur];
    std::copy(s.aptr.begin(), s.aptr.end(), A_ptr);
    if (A_idx) std::copy(s.aidx.begin(), s.aidx.end(), A_idx);
    return QEMB_OK;
}

// ------------------------------------------------------------------------------------------------
// fragment: kernels
// ------------------------------------------------------------------------------------------------
namespace {

int contract_1st(qemb_workspace *w, const qemb_tensor *t, int64_t p0, int64_t p1)
{
    Slot &s = w->slot[w->cur];
    REQ(t->nao == s.nao, "tensor nao != TA rows");
    REQ(w->ctx == t->ctx, "workspace and tensor live on different contexts");
    if (p1 < 0) p1 = t->naux;
    REQ(p0 >= 0 && p0 < p1 && p1 <= t->naux, "bad auxiliary slab");
    REQ(p0 % 16 == 0 && (p1 % 16 == 0 || p1 == t->naux), "auxiliary slab bounds must be multiples of 16");
    REQ((p0 == 0 && p1 == t->naux) || t->whitened, "an auxiliary slab needs a whitened tensor");
    w->ldp = t->ldp;
    w->naux = p1 - p0;
    w->mrows = round_up(p1, 16) - p0;
    TRY(build_plan1(w, t));
    TRY(w->Bint.reserve((size_t)w->ldp * (size_t)std::max<int64_t>(s.nA, 1) * 8));
    cudaStream_t cs = w->compute;
    CU(cudaStreamWaitEvent(cs, s.ta_ready, 0));
    CU(cudaStreamWaitEvent(cs, s.uploaded1, 0));
    GatherParams p;
    p.Asrc = t->data.as<double>() + p0;
    p.lda = t->ldp;
    p.M = (int)w->mrows;
    p.Mread = (int)(t->ldp - p0);
    p.acol = t->d_rcol.as<int>();
    p.brow = t->d_ridx.as<int>();
    p.TA = s.TA.as<double>();
    p.ldta = s.nao;
    p.bcol = s.pd.bcol1;
    p.ccol = s.pd.ccol1;
    p.C = w->Bint.as<double>();
    p.ldc = w->ldp;
    p.groups = s.pd.groups1;
    p.tiles = s.pd.tiles1;
    p.ntiles = s.pd.ntiles1;
    if (w->ctx->timing) {
        CU(cudaEventRecord(w->ev[0], cs));
        w->ev_valid[0] = true;
    }
    CU(launch_gather_gemm(p, w->ctx->num_sms, cs));
    w->ctx->launches += (p.ntiles > 0);
    if (w->ctx->timing) {
        CU(cudaEventRecord(w->ev[1], cs));
        w->ev_valid[1] = true;
    }
    CU(cudaEventRecord(s.consumed, cs)); // moved forward again by contract_2nd
    s.consumed_pending = true;
    w->have_B = true;
    return QEMB_OK;
}

int contract_2nd(qemb_workspace *w)
{
    Slot &s = w->slot[w->cur];
    REQ(w->have_B, "contract_1st has not run for this fragment");
    TRY(build_plan2(w));
    TRY(w->sym.reserve((size_t)w->ldp * (size_t)s.npair * 8));
    cudaStream_t cs = w->compute;
    CU(cudaStreamWaitEvent(cs, s.uploaded2, 0));
    GatherParams p;
    p.Asrc = w->Bint.as<double>();
    p.lda = w->ldp;
    p.M = (int)w->mrows;
    p.Mread = (int)w->ldp;
    p.acol = s.pd.acol2;
    p.brow = s.pd.brow2;
    p.TA = s.TA.as<double>();
    p.ldta = s.nao;
    p.bcol = s.pd.bcol2;
    p.ccol = s.pd.ccol2;
    p.C = w->sym.as<double>();
    p.ldc = w->ldp;
    p.groups = s.pd.groups2;
    p.tiles = s.pd.tiles2;
    p.ntiles = s.pd.ntiles2;
    if (w->ctx->timing) {
        CU(cudaEventRecord(w->ev2s, cs));
        w->ev2s_valid = true;
    }
    CU(launch_gather_gemm(p, w->ctx->num_sms, cs));
    w->ctx->launches += (p.ntiles > 0);
    if (w->ctx->timing) {
        CU(cudaEventRecord(w->ev[2], cs));
        w->ev_valid[2] = true;
    }
    // TA and the plan tables of this slot are free once K2a has run
    CU(cudaEventRecord(s.consumed, cs));
    s.consumed_pending = true;
    w->have_sym = true;
    w->solved = false;
    return QEMB_OK;
}

int solve(qemb_workspace *w, const qemb_metric *m)
{
    Slot &s = w->slot[w->cur];
    REQ(w->have_sym, "contract_2nd has not run for this fragment");
    REQ(!w->solved, "already solved");
    REQ(w->ctx == m->ctx, "workspace and metric live on different contexts");
    REQ(m->naux == w->naux && m->naux_pad == w->ldp, "naux mismatch between metric and fragment data");
    TrsmParams p;
    p.L = m->L.as<double>();
    p.ldl = m->naux_pad;
    p.Dinv = m->Dinv.as<double>();
    p.nblocks = m->nblocks;
    p.X = w->sym.as<double>();
    p.ldx = w->ldp;
    p.ncol = s.npair;
    cudaStream_t cs = w->compute;
    CU(launch_trsm_rows(p, (int)w->mrows, cs));
    w->ctx->launches += 1;
    if (w->ctx->timing) {
        CU(cudaEventRecord(w->ev[3], cs));
        w->ev_valid[3] = true;
    }
    s.work[2] = (double)m->naux * (double)m->naux * (double)s.npair;                                         // F_K2b
    s.work[6] = 8.0 * ((double)m->naux * ((double)m->naux + 1) / 2 + 2.0 * (double)m->naux * (double)s.npair); // Bytes_K2b
    w->solved = true;
    return QEMB_OK;
}

// G = X^T X into `dst` (device, ld)
int gram(qemb_workspace *w, double *dst, int64_t ld, int accumulate = 0)
{
    Slot &s = w->slot[w->cur];
    REQ(w->have_sym, "contract_2nd has not run for this fragment");
    const int nt = (int)((s.npair + GRAM_TILE - 1) / GRAM_TILE);
    GramParams p;
    TRY(get_gram_tiles(w, nt, &p.tiles, &p.ntiles));
    p.X = w->sym.as<double>();
    p.ldx = w->ldp;
    p.ktiles = (int)(w->mrows / 16);
    p.npair = s.npair;
    p.G = dst;
    p.ldg = ld;
    p.accumulate = accumulate;
    const int nsplit = accumulate ? 1 : gram_splits(p.ntiles, p.ktiles, w->ctx->num_sms);
    if (nsplit > 1) { // small fragment: fewer tiles than SMs, split the auxiliary index (deterministic two-pass sum)
        TRY(w->gram_partial.reserve((size_t)nsplit * (size_t)s.npair * (size_t)s.npair * 8));
        p.partial = w->gram_partial.as<double>();
    }
    cudaStream_t cs = w->compute;
    if (w->ctx->timing && !w->ev_valid[3]) { // no solve in between: K3 starts where K2a ended
        CU(cudaEventRecord(w->ev[3], cs));
        w->ev_valid[3] = true;
    }
    // opt-in int8-slice K3 (k3_int8.cu), plain and accumulating; slabs deep enough to overflow the exact int32 sums stay
    // FP64 (digits in [-128, 127], at most `slices` digit pairs per order: slices * rows * 128^2 < 2^31)
    int slices = w->ctx->k3_slices;
    w->k3_last_expo = 0;
    if (w->ctx->k3_tol > 0.0) {
        // guaranteed-tolerance mode: e_max of X first (one pass over X, one int back: this waits for K2a, the price of
        // deciding on the host which kernel instantiation runs), then the smallest S with bound * 2^(2 e_max) <= tol
        TRY(w->k3_expo.reserve(sizeof(int)));
        TRY(w->k3_expo_host.reserve(sizeof(int)));
        CU(launch_max_expo(p.X, p.ldx, (int)w->mrows, s.npair, w->k3_expo.as<int>(), cs));
        CU(cudaMemcpyAsync(w->k3_expo_host.p, w->k3_expo.p, sizeof(int), cudaMemcpyDeviceToHost, cs));
        CU(cudaStreamSynchronize(cs));
        w->ctx->launches += 1;
        const int emax = *static_cast<int *>(w->k3_expo_host.p);
        w->k3_last_expo = emax;
        slices = 0;
        for (int S = 5; S <= 8 && !slices; ++S)
            if (emax <= -(1 << 29) || gram_i8_bound(w->mrows, S) * std::ldexp(1.0, 2 * emax) <= w->ctx->k3_tol) slices = S;
    }
    w->k3_last_slices = 0;
    if (slices && (int64_t)slices * w->mrows * 16384 < (int64_t)2147483647) {
        w->k3_last_slices = slices;
        GramI8Params q;
        q.X = p.X;
        q.ldx = p.ldx;
        q.nrows = (int)w->mrows;
        q.npair = s.npair;
        q.G = dst;
        q.ldg = ld;
        q.slices = slices;
        q.accumulate = accumulate;
        TRY(w->gram_i8.reserve(gram_i8_scratch_bytes(q.nrows, q.npair, q.slices)));
        q.scratch = w->gram_i8.as<int8_t>();
        CU(launch_gram_int8(q, w->ctx->num_sms, cs));
        w->ctx->launches += 3;
    } else {
        CU(launch_gram(p, w->ctx->num_sms, cs));
        w->ctx->launches += nsplit > 1 ? 2 : 1;
    }
    if (w->ctx->timing) {
        CU(cudaEventRecord(w->ev[4], cs));
        w->ev_valid[4] = true;
    }
    s.work[3] = (double)w->naux * (double)s.npair * ((double)s.npair + 1);                 // F_K3 (symmetric count)
    s.work[7] = 8.0 * ((double)w->naux * (double)s.npair + (double)s.npair * (double)s.npair); // Bytes_K3
    w->G_last = dst;
    w->G_last_ld = ld;
    w->have_G = true;
    return QEMB_OK;
}

// G into the workspace's own (double-buffered) storage
int gram_ws(qemb_workspace *w)
{
    Slot &s = w->slot[w->cur];
    w->gi ^= 1;
    const int gi = w->gi;
    TRY(w->G[gi].reserve((size_t)s.npair * (size_t)s.npair * 8));
    if (w->g_copy_pending[gi]) { // the copy stream may still be reading this buffer
        CU(cudaStreamWaitEvent(w->compute, w->g_copied[gi], 0));
        w->g_copy_pending[gi] = false;
    }
    return gram(w, w->G[gi].as<double>(), s.npair);
}

void reset_events(qemb_workspace *w)
{
    for (int i = 0; i < 5; ++i) w->ev_valid[i] = false;
    w->ev2s_valid = false;
    w->evs_valid = false;
}

} // namespace

extern "C" int qemb_fragment_contract_1st(qemb_workspace *w, const qemb_tensor *t)
{
    REQ(w && t && w->started, "null argument or no fragment planned");
    DeviceGuard g(w->ctx->device);
    return contract_1st(w, t, 0, -1);
}

// (P | mu i) supplied by the caller instead of computed by K1: the reference lets a user construct a SemiSparse3DTensor
// from host data (pyi:41-58, cpp:308-352) and hand it to contract_with_TA_2nd_to_sym_dense (cpp:560-605)
extern "C" int qemb_fragment_set_B(qemb_workspace *w, int64_t naux, const double *B_host, int64_t ld)
{
    REQ(w && w->started, "null argument or no fragment planned (qemb_fragment_set_lists comes first)");
    REQ(naux > 0 && naux < ((int64_t)1 << 31) - 128, "naux out of range");
    Slot &s = w->slot[w->cur];
    REQ(s.nA == 0 || (B_host && ld >= naux), "null B or ld < naux");
    DeviceGuard g(w->ctx->device);
    w->naux = naux;
    w->mrows = round_up(naux, 16);
    w->ldp = round_up(naux, 128);
    const size_t bytes = (size_t)w->ldp * (size_t)std::max<int64_t>(s.nA, 1) * 8;
    TRY(w->Bint.reserve(bytes));
    cudaStream_t cs = w->compute;
    CU(cudaStreamWaitEvent(cs, s.ta_ready, 0)); // K2a reads TA; K1, which normally orders this, does not run
    CU(cudaMemsetAsync(w->Bint.p, 0, bytes, cs)); // rows [naux, ldp) must read as zero
    if (s.nA > 0) {
        CU(cudaMemcpy2DAsync(w->Bint.p, (size_t)w->ldp * 8, B_host, (size_t)ld * 8, (size_t)naux * 8, (size_t)s.nA,
                             cudaMemcpyHostToDevice, cs));
        w->ctx->h2d_bytes += naux * s.nA * 8;
    }
    CU(cudaStreamSynchronize(cs)); // B_host may be pageable and is the caller's again when this returns
    w->have_B = true;
    w->have_sym = w->solved = w->have_G = false;
    return QEMB_OK;
}

extern "C" int qemb_fragment_contract_2nd(qemb_workspace *w)
{
    REQ(w && w->started, "null argument or no fragment planned");
    DeviceGuard g(w->ctx->device);
    return contract_2nd(w);
}

extern "C" int qemb_fragment_solve(qemb_workspace *w, const qemb_metric *m)
{
    REQ(w && m && w->started, "null argument or no fragment planned");
    DeviceGuard g(w->ctx->device);
    return solve(w, m);
}

extern "C" int qemb_fragment_gram(qemb_workspace *w)
{
    REQ(w && w->started, "null argument or no fragment planned");
    DeviceGuard g(w->ctx->device);
    return gram_ws(w);
}

extern "C" int qemb_fragment_get_B(const qemb_workspace *w, double *host, int64_t ld)
{
    REQ(w && host && w->have_B, "no K1 result available");
    const Slot &s = w->slot[w->cur];
    REQ(ld >= w->naux, "ld < naux");
    if (s.nA == 0) return QEMB_OK;
    DeviceGuard g(w->ctx->device);
    CU(cudaStreamSynchronize(w->compute));
    CU(cudaMemcpy2D(host, (size_t)ld * 8, w->Bint.p, (size_t)w->ldp * 8, (size_t)w->naux * 8, (size_t)s.nA,
                    cudaMemcpyDeviceToHost));
    return QEMB_OK;
}

extern "C" int qemb_fragment_get_sym(const qemb_workspace *w, double *host, int64_t ld)
{
    REQ(w && host && w->have_sym, "no K2a result available");
    const Slot &s = w->slot[w->cur];
    REQ(ld >= w->naux, "ld < naux");
    DeviceGuard g(w->ctx->device);
    CU(cudaStreamSynchronize(w->compute));
    CU(cudaMemcpy2D(host, (size_t)ld * 8, w->sym.p, (size_t)w->ldp * 8, (size_t)w->naux * 8, (size_t)s.npair,
                    cudaMemcpyDeviceToHost));
    return QEMB_OK;
}

extern "C" int qemb_fragment_get_G(const qemb_workspace *w, double *host, int64_t ld)
{
    REQ(w && host && w->have_G, "no K3 result available");
    const Slot &s = w->slot[w->cur];
    REQ(ld >= s.npair, "ld < npair");
    DeviceGuard g(w->ctx->device);
    CU(cudaStreamSynchronize(w->compute));
    CU(cudaMemcpy2D(host, (size_t)ld * 8, w->G_last, (size_t)w->G_last_ld * 8, (size_t)s.npair * 8, (size_t)s.npair,
                    cudaMemcpyDeviceToHost));
    return QEMB_OK;
}

extern "C" int qemb_fragment_G_dev(const qemb_workspace *w, void **dev, int64_t *npair)
{
    REQ(w && dev && w->have_G, "no K3 result available");
    *dev = w->G_last;
    if (npair) *npair = w->slot[w->cur].npair;
    return QEMB_OK;
}

namespace {

struct MirrorCb {
    OutJob *job;
    double *out;
    int64_t ld, n, c0, c1;
    bool last;
    int64_t cs; // columns >= cs arrive whole over PCIe: mirror destinations stop there
};

// runs on a driver thread once the copy group [c0, c1) has landed: hand its mirror work to the pool
void CUDART_CB mirror_callback(void *arg)
{
    MirrorCb *cb = static_cast<MirrorCb *>(arg);
    OutJob *job = cb->job;
    const int64_t width = cb->c1 - cb->c0;
    const int64_t rows = std::max<int64_t>(0, cb->cs - cb->c1); // rows below the group's diagonal block whose
                                                                 // transposed image is not delivered by PCIe
    int64_t chunk = 0, ntasks = 0;
    if (rows > 0 && width > 0) {
        // ~2 MB per task, whole 256-row passes of the tiled mirror (host_pool.cpp)
        chunk = std::max<int64_t>(256, ((int64_t)(2 << 20) / (8 * width) + 255) / 256 * 256);
        ntasks = (rows + chunk - 1) / chunk;
    }
    {
        std::lock_guard<std::mutex> lk(job->m);
        job->remaining += ntasks - (cb->last ? 1 : 0);
        if (job->remaining == 0) job_finished(*job);
    }
    double *out = cb->out;
    const int64_t ld = cb->ld, n = cb->cs, c0 = cb->c0, c1 = cb->c1;
    for (int64_t t = 0; t < ntasks; ++t) {
        const int64_t r0 = c1 + t * chunk, r1 = std::min(n, r0 + chunk);
        host_pool().submit([=] {
            mirror_block(out, ld, c0, c1, r0, r1);
            std::lock_guard<std::mutex> lk(job->m);
            if (--job->remaining == 0) job_finished(*job);
        });
    }
    delete cb;
}

// device G (n x n, ld n) -> host `out` (ld_out) on the copy stream
int send_G(qemb_workspace *w, OutJob &job, const double *G, int64_t n, double *out, int64_t ld_out)
{
    cudaStream_t cs = w->ctx->copy;
    if (w->ctx->out_mode == 0 || n < 1024) {
        w->ctx->d2h_bytes += n * n * 8;
        CU(cudaMemcpy2DAsync(out, (size_t)ld_out * 8, G, (size_t)n * 8, (size_t)n * 8, (size_t)n, cudaMemcpyDeviceToHost, cs));
        return QEMB_OK;
    }
    // column groups of ~out_group_bytes; group [c0, c1) sends rows [c0, n) (its diagonal block travels whole), the
    // host mirrors rows [c1, n) of it into G[c0:c1, c1:n] as soon as the group has landed
    {
        std::lock_guard<std::mutex> lk(job.m);
        job.remaining = 1;
        job.active = true;
        job.counted = true;
        job.ctx = w->ctx;
        std::lock_guard<std::mutex> lk2(w->ctx->out_mutex);
        w->ctx->out_pending += 1;
    }
    const int64_t target = w->ctx->out_group_bytes;
    // column split: [0, csplit) travel as lower triangle + host mirror, [csplit, n) whole
    int64_t csplit = n;
    if (w->ctx->full_frac > 0.0) csplit = std::max<int64_t>(0, (int64_t)((1.0 - w->ctx->full_frac) * (double)n) / 256 * 256);
    for (int64_t c0 = 0; c0 < n;) {
        const bool whole = c0 >= csplit;
        const int64_t r_first = whole ? 0 : c0;
        const int64_t rows = n - r_first;
        int64_t width = std::max<int64_t>(256, (target / (8 * rows) + 255) / 256 * 256);
        int64_t c1 = std::min(n, c0 + width);
        if (!whole && c1 > csplit) c1 = csplit; // groups do not straddle the split
        if (n - c1 < 256 && (whole || csplit == n)) c1 = n;
        w->ctx->d2h_bytes += rows * 8 * (c1 - c0);
        cudaError_t e = cudaMemcpy2DAsync(out + r_first + ld_out * c0, (size_t)ld_out * 8, G + r_first + n * c0, (size_t)n * 8,
                                          (size_t)rows * 8, (size_t)(c1 - c0), cudaMemcpyDeviceToHost, cs);
        MirrorCb *cb = nullptr;
        if (e == cudaSuccess) {
            cb = new MirrorCb{&job, out, ld_out, n, c0, c1, c1 == n, whole ? c1 : csplit};
            e = cudaLaunchHostFunc(cs, mirror_callback, cb);
        }
        if (e != cudaSuccess) {
            // no further callback will drop the sentinel: drop it here so that wait_job / qemb_workspace_destroy do
            // not block forever (callbacks already queued only add and remove their own mirror tasks)
            delete cb;
            {
                std::lock_guard<std::mutex> lk(job.m);
                if (--job.remaining == 0) job_finished(job);
            }
            CU(e);
        }
        c0 = c1;
    }
    return QEMB_OK;
}

} // namespace

static int transform_impl(qemb_workspace *w, const qemb_tensor *t, const qemb_metric *m, const qemb_screen *sc,
                          const double *TA, int64_t ld, int64_t nao, int64_t nmo, double eps, int64_t p_begin,
                          int64_t p_end, double *out, int64_t ld_out, int out_is_device, int64_t *ticket)
{
    if (ticket) *ticket = -1;
    REQ(w && t && sc && TA, "null argument");
    REQ(t->whitened || m, "a metric is required unless the tensor is whitened");
    REQ(!out || ld_out >= nmo * (nmo + 1) / 2, "ld_out < npair");
    DeviceGuard g(w->ctx->device);
    reset_events(w);
    {
        NvtxRange nv("qemb:screen");
        TRY(qemb_fragment_screen(w, sc, TA, ld, nao, nmo, eps, nullptr));
    }
    {
        NvtxRange nv("qemb:K1");
        TRY(contract_1st(w, t, p_begin, p_end));
    }
    {
        NvtxRange nv("qemb:K2a");
        TRY(contract_2nd(w));
    }
    if (!t->whitened) {
        NvtxRange nv("qemb:K2b");
        TRY(solve(w, m));
    }
    NvtxRange nv3("qemb:K3+output");
    Slot &s = w->slot[w->cur];
    if (out && out_is_device) return gram(w, out, ld_out, out_is_device == 2);
    TRY(gram_ws(w));
    if (!out) return QEMB_OK; // resident: consumers (qemb_fragment_jk, qemb_fragment_get_G) read it in place
    const int gi = w->gi;
    // hand G to the copy stream; the compute stream moves on to the next fragment
    cudaEvent_t done;
    CU(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
    CU(cudaEventRecord(done, w->compute));
    CU(cudaStreamWaitEvent(w->ctx->copy, done, 0));
    CU(cudaEventDestroy(done));
    w->ticket += 1; // ticket k uses G buffer / event k & 1 (gi toggles once per host-output transform)
    OutJob &job = w->jobs[w->ticket % OUT_RING];
    wait_job(job); // a ticket issued OUT_RING transforms ago that nobody waited for
    TRY(send_G(w, job, w->G[gi].as<double>(), s.npair, out, ld_out));
    CU(cudaEventRecord(w->g_copied[gi], w->ctx->copy)); // device side: K3 of the fragment after next reuses G[gi]
    CU(cudaEventRecord(w->out_done[w->ticket % OUT_RING], w->ctx->copy)); // host side: this ticket only
    w->g_copy_pending[gi] = true;
    w->ticket_gi[w->ticket & 1] = gi;
    if (ticket) *ticket = w->ticket;
    return QEMB_OK;
}

extern "C" int qemb_fragment_transform(qemb_workspace *w, const qemb_tensor *t, const qemb_metric *m,
                                       const qemb_screen *sc, const double *TA, int64_t ld, int64_t nao, int64_t nmo,
                                       double eps, int64_t p_begin, int64_t p_end, double *out, int64_t ld_out,
                                       int out_is_device, int64_t *ticket)
{
    REQ(w, "null workspace");
    NvtxRange nv("qemb:fragment_transform");
    // log level <= Info (hpp:170-204, the reference's Timer prints "[TIMER] <function> finished in <t> s" per step):
    // per-kernel CUDA-event times of this fragment.  Reading them waits for the fragment, i.e. Info-level logging
    // serialises the fragment pipeline -- like the reference's std::cout lines it is a diagnostic mode.
    const bool info = log_info();
    const bool saved = w->ctx->timing;
    if (info) w->ctx->timing = true;
    const int rc = transform_impl(w, t, m, sc, TA, ld, nao, nmo, eps, p_begin, p_end, out, ld_out, out_is_device, ticket);
    w->ctx->timing = saved;
    if (info && rc == QEMB_OK) {
        float ms[5];
        if (qemb_fragment_times(w, ms) == QEMB_OK) {
            static const char *names[5] = {"get_AO_per_MO", "contract_with_TA_1st", "contract_with_TA_2nd_to_sym_dense",
                                           "eval_via_cholesky:trsm", "eval_via_cholesky:gram"};
            for (int i = 0; i < 5; ++i)
                if (i != 3 || ms[3] > 0.f) std::printf("[TIMER] %s finished in %g s\n", names[i], (double)ms[i] * 1e-3);
            std::fflush(stdout);
        }
    }
    return rc;
}

// J/K of a symmetric density from the (ij|kl) the last transform left on the device (jk.cu).
static int jk_impl(qemb_workspace *w, const double *D, int64_t ldd, double *vj, double *vk, int64_t ldo, bool sync)
{
    REQ(w && D && vj && vk && w->have_G, "null argument or no (ij|kl) on the device");
    const Slot &s = w->slot[w->cur];
    const int64_t nmo = s.nmo, npair = s.npair;
    REQ(ldd >= nmo && ldo >= nmo, "leading dimension < nmo");
    REQ(nmo <= jk_max_nmo(), "nmo too large for the J/K kernel");
    double dmax = 0.0, amax = 0.0;
    for (int64_t l = 0; l < nmo; ++l)
        for (int64_t k = 0; k < nmo; ++k) {
            dmax = std::max(dmax, std::fabs(D[k + ldd * l]));
            amax = std::max(amax, std::fabs(D[k + ldd * l] - D[l + ldd * k]));
        }
    REQ(amax <= 1e-12 * std::max(dmax, 1e-300) || amax == 0.0, "density matrix must be symmetric (hermi=1)");
    DeviceGuard g(w->ctx->device);
    NvtxRange nv("qemb:jk");
    cudaStream_t cs = w->compute;
    // chunk table: whole packed rows, at most jk_chunk_doubles() doubles per chunk
    std::vector<int> crow{0};
    {
        int64_t used = 0;
        for (int64_t k = 0; k < nmo; ++k) {
            if (used + k + 1 > jk_chunk_doubles()) {
                crow.push_back((int)k);
                used = 0;
            }
            used += k + 1;
        }
        crow.push_back((int)nmo);
    }
    const int nch = (int)crow.size() - 1;
    // single-pass streaming kernel (jk_stream.cu) when the fragment fits it; QEMB_JK_KERNEL=rows forces jk.cu
    int nst = 0, ring_doubles = 0, stream_smem = 0;
    {
        const char *force = std::getenv("QEMB_JK_KERNEL");
        const bool want_rows = force && std::strcmp(force, "rows") == 0;
        if (!want_rows && nmo <= jk_stream_max_nmo() && (reinterpret_cast<uintptr_t>(w->G_last) & 15) == 0)
            nst = jk_stream_plan((int)nmo, &ring_doubles, &stream_smem);
    }
    const bool stream = nst > 0;
    // inputs: D (nmo^2) | dpk (npair) | crow (jk.cu chunk table)
    // (dpk 16-byte aligned and padded by two doubles: jk_stream.cu copies whole rows of it with 16-byte bulk copies)
    const size_t oD = 0, oP = ((size_t)nmo * nmo * 8 + 15) / 16 * 16, oC = oP + ((size_t)npair + 2) * 8;
    const size_t in_bytes = oC + crow.size() * 4;
    // host staging in pinned memory, two slots: an enqueued (asynchronous) call returns while its copy is in flight
    const int hk = w->jk_stage_k;
    w->jk_stage_k ^= 1;
    if (!w->jk_stage_free[hk]) CU(cudaEventCreateWithFlags(&w->jk_stage_free[hk], cudaEventDisableTiming));
    CU(cudaEventSynchronize(w->jk_stage_free[hk]));
    TRY(w->jk_stage[hk].reserve(in_bytes + 8));
    double *hin = static_cast<double *>(w->jk_stage[hk].p);
    std::memset(hin, 0, in_bytes + 8);
    for (int64_t l = 0; l < nmo; ++l)
        for (int64_t k = 0; k < nmo; ++k) hin[k + nmo * l] = D[k + ldd * l];
    for (int64_t k = 0; k < nmo; ++k)
        for (int64_t l = 0; l <= k; ++l) hin[oP / 8 + k * (k + 1) / 2 + l] = (k == l ? 1.0 : 2.0) * D[k + ldd * l];
    std::memcpy(reinterpret_cast<char *>(hin) + oC, crow.data(), crow.size() * 4);
    TRY(w->jk_in.reserve(in_bytes));
    const int njp1 = stream ? 1 : jk_warps();
    const int64_t njp = npair * njp1;
    TRY(w->jk_scratch.reserve(((size_t)njp + 2 * (size_t)npair * nmo) * 8));
    TRY(w->jk_out.reserve(2 * (size_t)nmo * nmo * 8));
    CU(cudaMemcpyAsync(w->jk_in.p, hin, in_bytes, cudaMemcpyHostToDevice, cs));
    CU(cudaEventRecord(w->jk_stage_free[hk], cs));
    w->ctx->h2d_bytes += (int64_t)in_bytes;
    JkParams p;
    p.G = w->G_last;
    p.ldg = w->G_last_ld;
    p.npair = npair;
    p.nmo = (int)nmo;
    p.D = reinterpret_cast<const double *>(static_cast<char *>(w->jk_in.p) + oD);
    p.dpk = reinterpret_cast<const double *>(static_cast<char *>(w->jk_in.p) + oP);
    p.crow = reinterpret_cast<const int *>(static_cast<char *>(w->jk_in.p) + oC);
    p.nchunks = nch;
    p.nstages = nst;
    p.ring_doubles = ring_doubles;
    p.njp = njp1;
    p.Jp = w->jk_scratch.as<double>();
    p.R1 = p.Jp + njp;
    p.R2 = p.R1 + npair * nmo;
    double *Jd = w->jk_out.as<double>(), *Kd = Jd + nmo * nmo;
    for (auto &e : w->jk_ev)
        if (!e) CU(cudaEventCreate(&e));
    CU(cudaEventRecord(w->jk_ev[0], cs));
    if (stream) {
        CU(launch_jk_stream(p, stream_smem, w->ctx->num_sms, cs));
        CU(launch_jk_reduce(p, Jd, Kd, nmo, cs));
    } else {
        CU(launch_jk(p, Jd, Kd, nmo, w->ctx->num_sms, cs));
    }
    CU(cudaEventRecord(w->jk_ev[1], cs));
    w->ctx->launches += 2;
    CU(cudaMemcpy2DAsync(vj, (size_t)ldo * 8, Jd, (size_t)nmo * 8, (size_t)nmo * 8, (size_t)nmo, cudaMemcpyDeviceToHost, cs));
    CU(cudaMemcpy2DAsync(vk, (size_t)ldo * 8, Kd, (size_t)nmo * 8, (size_t)nmo * 8, (size_t)nmo, cudaMemcpyDeviceToHost, cs));
    w->ctx->d2h_bytes += 2 * nmo * nmo * 8;
    w->jk_ms = -1.f;
    if (sync) {
        CU(cudaStreamSynchronize(cs));
        CU(cudaEventElapsedTime(&w->jk_ms, w->jk_ev[0], w->jk_ev[1]));
    }
    return QEMB_OK;
}

extern "C" int qemb_fragment_jk(qemb_workspace *w, const double *D, int64_t ldd, double *vj, double *vk, int64_t ldo)
{
    return jk_impl(w, D, ldd, vj, vk, ldo, true);
}

// Same, enqueued: returns once the density has been staged; vj / vk (pinned host memory for true asynchrony) are
// complete after qemb_ctx_synchronize.  The all-fragment consumer pipeline (driver.build_consumed) calls this right
// behind every resident transform, so that only nmo x nmo matrices ever leave the device and the compute stream
// never idles between fragments.
extern "C" int qemb_fragment_jk_enqueue(qemb_workspace *w, const double *D, int64_t ldd, double *vj, double *vk,
                                        int64_t ldo)
{
    return jk_impl(w, D, ldd, vj, vk, ldo, false);
}

// device time of the last qemb_fragment_jk (both kernels) and its algorithmic bytes (8 npair^2: G read once)
extern "C" int qemb_fragment_jk_time(const qemb_workspace *w, float *ms, double *bytes)
{
    REQ(w && ms, "null argument");
    if (w->jk_ms < 0.f && w->jk_ev[0] && w->jk_ev[1]) { // last call was enqueued: its events complete in stream order
        DeviceGuard g(w->ctx->device);
        CU(cudaEventSynchronize(w->jk_ev[1]));
        float ms_ = 0.f;
        CU(cudaEventElapsedTime(&ms_, w->jk_ev[0], w->jk_ev[1]));
        const_cast<qemb_workspace *>(w)->jk_ms = ms_;
    }
    *ms = w->jk_ms;
    if (bytes) {
        const double np = (double)w->slot[w->cur].npair;
        *bytes = 8.0 * np * np;
    }
    return QEMB_OK;
}

// ------------------------------------------------------------------------------------------------
// peer memory (aux-sharded fragments: K3 adds its tiles straight into the owner rank's matrix over NVLink)
// ------------------------------------------------------------------------------------------------
extern "C" int qemb_ipc_alloc(qemb_ctx *c, int64_t bytes, void **dev, unsigned char *handle64)
{
    REQ(c && dev && handle64 && bytes > 0, "bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    DeviceGuard g(c->device);
    *dev = nullptr;
    cudaError_t e = cudaMalloc(dev, (size_t)bytes);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return fail(QEMB_ENOMEM, "cudaMalloc of " + std::to_string(bytes) + " bytes failed");
    }
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, *dev));
    std::memcpy(handle64, &h, 64);
    return QEMB_OK;
}

extern "C" int qemb_ipc_open(qemb_ctx *c, const unsigned char *handle64, void **dev)
{
    REQ(c && dev && handle64, "null argument");
    DeviceGuard g(c->device);
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, 64);
    CU(cudaIpcOpenMemHandle(dev, h, cudaIpcMemLazyEnablePeerAccess));
    return QEMB_OK;
}

extern "C" int qemb_ipc_close(qemb_ctx *c, void *dev)
{
    REQ(c && dev, "null argument");
    DeviceGuard g(c->device);
    TRY(sync_workspaces(c));
    CU(cudaStreamSynchronize(c->compute));
    CU(cudaIpcCloseMemHandle(dev));
    return QEMB_OK;
}

extern "C" int qemb_ipc_free(qemb_ctx *c, void *dev)
{
    REQ(c && dev, "null argument");
    DeviceGuard g(c->device);
    TRY(sync_workspaces(c));
    CU(cudaStreamSynchronize(c->compute));
    CU(cudaFree(dev));
    return QEMB_OK;
}

extern "C" int qemb_device_zero(qemb_ctx *c, void *dev, int64_t bytes)
{
    REQ(c && dev && bytes >= 0, "bad argument");
    DeviceGuard g(c->device);
    TRY(join_workspaces(c));
    CU(cudaMemsetAsync(dev, 0, (size_t)bytes, c->compute));
    return QEMB_OK;
}

extern "C" int qemb_device_mirror_lower(qemb_ctx *c, double *G, int64_t ld, int64_t n)
{
    REQ(c && G && ld >= n && n >= 0, "bad argument");
    DeviceGuard g(c->device);
    TRY(join_workspaces(c));
    CU(launch_mirror_lower(G, ld, n, c->compute));
    c->launches += 1;
    return QEMB_OK;
}

extern "C" int qemb_device_to_host(qemb_ctx *c, void *host, const void *dev, int64_t bytes)
{
    REQ(c && host && dev && bytes >= 0, "bad argument");
    DeviceGuard g(c->device);
    TRY(join_workspaces(c));
    CU(cudaMemcpyAsync(host, dev, (size_t)bytes, cudaMemcpyDefault, c->compute)); // UVA: `host` may be any pointer
    c->d2h_bytes += bytes;
    CU(cudaStreamSynchronize(c->compute));
    return QEMB_OK;
}

// Self-tuning, second stage: the model of autotune_model picks a split f*, but the rates it rests on move with the
// contention of a real delivery (the copy engine loses more to the mirror's DRAM traffic in a build than in a short
// probe; measured on the 2-GPU box: model 1.29 U / 42.5 GB/s, build 32 GB/s).  So the candidates -- full squares, the
// pure triangle, the split f* -- are RUN: a test matrix of n x n doubles is delivered `reps` times back to back through
// the real machinery (column groups on the copy stream, completion callbacks, mirror tasks on the host pool, two host
// buffers in rotation) and the fastest wins.  All ranks of a node call this at the same moment and walk the candidates
// in the same order, so every candidate is timed under the contention it would see.
// Which candidate wins -- 0 full squares, 1 pure triangle, 2 a split sending the fraction upper[2] of the upper triangle
// over PCIe -- given the seconds their trial deliveries took (<= 0: not run) and the measured rates.  A trial is four
// deliveries of a matrix a tenth of a fragment: candidates it puts within 5 % of the fastest are a TIE (one N = 1 box:
// 0.03809 s for full squares, 0.03829 s for the triangle; the builds took 2.81 s and 2.30 s).  Ties go to the candidate
// the rate model makes cheapest per byte of upper triangle: 2 / P0 for full squares, max((1 + f) / P, (1 - f) / M) for
// a split f (f = 0: the pure triangle).  On every box recorded under profiles/ this is the mode that won the build
// (tests/test_driver_host.py).  Returns -1 when nothing was timed.
static int autotune_pick(const double *secs, const double *upper, double P0, double P, double M)
{
    int fastest = -1;
    for (int k = 0; k < 3; ++k)
        if (secs[k] > 0 && (fastest < 0 || secs[k] < secs[fastest])) fastest = k;
    if (fastest < 0 || !(P0 > 0 && P > 0 && M > 0)) return fastest;
    int pick = -1;
    double best_cost = 0.0;
    for (int k = 0; k < 3; ++k) {
        if (secs[k] <= 0 || secs[k] > 1.05 * secs[fastest]) continue;
        const double cost = k == 0 ? 2.0 / P0 : std::max((1.0 + upper[k]) / P, (1.0 - upper[k]) / M);
        if (pick < 0 || cost < best_cost) {
            pick = k;
            best_cost = cost;
        }
    }
    return pick;
}

// test hooks for the host-side schedules of K3 and of the gather GEMMs (host only, no device):
//   tiles: the lower-triangular tile walk of K3 (bands of QEMB_K3_GROUP row tiles), (ti, tj) pairs, returns the count
//   splits: k splits of a small K3 and the 16-row k tiles per split launch_gram derives from it
//   balance: the persistent-kernel order of a gather-GEMM tile list with the given costs (a permutation of 0 .. n-1)
extern "C" int64_t qemb_test_gram_tiles(int nt, int32_t *ti_tj)
{
    if (nt < 0 || (nt > 0 && !ti_tj)) return fail(QEMB_EINVAL, "bad argument");
    std::vector<int2> h((size_t)nt * (nt + 1) / 2 + 1);
    const int64_t n = build_gram_tiles(nt, h.data());
    for (int64_t k = 0; k < n; ++k) {
        ti_tj[2 * k] = h[(size_t)k].x;
        ti_tj[2 * k + 1] = h[(size_t)k].y;
    }
    return n;
}

extern "C" int qemb_test_gram_splits(int ntiles, int ktiles, int num_sms, int *nsplit, int *kchunk)
{
    REQ(nsplit && kchunk && ntiles >= 0 && ktiles >= 0 && num_sms > 0, "bad argument");
    *nsplit = gram_splits(ntiles, ktiles, num_sms);
    *kchunk = (ktiles + *nsplit - 1) / *nsplit;
    return QEMB_OK;
}

extern "C" int qemb_test_balance_tiles(int64_t n, const float *cost, int num_sms, int32_t *order)
{
    REQ(n >= 0 && num_sms > 0 && (n == 0 || (cost && order)), "bad argument");
    std::vector<int4> tiles((size_t)n);
    for (int64_t i = 0; i < n; ++i) tiles[(size_t)i] = make_int4((int)i, 0, 0, 0);
    balance_tiles(tiles.data(), std::vector<float>(cost, cost + n), n, num_sms);
    for (int64_t i = 0; i < n; ++i) order[i] = tiles[(size_t)i].x;
    return QEMB_OK;
}

// test hook for the rule above (host only, no device): secs3 / upper3 = the three candidates
extern "C" int qemb_autotune_pick(const double *secs3, const double *upper3, double d2h_alone_gbs, double d2h_gbs,
                                  double mirror_gbs, int *pick)
{
    REQ(secs3 && upper3 && pick, "null argument");
    *pick = autotune_pick(secs3, upper3, d2h_alone_gbs, d2h_gbs, mirror_gbs);
    return QEMB_OK;
}

extern "C" int qemb_ctx_autotune_output(qemb_ctx *c, double *d2h_gbs, double *mirror_gbs, double *upper_fraction_pcie)
{
    double P = 0, M = 0, f = 0;
    TRY(autotune_model(c, &P, &M, &f));
    if (d2h_gbs) *d2h_gbs = P;
    if (mirror_gbs) *mirror_gbs = M;
    if (upper_fraction_pcie) *upper_fraction_pcie = f; // the model's choice stands unless the trials below complete
    DeviceGuard g(c->device);
    const int model_mode = c->out_mode;
    const double model_frac = c->full_frac;
    const int64_t n = 8192; // 0.5 GB per delivery
    const int reps = 4;
    const size_t bytes = (size_t)n * n * 8;
    struct Cand {
        int mode;
        double frac, upper, secs;
    } cand[3] = {{0, 0.0, 1.0, -1.0}, {1, 0.0, 0.0, -1.0}, {1, 0.0, f, -1.0}};
    int ncand = 2;
    if (f > 0.03 && f < 0.9) {
        cand[2].frac = 1.0 - std::sqrt(1.0 - f);
        ncand = 3;
    }
    for (int k = 0; k < 3; ++k) c->tune_secs[k] = -1.0;
    qemb_workspace *w = nullptr;
    void *dv = nullptr, *hb[2] = {nullptr, nullptr};
    auto cleanup = [&] {
        if (w) qemb_workspace_destroy(w);
        if (dv) cudaFree(dv);
        for (auto &h : hb)
            if (h) cudaFreeHost(h);
    };
    const int rc = qemb_workspace_create(c, &w);
    cudaError_t e = rc == QEMB_OK ? cudaMalloc(&dv, bytes) : cudaErrorUnknown;
    for (int i = 0; i < 2 && e == cudaSuccess; ++i) e = cudaMallocHost(&hb[i], bytes);
    if (e == cudaSuccess) e = cudaMemsetAsync(dv, 0, bytes, c->copy);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->copy);
    if (rc != QEMB_OK || e != cudaSuccess) {
        (void)cudaGetLastError();
        cleanup();
        return QEMB_OK;
    }
    for (int k = 0; k < ncand; ++k) {
        c->out_mode = cand[k].mode;
        c->full_frac = cand[k].frac;
        bool ok = true;
        std::chrono::steady_clock::time_point t0;
        for (int r = 0; r < reps + 1 && ok; ++r) { // the first delivery warms the path up and is not timed
            if (r == 1) {
                ok = qemb_fragment_wait_output(w, w->ticket) == QEMB_OK;
                t0 = std::chrono::steady_clock::now();
            }
            if (r >= 3) ok = ok && qemb_fragment_wait_output(w, w->ticket - 1) == QEMB_OK; // host buffer r & 1 is free again
            w->ticket += 1;
            OutJob &job = w->jobs[w->ticket % OUT_RING];
            wait_job(job);
            ok = ok && send_G(w, job, static_cast<double *>(dv), n, static_cast<double *>(hb[r & 1]), n) == QEMB_OK;
            ok = ok && cudaEventRecord(w->out_done[w->ticket % OUT_RING], c->copy) == cudaSuccess;
        }
        ok = ok && qemb_fragment_wait_output(w, w->ticket - 1) == QEMB_OK && qemb_fragment_wait_output(w, w->ticket) == QEMB_OK;
        if (ok) cand[k].secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        qemb_ctx_synchronize(c);
    }
    int best = -1;
    for (int k = 0; k < ncand; ++k)
        if (cand[k].secs > 0 && (best < 0 || cand[k].secs < cand[best].secs)) best = k;
    if (best >= 0) {
        double secs[3], upper[3];
        for (int k = 0; k < 3; ++k) {
            secs[k] = k < ncand ? cand[k].secs : -1.0;
            upper[k] = cand[k].upper;
        }
        const int pick = autotune_pick(secs, upper, c->tune_p0, P, M);
        if (pick >= 0) best = pick;
    }
    cleanup();
    if (best < 0) { // nothing could be timed: keep the model's choice
        (void)cudaGetLastError();
        c->out_mode = model_mode;
        c->full_frac = model_frac;
        return QEMB_OK;
    }
    c->out_mode = cand[best].mode;
    c->full_frac = cand[best].frac;
    for (int k = 0; k < ncand; ++k) c->tune_secs[k] = cand[k].secs;
    if (upper_fraction_pcie) *upper_fraction_pcie = cand[best].upper;
    return QEMB_OK;
}

// seconds the candidates of the last qemb_ctx_autotune_output took for their test deliveries (4 x 0.5 GB): full
// squares, pure triangle, model split (-1: not run)
extern "C" int qemb_ctx_autotune_trials(const qemb_ctx *c, double *secs3)
{
    REQ(c && secs3, "null argument");
    for (int k = 0; k < 3; ++k) secs3[k] = c->tune_secs[k];
    return QEMB_OK;
}

// how many host-output transforms may be in flight per workspace before the oldest ticket's completion state is
// recycled (callers keep at most this many pinned output buffers pending)
extern "C" int qemb_output_ring_depth(void) { return OUT_RING - 1; }

extern "C" int qemb_fragment_wait_output(qemb_workspace *w, int64_t ticket)
{
    REQ(w, "null workspace");
    REQ(ticket >= 1 && ticket <= w->ticket, "unknown ticket");
    REQ(ticket > w->ticket - OUT_RING, "ticket too old: its completion state has been recycled");
    // one event per ticket of the ring: waiting for ticket k must not wait for the copies of a later ticket
    // (an event shared by parity did exactly that with three host buffers: 145 ms per fragment instead of 85)
    DeviceGuard g(w->ctx->device);
    CU(cudaEventSynchronize(w->out_done[ticket % OUT_RING]));
    wait_job(w->jobs[ticket % OUT_RING]); // host mirror tasks of a triangle transfer
    return QEMB_OK;
}

extern "C" int qemb_fragment_work(const qemb_workspace *w, const qemb_tensor *, double *out8)
{
    REQ(w && out8 && w->started, "no fragment planned");
    std::copy(w->slot[w->cur].work, w->slot[w->cur].work + 8, out8);
    return QEMB_OK;
}

extern "C" int qemb_fragment_times(const qemb_workspace *w, float *ms5)
{
    REQ(w && ms5, "null argument");
    DeviceGuard g(w->ctx->device);
    CU(cudaStreamSynchronize(w->prep));
    CU(cudaStreamSynchronize(w->compute));
    for (int i = 0; i < 5; ++i) ms5[i] = 0.f;
    if (w->evs_valid) CU(cudaEventElapsedTime(&ms5[0], w->evs[0], w->evs[1]));
    for (int i = 0; i < 4; ++i)
        if (w->ev_valid[i] && w->ev_valid[i + 1]) CU(cudaEventElapsedTime(&ms5[i + 1], w->ev[i], w->ev[i + 1]));
    if (w->ev2s_valid && w->ev_valid[2]) CU(cudaEventElapsedTime(&ms5[2], w->ev2s, w->ev[2]));
    return QEMB_OK;
}

// ------------------------------------------------------------------------------------------------
// DMMA throughput probe: the gram kernel on an L2-resident problem
// ------------------------------------------------------------------------------------------------
extern "C" int qemb_probe_dmma_tflops(qemb_ctx *ctx, int64_t n, int64_t k, int iters, double *tflops)
{
    REQ(ctx && tflops && n > 0 && k > 0 && iters > 0, "bad argument");
    DeviceGuard g(ctx->device);
    const int64_t ldx = round_up(k, 128);
    DevBuf X, G, T;
    TRY(X.reserve((size_t)ldx * n * 8));
    TRY(G.reserve((size_t)n * n * 8));
    const int nt = (int)((n + GRAM_TILE - 1) / GRAM_TILE);
    std::vector<int2> h((size_t)nt * (nt + 1) / 2);
    const int64_t ntiles = build_gram_tiles(nt, h.data());
    TRY(T.reserve(h.size() * sizeof(int2)));
    cudaStream_t s = ctx->compute;
    CU(cudaMemcpyAsync(T.p, h.data(), h.size() * sizeof(int2), cudaMemcpyHostToDevice, s));
    CU(launch_fill(X.as<double>(), ldx * n, 1e-3, s));
    GramParams p;
    p.X = X.as<double>();
    p.ldx = ldx;
    p.ktiles = (int)(round_up(k, 16) / 16);
    p.npair = n;
    p.G = G.as<double>();
    p.ldg = n;
    p.tiles = T.as<int2>();
    p.ntiles = (int)ntiles;
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    CU(launch_gram(p, ctx->num_sms, s));
    CU(cudaEventRecord(e0, s));
    for (int i = 0; i < iters; ++i) CU(launch_gram(p, ctx->num_sms, s));
    CU(cudaEventRecord(e1, s));
    CU(cudaStreamSynchronize(s));
    ctx->launches += iters + 2;
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    // executed flops: every tile is a full 128 x 128 x k16 block
    const double fl = 2.0 * 128.0 * 128.0 * (double)(p.ktiles * 16) * (double)ntiles * iters;
    *tflops = fl / (ms * 1e-3) / 1e12;
    X.release();
    G.release();
    T.release();
    return QEMB_OK;
}
