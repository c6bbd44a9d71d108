// gram_noalloc.cu -- qemb_gram_noalloc: the split-K form of K3 (G = X^T X, k3_gram.cu) for callers that must not
// allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  It has NO counterpart
// in the reference.
//
// The workspace path (qemb_fragment_gram) keeps the split-K partial matrices and the tile table in cudaMalloc'ed
// buffers that grow on demand.  Here
//   * the partial matrices live in ONE file-scope `static __device__` array of a fixed size (NA_SCRATCH_BYTES, 32 MiB,
//     held by every GPU that loads this module until the process exits),
//   * the tile (ti, tj) of a work item is decoded from its index inside the kernel (no tile table),
//   * the tensor map reads X with its true row count (TMA zero-fills the k tail and the rows past npair), so the caller's
//     X needs no zero padding,
// and the entry point calls no cudaMalloc / cudaFree / allocating library routine: argument checks, one tensor-map
// encode on the host, two kernel launches on the caller's stream.
//
// In host code the name `g_na_scratch` is only a host-side stand-in for the device array, not a device address; the
// address a kernel may dereference is obtained ONCE PER DEVICE, with that device current, by cudaGetSymbolAddress
// (na_setup) and handed to the kernels as an ordinary pointer argument.
//
// Every call shares the array and finds the previous call's contents in it.  Nothing has to start at zero: every split
// STORES (never accumulates) every entry of its npair x npair partial matrix -- lower tiles directly, upper tiles as
// mirrors, a split whose k range is empty stores zeros -- before the reduction kernel of the same call, which follows on
// the same stream, reads them.  Calls that overlap on the GPU (two streams, two host threads, two processes sharing
// nothing but each its own copy -- the array is per process and device) RACE on the array: serialise them.
#include <cuda.h>

#include <mutex>
#include <string>

#include "../../include/qemb_b200.h"
#include "dmma_core.cuh"
#include "kernels.cuh"

namespace qemb {

int api_fail(int code, const char *msg); // api.cu: sets qemb_last_error() of the calling thread, returns code

namespace {

constexpr int64_t NA_SCRATCH_BYTES = 32ll << 20;
constexpr int64_t NA_SCRATCH_DOUBLES = NA_SCRATCH_BYTES / 8;
constexpr int64_t NA_MAX_NPAIR = 2048; // 2048^2 doubles = the whole array (one split)

static __device__ __align__(256) double g_na_scratch[NA_SCRATCH_DOUBLES];

constexpr int TM = 128, TN = 128;
constexpr int WM = 64, WN = 32;
constexpr int NCONSUMER_WARPS = 8;
constexpr int NTHREADS = 32 * (NCONSUMER_WARPS + 4);
constexpr int STAGES = 6;
constexpr int TILE_BYTES = 128 * BK * 8;
constexpr int STAGE_BYTES = 2 * TILE_BYTES;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 2 * STAGES * 8;

__device__ __forceinline__ void na_mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void na_mbar_expect_tx(uint64_t *bar, int bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void na_mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void na_mbar_wait(uint64_t *bar, int parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "NA_WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra NA_DONE;\n"
        "bra NA_WAIT_LOOP;\n"
        "NA_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void na_tma_load_2d(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_u32(dst)),
        "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}

struct NaParams {
    int ktiles;  // 16-wide k steps (the last one may reach past naux: zero-filled by TMA)
    int nt;      // row tiles of G
    int ntiles;  // nt (nt + 1) / 2 lower tiles
    int nsplit, kchunk;
    int64_t npair;
    double *partial; // nsplit x npair x npair, inside the static scratch array
};

// lower tile number q -> (ti >= tj), column by column (nt <= 16: a loop is cheaper than a square root)
__device__ __forceinline__ void na_tile(int q, int nt, int &ti, int &tj)
{
    int j = 0;
    while (q >= nt - j) {
        q -= nt - j;
        ++j;
    }
    tj = j;
    ti = j + q;
}

// The mainloop of gram_kernel (k3_gram.cu: TMA producer warp, 8 DMMA consumer warps, deferred stage release); always
// split-K into the scratch matrices, tiles decoded in place.
__global__ void __launch_bounds__(NTHREADS, 1) gram_noalloc_kernel(const __grid_constant__ CUtensorMap tmap, NaParams p)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
    uint64_t *empty = full + STAGES;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            na_mbar_init(&full[s], 1);
            na_mbar_init(&empty[s], NCONSUMER_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int nwork = p.ntiles * p.nsplit;

    if (warp >= NCONSUMER_WARPS) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == NCONSUMER_WARPS && lane == 0) {
            int s = 0, phase = 0;
            for (int w = blockIdx.x; w < nwork; w += gridDim.x) {
                int ti, tj;
                na_tile(w / p.nsplit, p.nt, ti, tj);
                const int m0 = ti * TM, n0 = tj * TN;
                const int kbeg = (w % p.nsplit) * p.kchunk, kend = min(p.ktiles, kbeg + p.kchunk);
                for (int kt = kbeg; kt < kend; ++kt) {
                    na_mbar_wait(&empty[s], phase ^ 1);
                    na_mbar_expect_tx(&full[s], STAGE_BYTES);
                    uint8_t *As = smem + s * STAGE_BYTES;
                    na_tma_load_2d(As, &tmap, kt * BK, m0, &full[s]);
                    na_tma_load_2d(As + TILE_BYTES, &tmap, kt * BK, n0, &full[s]);
                    if (++s == STAGES) {
                        s = 0;
                        phase ^= 1;
                    }
                }
            }
        }
        return;
    }

    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const int wm0 = (warp & 1) * WM, wn0 = (warp >> 1) * WN;
    const int g = lane >> 2, t = lane & 3;
    int koff[BK / 4];
#pragma unroll
    for (int ks = 0; ks < BK / 4; ++ks) koff[ks] = (((4 * (t >> 1) + ks) ^ g) << 1) + (t & 1);

    int s = 0, phase = 0, held = -1;
    for (int w = blockIdx.x; w < nwork; w += gridDim.x) {
        int ti, tj;
        na_tile(w / p.nsplit, p.nt, ti, tj);
        const int64_t m0 = (int64_t)ti * TM, n0 = (int64_t)tj * TN;
        const int split = w % p.nsplit;
        const int kbeg = split * p.kchunk, kend = min(p.ktiles, kbeg + p.kchunk);
        double *__restrict__ Gout = p.partial + (int64_t)split * p.npair * p.npair;
        WarpAcc<WM, WN> acc;
        acc.zero();
        for (int kt = kbeg; kt < kend; ++kt) {
            na_mbar_wait(&full[s], phase);
            if (held >= 0) { // release of the PREVIOUS stage: all its DMMAs have been issued (see k3_gram.cu)
                __syncwarp();
                if (lane == 0) na_mbar_arrive(&empty[held]);
            }
            held = s;
            const double *As = reinterpret_cast<const double *>(smem + s * STAGE_BYTES);
            const double *Bs = As + TILE_BYTES / 8;
#pragma unroll
            for (int ks = 0; ks < BK / 4; ++ks) {
                double a[WM / 8], b[WN / 8];
#pragma unroll
                for (int i = 0; i < WM / 8; ++i) a[i] = As[(wm0 + 8 * i + g) * BK + koff[ks]];
#pragma unroll
                for (int j = 0; j < WN / 8; ++j) b[j] = Bs[(wn0 + 8 * j + g) * BK + koff[ks]];
#pragma unroll
                for (int i = 0; i < WM / 8; ++i)
#pragma unroll
                    for (int j = 0; j < WN / 8; ++j) dmma884(acc.c[i][j][0], acc.c[i][j][1], a[i], b[j]);
            }
            if (++s == STAGES) {
                s = 0;
                phase ^= 1;
            }
        }
        const bool offdiag = ti != tj;
#pragma unroll
        for (int i = 0; i < WM / 8; ++i) {
            const int64_t row = m0 + wm0 + 8 * i + g;
            if (row >= p.npair) continue;
#pragma unroll
            for (int j = 0; j < WN / 8; ++j) {
                const int64_t col = n0 + wn0 + 8 * j + 2 * t;
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    if (col + e < p.npair) {
                        const double v = acc.c[i][j][e];
                        Gout[row + p.npair * (col + e)] = v;
                        if (offdiag) Gout[(col + e) + p.npair * row] = v;
                    }
                }
            }
        }
    }
}

// G[:, c] = sum_s partial[s][:, c], s = 0 .. nsplit-1 in that order (deterministic)
__global__ void gram_noalloc_reduce_kernel(const double *__restrict__ P, int nsplit, int64_t npair, double *__restrict__ G,
                                           int64_t ldg)
{
    const int64_t total = npair * npair;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        double v = 0.0;
        for (int s2 = 0; s2 < nsplit; ++s2) v += P[(int64_t)s2 * total + i];
        const int64_t c = i / npair, r = i - c * npair;
        G[r + ldg * c] = v;
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// one-off set-up per device (the first call on a device; under graph capture, make that call before the capture)
struct NaDevice {
    bool ready = false;
    double *scratch = nullptr; // device address of g_na_scratch on this device
    int num_sms = 0;
};
constexpr int NA_MAX_DEVICES = 64;
NaDevice g_na_dev[NA_MAX_DEVICES];
EncodeTiledFn g_na_encode = nullptr;
std::mutex g_na_mutex;

int na_fail_cuda(const char *what, cudaError_t e)
{
    const std::string msg = std::string("qemb_gram_noalloc: ") + what + ": " + cudaGetErrorString(e);
    return api_fail(e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? QEMB_ENODEV : QEMB_ECUDA, msg.c_str());
}
#define NA_CU(call)                                                                                     \
    do {                                                                                                \
        cudaError_t e__ = (call);                                                                       \
        if (e__ != cudaSuccess) return na_fail_cuda(#call, e__);                                        \
    } while (0)
#define NA_REQ(cond, msg)                                                                               \
    do {                                                                                                \
        if (!(cond)) return api_fail(QEMB_EINVAL, "qemb_gram_noalloc: " msg " [" #cond "]");           \
    } while (0)

// `dev` is the CURRENT device of the calling thread
int na_setup(int dev, NaDevice **out)
{
    std::lock_guard<std::mutex> lock(g_na_mutex);
    NaDevice &d = g_na_dev[dev];
    if (!d.ready) {
        int major = 0;
        NA_CU(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        if (major != 10) return api_fail(QEMB_ENODEV, "qemb_gram_noalloc: the current device is not sm_100 (no CPU fallback)");
        NA_CU(cudaDeviceGetAttribute(&d.num_sms, cudaDevAttrMultiProcessorCount, dev));
        void *ptr = nullptr;
        NA_CU(cudaGetSymbolAddress(&ptr, g_na_scratch)); // the symbol's address ON THE CURRENT DEVICE
        if (!ptr) return api_fail(QEMB_ECUDA, "qemb_gram_noalloc: cudaGetSymbolAddress returned a null address");
        d.scratch = static_cast<double *>(ptr);
        NA_CU(cudaFuncSetAttribute(gram_noalloc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        if (!g_na_encode) {
            void *fn = nullptr;
            cudaDriverEntryPointQueryResult q;
            NA_CU(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
            if (q != cudaDriverEntryPointSuccess || !fn)
                return api_fail(QEMB_ECUDA, "qemb_gram_noalloc: cuTensorMapEncodeTiled is not available in this driver");
            g_na_encode = reinterpret_cast<EncodeTiledFn>(fn);
        }
        d.ready = true;
    }
    *out = &d;
    return QEMB_OK;
}

} // namespace
} // namespace qemb

using namespace qemb;

extern "C" int qemb_gram_noalloc_limits(int64_t *max_npair, int64_t *scratch_bytes)
{
    if (max_npair) *max_npair = NA_MAX_NPAIR;
    if (scratch_bytes) *scratch_bytes = NA_SCRATCH_BYTES;
    return QEMB_OK;
}

extern "C" int qemb_gram_noalloc(const double *X_dev, int64_t ldx, int64_t naux, int64_t npair, double *G_dev, int64_t ldg,
                                 void *stream)
{
    NA_REQ(X_dev && G_dev, "null device pointer");
    NA_REQ(naux >= 1 && naux <= (int64_t)1 << 30, "naux out of range");
    NA_REQ(npair >= 1, "npair < 1");
    if (npair > NA_MAX_NPAIR) {
        const std::string msg = "qemb_gram_noalloc: npair = " + std::to_string(npair) + " exceeds the maximum of " +
                                std::to_string(NA_MAX_NPAIR) + " this entry point's fixed " +
                                std::to_string(NA_SCRATCH_BYTES >> 20) +
                                " MiB scratch array holds (use qemb_fragment_gram / qemb_fragment_transform, which allocate)";
        return api_fail(QEMB_EINVAL, msg.c_str());
    }
    NA_REQ(ldx >= naux, "ldx < naux");
    NA_REQ(ldx % 2 == 0 && reinterpret_cast<uintptr_t>(X_dev) % 16 == 0,
           "X must be 16-byte aligned with an even leading dimension (TMA global stride)");
    NA_REQ(ldg >= npair, "ldg < npair");

    int dev = -1;
    NA_CU(cudaGetDevice(&dev));
    NA_REQ(dev >= 0 && dev < NA_MAX_DEVICES, "device ordinal out of range");
    NaDevice *d = nullptr;
    if (int rc = na_setup(dev, &d)) return rc;

    cudaPointerAttributes ax, ag;
    NA_CU(cudaPointerGetAttributes(&ax, X_dev));
    NA_CU(cudaPointerGetAttributes(&ag, G_dev));
    NA_REQ(ax.type == cudaMemoryTypeDevice && ax.device == dev, "X is not device memory of the current device");
    NA_REQ((ag.type == cudaMemoryTypeDevice && ag.device == dev) || ag.type == cudaMemoryTypeManaged,
           "G is not device memory of the current device");

    NaParams p;
    p.ktiles = (int)((naux + BK - 1) / BK);
    p.nt = (int)((npair + TM - 1) / TM);
    p.ntiles = p.nt * (p.nt + 1) / 2;
    p.npair = npair;
    p.partial = d->scratch;
    // as many splits of the auxiliary index as the workspace path would take, capped by what the array holds
    int nsplit = gram_splits(p.ntiles, p.ktiles, d->num_sms);
    const int64_t fit = NA_SCRATCH_DOUBLES / (npair * npair);
    if (nsplit > fit) nsplit = (int)fit;
    if (nsplit < 1) nsplit = 1;
    p.kchunk = (p.ktiles + nsplit - 1) / nsplit;
    p.nsplit = (p.ktiles + p.kchunk - 1) / p.kchunk; // no empty split

    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {(cuuint64_t)naux, (cuuint64_t)npair};
    const cuuint64_t gstride[1] = {(cuuint64_t)ldx * 8};
    const cuuint32_t box[2] = {BK, 128};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = g_na_encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double *>(X_dev), gdim, gstride,
                                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return api_fail(QEMB_ECUDA, ("qemb_gram_noalloc: cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r)).c_str());

    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int nwork = p.ntiles * p.nsplit;
    const int grid = nwork < d->num_sms ? nwork : d->num_sms;
    gram_noalloc_kernel<<<grid, NTHREADS, SMEM_BYTES, s>>>(tmap, p);
    NA_CU(cudaGetLastError());
    const int64_t blocks = (npair * npair + 255) / 256;
    const int64_t cap = (int64_t)d->num_sms * 8;
    gram_noalloc_reduce_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, s>>>(p.partial, p.nsplit, npair, G_dev, ldg);
    NA_CU(cudaGetLastError());
    return QEMB_OK;
}
