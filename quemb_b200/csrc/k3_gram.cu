// k3_gram.cu -- K3: (ij|kl) = sum_P X[P,ij] X[P,kl]      (reference: eval_via_cholesky step 2,
// src/quemb/molbe/_cpp/eri_sparse_DF.cpp:620, and the cuBLAS Dsyrk + host mirror of its GPU variant,
// cpp:684-708).
//
// Symmetric FP64 tensor-core contraction.  X is (ldx x npair) column-major, i.e. every pair column is
// k-contiguous -- exactly the "row.col" operand order of DMMA for both operands.  Only the nt(nt+1)/2
// lower 128x128 tiles are computed (naux*npair*(npair+1) flops instead of the reference CPU path's full
// 2*naux*npair^2 GEMM); every off-diagonal tile is also stored transposed, so the full npair x npair
// square the consumers read (pfrag.py:383, mbe.py:1155) is emitted directly in s4 pair order.
//
// Data movement (measured on B200, npair 24310, naux 2716: 35.9 TFLOP/s = 96 % of the 37.2 TFLOP/s DMMA
// peak, cuBLAS DGEMM 35.5; the cp.async + bar.sync version of the same tiling reached 29-31):
//   * one producer warp issues TMA tile loads (cp.async.bulk.tensor.2d, 128B swizzle) of the two
//     128-row x 16-k operand slabs of X into a ring of shared-memory stages;
//   * eight consumer warps (64x32 register tiles of DMMA.8x8x4 accumulators) wait on the stage's "full"
//     mbarrier, read their fragments with 64-bit LDS, and release the stage through its "empty"
//     mbarrier -- there is NO CTA-wide barrier in the main loop, so a warp that waits for shared memory
//     never holds back the other warp of its scheduler (the cp.async version idles the DMMA pipe ~16 %
//     of the time on bar.sync + LDS + cp.async-wait stalls that hit all warps at once; ncu, profiles/);
//   * the kernel is persistent (one CTA per SM): the producer runs ahead into the next tile while the
//     consumers store the finished one, which hides the pipeline fill and the epilogue.
//
// Bank conflicts: with the 128B swizzle the 16-byte chunk c of row r sits at chunk c ^ (r & 7).  The DMMA
// fragment pattern (lane = 4g + t reads row g, k = t) would then be 2-way conflicted, so the k index
// inside a 16-wide stage is permuted: k-step ks, lane t reads k = 8*(t>>1) + 2*ks + (t&1).  Both
// operands use the same permutation, so the sum over k is unchanged; the 16 lanes of an LDS.64 phase hit
// 16 distinct bank pairs.
#include <cuda.h>

#include <cstdlib>

#include "dmma_core.cuh"
#include "kernels.cuh"

namespace qemb {

namespace {
constexpr int TM = 128, TN = 128;
constexpr int WM = 64, WN = 32;
constexpr int NCONSUMER_WARPS = 8;
constexpr int NTHREADS = 32 * (NCONSUMER_WARPS + 4); // + one producer warpgroup (register file is split per
                                                     // scheduler: a 9th warp would cap everybody at 168 regs)
constexpr int STAGES = 6;
constexpr int TILE_BYTES = 128 * BK * 8; // 16 KB per operand slab
constexpr int STAGE_BYTES = 2 * TILE_BYTES;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /* alignment slack */ + 2 * STAGES * 8;

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, int bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, int parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_u32(dst)),
        "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}
// same, with an L2 eviction-priority hint (createpolicy): the row panels of the resident band of tiles are wanted
// again by the next wave of CTAs (evict_last), the column panels stream through once per band (evict_first)
__device__ __forceinline__ void tma_load_2d_hint(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar,
                                                 uint64_t policy)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;"
        ::"r"(smem_u32(dst)),
        "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
} // namespace

struct GramTmaParams {
    int ktiles; // 16-wide k steps
    int64_t npair;
    double *G;
    int64_t ldg;
    const int2 *tiles;
    int ntiles;
    // split-K for small npair (fewer tiles than SMs): work item w = (tile w / nsplit, k range of split w % nsplit),
    // split s writes its partial tile into G + s * split_stride; gram_reduce_kernel adds them up in a fixed order
    int nsplit, kchunk;
    int64_t split_stride;
    // 1: the lower tiles are ADDED (system-scope red.add.f64) into G, which may live in a peer GPU's memory (NVLink):
    // the reduce-sum of an aux-sharded fragment happens in the epilogue, tile by tile, while the other tiles compute
    int accumulate;
    int hint; // L2 hints of the operand loads: bits 0-1: 0 none, 1 row panel evict_last, 2 + column panel evict_first;
              // bit 2 (value 4): alternate the k direction from tile to tile
};

__global__ void __launch_bounds__(NTHREADS, 1)
gram_kernel(const __grid_constant__ CUtensorMap tmap, GramTmaParams p)
{
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment: the swizzle phase of a row is taken from address bits 7..9
    uint8_t *smem = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
    uint64_t *empty = full + STAGES;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NCONSUMER_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp >= NCONSUMER_WARPS) {
        // ---------------- producer warpgroup: gives its registers to the consumers ----------------
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == NCONSUMER_WARPS && lane == 0) {
            int s = 0, phase = 0;
            uint64_t pol_last, pol_first;
            asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_last));
            asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_first));
            for (int w = blockIdx.x; w < p.ntiles * p.nsplit; w += gridDim.x) {
                const int2 tt = p.tiles[w / p.nsplit];
                const int m0 = tt.x * TM, n0 = tt.y * TN;
                const int kbeg = (w % p.nsplit) * p.kchunk, kend = min(p.ktiles, kbeg + p.kchunk);
                // hint bit 2 (experiment, off by default): every other tile of a CTA walks k downwards.  The CTAs of a
                // wave advance through k together, so when a wave ends L2 holds the k TAILS of its panels; a next wave
                // that starts at k = 0 finds none of them (cyclic LRU thrash), one that starts at the tail re-uses
                // the band's row panels as far as they survived.  Only the producer's coordinates change: the
                // consumers accumulate stages in arrival order (sum order of such a tile reversed, still deterministic).
                const bool rev = (p.hint & 4) && (((w - (int)blockIdx.x) / (int)gridDim.x) & 1);
                for (int kq = kbeg; kq < kend; ++kq) {
                    const int kt = rev ? kend - 1 - (kq - kbeg) : kq;
                    mbar_wait(&empty[s], phase ^ 1);
                    mbar_expect_tx(&full[s], STAGE_BYTES);
                    uint8_t *As = smem + s * STAGE_BYTES;
                    if (p.hint & 3)
                        tma_load_2d_hint(As, &tmap, kt * BK, m0, &full[s], pol_last);
                    else
                        tma_load_2d(As, &tmap, kt * BK, m0, &full[s]);
                    if ((p.hint & 3) >= 2)
                        tma_load_2d_hint(As + TILE_BYTES, &tmap, kt * BK, n0, &full[s], pol_first);
                    else
                        tma_load_2d(As + TILE_BYTES, &tmap, kt * BK, n0, &full[s]);
                    if (++s == STAGES) {
                        s = 0;
                        phase ^= 1;
                    }
                }
            }
        }
        return;
    }

    // ---------------- consumers ----------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const int wm0 = (warp & 1) * WM, wn0 = (warp >> 1) * WN;
    const int g = lane >> 2, t = lane & 3;
    int koff[BK / 4]; // swizzled position (in doubles) of this lane's k element inside a 16-double row
#pragma unroll
    for (int ks = 0; ks < BK / 4; ++ks) koff[ks] = (((4 * (t >> 1) + ks) ^ g) << 1) + (t & 1);

    // Stage release is DEFERRED by one stage: the "empty" arrive of stage s-1 is issued after the wait for stage s.
    // ptxas schedules an arrive that directly follows the stage body right behind the last LDS *issue*, ahead of
    // the DMMAs that consume them (seen in the SASS), i.e. before those loads are known to have completed; rows
    // of the refill (TMA, async proxy) then sporadically reached a slow LDS: one 128-byte row of one warp wrong
    // in ~1 of 10^4 tiles.  Behind the wait loop of the next stage every DMMA of stage s-1 has been issued, which
    // needs its LDS operands in registers, so the release is safe by construction and costs no stall.
    int s = 0, phase = 0, held = -1;
    for (int w = blockIdx.x; w < p.ntiles * p.nsplit; w += gridDim.x) {
        const int2 tt = p.tiles[w / p.nsplit];
        const int64_t m0 = (int64_t)tt.x * TM, n0 = (int64_t)tt.y * TN;
        const int split = w % p.nsplit;
        const int kbeg = split * p.kchunk, kend = min(p.ktiles, kbeg + p.kchunk);
        double *__restrict__ Gout = p.G + (int64_t)split * p.split_stride;
        WarpAcc<WM, WN> acc;
        acc.zero();
        for (int kt = kbeg; kt < kend; ++kt) {
            mbar_wait(&full[s], phase);
            if (held >= 0) {
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[held]);
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
        // epilogue: direct tile + mirrored tile (diagonal tiles are complete by themselves)
        const bool offdiag = tt.x != tt.y;
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
                        if (p.accumulate) {
                            atomicAdd_system(&Gout[row + p.ldg * (col + e)], v);
                        } else {
                            // streaming stores (evict-first): the 8 npair^2 bytes of G pass through L2 once and must
                            // not push out the X panels the resident band of tiles keeps re-reading
                            __stcs(&Gout[row + p.ldg * (col + e)], v);
                            if (offdiag) __stcs(&Gout[(col + e) + p.ldg * row], v);
                        }
                    }
                }
            }
        }
    }
}

// G[:, c] = sum_s P[s][:, c] in the fixed order s = 0 .. nsplit-1 (deterministic)
__global__ void gram_reduce_kernel(const double *__restrict__ P, int64_t stride, int nsplit, int64_t npair,
                                   double *__restrict__ G, int64_t ldg)
{
    const int64_t total = npair * npair;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        double v = 0.0;
        for (int s2 = 0; s2 < nsplit; ++s2) v += P[(int64_t)s2 * stride + i];
        const int64_t c = i / npair, r = i - c * npair;
        G[r + ldg * c] = v;
    }
}

// G[c, r] = G[r, c] for r > c, 32 x 32 blocks through shared memory (after an accumulated, lower-only K3)
__global__ void mirror_lower_kernel(double *__restrict__ G, int64_t ld, int64_t n, int nb)
{
    __shared__ double t[32][33];
    // block index -> (bi >= bj) of the lower triangle of 32-blocks
    const int64_t b = blockIdx.x;
    int bi = (int)((sqrt(8.0 * (double)b + 1.0) - 1.0) * 0.5);
    while ((int64_t)(bi + 1) * (bi + 2) / 2 <= b) ++bi;
    while ((int64_t)bi * (bi + 1) / 2 > b) --bi;
    const int bj = (int)(b - (int64_t)bi * (bi + 1) / 2);
    if (bi >= nb) return;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5; // 32 x 8
    for (int c = ty; c < 32; c += 8) {
        const int64_t row = (int64_t)bi * 32 + tx, col = (int64_t)bj * 32 + c;
        t[c][tx] = (row < n && col < n) ? G[row + ld * col] : 0.0;
    }
    __syncthreads();
    for (int c = ty; c < 32; c += 8) {
        // destination element (row' = bj*32 + tx, col' = bi*32 + c) = source (bi*32 + c, bj*32 + tx) = t[tx][c]
        const int64_t row = (int64_t)bj * 32 + tx, col = (int64_t)bi * 32 + c;
        if (row < n && col < n && col > row) G[row + ld * col] = t[tx][c];
    }
}

cudaError_t launch_mirror_lower(double *G, int64_t ld, int64_t n, cudaStream_t s)
{
    if (n <= 1) return cudaSuccess;
    const int nb = (int)((n + 31) / 32);
    const int64_t blocks = (int64_t)nb * (nb + 1) / 2;
    mirror_lower_kernel<<<(unsigned)blocks, 256, 0, s>>>(G, ld, n, nb);
    return cudaGetLastError();
}

// ---- host ---------------------------------------------------------------------------------------------
namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode()
{
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}
} // namespace

// k splits worth using for `ntiles` output tiles: none once the tiles alone fill the SMs; otherwise enough work
// items for ~2 per SM, each at least 4 k tiles deep
int gram_splits(int ntiles, int ktiles, int num_sms)
{
    if (ntiles <= 0 || ntiles >= num_sms || ktiles < 8) return 1;
    int want = (2 * num_sms + ntiles - 1) / ntiles;
    want = want < ktiles / 4 ? want : ktiles / 4;
    if (want <= 1) return 1;
    const int kchunk = (ktiles + want - 1) / want;
    return (ktiles + kchunk - 1) / kchunk;
}

cudaError_t launch_gram(const GramParams &p, int num_sms, cudaStream_t s)
{
    if (p.ntiles <= 0) return cudaSuccess;
    EncodeTiledFn enc = get_encode();
    if (!enc) return cudaErrorNotSupported;
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {(cuuint64_t)p.ktiles * BK, (cuuint64_t)p.npair};
    const cuuint64_t gstride[1] = {(cuuint64_t)p.ldx * 8};
    const cuuint32_t box[2] = {BK, 128};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double *>(p.X), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    const int nsplit = (p.partial && !p.accumulate) ? gram_splits(p.ntiles, p.ktiles, num_sms) : 1;
    static const int hint = [] {
        // default 1: row panels evict_last (8.10 instead of 8.48 GB of DRAM reads at npair 25425, same speed); hinting
        // the column panels evict_first makes it worse (13 GB) -- profiles/k3_hint_r2.txt
        const char *e = std::getenv("QEMB_K3_HINT");
        return e ? std::atoi(e) : 1;
    }();
    GramTmaParams kp{p.ktiles, p.npair, p.G, p.ldg, p.tiles, p.ntiles, 1, p.ktiles, 0, p.accumulate, hint};
    if (nsplit > 1) {
        kp.G = p.partial;
        kp.ldg = p.npair;
        kp.nsplit = nsplit;
        kp.kchunk = (p.ktiles + nsplit - 1) / nsplit;
        kp.split_stride = p.npair * p.npair;
    }
    const int nwork = p.ntiles * kp.nsplit;
    const int grid = nwork < num_sms ? nwork : num_sms;
    gram_kernel<<<grid, NTHREADS, SMEM_BYTES, s>>>(tmap, kp);
    e = cudaGetLastError();
    if (e != cudaSuccess || nsplit == 1) return e;
    const int64_t blocks = (p.npair * p.npair + 255) / 256;
    gram_reduce_kernel<<<(unsigned)(blocks < 148 * 8 ? blocks : 148 * 8), 256, 0, s>>>(p.partial, kp.split_stride, nsplit,
                                                                                    p.npair, p.G, p.ldg);
    return cudaGetLastError();
}

// Lower-triangular tile list in bands of GROUP row tiles: inside a band the column tile is the slow
// index, so the resident wave touches <= GROUP row panels and ~148/GROUP column panels.  DRAM reads of X are
// ~ nt^2 / (2 GROUP) panels as long as the GROUP row panels of a band (GROUP x 128 x naux x 8 bytes) stay in L2.
int gram_tile_group()
{
    static const int group = [] {
        int g = 12;
        if (const char *e = std::getenv("QEMB_K3_GROUP")) g = std::atoi(e);
        return g < 1 ? 1 : (g > 148 ? 148 : g);
    }();
    return group;
}

int64_t build_gram_tiles(int nt, int2 *out)
{
    const int GROUP = gram_tile_group();
    int64_t n = 0;
    for (int i0 = 0; i0 < nt; i0 += GROUP) {
        const int i1 = (i0 + GROUP < nt) ? i0 + GROUP : nt;
        for (int j = 0; j < i1; ++j)
            for (int i = (j > i0 ? j : i0); i < i1; ++i) out[n++] = make_int2(i, j);
    }
    return n;
}

} // namespace qemb
