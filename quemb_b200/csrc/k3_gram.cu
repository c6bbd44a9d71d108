// k3_gram.cu -- K3: (ij|kl) = sum_P X[P,ij] X[P,kl]      (reference: eval_via_cholesky step 2,
// src/quemb/molbe/_cpp/eri_sparse_DF.cpp:620, and the cuBLAS Dsyrk + host mirror of its GPU variant,
// cpp:684-708).
//
// Design: symmetric FP64 tensor-core contraction.  X is (ldx x npair) column-major, i.e. every pair
// column is k-contiguous -- exactly the "row.col" operand order of DMMA for both operands, so the
// tile loads are plain 128-byte row segments.  Only the nt(nt+1)/2 lower tiles are computed
// (naux*npair*(npair+1) flops instead of the reference CPU path's full 2*naux*npair^2 GEMM); every
// off-diagonal tile is also stored transposed, so the full npair x npair square the consumers read
// (pfrag.py:383, mbe.py:1155) is emitted directly in s4 pair order with no host mirror pass.
//
// CTA tile 128x128x16, 8 warps (2 x 4), warp tile 64x32 = 32 DMMA.8x8x4 accumulators (128 regs),
// STAGES-deep cp.async ring (40 KB per stage).  Tiles are ordered in bands of GROUP row-tiles so that
// the 148 concurrently resident tiles share ~2*sqrt(148) X panels in the 126 MB L2.
#include "dmma_core.cuh"
#include "kernels.cuh"

namespace qemb {

namespace {
constexpr int TM = GRAM_TILE, TN = GRAM_TILE;
constexpr int STAGES = 4;
constexpr int STAGE_DOUBLES = (TM + TN) * PITCH_K;
constexpr int SMEM_BYTES = STAGES * STAGE_DOUBLES * 8;
constexpr int NTHREADS = 256;
constexpr int WM = 64, WN = 32;

__device__ __forceinline__ void load_rows_kmajor(double *dst, const double *__restrict__ X, int64_t ldx,
                                                 int64_t row0, int64_t nrows_total, int k0, int tid)
{
    // 128 rows x 16 k = 128 x 8 chunks of 16 B; 8 consecutive threads fetch one 128-byte row segment
#pragma unroll
    for (int r = 0; r < (TM * 8) / NTHREADS; ++r) {
        const int idx = tid + NTHREADS * r;
        const int row = idx >> 3, ch = idx & 7;
        const int64_t grow = row0 + row;
        const bool ok = grow < nrows_total;
        const double *src = X + (ok ? grow : 0) * ldx + k0 + ch * 2;
        cp_async16(dst + row * PITCH_K + ch * 2, src, ok);
    }
}
} // namespace

__global__ void __launch_bounds__(NTHREADS, 1) gram_kernel(GramParams p)
{
    extern __shared__ __align__(16) double smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm0 = (warp & 1) * WM, wn0 = (warp >> 1) * WN;

    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const int2 tt = p.tiles[tile];
        const int64_t m0 = (int64_t)tt.x * TM, n0 = (int64_t)tt.y * TN;

        WarpAcc<WM, WN> acc;
        acc.zero();

        auto load_stage = [&](int s, int kt) {
            double *As = smem + s * STAGE_DOUBLES;
            double *Bs = As + TM * PITCH_K;
            load_rows_kmajor(As, p.X, p.ldx, m0, p.npair, kt * BK, tid);
            load_rows_kmajor(Bs, p.X, p.ldx, n0, p.npair, kt * BK, tid);
        };

#pragma unroll
        for (int s = 0; s < STAGES - 1; ++s) {
            if (s < p.ktiles) load_stage(s, s);
            cp_async_commit();
        }
        for (int kt = 0; kt < p.ktiles; ++kt) {
            cp_async_wait<STAGES - 2>();
            __syncthreads();
            const int nk = kt + STAGES - 1;
            if (nk < p.ktiles) load_stage(nk % STAGES, nk);
            cp_async_commit();
            const double *As = smem + (kt % STAGES) * STAGE_DOUBLES;
            const double *Bs = As + TM * PITCH_K;
            warp_mma_stage<WM, WN, true, PITCH_K, PITCH_K>(acc, As, Bs, wm0, wn0, lane);
        }
        cp_async_wait<0>();
        __syncthreads(); // all warps done with the ring before the next tile's prologue refills it

        // epilogue: direct tile + mirrored tile (skipped on the diagonal, where the tile is complete)
        const int g = lane >> 2, t = lane & 3;
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
                        p.G[row + p.ldg * (col + e)] = v;
                        if (offdiag) p.G[(col + e) + p.ldg * row] = v;
                    }
                }
            }
        }
    }
}

cudaError_t launch_gram(const GramParams &p, cudaStream_t s)
{
    // per device, so set on every launch (a few hundred ns)
    cudaError_t e = cudaFuncSetAttribute(gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    if (p.ntiles <= 0) return cudaSuccess;
    gram_kernel<<<p.ntiles, NTHREADS, SMEM_BYTES, s>>>(p);
    return cudaGetLastError();
}

// Lower-triangular tile list in bands of GROUP row tiles: inside a band the column tile is the slow
// index, so the resident wave touches <= GROUP row panels and ~148/GROUP column panels.
int64_t build_gram_tiles(int nt, int2 *out)
{
    constexpr int GROUP = 12;
    int64_t n = 0;
    for (int i0 = 0; i0 < nt; i0 += GROUP) {
        const int i1 = (i0 + GROUP < nt) ? i0 + GROUP : nt;
        for (int j = 0; j < i1; ++j)
            for (int i = (j > i0 ? j : i0); i < i1; ++i) out[n++] = make_int2(i, j);
    }
    return n;
}

} // namespace qemb
