// gram_impl.cuh -- templated body of K3 (G = X^T X, lower tiles + mirror), shared by the library
// (k3_gram.cu instantiates the tuned configuration) and by scripts/gram_variants.cu (tile-shape sweep).
#pragma once
#include "dmma_core.cuh"
#include "kernels.cuh"

namespace qemb {

template <int TM_, int TN_, int WM_, int WN_, int BK_, int STAGES_, int MINB_>
struct GramCfg {
    static constexpr int TM = TM_, TN = TN_, WM = WM_, WN = WN_, KB = BK_, STAGES = STAGES_, MINB = MINB_;
    static constexpr int WARPS_M = TM / WM, WARPS_N = TN / WN;
    static constexpr int NTHREADS = 32 * WARPS_M * WARPS_N;
    static constexpr int PK = KB + 4; // [row][k] pitch: (g*PK + t) mod 16 = 4g + t for KB = 16, 32
    static constexpr int STAGE_DOUBLES = (TM + TN) * PK;
    static constexpr int SMEM_BYTES = STAGES * STAGE_DOUBLES * 8;
    static_assert(KB % 16 == 0 && (TM * (KB / 2)) % NTHREADS == 0 && (TN * (KB / 2)) % NTHREADS == 0, "tile/threads");
};

template <class C>
__device__ __forceinline__ void gram_load_rows(double *dst, const double *__restrict__ X, int64_t ldx, int64_t row0,
                                               int64_t nrows_total, int k0, int klimit, int tid, int rows)
{
    constexpr int CH = C::KB / 2; // 16-byte chunks per row
    for (int idx = tid; idx < rows * CH; idx += C::NTHREADS) {
        const int row = idx / CH, ch = idx % CH;
        const int64_t grow = row0 + row;
        const bool ok = grow < nrows_total && (k0 + ch * 2) < klimit; // k tail (KB = 32) reads as zero
        cp_async16(dst + row * C::PK + ch * 2, X + (ok ? grow : 0) * ldx + (ok ? k0 + ch * 2 : 0), ok);
    }
}

// tiles: (ti, tj) = (row tile of TM rows, column tile of TN columns); every tile that intersects the lower
// triangle is listed.  Elements with col > row are not stored (they are the mirror of another tile's).
template <class C>
__global__ void __launch_bounds__(C::NTHREADS, C::MINB) gram_kernel_t(GramParams p)
{
    extern __shared__ __align__(16) double smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm0 = (warp % C::WARPS_M) * C::WM, wn0 = (warp / C::WARPS_M) * C::WN;

    const int2 tt = p.tiles[blockIdx.x];
    const int64_t m0 = (int64_t)tt.x * C::TM, n0 = (int64_t)tt.y * C::TN;

    WarpAcc<C::WM, C::WN> acc;
    acc.zero();

    auto load_stage = [&](int s, int kt) {
        double *As = smem + s * C::STAGE_DOUBLES;
        double *Bs = As + C::TM * C::PK;
        gram_load_rows<C>(As, p.X, p.ldx, m0, p.npair, kt * C::KB, p.ktiles * BK, tid, C::TM);
        gram_load_rows<C>(Bs, p.X, p.ldx, n0, p.npair, kt * C::KB, p.ktiles * BK, tid, C::TN);
    };

    const int ktiles = (p.ktiles * BK + C::KB - 1) / C::KB; // p.ktiles counts 16-row tiles; rows are zero padded to 128
#pragma unroll
    for (int s = 0; s < C::STAGES - 1; ++s) {
        if (s < ktiles) load_stage(s, s);
        cp_async_commit();
    }
    for (int kt = 0; kt < ktiles; ++kt) {
        cp_async_wait<C::STAGES - 2>();
        __syncthreads();
        const int nk = kt + C::STAGES - 1;
        if (nk < ktiles) load_stage(nk % C::STAGES, nk);
        cp_async_commit();
        const double *As = smem + (kt % C::STAGES) * C::STAGE_DOUBLES;
        const double *Bs = As + C::TM * C::PK;
#pragma unroll
        for (int kk = 0; kk < C::KB / BK; ++kk)
            warp_mma_stage<C::WM, C::WN, true, C::PK, C::PK>(acc, As + kk * BK, Bs + kk * BK, wm0, wn0, lane);
    }
    cp_async_wait<0>();

    // epilogue: lower-triangle elements + their mirror
    const int g = lane >> 2, t = lane & 3;
    const bool crosses = (n0 + C::TN - 1) > m0; // tile touches or crosses the diagonal
#pragma unroll
    for (int i = 0; i < C::WM / 8; ++i) {
        const int64_t row = m0 + wm0 + 8 * i + g;
        if (row >= p.npair) continue;
#pragma unroll
        for (int j = 0; j < C::WN / 8; ++j) {
            const int64_t col = n0 + wn0 + 8 * j + 2 * t;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int64_t c = col + e;
                if (c >= p.npair || (crosses && c > row)) continue;
                const double v = acc.c[i][j][e];
                p.G[row + p.ldg * c] = v;
                if (c != row) p.G[c + p.ldg * row] = v;
            }
        }
    }
}

// host: tile list for a TM x TN tiling of the lower triangle, in bands of `group` row tiles (column tile
// slow, row tile fast inside a band) so that a resident wave shares few X panels in L2
template <class C>
inline int64_t gram_build_tiles(int64_t npair, int2 *out, int group = 12)
{
    const int ntm = (int)((npair + C::TM - 1) / C::TM), ntn = (int)((npair + C::TN - 1) / C::TN);
    int64_t n = 0;
    for (int i0 = 0; i0 < ntm; i0 += group) {
        const int i1 = (i0 + group < ntm) ? i0 + group : ntm;
        for (int j = 0; j < ntn; ++j)
            for (int i = i0; i < i1; ++i) {
                const int64_t row_hi = (int64_t)(i + 1) * C::TM - 1; // last row of the tile
                if ((int64_t)j * C::TN > row_hi) continue;          // entirely above the diagonal
                if (out) out[n] = make_int2(i, j);
                ++n;
            }
    }
    return n;
}

template <class C>
inline cudaError_t gram_launch(const GramParams &p, cudaStream_t s)
{
    cudaError_t e = cudaFuncSetAttribute(gram_kernel_t<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    if (p.ntiles <= 0) return cudaSuccess;
    gram_kernel_t<C><<<p.ntiles, C::NTHREADS, C::SMEM_BYTES, s>>>(p);
    return cudaGetLastError();
}

} // namespace qemb
