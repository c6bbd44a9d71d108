// trsm.cu -- K2b: X = L^-1 B, left side, lower, non-unit, in place      (reference: eval_via_cholesky
// step 1, src/quemb/molbe/_cpp/eri_sparse_DF.cpp:615 (Eigen triangularView::solve) and cublasDtrsm
// of its GPU variant, cpp:667-678).
//
// Blocked forward substitution on FP64 tensor cores.  The diagonal 128x128 blocks of L are inverted
// once per metric (invert_diag_blocks_kernel); a CTA then owns a panel of 64 right-hand-side columns
// and sweeps the block rows kb = 0..nb-1:
//     acc   = L[kb, 0:kb] * X[0:kb, panel]                (DMMA, cp.async ring; L is L2 resident)
//     T     = B[kb, panel] - acc          -> shared memory
//     X[kb] = Dinv[kb] * T                                (DMMA, B operand straight from shared memory)
// Panels are independent, so there is no inter-CTA dependency and one launch solves everything.
// Only diagonal blocks are inverted (like cuBLAS/MAGMA trsm), never the whole factor, which keeps the
// backward stability of substitution for the ill-conditioned Coulomb metric.
#include <algorithm>
#include "dmma_core.cuh"
#include "kernels.cuh"

namespace qemb {

namespace {
constexpr int NB = TRSM_BLOCK;  // 128 rows per block row
constexpr int PN = TRSM_PANEL;  // 64 columns per CTA
constexpr int PA = NB + 4;      // As[k][m]
constexpr int PT = NB + 4;      // Ts[n][k]
constexpr int STAGES = 3;
constexpr int A_DOUBLES = BK * PA;
constexpr int B_DOUBLES = PN * PITCH_K;
constexpr int STAGE_DOUBLES = A_DOUBLES + B_DOUBLES;
constexpr int T_DOUBLES = PN * PT;
constexpr int SMEM_BYTES = (STAGES * STAGE_DOUBLES + T_DOUBLES) * 8;
constexpr int NTHREADS = 256;
constexpr int WM = 32, WN = 32; // 4 x 2 warps
} // namespace

struct TrsmKParams {
    TrsmParams p;
    int mrows; // valid rows (multiple of 16); rows beyond read as zero and are not written
};

__global__ void __launch_bounds__(NTHREADS, 1) trsm_panel_kernel(TrsmKParams kp)
{
    const TrsmParams &p = kp.p;
    extern __shared__ __align__(16) double smem[];
    double *Ts = smem + STAGES * STAGE_DOUBLES;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm0 = (warp & 3) * WM, wn0 = (warp >> 2) * WN;
    const int g = lane >> 2, t = lane & 3;
    const int64_t n0 = (int64_t)blockIdx.x * PN;

    // A operand: 16 columns of a 128-row block of a column-major matrix -> As[k][m]
    auto load_A = [&](double *As, const double *src /* top-left of the 128 x 16 slab */, int64_t ld) {
#pragma unroll
        for (int r = 0; r < (BK * NB / 2) / NTHREADS; ++r) {
            const int idx = tid + NTHREADS * r;
            const int k = idx >> 6, ch = idx & 63;
            cp_async16(As + k * PA + ch * 2, src + (int64_t)k * ld + ch * 2, true);
        }
    };
    // B operand: rows [k0, k0+16) of the panel columns -> Bs[n][k]
    auto load_B = [&](double *Bs, int k0) {
#pragma unroll
        for (int r = 0; r < (PN * 8) / NTHREADS; ++r) {
            const int idx = tid + NTHREADS * r;
            const int n = idx >> 3, ch = idx & 7;
            const bool ok = (n0 + n) < p.ncol;
            cp_async16(Bs + n * PITCH_K + ch * 2, p.X + (ok ? (n0 + n) : 0) * p.ldx + k0 + ch * 2, ok);
        }
    };

    for (int kb = 0; kb < p.nblocks; ++kb) {
        const int64_t r0 = (int64_t)kb * NB;
        WarpAcc<WM, WN> acc;
        acc.zero();

        // ---- phase 1: acc = L[kb, 0:kb] * X[0:kb, panel] ----
        const int ktiles = kb * (NB / BK);
        auto load_stage = [&](int s, int kt) {
            double *As = smem + s * STAGE_DOUBLES;
            load_A(As, p.L + (int64_t)kt * BK * p.ldl + r0, p.ldl);
            load_B(As + A_DOUBLES, kt * BK);
        };
#pragma unroll
        for (int s = 0; s < STAGES - 1; ++s) {
            if (s < ktiles) load_stage(s, s);
            cp_async_commit();
        }
        for (int kt = 0; kt < ktiles; ++kt) {
            cp_async_wait<STAGES - 2>();
            __syncthreads();
            const int nk = kt + STAGES - 1;
            if (nk < ktiles) load_stage(nk % STAGES, nk);
            cp_async_commit();
            const double *As = smem + (kt % STAGES) * STAGE_DOUBLES;
            warp_mma_stage<WM, WN, false, PA, PITCH_K>(acc, As, As + A_DOUBLES, wm0, wn0, lane);
        }
        cp_async_wait<0>();
        __syncthreads();

        // prefetch the first Dinv stages while T is formed
        const double *Dk = p.Dinv + (int64_t)kb * NB * NB;
        constexpr int KT2 = NB / BK;
#pragma unroll
        for (int s = 0; s < STAGES - 1; ++s) {
            load_A(smem + s * STAGE_DOUBLES, Dk + (int64_t)s * BK * NB, NB);
            cp_async_commit();
        }

        // ---- phase 2: T = B[kb, panel] - acc -> Ts[n][row] ----
#pragma unroll
        for (int j = 0; j < WN / 8; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int n = wn0 + 8 * j + 2 * t + e;
                const bool nok = (n0 + n) < p.ncol;
                const double *src = p.X + (nok ? (n0 + n) : 0) * p.ldx + r0;
#pragma unroll
                for (int i = 0; i < WM / 8; ++i) {
                    const int row = wm0 + 8 * i + g;
                    const double b = (nok && (r0 + row) < kp.mrows) ? src[row] : 0.0;
                    Ts[n * PT + row] = b - acc.c[i][j][e];
                }
            }
        acc.zero();

        // ---- phase 3: X[kb] = Dinv[kb] * T ----
        for (int kt = 0; kt < KT2; ++kt) {
            cp_async_wait<STAGES - 2>();
            __syncthreads(); // also publishes Ts on the first pass
            const int nk = kt + STAGES - 1;
            if (nk < KT2) load_A(smem + (nk % STAGES) * STAGE_DOUBLES, Dk + (int64_t)nk * BK * NB, NB);
            cp_async_commit();
            const double *As = smem + (kt % STAGES) * STAGE_DOUBLES;
            warp_mma_stage<WM, WN, false, PA, PT>(acc, As, Ts + kt * BK, wm0, wn0, lane);
        }
        cp_async_wait<0>();

#pragma unroll
        for (int j = 0; j < WN / 8; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int n = wn0 + 8 * j + 2 * t + e;
                if ((n0 + n) >= p.ncol) continue;
                double *dst = p.X + (n0 + n) * p.ldx + r0;
#pragma unroll
                for (int i = 0; i < WM / 8; ++i) {
                    const int row = wm0 + 8 * i + g;
                    if ((r0 + row) < kp.mrows) dst[row] = acc.c[i][j][e];
                }
            }
        __threadfence(); // the next block row reads these rows back through cp.async (L2)
        __syncthreads();
    }
}

// inverse of every 128x128 lower-triangular diagonal block; thread j builds column j by forward
// substitution (uniform loops -> L is read as a warp broadcast out of L1)
__global__ void __launch_bounds__(NB, 1) invert_diag_blocks_kernel(const double *__restrict__ L, int64_t ldl,
                                                                    double *__restrict__ Dinv)
{
    extern __shared__ __align__(16) double xs[]; // xs[c][j]
    const int b = blockIdx.x, j = threadIdx.x;
    const double *Lb = L + (int64_t)b * NB * ldl + (int64_t)b * NB;
    for (int r = 0; r < NB; ++r) {
        double s = (r == j) ? 1.0 : 0.0;
        for (int c = 0; c < r; ++c) s -= Lb[r + (int64_t)c * ldl] * xs[c * NB + j];
        s = (r < j) ? 0.0 : s / Lb[r + (int64_t)r * ldl];
        xs[r * NB + j] = s;
    }
    double *D = Dinv + (int64_t)b * NB * NB;
    for (int r = 0; r < NB; ++r) D[r + j * NB] = xs[r * NB + j];
}

cudaError_t launch_invert_diag_blocks(const double *L, int64_t ldl, int nblocks, double *Dinv, cudaStream_t s)
{
    const int bytes = NB * NB * 8;
    cudaError_t e =
        cudaFuncSetAttribute(invert_diag_blocks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    if (nblocks <= 0) return cudaSuccess;
    invert_diag_blocks_kernel<<<nblocks, NB, bytes, s>>>(L, ldl, Dinv);
    return cudaGetLastError();
}

cudaError_t launch_trsm_rows(const TrsmParams &p, int mrows, cudaStream_t s)
{
    cudaError_t e =
        cudaFuncSetAttribute(trsm_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    if (p.ncol <= 0 || p.nblocks <= 0) return cudaSuccess;
    TrsmKParams kp{p, mrows};
    // forward substitution never needs block rows below the last valid row: a slab that ends at row p_end costs
    // p_end^2 ncol flops, not naux^2 ncol (aux-sharded staging: rank r solves only down to its own slab)
    kp.p.nblocks = std::min(p.nblocks, (mrows + NB - 1) / NB);
    if (kp.p.nblocks <= 0) return cudaSuccess;
    const unsigned grid = (unsigned)((p.ncol + PN - 1) / PN);
    trsm_panel_kernel<<<grid, NTHREADS, SMEM_BYTES, s>>>(kp);
    return cudaGetLastError();
}

cudaError_t launch_trsm(const TrsmParams &p, cudaStream_t s)
{
    return launch_trsm_rows(p, p.nblocks * NB, s);
}

} // namespace qemb
