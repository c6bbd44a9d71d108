// gather_gemm.cu -- K1 and K2a as ONE grouped, gathered FP64 tensor-core GEMM.
//
//   C[:, ccol[n]] = sum_{k<K} Asrc[:, acol[k]] * TA[brow[k], bcol[n]]          n < N, per group
//
// K1 (contract_with_TA_1st, src/quemb/molbe/_cpp/eri_sparse_DF.cpp:484-532), one group per AO mu:
//     k runs over R(mu) (acol = column of the pair (mu,nu) in the unique data, brow = nu),
//     n runs over the MOs i with mu in A(i) (bcol = i, ccol = off(i,mu)).
//     The reference streams every naux-column of (P|mu nu) once per (i, mu, nu) axpy with no reuse
//     across i (cpp:519-526); here the gathered columns are staged once per (mu, 128-row slab) in
//     shared memory and reused by all MOs that reach mu.
// K2a (contract_with_TA_2nd_to_sym_dense, cpp:560-605), one group per MO a (= the SMALLER index of
//     the pair, cpp:595 + indexers.hpp:95):
//     k runs over A(a) (acol = off(a,mu) -- contiguous, brow = mu), n over b = a..nmo-1
//     (bcol = b, ccol = T(b)+a).
//
// The auxiliary index P is the GEMM's m dimension: it is contiguous in HBM in the source columns and
// in the output columns, so A-tiles are 1 KB contiguous runs (128 P x 8 B) fetched with 16-byte
// cp.async; the TA tile is an element-wise double gather (8-byte cp.async) out of the L2-resident TA.
// CTA tile 128(P) x 128(n) x 16, 8 warps, warp tile 64x32, 4-stage ring.
#include "dmma_core.cuh"
#include "kernels.cuh"

namespace qemb {

namespace {
constexpr int TM = GATHER_TILE_M, TN = GATHER_TILE_N;
constexpr int PA = TM + 4; // As[k][m]
constexpr int STAGES = 4;
constexpr int A_DOUBLES = BK * PA;
constexpr int B_DOUBLES = TN * PITCH_K;
constexpr int STAGE_DOUBLES = A_DOUBLES + B_DOUBLES;
constexpr int SMEM_BYTES = STAGES * STAGE_DOUBLES * 8 + 2 * TN * 4;
constexpr int NTHREADS = 256;
constexpr int WM = 64, WN = 32;
} // namespace

__global__ void __launch_bounds__(NTHREADS, 1) gather_gemm_kernel(GatherParams p)
{
    extern __shared__ __align__(16) double smem[];
    int *bcol_s = reinterpret_cast<int *>(smem + STAGES * STAGE_DOUBLES);
    int *ccol_s = bcol_s + TN;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm0 = (warp & 1) * WM, wn0 = (warp >> 1) * WN;

    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const int4 tt = p.tiles[tile];
        const GatherGroup grp = p.groups[tt.x];
        const int m0 = tt.y * TM, n0 = tt.z * TN;
        const int K = grp.K;
        const int nvalid = min(TN, grp.N - n0);
        const int *acol = p.acol + grp.a_off;
        const int *brow = p.brow + grp.a_off;
        const int ktiles = (K + BK - 1) / BK;

        __syncthreads(); // previous tile's epilogue has finished reading ccol_s
        if (tid < TN) {
            const bool ok = tid < nvalid;
            bcol_s[tid] = ok ? p.bcol[grp.b_off + n0 + tid] : 0;
            ccol_s[tid] = ok ? p.ccol[grp.b_off + n0 + tid] : -1;
        }
        __syncthreads();

        WarpAcc<WM, WN> acc;
        acc.zero();

        auto load_stage = [&](int s, int kt) {
            double *As = smem + s * STAGE_DOUBLES;
            double *Bs = As + A_DOUBLES;
            const int k0 = kt * BK;
            // A: 16 gathered columns x 128 rows: 64 chunks of 16 B per column, a warp reads 512 B contiguous
#pragma unroll
            for (int r = 0; r < (BK * TM / 2) / NTHREADS; ++r) {
                const int idx = tid + NTHREADS * r;
                const int k = idx >> 6, ch = idx & 63;
                const bool kok = (k0 + k) < K;
                const int64_t col = kok ? acol[k0 + k] : 0;
                const int row = m0 + ch * 2;
                const bool ok = kok && row < p.M;
                cp_async16(As + k * PA + ch * 2, p.Asrc + col * p.lda + (ok ? row : 0), ok);
            }
            // B: TA[brow[k], bcol[n]] -> Bs[n][k]; a thread keeps one k and walks n
            {
                const int k = tid & 15;
                const bool kok = (k0 + k) < K;
                const int64_t rr = kok ? brow[k0 + k] : 0;
#pragma unroll
                for (int r = 0; r < (TN * BK) / NTHREADS; ++r) {
                    const int n = (tid >> 4) + 16 * r;
                    const bool ok = kok && n < nvalid;
                    cp_async8(Bs + n * PITCH_K + k, p.TA + rr + p.ldta * (int64_t)bcol_s[n], ok);
                }
            }
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
            const double *Bs = As + A_DOUBLES;
            warp_mma_stage<WM, WN, false, PA, PITCH_K>(acc, As, Bs, wm0, wn0, lane);
        }
        cp_async_wait<0>();

        // epilogue: scatter the columns (K == 0 groups write zeros, like the reference's zero init)
        const int g = lane >> 2, t = lane & 3;
#pragma unroll
        for (int j = 0; j < WN / 8; ++j) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int n = wn0 + 8 * j + 2 * t + e;
                const int cc = ccol_s[n];
                if (cc < 0) continue;
                double *dst = p.C + (int64_t)cc * p.ldc;
#pragma unroll
                for (int i = 0; i < WM / 8; ++i) {
                    const int row = m0 + wm0 + 8 * i + g;
                    if (row < p.M) dst[row] = acc.c[i][j][e];
                }
            }
        }
    }
}

cudaError_t launch_gather_gemm(const GatherParams &p, cudaStream_t s)
{
    cudaError_t e =
        cudaFuncSetAttribute(gather_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    if (p.ntiles <= 0) return cudaSuccess;
    gather_gemm_kernel<<<p.ntiles, NTHREADS, SMEM_BYTES, s>>>(p);
    return cudaGetLastError();
}

} // namespace qemb
