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
// The auxiliary index P is the GEMM's m dimension: it is contiguous in HBM in the source columns and in
// the output columns.  Persistent kernel, one CTA per SM, warp specialised like K3 (k3_gram.cu):
//   * producer warpgroup (4 warps, 56 registers each after setmaxnreg; the consumers take 224):
//       - A tile: 16 gathered columns x 128 rows per stage = 16 bulk copies of 1 KB (cp.async.bulk, the 1-D
//         TMA path, one per lane) into As[k][m] (pitch 132), completing on the stage's "full" mbarrier;
//       - B tile: the element-wise double gather TA[brow[k], bcol[n]] out of the L1/L2-resident TA with plain
//         loads + st.shared into Bs[n][k] (pitch 20); the four warps arrive on the same mbarrier.
//     The producers run up to STAGES stages ahead, across tile boundaries, so index loads, the gather
//     latency, the pipeline fill of the next tile and the epilogue of the previous one are all hidden.
//   * eight consumer warps (64 x 32 register tiles of DMMA.8x8x4 accumulators) wait on "full", read
//     fragments with conflict-free 64-bit LDS and release the stage through "empty".  No CTA-wide barrier.
//   * n quantisation: the 16 column blocks (of 8) of a tile are dealt round-robin to the four n-warps, and a
//     warp skips the blocks beyond the group's N, so a ragged tile costs ceil(N/32)*32 columns on every
//     warp instead of 128 (K2a groups have N = 1..nmo: 61 % -> 87 % useful work at nmo = 220).
#include "dmma_core.cuh"
#include "kernels.cuh"

namespace qemb {

namespace {
constexpr int TM = GATHER_TILE_M, TN = GATHER_TILE_N;
constexpr int PA = TM + 4; // As[k][m]
constexpr int PB = PITCH_K; // Bs[n][k]
constexpr int STAGES = 5;
constexpr int A_BYTES = BK * PA * 8;
constexpr int B_BYTES = TN * PB * 8;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 2 * STAGES * 8 + 16;
constexpr int NCONSUMER_WARPS = 8;
constexpr int NPRODUCER_WARPS = 4;
constexpr int NTHREADS = 32 * (NCONSUMER_WARPS + NPRODUCER_WARPS);
constexpr int WM = 64, WN = 32;

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
// raises the transaction count of the current phase without arriving
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, int bytes)
{
    asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
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
        "GG_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra GG_DONE;\n"
        "bra GG_WAIT;\n"
        "GG_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_load(void *dst, const void *src, int bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
} // namespace

__global__ void __launch_bounds__(NTHREADS, 1) gather_gemm_kernel(GatherParams p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
    uint64_t *empty = full + STAGES;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], NPRODUCER_WARPS);
            mbar_init(&empty[s], NCONSUMER_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp >= NCONSUMER_WARPS) {
        // ------------------------------------ producers ------------------------------------------
        asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
        const int pw = warp - NCONSUMER_WARPS; // 0..3
        const int ptid = tid - 32 * NCONSUMER_WARPS; // 0..127
        const int kk = ptid & 15;                    // this thread's k inside a stage (B gather)
        const int nb = ptid >> 4;                    // first n (B gather): n = nb + 8 r
        int s = 0, phase = 0;
        for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
            const int4 tt = p.tiles[tile];
            const GatherGroup grp = p.groups[tt.x];
            const int m0 = tt.y * TM, n0 = tt.z * TN;
            const int K = grp.K;
            const int nvalid = min(TN, grp.N - n0);
            const int ktiles = (K + BK - 1) / BK;
            const int rows = min(TM, p.Mread - m0); // rows of the A tile that exist in the source columns
            const int *acol = p.acol + grp.a_off;
            const int *brow = p.brow + grp.a_off;
            const int *bcol = p.bcol + grp.b_off + n0;
            // the MO columns this thread gathers from (fixed for the tile)
            int boff[TN / 8]; // nmo * nao < 2^31
#pragma unroll
            for (int r = 0; r < TN / 8; ++r) {
                const int n = nb + 8 * r;
                boff[r] = n < nvalid ? bcol[n] * (int)p.ldta : -1;
            }
            for (int kt = 0; kt < ktiles; ++kt) {
                const int k0 = kt * BK;
                const int nk = min(BK, K - k0);
                // index loads before the wait: they overlap the wait for the stage
                const int64_t rr = kk < nk ? brow[k0 + kk] : -1;
                int64_t ac = -1;
                if (pw == 0 && lane < nk) ac = acol[k0 + lane];
                mbar_wait(&empty[s], phase ^ 1);
                uint8_t *stage = smem + s * STAGE_BYTES;
                double *As = reinterpret_cast<double *>(stage);
                double *Bs = reinterpret_cast<double *>(stage + A_BYTES);
                if (pw == 0) {
                    // k tail: unused columns must read as zero (finite times zero), rows beyond `rows` are never stored
                    if (nk < BK) {
                        for (int c = nk; c < BK; ++c)
                            for (int m = lane; m < TM; m += 32) As[c * PA + m] = 0.0;
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        __syncwarp();
                    }
                    if (lane == 0) mbar_expect_tx(&full[s], nk * rows * 8);
                    __syncwarp();
                    if (lane < nk) bulk_load(As + lane * PA, p.Asrc + ac * p.lda + m0, rows * 8, &full[s]);
                }
                // B gather: Bs[n][k] = TA[brow[k], bcol[n]]
#pragma unroll
                for (int r = 0; r < TN / 8; ++r) {
                    const int n = nb + 8 * r;
                    double v = 0.0;
                    if (rr >= 0 && boff[r] >= 0) v = __ldg(p.TA + rr + boff[r]);
                    Bs[n * PB + kk] = v;
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&full[s]);
                if (++s == STAGES) {
                    s = 0;
                    phase ^= 1;
                }
            }
        }
        return;
    }

    // -------------------------------------- consumers ------------------------------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
    const int wmi = warp & 1, wni = warp >> 1;
    const int wm0 = wmi * WM;
    const int g = lane >> 2, t = lane & 3;

    int s = 0, phase = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const int4 tt = p.tiles[tile];
        const GatherGroup grp = p.groups[tt.x];
        const int m0 = tt.y * TM, n0 = tt.z * TN;
        const int nvalid = min(TN, grp.N - n0);
        const int ktiles = (grp.K + BK - 1) / BK;
        // column blocks (of 8) are dealt round-robin to the four n-warps: block index = 4 j + wni
        const int nblocks = (nvalid + 7) >> 3;
        const int jv = nblocks > wni ? (nblocks - wni + 3) >> 2 : 0; // valid j of this warp
        const int mvalid = min(TM, p.M - m0) - wm0;                  // valid rows of this warp's half
        const int iv = mvalid > 0 ? min(WM / 8, (mvalid + 7) >> 3) : 0;
        // output columns of this thread: n = 8 (4 j + wni) + 2 t + e
        int cc[WN / 8][2];
#pragma unroll
        for (int j = 0; j < WN / 8; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int n = 8 * (4 * j + wni) + 2 * t + e;
                cc[j][e] = n < nvalid ? p.ccol[grp.b_off + n0 + n] : -1;
            }

        WarpAcc<WM, WN> acc;
        acc.zero();
        for (int kt = 0; kt < ktiles; ++kt) {
            mbar_wait(&full[s], phase);
            const double *As = reinterpret_cast<const double *>(smem + s * STAGE_BYTES);
            const double *Bs = reinterpret_cast<const double *>(smem + s * STAGE_BYTES + A_BYTES);
#pragma unroll
            for (int ks = 0; ks < BK / 4; ++ks) {
                double a[WM / 8], b[WN / 8];
#pragma unroll
                for (int i = 0; i < WM / 8; ++i) a[i] = As[(4 * ks + t) * PA + wm0 + 8 * i + g];
#pragma unroll
                for (int j = 0; j < WN / 8; ++j) b[j] = Bs[(8 * (4 * j + wni) + g) * PB + 4 * ks + t];
#pragma unroll
                for (int j = 0; j < WN / 8; ++j)
                    if (j < jv) {
#pragma unroll
                        for (int i = 0; i < WM / 8; ++i)
                            if (i < iv) dmma884(acc.c[i][j][0], acc.c[i][j][1], a[i], b[j]);
                    }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
            if (++s == STAGES) {
                s = 0;
                phase ^= 1;
            }
        }

        // epilogue: scatter the columns (K == 0 groups write zeros, like the reference's zero init)
#pragma unroll
        for (int j = 0; j < WN / 8; ++j) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                if (cc[j][e] < 0) continue;
                double *dst = p.C + (int64_t)cc[j][e] * p.ldc;
#pragma unroll
                for (int i = 0; i < WM / 8; ++i) {
                    const int row = m0 + wm0 + 8 * i + g;
                    if (row < p.M) dst[row] = acc.c[i][j][e];
                }
            }
        }
    }
}

cudaError_t launch_gather_gemm(const GatherParams &p, int num_sms, cudaStream_t s)
{
    cudaError_t e =
        cudaFuncSetAttribute(gather_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    if (p.ntiles <= 0) return cudaSuccess;
    const int grid = p.ntiles < num_sms ? p.ntiles : num_sms;
    gather_gemm_kernel<<<grid, NTHREADS, SMEM_BYTES, s>>>(p);
    return cudaGetLastError();
}

} // namespace qemb
