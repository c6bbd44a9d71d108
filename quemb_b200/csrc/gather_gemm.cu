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
//   * producer warpgroup (4 warps, 88 registers each after setmaxnreg; the consumers take 208):
//       - A tile: 16 gathered columns x 128 rows per stage = 16 bulk copies of 1 KB (cp.async.bulk, the 1-D
//         TMA path, one per lane) into As[k][m] (pitch 132), completing on the stage's "full" mbarrier;
//       - B tile: the element-wise double gather TA[brow[k], bcol[n]] out of the L1/L2-resident TA with plain
//         loads + st.shared into Bs[n][k] (pitch 20); the four warps arrive on the same mbarrier.
//     The producers run up to STAGES stages ahead, across tile boundaries, so index loads, the gather
//     latency, the pipeline fill of the next tile and the epilogue of the previous one are all hidden.
//   * eight consumer warps (32 x 64 register tiles of DMMA.8x8x4 accumulators, 4 m-warps x 2 n-warps) wait on
//     "full", read fragments with conflict-free 64-bit LDS and release the stage through "empty".  No
//     CTA-wide barrier.
//   * n quantisation: the 16 column blocks (of 8) of a tile alternate between the two n-warps, which sit on the
//     same scheduler; a warp runs the stage body specialised for its number of live blocks (8 code paths,
//     warp-uniform branch), so a ragged tile costs ceil(N/8)*8 columns of tensor-pipe time instead of 128
//     (K2a groups have N = 1..nmo: 61 % -> 96 % useful work at nmo = 220; K1: 61 % -> 93 %).
#include <cstdlib>

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
constexpr int WM = 32, WN = 64; // consumer warp tile: 4 m-warps x 2 n-warps

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

// One BK = 16 stage of a consumer warp with JV live column blocks of 8.  (A predicated form -- `if (j < jv)`
// around every DMMA -- costs a full stage whatever jv is: a predicated-off DMMA still takes its slot in the
// pipe; measured, K1 4.3 ms with and without the skips.)  JV is warp-uniform, so the dispatch is a plain branch.
template <int JV>
__device__ __forceinline__ void consumer_stage(WarpAcc<WM, WN> &acc, const double *__restrict__ As,
                                               const double *__restrict__ Bs, int wm0, int wni, int g, int t)
{
#pragma unroll
    for (int ks = 0; ks < BK / 4; ++ks) {
        double a[WM / 8], b[JV];
#pragma unroll
        for (int i = 0; i < WM / 8; ++i) a[i] = As[(4 * ks + t) * PA + wm0 + 8 * i + g];
#pragma unroll
        for (int j = 0; j < JV; ++j) b[j] = Bs[(8 * (2 * j + wni) + g) * PB + 4 * ks + t];
#pragma unroll
        for (int j = 0; j < JV; ++j)
#pragma unroll
            for (int i = 0; i < WM / 8; ++i) dmma884(acc.c[i][j][0], acc.c[i][j][1], a[i], b[j]);
    }
}

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
        asm volatile("setmaxnreg.dec.sync.aligned.u32 88;");
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
            // software pipeline: the indices of stage kt+1 and the TA values of stage kt are loaded BEFORE the
            // wait for the stage's buffer, so their latency overlaps the wait (or, when the consumers are
            // starved, only one load latency + the stores sits between two stages)
            const int nfill = ((nvalid + 7) >> 3) << 3; // consumers never use column blocks beyond this
            int rr = (kk < K) ? brow[kk] : -1;
            int ac = (pw == 0 && lane < min(BK, K)) ? acol[lane] : -1;
            for (int kt = 0; kt < ktiles; ++kt) {
                const int k0 = kt * BK;
                const int nk = min(BK, K - k0);
                double v[TN / 8];
#pragma unroll
                for (int r = 0; r < TN / 8; ++r) {
                    v[r] = 0.0;
                    if (rr >= 0 && boff[r] >= 0 && !(p.debug & 1)) v[r] = __ldg(p.TA + rr + boff[r]);
                }
                const int k1 = k0 + BK;
                const int rr_next = (k1 + kk < K) ? brow[k1 + kk] : -1;
                int ac_next = -1;
                if (pw == 0 && k1 + lane < K && lane < BK) ac_next = acol[k1 + lane];
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
                    if (lane == 0) mbar_expect_tx(&full[s], (p.debug & 2) ? 0 : nk * rows * 8);
                    __syncwarp();
                    if (lane < nk && !(p.debug & 2)) bulk_load(As + lane * PA, p.Asrc + (int64_t)ac * p.lda + m0, rows * 8, &full[s]);
                }
                // B tile: Bs[n][k] = TA[brow[k], bcol[n]]
#pragma unroll
                for (int r = 0; r < TN / 8; ++r) {
                    const int n = nb + 8 * r;
                    if (n < nfill) Bs[n * PB + kk] = v[r];
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&full[s]);
                rr = rr_next;
                ac = ac_next;
                if (++s == STAGES) {
                    s = 0;
                    phase ^= 1;
                }
            }
        }
        return;
    }

    // -------------------------------------- consumers ------------------------------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
    // warps w and w + 4 share a scheduler: they take the same rows and the two column-block parities, so every
    // scheduler sees ceil(N/8) blocks in total whatever N is
    const int wm0 = (warp & 3) * WM, wni = warp >> 2;
    const int g = lane >> 2, t = lane & 3;

    // the "empty" arrive of a stage is deferred until after the wait for the next one (see k3_gram.cu: an arrive
    // right behind the stage body is scheduled ahead of the DMMAs, i.e. before the stage's LDS are known to be done)
    int s = 0, phase = 0, held = -1;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const int4 tt = p.tiles[tile];
        const GatherGroup grp = p.groups[tt.x];
        const int m0 = tt.y * TM, n0 = tt.z * TN;
        const int nvalid = min(TN, grp.N - n0);
        const int ktiles = (grp.K + BK - 1) / BK;
        const int nblocks = (nvalid + 7) >> 3;      // live column blocks of 8
        const int jv = (nblocks - wni + 1) >> 1;    // ... of which this warp owns blocks 2 j + wni

        WarpAcc<WM, WN> acc;
        acc.zero();
        for (int kt = 0; kt < ktiles; ++kt) {
            mbar_wait(&full[s], phase);
            if (held >= 0) {
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[held]);
            }
            held = s;
            const double *As = reinterpret_cast<const double *>(smem + s * STAGE_BYTES);
            const double *Bs = reinterpret_cast<const double *>(smem + s * STAGE_BYTES + A_BYTES);
            switch (jv) {
#define GG_CASE(J) case J: consumer_stage<J>(acc, As, Bs, wm0, wni, g, t); break;
                GG_CASE(8) GG_CASE(7) GG_CASE(6) GG_CASE(5) GG_CASE(4) GG_CASE(3) GG_CASE(2) GG_CASE(1)
#undef GG_CASE
            default: break;
            }
            if (++s == STAGES) {
                s = 0;
                phase ^= 1;
            }
        }

        // epilogue: scatter the columns (K == 0 groups write zeros, like the reference's zero init); rows beyond
        // M were computed on whatever the A stage held and are dropped here
        const int *ccol = p.ccol + grp.b_off + n0;
#pragma unroll
        for (int j = 0; j < WN / 8; ++j) {
            if (j >= jv) break;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int n = 8 * (2 * j + wni) + 2 * t + e;
                if (n >= nvalid) continue;
                double *dst = p.C + (int64_t)ccol[n] * p.ldc;
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
    GatherParams q = p;
    static const int dbg = getenv("QEMB_GG_DEBUG") ? atoi(getenv("QEMB_GG_DEBUG")) : 0;
    q.debug = dbg;
    gather_gemm_kernel<<<grid, NTHREADS, SMEM_BYTES, s>>>(q);
    return cudaGetLastError();
}

} // namespace qemb
