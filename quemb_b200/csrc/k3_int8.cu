// k3_int8.cu -- OPT-IN K3 (G = X^T X) through 8-bit integer slices on the int8 tensor cores (tcgen05.mma kind::i8,
// SASS UTCIMMA, int32 accumulators in TMEM); off by default (qemb_ctx_set_k3_int8_slices / QEMB_K3_INT8 = 5 .. 8).
//
// Same result as k3_gram.cu (reference: G = X^T X in FP64, cpp:620) up to the dropped slices:
//   X[:, j] = 2^e_j * ( sum_{a<S} 2^-(8a+6) D_a[:, j] + R ),  |R| <= 2^-(8S-1),  D_a int8: balanced base-256 digits of the
//             column rounded to 8S-2 bits below its exponent (k3_int8_digits.h); the split is error free
//   G_ij    = 2^(e_i+e_j) * sum_{d<S} 2^-(8d+12) * sum_{a+b=d} <D_a[:, i], D_b[:, j]>     (inner sums exact in int32)
// History: round 1 measured a truncating base-128 slicer on the GPU (S = 7 -> 1.3e-12 of max|G| at 1.9x the FP64 tensor
// peak, S = 8 -> 1.1e-14 at 1.3x; profiles/k3_i8_proto_r1.txt); rounded, balanced base-256 digits give more per slice
// product: S = 6 -> 2e-13, S = 7 -> 3e-16 on the same operands (profiles/k3_int8_parity_study_r1.txt).
// The FP64 DMMA kernel stays the default and the parity reference.
//
// slice_kernel : one CTA per column of X -> exponent e_j and the S slices, K-major per pair:
//                D[pair][a * naux_pad + P]  (one row of S * naux_pad bytes per pair, pad pairs zero)
// gemm_kernel  : persistent, one CTA per SM, warp specialised:
//                  warp 0   TMA producer (SWIZZLE_128B boxes of 128 bytes of K): per 128-deep k chunk the S slices of
//                           the B tile (64 pairs) and, through a ring, the S slices of the A tile (128 pairs)
//                  warp 1   one thread issues tcgen05.mma 128x64x32: slice a of A times slices b <= S-1-a of B into the
//                           accumulator of order d = a + b; S accumulators of 64 columns (<= 512 TMEM columns), so
//                           every slice tile loaded once serves up to S products
//                  warps 2-5 epilogue: tcgen05.ld of the S accumulators, FP64 recombination, scaling, store + mirror
#include "common.cuh"
#include "k3_int8_digits.h"
#include "kernels.cuh"
#include <cuda.h>

#include <cmath>

namespace qemb {

namespace {
constexpr int TM = 128;       // A tile: pairs i (TMEM lanes)
constexpr int TN = 64;        // B tile: pairs j (TMEM columns per accumulator)
constexpr int KC = 128;       // bytes (= int8 elements) of K per stage = one swizzle row
constexpr int A_RING = 4;
constexpr int A_BYTES = TM * KC, B_BYTES = TN * KC;
constexpr int NTHREADS_I8 = 32 * 6;
constexpr int EXPO_POISON = -(1 << 30); // exponent entry of a pair whose column of X holds a NaN or an Inf
template <int S>
constexpr int smem_bytes_i8() { return A_RING * A_BYTES + 2 * S * B_BYTES + 1024; }

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t *b, int bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *b, int parity)
{
    asm volatile("{\n.reg .pred p;\nKW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra KD;\nbra KW;\nKD:\n}\n" ::"r"(s32(b)),
                 "r"(parity)
                 : "memory");
}
__device__ __forceinline__ void tma_load(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(s32(dst)),
                 "l"(map), "r"(c0), "r"(c1), "r"(s32(bar))
                 : "memory");
}
__device__ __forceinline__ void umma(uint32_t dcol, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(dcol),
                 "l"(da), "l"(db), "r"(idesc), "r"(acc)
                 : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(bar)) : "memory");
}
// K-major, SWIZZLE_128B operand: rows of 128 bytes, 8-row groups 1024 bytes apart
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr)
{
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}

// ---------------------------------------------------------------------------------------------------------------
template <int S>
__global__ void slice_kernel(const double *X, int64_t ldx, int naux, int naux_pad, int64_t npair, int8_t *Dcat, int *expo)
{
    const int j = blockIdx.x;
    int8_t *row = Dcat + (int64_t)j * S * naux_pad;
    if (j >= npair) { // pad pairs of the last tile: zero rows
        for (int p = threadIdx.x; p < S * naux_pad; p += blockDim.x) row[p] = 0;
        if (threadIdx.x == 0) expo[j] = 0;
        return;
    }
    const double *x = X + ldx * j;
    __shared__ double red[32];
    double m = 0.0;
    int bad = 0; // NaN or Inf in this column: the FP64 product would poison row and column j of G, so does this path
    for (int p = threadIdx.x; p < naux; p += blockDim.x) {
        const double a = fabs(x[p]);
        m = fmax(m, a);
        bad |= !(a <= 1.7976931348623157e308);
    }
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    bad = __syncthreads_or(bad);
    m = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) m = fmax(m, red[w]);
    int e = 0;
    if (m > 0.0 && !bad) {
        frexp(m, &e); // m = f * 2^e, f in [0.5, 1)  ->  |x| * 2^-e < 1
    }
    if (threadIdx.x == 0) expo[j] = bad ? EXPO_POISON : e;
    for (int p = threadIdx.x; p < naux_pad; p += blockDim.x) i8_digits<S>((p < naux && !bad) ? x[p] : 0.0, e, row + p, naux_pad);
}

// ---------------------------------------------------------------------------------------------------------------
struct GemmParams {
    int64_t npair;
    int nkc;           // pairs, k chunks of 128 per slice (naux_pad / 128)
    int naux_pad;
    const int *expo;
    double *G;
    int64_t ldg;
    const int2 *tiles;        // (I, J): A tile rows 128 I .., B tile rows 64 J ..; lower triangle incl. diagonal blocks
    int ntiles;
    int accumulate;           // 1: add the lower triangle into G (possibly peer memory), no mirror
};

template <int S>
__global__ void __launch_bounds__(NTHREADS_I8, 1) gemm_kernel_i8(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, GemmParams p)
{
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t *sA = smem;                        // A_RING x 16 KB
    uint8_t *sB = smem + A_RING * A_BYTES;     // 2 x S x 8 KB
    __shared__ uint64_t fullA[A_RING], emptyA[A_RING], fullB[2], emptyB[2], accFull, accEmpty;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int s = 0; s < A_RING; ++s) {
            mbar_init(&fullA[s], 1);
            mbar_init(&emptyA[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&fullB[s], 1);
            mbar_init(&emptyB[s], 1);
        }
        mbar_init(&accFull, 1);
        mbar_init(&accEmpty, 4); // one arrive per epilogue warp
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tmem_base)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tacc = tmem_base;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int sa = 0, pa = 0, sb = 0, pb = 0;
            for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x) {
                const int2 tile = p.tiles[t];
                for (int kc = 0; kc < p.nkc; ++kc) {
                    mbar_wait(&emptyB[sb], pb ^ 1);
                    mbar_expect(&fullB[sb], S * B_BYTES);
                    for (int b = 0; b < S; ++b)
                        tma_load(sB + (sb * S + b) * B_BYTES, &mapB, b * p.naux_pad + kc * KC, tile.y * TN, &fullB[sb]);
                    if (++sb == 2) {
                        sb = 0;
                        pb ^= 1;
                    }
                    for (int a = 0; a < S; ++a) {
                        mbar_wait(&emptyA[sa], pa ^ 1);
                        mbar_expect(&fullA[sa], A_BYTES);
                        tma_load(sA + sa * A_BYTES, &mapA, a * p.naux_pad + kc * KC, tile.x * TM, &fullA[sa]);
                        if (++sa == A_RING) {
                            sa = 0;
                            pa ^= 1;
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
            int sa = 0, pa = 0, sb = 0, pb = 0, pacc = 0;
            for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x) {
                mbar_wait(&accEmpty, pacc ^ 1); // epilogue of the previous tile has drained TMEM
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int kc = 0; kc < p.nkc; ++kc) {
                    mbar_wait(&fullB[sb], pb);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    for (int a = 0; a < S; ++a) {
                        mbar_wait(&fullA[sa], pa);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const uint64_t da0 = make_desc(s32(sA + sa * A_BYTES));
                        const uint64_t db0 = make_desc(s32(sB + sb * S * B_BYTES));
                        const bool first = kc == 0 && a == 0;
                        for (int b = 0; b + a < S; ++b) {
                            const uint64_t db = db0 + (uint64_t)(b * (B_BYTES >> 4)); // next slice: start address + 8 KB
                            const uint32_t dcol = tacc + (uint32_t)((a + b) * TN);
                            umma(dcol, da0, db, idesc, first ? 0u : 1u);
                            umma(dcol, da0 + 2, db + 2, idesc, 1u); // + 32 bytes of K inside the swizzle row
                            umma(dcol, da0 + 4, db + 4, idesc, 1u);
                            umma(dcol, da0 + 6, db + 6, idesc, 1u);
                        }
                        umma_commit(&emptyA[sa]); // slot free once the MMAs above have read it
                        if (++sa == A_RING) {
                            sa = 0;
                            pa ^= 1;
                        }
                    }
                    umma_commit(&emptyB[sb]);
                    if (++sb == 2) {
                        sb = 0;
                        pb ^= 1;
                    }
                }
                umma_commit(&accFull);
                pacc ^= 1;
            }
        }
    } else {
        // ------------------------------------------------------------------ epilogue (warps 2..5)
        const int lbase = 32 * (warp & 3); // TMEM lanes this warp may read
        int pacc = 0;
        for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x) {
            const int2 tile = p.tiles[t];
            mbar_wait(&accFull, pacc);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int64_t i = (int64_t)tile.x * TM + lbase + lane;
            const int ei = i < p.npair ? p.expo[i] : 0;
            for (int c0 = 0; c0 < TN; c0 += 8) {
                double g[8];
                uint32_t v[S][8];
#pragma unroll
                for (int d = 0; d < S; ++d) { // all S accumulators of these 8 columns in flight, one wait
                    const uint32_t addr = tacc + ((uint32_t)lbase << 16) + (uint32_t)(d * TN + c0);
                    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                 : "=r"(v[d][0]), "=r"(v[d][1]), "=r"(v[d][2]), "=r"(v[d][3]), "=r"(v[d][4]), "=r"(v[d][5]), "=r"(v[d][6]),
                                   "=r"(v[d][7])
                                 : "r"(addr));
                }
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int q = 0; q < 8; ++q) g[q] = 0.0;
#pragma unroll
                for (int d = S - 1; d >= 0; --d) { // small terms first
                    const double w = ldexp(1.0, -(8 * d + 12));
#pragma unroll
                    for (int q = 0; q < 8; ++q) g[q] = fma((double)(int32_t)v[d][q], w, g[q]);
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const int64_t j = (int64_t)tile.y * TN + c0 + q;
                    if (i < p.npair && j < p.npair && j <= i) {
                        const int ej = p.expo[j];
                        const double val = (ei == EXPO_POISON || ej == EXPO_POISON) ? __longlong_as_double(0x7ff8000000000000ll)
                                                                                    : ldexp(g[q], ei + ej);
                        if (p.accumulate) {
                            atomicAdd_system(&p.G[i + p.ldg * j], val);
                        } else {
                            p.G[i + p.ldg * j] = val;
                            p.G[j + p.ldg * i] = val;
                        }
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&accEmpty);
            pacc ^= 1;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tacc), "r"(512));
}


// tile list on the device: row block I (128 pairs) times column blocks J (64 pairs) up to the diagonal
__global__ void i8_tiles_kernel(int2 *tiles, int nI, int nJ)
{
    const int I = blockIdx.x * blockDim.x + threadIdx.x;
    if (I >= nI) return;
    // row blocks before I hold 2 i + 2 column blocks each (none of them is clipped by nJ): I (I + 1) tiles
    const int off = I * (I + 1);
    const int cnt = 2 * I + 2 < nJ ? 2 * I + 2 : nJ;
    for (int J = 0; J < cnt; ++J) tiles[off + J] = make_int2(I, J);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode_i8()
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

template <int S>
cudaError_t launch_i8(const GramI8Params &p, int num_sms, cudaStream_t s)
{
    EncodeTiledFn enc = get_encode_i8();
    if (!enc) return cudaErrorNotSupported;
    const int naux_pad = gram_i8_naux_pad(p.nrows);
    const int64_t npair_pad = (p.npair + TM - 1) / TM * TM;
    const int nI = (int)(npair_pad / TM), nJ = (int)((p.npair + TN - 1) / TN);
    int8_t *D = p.scratch;
    int *expo = reinterpret_cast<int *>(p.scratch + (size_t)npair_pad * S * naux_pad);
    int2 *tiles = reinterpret_cast<int2 *>(expo + npair_pad);
    int ntiles = 0;
    for (int I = 0; I < nI; ++I) ntiles += 2 * I + 2 < nJ ? 2 * I + 2 : nJ;
    CUtensorMap maps[2];
    for (int m = 0; m < 2; ++m) {
        const cuuint64_t dims[2] = {(cuuint64_t)S * naux_pad, (cuuint64_t)npair_pad};
        const cuuint64_t strides[1] = {(cuuint64_t)S * naux_pad};
        const cuuint32_t box[2] = {KC, (cuuint32_t)(m == 0 ? TM : TN)};
        const cuuint32_t estr[2] = {1, 1};
        if (enc(&maps[m], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, D, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return cudaErrorInvalidValue;
    }
    slice_kernel<S><<<(unsigned)npair_pad, 256, 0, s>>>(p.X, p.ldx, p.nrows, naux_pad, p.npair, D, expo);
    i8_tiles_kernel<<<(nI + 127) / 128, 128, 0, s>>>(tiles, nI, nJ);
    cudaError_t e = cudaFuncSetAttribute(gemm_kernel_i8<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes_i8<S>());
    if (e != cudaSuccess) return e;
    GemmParams gp{p.npair, naux_pad / KC, naux_pad, expo, p.G, p.ldg, tiles, ntiles, p.accumulate};
    const int grid = ntiles < num_sms ? ntiles : num_sms;
    gemm_kernel_i8<S><<<grid, NTHREADS_I8, smem_bytes_i8<S>(), s>>>(maps[0], maps[1], gp);
    return cudaGetLastError();
}
} // namespace

int gram_i8_naux_pad(int nrows) { return (nrows + KC - 1) / KC * KC; }

size_t gram_i8_scratch_bytes(int nrows, int64_t npair, int slices)
{
    const int64_t npair_pad = (npair + TM - 1) / TM * TM;
    const int64_t nI = npair_pad / TM;
    return (size_t)npair_pad * slices * gram_i8_naux_pad(nrows) + (size_t)npair_pad * 4 + (size_t)(nI * (nI + 1) + 2 * nI + 8) * sizeof(int2);
}

// ---- guaranteed-tolerance slice choice ------------------------------------------------------------------------
// Largest binary exponent over the columns of X (frexp of the column maximum, the slicer's own rule; columns holding
// a NaN / Inf or only zeros do not count).  With e_max known BEFORE slicing, gram_i8_bound(naux, S) * 2^(2 e_max) is a
// worst-case bound on |G_int8 - X^T X| for every entry, so the smallest S that meets a caller tolerance can be picked
// (and the FP64 kernel when none does).
__global__ void max_expo_kernel(const double *X, int64_t ldx, int naux, int64_t npair, int *out)
{
    __shared__ double red[32];
    for (int64_t j = blockIdx.x; j < npair; j += gridDim.x) {
        const double *x = X + ldx * j;
        double m = 0.0;
        int bad = 0;
        for (int p = threadIdx.x; p < naux; p += blockDim.x) {
            const double a = fabs(x[p]);
            m = fmax(m, a);
            bad |= !(a <= 1.7976931348623157e308);
        }
        for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
        bad = __syncthreads_or(bad);
        if (threadIdx.x == 0) {
            m = 0.0;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) m = fmax(m, red[w]);
            if (m > 0.0 && !bad) {
                int e = 0;
                frexp(m, &e);
                atomicMax(out, e);
            }
        }
        __syncthreads();
    }
}

cudaError_t launch_max_expo(const double *X, int64_t ldx, int naux, int64_t npair, int *out_dev, cudaStream_t s)
{
    const int init = -(1 << 30);
    cudaError_t e = cudaMemcpyAsync(out_dev, &init, sizeof(int), cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess || npair <= 0) return e;
    max_expo_kernel<<<(unsigned)(npair < 148 * 8 ? npair : 148 * 8), 256, 0, s>>>(X, ldx, naux, npair, out_dev);
    return cudaGetLastError();
}

// worst case of |G_int8 - X^T X| / 2^(e_i + e_j): remainders of both operands (|R| <= 2^-(8S-1) after rounding to
// 8S-2 bits) plus the slice orders d >= S that are never formed (digits in [-128, 127]); oracle/k3_int8_oracle.py
// `truncation_bound` is the same sum
double gram_i8_bound(int64_t naux, int slices)
{
    const double rem = std::ldexp(1.0, -(8 * slices - 1));
    double dropped = 0.0;
    for (int d = slices; d < 2 * slices - 1; ++d) {
        const int lo = d - slices + 1 > 0 ? d - slices + 1 : 0, hi = d < slices - 1 ? d : slices - 1;
        dropped += (double)(hi - lo + 1) * 16384.0 * std::ldexp(1.0, -(8 * d + 12));
    }
    return (double)naux * (2.0 * rem * (1.0 + rem) + rem * rem + dropped);
}

cudaError_t launch_gram_int8(const GramI8Params &p, int num_sms, cudaStream_t s)
{
    if (p.npair <= 0) return cudaSuccess;
    if (p.slices == 5) return launch_i8<5>(p, num_sms, s);
    if (p.slices == 6) return launch_i8<6>(p, num_sms, s);
    if (p.slices == 7) return launch_i8<7>(p, num_sms, s);
    if (p.slices == 8) return launch_i8<8>(p, num_sms, s);
    return cudaErrorInvalidValue;
}

} // namespace qemb

// Host twin of slice_kernel (same exponent rule, same digit routine): D[(c * slices + a) * nrows + p], expo[c].
// For the CPU test suite (the device slicer against oracle/k3_int8_oracle.py without a GPU); not on the product path.
extern "C" int qemb_k3_int8_slice_host(const double *X, int64_t ldx, int64_t nrows, int64_t ncols, int slices, int8_t *D, int *expo)
{
    if (!X || !D || !expo || nrows < 0 || ncols < 0 || ldx < nrows || slices < 5 || slices > 8) return 1;
    for (int64_t c = 0; c < ncols; ++c) {
        const double *x = X + ldx * c;
        double m = 0.0;
        bool bad = false;
        for (int64_t p = 0; p < nrows; ++p) {
            const double a = fabs(x[p]);
            m = fmax(m, a);
            bad |= !(a <= 1.7976931348623157e308);
        }
        int e = 0;
        if (m > 0.0 && !bad) frexp(m, &e);
        expo[c] = bad ? qemb::EXPO_POISON : e;
        int8_t *row = D + c * slices * nrows;
        for (int64_t p = 0; p < nrows; ++p) {
            const double v = bad ? 0.0 : x[p];
            switch (slices) {
            case 5: qemb::i8_digits<5>(v, e, row + p, nrows); break;
            case 6: qemb::i8_digits<6>(v, e, row + p, nrows); break;
            case 7: qemb::i8_digits<7>(v, e, row + p, nrows); break;
            default: qemb::i8_digits<8>(v, e, row + p, nrows); break;
            }
        }
    }
    return 0;
}
