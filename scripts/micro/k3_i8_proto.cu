// k3_i8_proto.cu -- PROTOTYPE (not product code) of DESIGN section 8, item 1: K3 (G = X^T X in FP64) through 7-bit
// integer slices on the int8 tensor cores of sm_100a (tcgen05.mma kind::i8, accumulators in TMEM).
//
//   X[:, j] = 2^e_j * sum_a 2^-(7a+6) D_a[:, j]  (+ remainder below 2^-(7S-1)),  D_a int8, error free
//   G_ij    = 2^(e_i+e_j) * sum_{d<S} 2^-(7d+12) * sum_{a+b=d} <D_a[:, i], D_b[:, j]>     (inner sums exact in int32)
//
// slice_kernel : one CTA per column of X -> exponent e_j and the S slices, stored K-major per pair:
//                Dcat[pair][a * naux_pad + P]  (one row of S * naux_pad bytes per pair)
// gemm_kernel  : persistent, one CTA per SM, warp specialised:
//                  warp 0   TMA producer (SWIZZLE_128B boxes of 128 bytes of K): per 128-deep k chunk the S slices of
//                           the B tile (64 pairs) and, one after the other through a ring, the S slices of the A tile
//                           (128 pairs)
//                  warp 1   one thread issues tcgen05.mma 128x64x32: slice a of A times slices b <= S-1-a of B into the
//                           accumulator of order d = a + b; S accumulators of 64 columns = 448 of the 512 TMEM columns,
//                           so every slice tile loaded once serves up to S products
//                  warps 2-5 epilogue: tcgen05.ld of the S accumulators, FP64 recombination, scaling, store (+ mirror)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o k3_i8_proto k3_i8_proto.cu -lcuda && ./k3_i8_proto [npair] [naux]
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>
#include <vector>

#ifndef S_DEF
#define S_DEF 7
#endif
constexpr int S = S_DEF;      // slices (S * TN <= 512 TMEM columns)
constexpr int TM = 128;       // A tile: pairs i (TMEM lanes)
constexpr int TN = 64;        // B tile: pairs j (TMEM columns per accumulator)
constexpr int KC = 128;       // bytes (= int8 elements) of K per stage = one swizzle row
#ifndef A_RING_DEF
#define A_RING_DEF 4
#endif
constexpr int A_RING = A_RING_DEF;
constexpr int A_BYTES = TM * KC, B_BYTES = TN * KC;
constexpr int SMEM_BYTES = A_RING * A_BYTES + 2 * S * B_BYTES + 1024;
constexpr int NTHREADS = 32 * 6;

#define CK(x)                                                                                                          \
    do {                                                                                                               \
        cudaError_t e_ = (x);                                                                                          \
        if (e_ != cudaSuccess) {                                                                                       \
            printf("%s failed: %s\n", #x, cudaGetErrorString(e_));                                                     \
            exit(1);                                                                                                   \
        }                                                                                                              \
    } while (0)

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
__global__ void slice_kernel(const double *X, int64_t ldx, int naux, int naux_pad, int8_t *Dcat, int *expo)
{
    const int j = blockIdx.x;
    const double *x = X + ldx * j;
    __shared__ double red[32];
    double m = 0.0;
    for (int p = threadIdx.x; p < naux; p += blockDim.x) m = fmax(m, fabs(x[p]));
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    m = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) m = fmax(m, red[w]);
    int e = 0;
    if (m > 0.0) {
        frexp(m, &e); // m = f * 2^e, f in [0.5, 1)  ->  |x| * 2^-e < 1
    }
    if (threadIdx.x == 0) expo[j] = e;
    int8_t *row = Dcat + (int64_t)j * S * naux_pad;
    for (int p = threadIdx.x; p < naux_pad; p += blockDim.x) {
        double r = p < naux ? ldexp(x[p], -e) : 0.0; // exact
        double scale = 64.0;
#pragma unroll
        for (int a = 0; a < S; ++a) {
            const double t = trunc(r * scale);
            r = r * scale - t; // exact
            row[a * naux_pad + p] = (int8_t)t;
            scale = 128.0;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
struct GemmParams {
    int npair, nkc;           // pairs, k chunks of 128 per slice (naux_pad / 128)
    int naux_pad;
    const int *expo;
    double *G;
    int64_t ldg;
    const int2 *tiles;        // (I, J): A tile rows 128 I .., B tile rows 64 J ..; lower triangle incl. diagonal blocks
    int ntiles;
};

__global__ void __launch_bounds__(NTHREADS, 1) gemm_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, GemmParams p)
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
            const int i = tile.x * TM + lbase + lane;
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
                    const double w = ldexp(1.0, -(7 * d + 12));
#pragma unroll
                    for (int q = 0; q < 8; ++q) g[q] = fma((double)(int32_t)v[d][q], w, g[q]);
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const int j = tile.y * TN + c0 + q;
                    if (i < p.npair && j < p.npair && j <= i) {
                        const double val = ldexp(g[q], ei + p.expo[j]);
                        p.G[i + p.ldg * j] = val;
#ifndef NO_MIRROR
                        p.G[j + p.ldg * i] = val;
#endif
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

// ---------------------------------------------------------------------------------------------------------------
static CUtensorMap make_map(const int8_t *base, int64_t krow_bytes, int64_t rows, int box_rows)
{
    CUtensorMap m;
    cuuint64_t dims[2] = {(cuuint64_t)krow_bytes, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)krow_bytes};
    cuuint32_t box[2] = {KC, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = cuTensorMapEncodeTiled(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void *)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        printf("cuTensorMapEncodeTiled failed: %d\n", (int)r);
        exit(1);
    }
    return m;
}

int main(int argc, char **argv)
{
    const int npair = argc > 1 ? atoi(argv[1]) : 2000, naux = argc > 2 ? atoi(argv[2]) : 700;
    const int naux_pad = (naux + KC - 1) / KC * KC;
    const int npair_pad = (npair + TM - 1) / TM * TM;
    printf("npair %d naux %d (padded %d x %d), %d slices\n", npair, naux, npair_pad, naux_pad, S);
    CK(cudaFree(0));
    // K3-like operand: column norms over 2.6 decades, rows decaying
    std::vector<double> hX((size_t)naux * npair);
    srand(11);
    auto rnd = []() { return (rand() + 0.5) / ((double)RAND_MAX + 1.0); };
    std::vector<double> colw(npair), roww(naux);
    for (auto &c : colw) c = exp(-6.0 * rnd());
    for (auto &r : roww) r = exp(-4.0 * rnd());
    for (int j = 0; j < npair; ++j)
        for (int q = 0; q < naux; ++q) {
            const double u = sqrt(-2.0 * log(rnd())) * cos(6.283185307179586 * rnd());
            hX[(size_t)j * naux + q] = u * colw[j] * roww[q];
        }
    double *dX, *dG;
    int8_t *dD;
    int *dE;
    CK(cudaMalloc(&dX, hX.size() * 8));
    CK(cudaMemcpy(dX, hX.data(), hX.size() * 8, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dD, (size_t)npair_pad * S * naux_pad));
    CK(cudaMemset(dD, 0, (size_t)npair_pad * S * naux_pad));
    CK(cudaMalloc(&dE, npair_pad * 4));
    CK(cudaMemset(dE, 0, npair_pad * 4));
    CK(cudaMalloc(&dG, (size_t)npair * npair * 8));
    CK(cudaMemset(dG, 0xff, (size_t)npair * npair * 8));

    std::vector<int2> tiles;
    for (int I = 0; I < npair_pad / TM; ++I)
        for (int J = 0; J * TN <= I * TM + TM - 1 && J * TN < npair; ++J) tiles.push_back(make_int2(I, J));
    int2 *dT;
    CK(cudaMalloc(&dT, tiles.size() * sizeof(int2)));
    CK(cudaMemcpy(dT, tiles.data(), tiles.size() * sizeof(int2), cudaMemcpyHostToDevice));

    const CUtensorMap mapA = make_map(dD, (int64_t)S * naux_pad, npair_pad, TM), mapB = make_map(dD, (int64_t)S * naux_pad, npair_pad, TN);
    GemmParams p{npair, naux_pad / KC, naux_pad, dE, dG, npair, dT, (int)tiles.size()};
    cudaDeviceProp pr;
    CK(cudaGetDeviceProperties(&pr, 0));
    CK(cudaFuncSetAttribute(gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    cudaEvent_t e0, e1, e2;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventCreate(&e2);
    const int grid = (int)tiles.size() < pr.multiProcessorCount ? (int)tiles.size() : pr.multiProcessorCount;
    float ms_s = 0, ms_g = 0;
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        slice_kernel<<<npair, 256>>>(dX, naux, naux, naux_pad, dD, dE);
        cudaEventRecord(e1);
        gemm_kernel<<<grid, NTHREADS, SMEM_BYTES>>>(mapA, mapB, p);
        cudaEventRecord(e2);
        CK(cudaDeviceSynchronize());
        cudaEventElapsedTime(&ms_s, e0, e1);
        cudaEventElapsedTime(&ms_g, e1, e2);
    }
    const double ops = 2.0 * 128.0 * 64.0 * (double)tiles.size() * (S * (S + 1) / 2) * naux_pad;
    printf("slice %.3f ms, int8 gemm + recombination %.3f ms (%zu tiles, %.2f Pop/s); FP64 DMMA at 35.4 TF/s would need %.3f ms for the lower triangle\n",
           ms_s, ms_g, tiles.size(), ops / ms_g / 1e12, (double)npair * npair * naux / 35.4e12 * 1e3);

    // check sampled entries against long double on the host
    std::vector<double> hG((size_t)npair * npair);
    CK(cudaMemcpy(hG.data(), dG, hG.size() * 8, cudaMemcpyDeviceToHost));
    double gmax = 0, emax = 0;
    for (int s = 0; s < 4000; ++s) {
        const int i = rand() % npair, j = rand() % npair;
        long double r = 0;
        for (int q = 0; q < naux; ++q) r += (long double)hX[(size_t)i * naux + q] * hX[(size_t)j * naux + q];
        gmax = fmax(gmax, fabs((double)r));
        emax = fmax(emax, fabs(hG[i + (size_t)npair * j] - (double)r));
    }
    printf("4000 sampled entries: max |G_int8 - G_exact| = %.3e (%.2e of the largest sampled |G| %.3e)\n", emax, emax / gmax, gmax);
    return emax / gmax > 1e-10;
}
