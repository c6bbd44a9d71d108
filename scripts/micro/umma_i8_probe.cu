// umma_i8_probe.cu -- smallest possible check of the pieces an int8-slice K3 needs:
// tcgen05.mma kind::i8 (SASS UTCIMMA) with K-major 128-byte-swizzled shared-memory operands, an int32 accumulator in
// TMEM, tcgen05.commit -> mbarrier, tcgen05.ld epilogue.  One CTA, one 128 x 128 x 128 tile, operands staged by plain
// stores in the layout a SWIZZLE_128B TMA box would produce; result compared with the host.  Not product code.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_i8_probe umma_i8_probe.cu && ./umma_i8_probe
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int M = 128, N = 128, K = 128; // K in int8 elements = one 128-byte swizzle row

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, SWIZZLE_128B: row r at r * 128 bytes, 16-byte chunk c stored at chunk c ^ (r & 7); 8-row groups 1024 bytes apart
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);   // start address
    d |= (uint64_t)1 << 16;                   // leading byte offset (unused for swizzled K-major): 1
    d |= (uint64_t)(1024 >> 4) << 32;         // stride byte offset between 8-row groups
    d |= (uint64_t)1 << 46;                   // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                   // SWIZZLE_128B
    return d;
}

__global__ void __launch_bounds__(128, 1) probe(const int8_t *A, const int8_t *B, int32_t *C)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *sA = smem, *sB = smem + M * 128;
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    for (int idx = tid; idx < M * 8; idx += 128) {
        const int r = idx >> 3, c = idx & 7;
        *reinterpret_cast<int4 *>(sA + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const int4 *>(A + r * K + c * 16);
    }
    for (int idx = tid; idx < N * 8; idx += 128) {
        const int r = idx >> 3, c = idx & 7;
        *reinterpret_cast<int4 *>(sB + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const int4 *>(B + r * K + c * 16);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy stores -> visible to the tensor core
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tmem_base)), "r"(128));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tacc = tmem_base;

    if (tid == 0) {
        // instruction descriptor: D = S32, A = B = signed int8, both K-major, N >> 3, M >> 4
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        for (int k = 0; k < K / 32; ++k) { // one instruction = 32 int8 along K = 32 bytes inside the swizzle row
            const uint64_t da = make_desc(s32(sA) + 32 * k), db = make_desc(s32(sB) + 32 * k);
            const uint32_t acc = k > 0;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tacc),
                         "l"(da), "l"(db), "r"(idesc), "r"(acc)
                         : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)) : "memory");
    }
    // everybody waits for the accumulator
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(s32(&bar)) : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // epilogue: warp w owns TMEM lanes 32 w .. 32 w + 31 = rows of the tile; 8 columns per load
    for (int c0 = 0; c0 < N; c0 += 8) {
        uint32_t v[8];
        const uint32_t addr = tacc + ((uint32_t)(32 * warp) << 16) + (uint32_t)c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                     : "r"(addr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int j = 0; j < 8; ++j) C[(32 * warp + lane) * N + c0 + j] = (int32_t)v[j];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tacc), "r"(128));
}

// issue-rate probe: every SM issues `iters` x (K/32) UMMAs of shape 128 x NN x 32 on resident operands (no loads)
template <int NN>
__global__ void __launch_bounds__(128, 1) rate(int iters)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (M + 256) * 128 / 16; i += 128) reinterpret_cast<int4 *>(smem)[i] = make_int4(0x01010101, 0x01010101, 0x01010101, 0x01010101);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tmem_base)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tacc = tmem_base;
    if (tid == 0) {
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NN >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint32_t a0 = s32(smem), b0 = s32(smem + M * 128);
        for (int it = 0; it < iters; ++it) {
            const uint32_t acc_col = (it & 1) * NN; // alternate two accumulators
            for (int k = 0; k < K / 32; ++k) {
                const uint64_t da = make_desc(a0 + 32 * k), db = make_desc(b0 + 32 * k);
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tacc + acc_col),
                             "l"(da), "l"(db), "r"(idesc), "r"(1u)
                             : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)) : "memory");
    }
    asm volatile("{\n.reg .pred p;\nW2:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D2;\nbra W2;\nD2:\n}\n" ::"r"(s32(&bar)) : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tacc), "r"(512));
}

template <int NN>
static void run_rate(int sms)
{
    const int smem = (M + 256) * 128 + 1024, iters = 20000;
    cudaFuncSetAttribute(rate<NN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    rate<NN><<<sms, 128, smem>>>(100);
    cudaEventRecord(e0);
    rate<NN><<<sms, 128, smem>>>(iters);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double ops = 2.0 * M * NN * K * (double)iters * sms;
    printf("UTCIMMA 128x%dx32, %d SMs, resident operands: %.3f ms, %.2f Pop/s (%s)\n", NN, sms, ms, ops / ms / 1e12, cudaGetErrorString(e));
}

int main()
{
    int8_t *hA = (int8_t *)malloc(M * K), *hB = (int8_t *)malloc(N * K);
    srand(3);
    for (int i = 0; i < M * K; ++i) hA[i] = (int8_t)(rand() % 255 - 127);
    for (int i = 0; i < N * K; ++i) hB[i] = (int8_t)(rand() % 255 - 127);
    int8_t *dA, *dB;
    int32_t *dC, *hC = (int32_t *)malloc(M * N * 4);
    cudaMalloc(&dA, M * K);
    cudaMalloc(&dB, N * K);
    cudaMalloc(&dC, M * N * 4);
    cudaMemcpy(dA, hA, M * K, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB, N * K, cudaMemcpyHostToDevice);
    cudaMemset(dC, 0xff, M * N * 4);
    const int smem = (M + N) * 128 + 1024;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    probe<<<1, 128, smem>>>(dA, dB, dC);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    cudaMemcpy(hC, dC, M * N * 4, cudaMemcpyDeviceToHost);
    long bad = 0;
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) {
            int32_t r = 0;
            for (int k = 0; k < K; ++k) r += (int32_t)hA[m * K + k] * (int32_t)hB[n * K + k];
            if (r != hC[m * N + n]) {
                if (bad < 5) printf("  C[%d][%d] = %d, expected %d\n", m, n, hC[m * N + n], r);
                ++bad;
            }
        }
    printf("tcgen05.mma kind::i8 128x128x128: %ld of %d entries differ -> %s\n", bad, M * N, bad ? "MISMATCH" : "bit-exact");
    if (bad) return 1;
    cudaDeviceProp pr;
    cudaGetDeviceProperties(&pr, 0);
    run_rate<64>(pr.multiProcessorCount);
    run_rate<128>(pr.multiProcessorCount);
    run_rate<256>(pr.multiProcessorCount);
    return 0;
}
