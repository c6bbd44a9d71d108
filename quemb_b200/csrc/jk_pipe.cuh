// jk_pipe.cuh -- mbarrier / bulk-copy helpers shared by the streaming J/K kernels (jk_stream.cu, jk_stream12.cu).
#pragma once
#include "common.cuh"
#include <type_traits>

namespace qemb {
namespace jkpipe {

template <int I, int N, class F>
__device__ __forceinline__ void static_for(F &&f)
{
    if constexpr (I < N) {
        f(std::integral_constant<int, I>{});
        static_for<I + 1, N>(f);
    }
}

__device__ __forceinline__ int tri32(int k) { return (k * (k + 1)) >> 1; }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, int bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, int parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "JS_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra JS_DONE;\n"
        "bra JS_WAIT;\n"
        "JS_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_load(void *dst, const void *src, int bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void pair_of(int64_t r, int &i, int &j)
{
    int a = (int)((sqrt(8.0 * (double)r + 1.0) - 1.0) * 0.5);
    while ((int64_t)a * (a + 1) / 2 + a + 1 <= r) ++a; // tri(a + 1) <= r
    while ((int64_t)a * (a + 1) / 2 > r) --a;
    i = a;
    j = (int)(r - (int64_t)a * (a + 1) / 2);
}

} // namespace jkpipe
} // namespace qemb
