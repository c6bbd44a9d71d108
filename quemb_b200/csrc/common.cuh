// common.cuh -- device helpers shared by the sm_100a kernels of libqemb_b200.
//
// FP64 on Blackwell (sm_100a): tcgen05.mma has no f64 kind, so the FP64 tensor path is the
// warp-level `mma.sync.aligned.m8n8k4.f64` (SASS: DMMA.8x8x4; the larger m16n8k{4,8,16} PTX shapes
// are split by ptxas into the same DMMA.8x8x4 on sm_100a, checked with cuobjdump).  Operands are
// staged global -> shared with cp.async (LDGSTS) in a multi-stage ring and read back with 64-bit
// LDS in bank-conflict-free padded layouts.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace qemb {

// ---- cp.async -------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// 16-byte copy, L2 only (.cg); src_bytes = 0 zero-fills the destination (used for all tails).
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem, bool pred)
{
    const int n = pred ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(smem_u32(smem)), "l"(gmem), "r"(n));
}
// 8-byte copy (.ca is the only variant for < 16 bytes)
__device__ __forceinline__ void cp_async8(void *smem, const void *gmem, bool pred)
{
    const int n = pred ? 8 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(smem_u32(smem)), "l"(gmem), "r"(n));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

// ---- DMMA -----------------------------------------------------------------------------------
// D(8x8) += A(8x4, row) * B(4x8, col).  lane = 4*g + t:
//   a  = A[g][t]      b = B[t][g]      c0,c1 = C[g][2t], C[g][2t+1]
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// ---- tile geometry shared by all GEMM-shaped kernels ----------------------------------------
constexpr int BK = 16;          // k-depth of one pipeline stage (4 DMMA k-steps)
constexpr int PITCH_K = BK + 4; // [row][k] layouts: 20 doubles -> 16 lanes of a 64-bit LDS phase hit 16 distinct bank pairs
// [k][row] layouts use pitch rows+4 (132 for 128 rows, 68 for 64 rows): (rows+4) % 16 == 4

__host__ __device__ inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

} // namespace qemb
