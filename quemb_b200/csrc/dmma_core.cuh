// dmma_core.cuh -- the warp-level FP64 tensor-core inner product used by every GEMM-shaped kernel.
//
// A CTA tile is computed by 8 warps; each warp owns a WM x WN sub-tile held in registers as
// (WM/8) x (WN/8) DMMA.8x8x4 accumulators (2 doubles per thread each).  One pipeline stage holds
// BK = 16 k-values = 4 DMMA k-steps.
//
// Shared-memory operand layouts (doubles):
//   A, "k-major rows"  As[m][k], pitch PITCH_K (=20)       -- operand whose k index is contiguous in HBM
//   A, "m-major cols"  As[k][m], pitch PA (= rows + 4)     -- operand whose m index is contiguous in HBM
//   B                  Bs[n][k], pitch PB                   -- k contiguous
// With lane = 4*g + t the fragment loads touch, per 16-lane LDS.64 phase, 16 distinct bank pairs:
//   [row][k]: (g*20 + t) mod 16 = 4g + t ;  [k][row]: (t*132 + g) mod 16 = 4t + g.
#pragma once
#include "common.cuh"

namespace qemb {

template <int WM, int WN>
struct WarpAcc {
    double c[WM / 8][WN / 8][2];
    __device__ __forceinline__ void zero()
    {
#pragma unroll
        for (int i = 0; i < WM / 8; ++i)
#pragma unroll
            for (int j = 0; j < WN / 8; ++j) c[i][j][0] = c[i][j][1] = 0.0;
    }
};

// One BK=16 stage.  A_ROWS_KMAJOR selects the A layout; As/Bs already point at the stage.
// wm0/wn0: first row / column of the warp tile inside the CTA tile.  kofs: first k inside Bs/As.
template <int WM, int WN, bool A_ROWS_KMAJOR, int PA, int PB>
__device__ __forceinline__ void warp_mma_stage(WarpAcc<WM, WN> &acc, const double *__restrict__ As,
                                               const double *__restrict__ Bs, int wm0, int wn0, int lane)
{
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int ks = 0; ks < BK / 4; ++ks) {
        double a[WM / 8], b[WN / 8];
#pragma unroll
        for (int i = 0; i < WM / 8; ++i) {
            if (A_ROWS_KMAJOR)
                a[i] = As[(wm0 + 8 * i + g) * PA + 4 * ks + t];
            else
                a[i] = As[(4 * ks + t) * PA + wm0 + 8 * i + g];
        }
#pragma unroll
        for (int j = 0; j < WN / 8; ++j) b[j] = Bs[(wn0 + 8 * j + g) * PB + 4 * ks + t];
#pragma unroll
        for (int i = 0; i < WM / 8; ++i)
#pragma unroll
            for (int j = 0; j < WN / 8; ++j) dmma884(acc.c[i][j][0], acc.c[i][j][1], a[i], b[j]);
    }
}

} // namespace qemb
