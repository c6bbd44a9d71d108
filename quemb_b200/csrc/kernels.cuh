// kernels.cuh -- parameter blocks and host launchers of the sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace qemb {

// ---- K3: G = X^T X (symmetric, lower tiles + mirror) ------------------------------------------
struct GramParams {
    const double *X;   // (ldx x npair) column-major, rows = auxiliary index (zero padded to a multiple of 16)
    int64_t ldx;
    int ktiles;        // number of 16-row k tiles of the slab
    int64_t npair;
    double *G;         // npair x npair, ld = ldg
    int64_t ldg;
    const int2 *tiles; // (ti, tj), ti >= tj, L2-friendly order
    int ntiles;
    double *partial = nullptr; // split-K scratch: gram_splits(...) x npair^2 doubles, or null (no split)
    int accumulate = 0;        // 1: add the lower tiles into G (system-scope atomics; G may be peer memory), no mirror
};
// upper triangle from the lower one, on the device
cudaError_t launch_mirror_lower(double *G, int64_t ld, int64_t n, cudaStream_t s);
// number of k splits launch_gram uses when `partial` is given (1 = no split, no scratch needed)
int gram_splits(int ntiles, int ktiles, int num_sms);
// TMA + mbarrier pipeline, persistent (one CTA per SM, grid = min(ntiles, num_sms))
cudaError_t launch_gram(const GramParams &p, int num_sms, cudaStream_t s);
// host: fills the lower-triangular tile list for nt row tiles; returns count
int64_t build_gram_tiles(int nt, int2 *out /* nt*(nt+1)/2 */);
constexpr int GRAM_TILE = 128;

// ---- K3, opt-in int8-slice variant (k3_int8.cu): same G, tcgen05 int8 tensor cores -------------------------
struct GramI8Params {
    const double *X;   // (ldx x npair) column-major, rows = auxiliary index, `nrows` valid (zero padded) rows
    int64_t ldx;
    int nrows;
    int64_t npair;
    double *G;         // npair x npair, ld = ldg (both triangles written)
    int64_t ldg;
    int slices;        // 6, 7 or 8 (about 3e-12 / 1e-14 / FP64 rounding level of max|G|)
    int accumulate = 0; // 1: the lower triangle is ADDED into G (system-scope red.add.f64; G may be a peer GPU's memory),
                        //    nothing is mirrored -- same contract as GramParams::accumulate of the FP64 kernel
    int8_t *scratch;   // gram_i8_scratch_bytes(...) bytes, 256-byte aligned
};
int gram_i8_naux_pad(int nrows);
size_t gram_i8_scratch_bytes(int nrows, int64_t npair, int slices);
cudaError_t launch_gram_int8(const GramI8Params &p, int num_sms, cudaStream_t s);
// guaranteed-tolerance mode: largest column exponent of X (device int, -2^30 when there is none) and the worst-case
// bound of the S-slice product per unit of 2^(e_i + e_j)
cudaError_t launch_max_expo(const double *X, int64_t ldx, int naux, int64_t npair, int *out_dev, cudaStream_t s);
double gram_i8_bound(int64_t naux, int slices);

// ---- K1 / K2a: grouped gather GEMM --------------------------------------------------------------
//   C[:, ccol[n]] = sum_k  Asrc[:, acol[k]] * TA[brow[k], bcol[n]]        (k < K, n < N of the group)
struct GatherGroup {
    int K, N;
    int a_off; // offset into acol / brow
    int b_off; // offset into bcol / ccol
};
struct GatherParams {
    const double *Asrc;
    int64_t lda;
    int M;     // rows (auxiliary slab) that are computed and stored
    int Mread; // rows that exist below Asrc in every source column (>= M, multiple of 16): tiles are fetched whole
    const int *acol;
    const int *brow;
    const double *TA;
    int64_t ldta;
    const int *bcol;
    const int *ccol;
    double *C;
    int64_t ldc;
    const GatherGroup *groups;
    const int4 *tiles; // (group, m tile, n tile, unused)
    int ntiles;
    int debug; // development only (QEMB_GG_DEBUG): 1 = skip the TA gather, 2 = skip the A copies; results are garbage
};
cudaError_t launch_gather_gemm(const GatherParams &p, int num_sms, cudaStream_t s);
constexpr int GATHER_TILE_M = 128;
constexpr int GATHER_TILE_N = 128;

// ---- K2b: X = L^-1 B, blocked, in place -----------------------------------------------------------
struct TrsmParams {
    const double *L;    // naux_pad x naux_pad column-major lower (pad: identity)
    int64_t ldl;
    const double *Dinv; // nb blocks of 128x128 column-major: inverse of the diagonal blocks
    int nblocks;        // = naux_pad / 128
    double *X;          // ldx x ncol column-major, rows >= nblocks*128 allocated
    int64_t ldx;
    int64_t ncol;
};
cudaError_t launch_trsm(const TrsmParams &p, cudaStream_t s);
// same, but rows >= mrows of X are treated as zero on input and left untouched on output
cudaError_t launch_trsm_rows(const TrsmParams &p, int mrows, cudaStream_t s);
cudaError_t launch_invert_diag_blocks(const double *L, int64_t ldl, int nblocks, double *Dinv, cudaStream_t s);
constexpr int TRSM_BLOCK = 128;
constexpr int TRSM_PANEL = 64;

// ---- J / K from the resident (ij|kl) (consumer of the path; jk.cu) -----------------------------------
struct JkParams {
    const double *G; // npair x npair, symmetric, s4 pair order
    int64_t ldg;
    int64_t npair;
    int nmo;
    const double *D;   // nmo x nmo symmetric density, column-major
    const double *dpk; // packed weights: D[k,l] * (k == l ? 1 : 2) at T(k)+l
    const int *crow;   // chunk table: packed rows [crow[c], crow[c+1]) form chunk c
    int nchunks;
    double *Jp;        // npair x njp partial sums of J
    int njp;           // jk_warps() (jk.cu) or 1 (jk_stream.cu)
    double *R1, *R2;   // npair x nmo (row r at r*nmo)
    // jk_stream.cu: stages per column and capacity of the shared-memory ring in doubles (jk_stream_plan)
    int nstages, ring_doubles;
};
cudaError_t launch_jk(const JkParams &p, double *J, double *K, int64_t ldo, int num_sms, cudaStream_t s);
int jk_max_nmo();
int jk_chunk_doubles();
int jk_warps();
// single-pass streaming kernel (jk_stream.cu), nmo <= jk_stream_max_nmo(), G 16-byte aligned
int jk_stream_max_nmo();
int jk_stream_plan(int nmo, int *ring_doubles, int *smem_bytes); // stages per column, 0 = does not fit
cudaError_t launch_jk_stream(const JkParams &p, int smem, int num_sms, cudaStream_t s);
cudaError_t launch_jk_reduce(const JkParams &p, double *J, double *K, int64_t ldo, cudaStream_t s);

// ---- screening -------------------------------------------------------------------------------------
// mask[mu + nao*i] = |sum_nu S[mu,nu] TA[nu,i]| >= eps  (left-to-right, no FMA contraction)
cudaError_t launch_screen(const double *S, int64_t lds, const double *TA, int64_t ldta, int nao, int nmo,
                          double eps, uint8_t *mask, cudaStream_t s);

// ---- small utilities --------------------------------------------------------------------------------
cudaError_t launch_fill(double *p, int64_t n, double v, cudaStream_t s);
// zero rows [r0, r1) of every column of a column-major matrix
cudaError_t launch_zero_rows(double *p, int64_t ld, int64_t r0, int64_t r1, int64_t ncol, cudaStream_t s);

cudaError_t launch_scatter_block(const double *block, const int *col, int npairs, int naux, double *M, int64_t ld,
                                 cudaStream_t s);
// *out += number of NaN entries in rows [0, nrows) of `ncol` columns (out: zeroed device counter)
cudaError_t launch_count_nan(const double *p, int64_t ld, int64_t nrows, int64_t ncol, unsigned long long *out,
                             cudaStream_t s);
cudaError_t launch_set_diag(double *p, int64_t ld, int64_t i0, int64_t i1, double v, cudaStream_t s);

} // namespace qemb
