// screen.cu -- MO->AO screening and small utility kernels.
//
// get_AO_per_MO (src/quemb/molbe/_cpp/eri_sparse_DF.cpp:443-465): X = |S_abs * TA|, keep X(mu,i) >= eps.
// The sparsity pattern must match the reference's, so the sum over nu runs strictly left to right with
// separately rounded multiply and add (__dmul_rn/__dadd_rn forbid FMA contraction): bit for bit the value the
// plain-C oracle (and any non-fast-math scalar loop) produces.  The reference itself forms S_abs * TA with a blocked,
// FMA, -ffast-math Eigen GEMM, i.e. in another summation order: an entry within rounding distance of eps can land on
// the other side there.  Such entries are counted (bit 1 of the mask, qemb_fragment_borderline) and the driver logs
// them; on every case of the test suite, full-size comparisons with the compiled reference included, the count is 0
// and the lists are identical.  Work is 2*nao^2*nmo flops -- tiny next
// to the contraction kernels -- so one thread per (mu, i) with coalesced S_abs reads is enough.
#include "kernels.cuh"

namespace qemb {

__global__ void __launch_bounds__(128) screen_kernel(const double *__restrict__ S, int64_t lds,
                                                     const double *__restrict__ TA, int64_t ldta, int nao,
                                                     double eps, uint8_t *__restrict__ mask)
{
    const int mu = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (mu >= nao) return;
    const double *ta = TA + (int64_t)i * ldta;
    double s = 0.0;
    for (int nu = 0; nu < nao; ++nu) s = __dadd_rn(s, __dmul_rn(S[mu + (int64_t)nu * lds], ta[nu]));
    // bit 0: kept; bit 1: borderline -- within 1e-12 of the threshold, where another summation order
    // (the reference is built with -ffast-math) may decide differently
    mask[mu + (int64_t)nao * i] = ((fabs(s) >= eps) ? 1 : 0) | ((fabs(fabs(s) - eps) < 1e-12) ? 2 : 0);
}

cudaError_t launch_screen(const double *S, int64_t lds, const double *TA, int64_t ldta, int nao, int nmo,
                          double eps, uint8_t *mask, cudaStream_t s)
{
    if (nao <= 0 || nmo <= 0) return cudaSuccess;
    dim3 grid((nao + 127) / 128, nmo);
    screen_kernel<<<grid, 128, 0, s>>>(S, lds, TA, ldta, nao, eps, mask);
    return cudaGetLastError();
}

__global__ void fill_kernel(double *p, int64_t n, double v)
{
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = v;
}
cudaError_t launch_fill(double *p, int64_t n, double v, cudaStream_t s)
{
    if (n <= 0) return cudaSuccess;
    const int64_t blocks = (n + 255) / 256;
    fill_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, s>>>(p, n, v);
    return cudaGetLastError();
}

__global__ void zero_rows_kernel(double *p, int64_t ld, int64_t r0, int64_t nr, int64_t ncol)
{
    const int64_t total = nr * ncol;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t c = i / nr, r = i - c * nr;
        p[c * ld + r0 + r] = 0.0;
    }
}
cudaError_t launch_zero_rows(double *p, int64_t ld, int64_t r0, int64_t r1, int64_t ncol, cudaStream_t s)
{
    const int64_t nr = r1 - r0;
    if (nr <= 0 || ncol <= 0) return cudaSuccess;
    const int64_t blocks = (nr * ncol + 255) / 256;
    zero_rows_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, s>>>(p, ld, r0, nr, ncol);
    return cudaGetLastError();
}

// (mu nu|P) staging: block[(i * n_j + j) * naux + P]  ->  M[P + ld * col[i * n_j + j]]  for col >= 0
__global__ void scatter_block_kernel(const double *__restrict__ block, const int *__restrict__ col, int npairs,
                                     int naux, double *__restrict__ M, int64_t ld)
{
    for (int pr = blockIdx.x; pr < npairs; pr += gridDim.x) {
        const int c = col[pr];
        if (c < 0) continue;
        const double *src = block + (int64_t)pr * naux;
        double *dst = M + (int64_t)c * ld;
        for (int P = threadIdx.x; P < naux; P += blockDim.x) dst[P] = src[P];
    }
}
cudaError_t launch_scatter_block(const double *block, const int *col, int npairs, int naux, double *M, int64_t ld,
                                 cudaStream_t s)
{
    if (npairs <= 0 || naux <= 0) return cudaSuccess;
    scatter_block_kernel<<<npairs < 148 * 8 ? npairs : 148 * 8, 256, 0, s>>>(block, col, npairs, naux, M, ld);
    return cudaGetLastError();
}

// number of NaN entries in rows [0, nrows) of the columns of a column-major matrix (the reference's
// `assert not np.isnan(result.unique_dense_data).any()` after staging, eri_sparse_DF.py:492)
__global__ void count_nan_kernel(const double *__restrict__ p, int64_t ld, int64_t nrows, int64_t ncol,
                                 unsigned long long *__restrict__ out)
{
    const int64_t total = nrows * ncol;
    unsigned long long n = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t c = i / nrows, r = i - c * nrows;
        const double v = p[c * ld + r];
        n += (v != v);
    }
    for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(out, n);
}
cudaError_t launch_count_nan(const double *p, int64_t ld, int64_t nrows, int64_t ncol, unsigned long long *out,
                             cudaStream_t s)
{
    if (nrows <= 0 || ncol <= 0) return cudaSuccess;
    count_nan_kernel<<<148 * 8, 256, 0, s>>>(p, ld, nrows, ncol, out);
    return cudaGetLastError();
}

__global__ void set_diag_kernel(double *p, int64_t ld, int64_t i0, int64_t i1, double v)
{
    const int64_t i = i0 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < i1) p[i + i * ld] = v;
}
cudaError_t launch_set_diag(double *p, int64_t ld, int64_t i0, int64_t i1, double v, cudaStream_t s)
{
    if (i1 <= i0) return cudaSuccess;
    set_diag_kernel<<<(unsigned)((i1 - i0 + 127) / 128), 128, 0, s>>>(p, ld, i0, i1, v);
    return cudaGetLastError();
}

} // namespace qemb
