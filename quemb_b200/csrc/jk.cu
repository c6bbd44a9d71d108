// jk.cu -- Coulomb / exchange matrices straight from the device-resident (ij|kl) of a fragment
// (the consumer right after the path).
//
// Reference consumers: get_veff (src/quemb/molbe/helper.py:28-69, pyscf scf.hf.dot_eri_dm on the s4 ERIs with
// the Schmidt-space density) in BE.initialize (src/quemb/molbe/mbe.py:1146-1170), and the two-electron part of
// update_ebe_hf (src/quemb/molbe/pfrag.py:327-390), which is sum_j D_ij (2 J_ij - K_ij)/... of the same J and K
// built with the fragment HF density.  Both need the full npair x npair matrix on the HOST today (8 npair^2
// bytes over PCIe + HDF5); with J/K built where G already lives only 2 nmo^2 doubles travel.
//
//   J[ij]  = sum_{kl} (ij|kl) D[k,l]            K[i,k] = sum_{jl} (ij|kl) D[j,l]          (D symmetric)
//
// One pass over G, HBM bound (8 npair^2 bytes read once; everything else is O(npair nmo)).  G is symmetric, so
// "row" ij of the s4 matrix is the contiguous column ij.  A CTA streams a column through shared memory in chunks
// of whole packed rows k (cp.async double buffer) and unpacks it implicitly: with V[k,l] = (ij|kl),
//     v_x[k] = sum_l V[k,l] x[l] = sum_{l<=k} P[T(k)+l] x[l]  +  sum_{m>k} P[T(m)+k] x[m]
//   pass A (first sum):  warp per packed row k, lanes over l, shuffle reduction
//   pass B (second sum): thread per column l of the packed triangle, running over the rows k > l of the chunk;
//                        consecutive threads read consecutive shared-memory words (conflict free)
// for x = D[:,j] (-> contribution of row (i,j) to K[i,:]) and x = D[:,i] (-> K[j,:]), plus J[ij] in pass A.
// No atomics: per-column results go to R1/R2 (npair x nmo) and a second kernel sums them in a fixed order, so
// the result is deterministic.
#include "common.cuh"
#include "kernels.cuh"

namespace qemb {

namespace {
constexpr int JK_THREADS = 512;
constexpr int JK_WARPS = JK_THREADS / 32;
constexpr int JK_LMAX = 2;                   // packed-triangle columns per thread -> nmo <= 1024
constexpr int JK_CHUNK = 12288;              // doubles per pipeline buffer (96 KB); >= nmo
constexpr int JK_NMO_MAX = JK_THREADS * JK_LMAX;

__device__ __forceinline__ int64_t tri(int64_t k) { return k * (k + 1) / 2; }
} // namespace

__global__ void __launch_bounds__(JK_THREADS, 1) jk_rows_kernel(JkParams p)
{
    extern __shared__ __align__(16) double jk_smem[];
    double *buf0 = jk_smem, *buf1 = jk_smem + JK_CHUNK;
    double *xi = jk_smem + 2 * JK_CHUNK; // D[:, i]
    double *xj = xi + JK_NMO_MAX;        // D[:, j]
    double *vai = xj + JK_NMO_MAX;       // pass A results
    double *vaj = vai + JK_NMO_MAX;
    __shared__ double jred[JK_WARPS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nmo = p.nmo, nch = p.nchunks;

    // chunk c of column r: packed rows [crow[c], crow[c+1])
    auto prefetch = [&](double *dst, int64_t r, int c) {
        const int k0 = p.crow[c], k1 = p.crow[c + 1];
        const int64_t o0 = tri(k0);
        const int n = (int)(tri(k1) - o0);
        const double *src = p.G + p.ldg * r + o0;
        for (int e = tid; e < n; e += JK_THREADS) cp_async8(dst + e, src + e, true);
    };

    int64_t r = blockIdx.x;
    if (r >= p.npair) return;
    int c = 0, n = 0;
    prefetch(buf0, r, 0);
    cp_async_commit();

    double bi[JK_LMAX], bj[JK_LMAX], jw = 0.0;
    int i = 0, j = 0;
    while (r < p.npair) {
        if (c == 0) { // new column: (i, j) with i >= j, density columns, accumulators
            i = (int)((sqrt(8.0 * (double)r + 1.0) - 1.0) * 0.5);
            while (tri(i + 1) <= r) ++i;
            while (tri(i) > r) --i;
            j = (int)(r - tri(i));
            for (int l = tid; l < nmo; l += JK_THREADS) {
                xi[l] = p.D[l + (int64_t)nmo * i];
                xj[l] = p.D[l + (int64_t)nmo * j];
            }
#pragma unroll
            for (int q = 0; q < JK_LMAX; ++q) bi[q] = bj[q] = 0.0;
            jw = 0.0;
        }
        // next item of this CTA
        int64_t rn = r;
        int cn = c + 1;
        if (cn == nch) {
            cn = 0;
            rn = r + gridDim.x;
        }
        double *cur = (n & 1) ? buf1 : buf0;
        if (rn < p.npair) {
            prefetch((n & 1) ? buf0 : buf1, rn, cn);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads(); // chunk landed; xi / xj visible

        const int k0 = p.crow[c], k1 = p.crow[c + 1];
        const int64_t o0 = tri(k0);
        // pass A: v[k] += sum_{l <= k} P[T(k)+l] x[l];  J += sum P[T(k)+l] dpk[T(k)+l]
        for (int k = k0 + warp; k < k1; k += JK_WARPS) {
            const double *row = cur + (tri(k) - o0);
            const double *dk = p.dpk + tri(k);
            double ai = 0.0, aj = 0.0;
            for (int l = lane; l <= k; l += 32) {
                const double g = row[l];
                ai = fma(g, xi[l], ai);
                aj = fma(g, xj[l], aj);
                jw = fma(g, __ldg(dk + l), jw);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                ai += __shfl_xor_sync(0xffffffffu, ai, o);
                aj += __shfl_xor_sync(0xffffffffu, aj, o);
            }
            if (lane == 0) {
                vai[k] = ai;
                vaj[k] = aj;
            }
        }
        // pass B: v[l] += sum_{k > l} P[T(k)+l] x[k]
#pragma unroll
        for (int q = 0; q < JK_LMAX; ++q) {
            const int l = tid + JK_THREADS * q;
            if (l < k1 - 1 && l < nmo) {
                double si = bi[q], sj = bj[q];
                int k = (l + 1 > k0 ? l + 1 : k0);
                int off = (int)(tri(k) - o0) + l;
#pragma unroll 4
                for (; k < k1; ++k) {
                    const double g = cur[off];
                    si = fma(g, xi[k], si);
                    sj = fma(g, xj[k], sj);
                    off += k + 1;
                }
                bi[q] = si;
                bj[q] = sj;
            }
        }
        __syncthreads(); // buffer may be refilled; pass A results visible

        if (c == nch - 1) { // column finished: R1[r,:] = V D[:,j],  R2[r,:] = V D[:,i],  Jp[r]
#pragma unroll
            for (int q = 0; q < JK_LMAX; ++q) {
                const int l = tid + JK_THREADS * q;
                if (l < nmo) {
                    p.R1[r * nmo + l] = vaj[l] + bj[q];
                    p.R2[r * nmo + l] = vai[l] + bi[q];
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) jw += __shfl_xor_sync(0xffffffffu, jw, o);
            if (lane == 0) jred[warp] = jw;
            __syncthreads();
            if (tid == 0) {
                double s = 0.0;
                for (int w = 0; w < JK_WARPS; ++w) s += jred[w];
                p.Jp[r] = s;
            }
            __syncthreads(); // xi / xj / va* / jred are rewritten by the next column
        }
        r = rn;
        c = cn;
        ++n;
    }
}

// K[a,k] = sum_{j<=a} R1[(a,j)][k] + sum_{i>a} R2[(i,a)][k]   (fixed order);   J[a,k] = Jp[T(max)+min]
__global__ void jk_reduce_kernel(JkParams p, double *__restrict__ J, double *__restrict__ K, int64_t ldo)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    const int a = blockIdx.y;
    const int nmo = p.nmo;
    if (k >= nmo) return;
    double s = 0.0;
    for (int j = 0; j <= a; ++j) s += p.R1[(tri(a) + j) * nmo + k];
    for (int i = a + 1; i < nmo; ++i) s += p.R2[(tri(i) + a) * nmo + k];
    K[a + ldo * k] = s;
    const int hi = a > k ? a : k, lo = a > k ? k : a;
    J[a + ldo * k] = p.Jp[tri(hi) + lo];
}

int jk_max_nmo() { return JK_NMO_MAX; }
int jk_chunk_doubles() { return JK_CHUNK; }

cudaError_t launch_jk(const JkParams &p, double *J, double *K, int64_t ldo, int num_sms, cudaStream_t s)
{
    if (p.nmo <= 0) return cudaSuccess;
    if (p.nmo > JK_NMO_MAX) return cudaErrorInvalidValue;
    const int smem = (2 * JK_CHUNK + 4 * JK_NMO_MAX) * 8;
    cudaError_t e = cudaFuncSetAttribute(jk_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    const int grid = p.npair < num_sms ? (int)p.npair : num_sms;
    jk_rows_kernel<<<grid, JK_THREADS, smem, s>>>(p);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    dim3 g2((p.nmo + 127) / 128, p.nmo);
    jk_reduce_kernel<<<g2, 128, 0, s>>>(p, J, K, ldo);
    return cudaGetLastError();
}

} // namespace qemb
