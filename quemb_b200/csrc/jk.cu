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
//   pass A (first sum):  a warp takes two packed rows, lanes over l, one transposed butterfly for the four sums
//   pass B (second sum): thread per column l of the packed triangle, running over the rows k > l of the chunk;
//                        consecutive threads read consecutive shared-memory words (conflict free)
// for x = D[:,j] (-> contribution of row (i,j) to K[i,:]) and x = D[:,i] (-> K[j,:]), plus J[ij] in pass B (the
// weights dpk are read coalesced, independent of the accumulators, so their latency overlaps).
// No atomics: per-column results go to R1/R2 (npair x nmo) and a second kernel sums them in a fixed order, so
// the result is deterministic.
#include "common.cuh"
#include "kernels.cuh"

namespace qemb {

namespace {
constexpr int JK_MAXWARPS = 16;
constexpr int JK_CHUNK = 12288;     // doubles per pipeline buffer (96 KB); >= nmo
constexpr int JK_BUF = JK_CHUNK + 2; // + slack for the alignment shift of a column that starts on an odd double
constexpr int JK_NMO_MAX = 768;

__device__ __forceinline__ int64_t tri(int64_t k) { return k * (k + 1) / 2; }

// sums v0..v3 over the warp with one transposed butterfly (6 shuffles): afterwards lanes 0 / 8 / 16 / 24 (and the
// lanes that share their bits 3,4) hold the totals of v0 / v1 / v2 / v3
__device__ __forceinline__ double reduce4(double v0, double v1, double v2, double v3, int lane)
{
    const bool h16 = lane & 16, h8 = lane & 8;
    double ka = h16 ? v2 : v0, kb = h16 ? v3 : v1;
    ka += __shfl_xor_sync(0xffffffffu, h16 ? v0 : v2, 16);
    kb += __shfl_xor_sync(0xffffffffu, h16 ? v1 : v3, 16);
    double kk = h8 ? kb : ka;
    kk += __shfl_xor_sync(0xffffffffu, h8 ? ka : kb, 8);
    kk += __shfl_xor_sync(0xffffffffu, kk, 4);
    kk += __shfl_xor_sync(0xffffffffu, kk, 2);
    kk += __shfl_xor_sync(0xffffffffu, kk, 1);
    return kk;
}
} // namespace

// One packed row k of the column in shared memory, handled by one warp: lane owns the columns l = lane + 32 m.
//   A term: a_i/a_j += P[T(k)+l] x[l]          (reduced over the lanes at the end of the row pair)
//   B term: vb[m]   += P[T(k)+l] x[k],  l < k  (stays in the lane's registers for the whole column of G)
//   J     : jw      += P[T(k)+l] dpk[T(k)+l]
template <int MMAX>
__device__ __forceinline__ void jk_row(const double *__restrict__ row, const double *__restrict__ dk, int k, int lane,
                                       const double2 *__restrict__ x2, double (&vbi)[MMAX], double (&vbj)[MMAX],
                                       double &ai, double &aj, double &jw)
{
    const double2 xk = x2[k];
#pragma unroll
    for (int m = 0; m < MMAX; ++m) {
        if (32 * m > k) break; // warp uniform
        const int l = lane + 32 * m;
        if (l <= k) {
            const double g = row[l];
            const double2 x = x2[l];
            ai = fma(g, x.x, ai);
            aj = fma(g, x.y, aj);
            jw = fma(g, __ldg(dk + l), jw);
            if (l < k) {
                vbi[m] = fma(g, xk.x, vbi[m]);
                vbj[m] = fma(g, xk.y, vbj[m]);
            }
        }
    }
}

template <int MMAX, int NTHREADS>
__global__ void __launch_bounds__(NTHREADS, 1) jk_rows_kernel(JkParams p)
{
    constexpr int NWARPS = NTHREADS / 32;
    constexpr int NP = 32 * MMAX; // padded nmo of the cross-warp reduction
    static_assert(NWARPS * NP <= JK_CHUNK, "cross-warp reduction must fit a chunk buffer");
    extern __shared__ __align__(16) double jk_smem[];
    double *buf0 = jk_smem, *buf1 = jk_smem + JK_BUF;
    double2 *x2 = reinterpret_cast<double2 *>(jk_smem + 2 * JK_BUF); // (D[l,i], D[l,j])
    double2 *va2 = x2 + JK_NMO_MAX;                                  // A-term results per packed row

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nmo = p.nmo, nch = p.nchunks;

    // chunk c of column r: packed rows [crow[c], crow[c+1]); stored at dst + shift so that 16-byte copies line up
    auto chunk_src = [&](int64_t r, int c) { return p.G + p.ldg * r + tri(p.crow[c]); };
    auto prefetch = [&](double *dst, int64_t r, int c) {
        const int k0 = p.crow[c], k1 = p.crow[c + 1];
        const int n = (int)(tri(k1) - tri(k0));
        const double *src = chunk_src(r, c);
        const int sh = (int)((reinterpret_cast<uintptr_t>(src) >> 3) & 1);
        double *d = dst + sh;
        if (tid == 0 && sh) cp_async8(d, src, true);
        const int np2 = (n - sh) >> 1;
        for (int q = tid; q < np2; q += NTHREADS) cp_async16(d + sh + 2 * q, src + sh + 2 * q, true);
        if (tid == 0 && ((n - sh) & 1)) cp_async8(d + n - 1, src + n - 1, true);
    };

    int64_t r = blockIdx.x;
    if (r >= p.npair) return;
    int c = 0, n = 0;
    prefetch(buf0, r, 0);
    cp_async_commit();

    double vbi[MMAX], vbj[MMAX], jw = 0.0;
    while (r < p.npair) {
        if (c == 0) { // new column: (i, j) with i >= j, density columns, accumulators
            int i = (int)((sqrt(8.0 * (double)r + 1.0) - 1.0) * 0.5);
            while (tri(i + 1) <= r) ++i;
            while (tri(i) > r) --i;
            const int j = (int)(r - tri(i));
            for (int l = tid; l < nmo; l += NTHREADS)
                x2[l] = make_double2(p.D[l + (int64_t)nmo * i], p.D[l + (int64_t)nmo * j]);
#pragma unroll
            for (int m = 0; m < MMAX; ++m) vbi[m] = vbj[m] = 0.0;
            jw = 0.0;
        }
        int64_t rn = r;
        int cn = c + 1;
        if (cn == nch) {
            cn = 0;
            rn = r + gridDim.x;
        }
        double *curbuf = (n & 1) ? buf1 : buf0;
        if (rn < p.npair) {
            prefetch((n & 1) ? buf0 : buf1, rn, cn);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads(); // chunk landed; x2 visible

        const int k0 = p.crow[c], k1 = p.crow[c + 1];
        const int64_t o0 = tri(k0);
        const double *cur = curbuf + ((reinterpret_cast<uintptr_t>(chunk_src(r, c)) >> 3) & 1);
        // rows of the chunk, two per warp and step (k, k + NWARPS): one transposed reduction for the four A sums
        for (int k = k0 + warp; k < k1; k += 2 * NWARPS) {
            const int kb = k + NWARPS;
            double a0i = 0.0, a0j = 0.0, a1i = 0.0, a1j = 0.0;
            jk_row<MMAX>(cur + (tri(k) - o0), p.dpk + tri(k), k, lane, x2, vbi, vbj, a0i, a0j, jw);
            if (kb < k1) jk_row<MMAX>(cur + (tri(kb) - o0), p.dpk + tri(kb), kb, lane, x2, vbi, vbj, a1i, a1j, jw);
            const double t = reduce4(a0i, a0j, a1i, a1j, lane);
            if ((lane & 7) == 0) {
                const bool h16 = lane & 16, h8 = lane & 8;
                if (!h16 || kb < k1) reinterpret_cast<double *>(va2 + (h16 ? kb : k))[h8 ? 1 : 0] = t;
            }
        }
        __syncthreads(); // the chunk buffer is free; A-term results visible

        if (c == nch - 1) {
            // column finished.  v_x[l] = va[l] + sum over the warps' private B-term accumulators, in a fixed order,
            // through the chunk buffer that was just consumed (the prefetch in flight targets the other one)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) jw += __shfl_xor_sync(0xffffffffu, jw, o);
            if (lane == 0) {
                p.Jp[r * JK_MAXWARPS + warp] = jw;
                if (NWARPS < JK_MAXWARPS) p.Jp[r * JK_MAXWARPS + warp + NWARPS] = 0.0;
            }
            double *red = curbuf;
#pragma unroll
            for (int pass = 0; pass < 2; ++pass) { // 0: x = D[:,i] -> R2,  1: x = D[:,j] -> R1
#pragma unroll
                for (int m = 0; m < MMAX; ++m) red[warp * NP + lane + 32 * m] = pass ? vbj[m] : vbi[m];
                __syncthreads();
                for (int l = tid; l < nmo; l += NTHREADS) {
                    double sacc = pass ? va2[l].y : va2[l].x;
#pragma unroll
                    for (int w = 0; w < NWARPS; ++w) sacc += red[w * NP + l];
                    (pass ? p.R1 : p.R2)[r * nmo + l] = sacc;
                }
                __syncthreads();
            }
        }
        r = rn;
        c = cn;
        ++n;
    }
}

// K[a,k] = sum_{j<=a} R1[(a,j)][k] + sum_{i>a} R2[(i,a)][k]   (fixed order);   J[a,k] = sum_w Jp[T(max)+min][w]
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
    const double *jp = p.Jp + (tri(hi) + lo) * JK_MAXWARPS;
    double t = 0.0;
    for (int w = 0; w < JK_MAXWARPS; ++w) t += jp[w];
    J[a + ldo * k] = t;
}

int jk_warps() { return JK_MAXWARPS; }
int jk_max_nmo() { return JK_NMO_MAX; }
int jk_chunk_doubles() { return JK_CHUNK; }

namespace {
template <int MMAX, int NTHREADS>
cudaError_t launch_rows(const JkParams &p, int num_sms, cudaStream_t s)
{
    const int smem = 2 * JK_BUF * 8 + 2 * JK_NMO_MAX * 16;
    cudaError_t e = cudaFuncSetAttribute(jk_rows_kernel<MMAX, NTHREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    const int grid = p.npair < num_sms ? (int)p.npair : num_sms;
    jk_rows_kernel<MMAX, NTHREADS><<<grid, NTHREADS, smem, s>>>(p);
    return cudaGetLastError();
}
} // namespace

cudaError_t launch_jk(const JkParams &p, double *J, double *K, int64_t ldo, int num_sms, cudaStream_t s)
{
    if (p.nmo <= 0) return cudaSuccess;
    if (p.nmo > JK_NMO_MAX) return cudaErrorInvalidValue;
    cudaError_t e;
    if (p.nmo <= 256)
        e = launch_rows<8, 512>(p, num_sms, s);
    else if (p.nmo <= 512)
        e = launch_rows<16, 512>(p, num_sms, s);
    else
        e = launch_rows<24, 256>(p, num_sms, s);
    if (e != cudaSuccess) return e;
    dim3 g2((p.nmo + 127) / 128, p.nmo);
    jk_reduce_kernel<<<g2, 128, 0, s>>>(p, J, K, ldo);
    return cudaGetLastError();
}

} // namespace qemb
