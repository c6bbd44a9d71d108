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
//   A role (first sum,  warps 0..7):  one THREAD per packed row k, running over l <= k.  T(k) mod 16 is a permutation
//                         over any 16 consecutive k, so the 16 lanes of an LDS.64 phase hit 16 distinct banks; the
//                         sums stay in the thread, no reduction
//   B role (second sum, warps 8..15): one thread per column l of the packed triangle, running over the rows k > l of
//                         the chunk; consecutive threads read consecutive words
//   J[ij] = P . dpk:      element-wise over the chunk by all threads (coalesced weights, independent loads)
// for x = D[:,j] (-> contribution of row (i,j) to K[i,:]) and x = D[:,i] (-> K[j,:]).  Both roles work on the same
// chunk at the same time and have the same amount of work (the two halves of the symmetric matrix).
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

} // namespace

template <int QMAX>
__global__ void __launch_bounds__(512, 1) jk_rows_kernel(JkParams p)
{
    constexpr int NTHREADS = 512, GA = 256, GB = 256; // A role: threads 0..255, B role: threads 256..511
    extern __shared__ __align__(16) double jk_smem[];
    double *buf0 = jk_smem, *buf1 = jk_smem + JK_BUF;
    double2 *x2 = reinterpret_cast<double2 *>(jk_smem + 2 * JK_BUF); // (D[l,i], D[l,j])
    double2 *va2 = x2 + JK_NMO_MAX;                                  // A-role results per packed row

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool roleB = tid >= GA;
    const int tb = tid - GA;
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

    double bi[QMAX], bj[QMAX], jw = 0.0;
    while (r < p.npair) {
        if (c == 0) { // new column: (i, j) with i >= j, density columns, accumulators
            int i = (int)((sqrt(8.0 * (double)r + 1.0) - 1.0) * 0.5);
            while (tri(i + 1) <= r) ++i;
            while (tri(i) > r) --i;
            const int j = (int)(r - tri(i));
            for (int l = tid; l < nmo; l += NTHREADS)
                x2[l] = make_double2(p.D[l + (int64_t)nmo * i], p.D[l + (int64_t)nmo * j]);
#pragma unroll
            for (int q = 0; q < QMAX; ++q) bi[q] = bj[q] = 0.0;
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
        const int o0 = (int)tri(k0);
        const double *cur = curbuf + ((reinterpret_cast<uintptr_t>(chunk_src(r, c)) >> 3) & 1);
        // J += P . dpk over the chunk, element-wise by all threads: coalesced weights, independent loads (a weight
        // load inside the B loop put an L2 latency on the critical path of every row)
        {
            const double *dk = p.dpk + o0;
            const int ne = (int)tri(k1) - o0;
            double j0 = 0.0, j1 = 0.0, j2 = 0.0, j3 = 0.0;
            int e = tid;
            for (; e + 3 * NTHREADS < ne; e += 4 * NTHREADS) {
                const double w0 = __ldg(dk + e), w1 = __ldg(dk + e + NTHREADS), w2 = __ldg(dk + e + 2 * NTHREADS),
                             w3 = __ldg(dk + e + 3 * NTHREADS);
                j0 = fma(cur[e], w0, j0);
                j1 = fma(cur[e + NTHREADS], w1, j1);
                j2 = fma(cur[e + 2 * NTHREADS], w2, j2);
                j3 = fma(cur[e + 3 * NTHREADS], w3, j3);
            }
            for (; e < ne; e += NTHREADS) j0 = fma(cur[e], __ldg(dk + e), j0);
            jw += (j0 + j1) + (j2 + j3);
        }
        if (!roleB) {
            // A role: va[k] = sum_{l <= k} P[T(k)+l] x[l]
            for (int k = k0 + tid; k < k1; k += GA) {
                const double *row = cur + ((k * (k + 1)) / 2 - o0);
                double ai = 0.0, aj = 0.0, ci = 0.0, cj = 0.0; // two dependency chains per sum
                int l = 0;
#pragma unroll 4
                for (; l + 1 <= k; l += 2) {
                    const double g0 = row[l], g1 = row[l + 1];
                    const double2 x0 = x2[l], x1 = x2[l + 1];
                    ai = fma(g0, x0.x, ai);
                    aj = fma(g0, x0.y, aj);
                    ci = fma(g1, x1.x, ci);
                    cj = fma(g1, x1.y, cj);
                }
                if (l <= k) {
                    const double g0 = row[l];
                    const double2 x0 = x2[l];
                    ai = fma(g0, x0.x, ai);
                    aj = fma(g0, x0.y, aj);
                }
                va2[k] = make_double2(ai + ci, aj + cj);
            }
        } else {
            // B role: vb[l] += sum_{k > l} P[T(k)+l] x[k]
#pragma unroll
            for (int q = 0; q < QMAX; ++q) {
                const int l = tb + GB * q;
                if (l < k1 && l < nmo) {
                    double si = bi[q], sj = bj[q], ti = 0.0, tj = 0.0; // two dependency chains per sum
                    int k = (l + 1 > k0 ? l + 1 : k0);
                    int off = (k * (k + 1)) / 2 - o0 + l;
#pragma unroll 4
                    for (; k + 1 < k1; k += 2) {
                        const double g0 = cur[off], g1 = cur[off + k + 1];
                        const double2 x0 = x2[k], x1 = x2[k + 1];
                        si = fma(g0, x0.x, si);
                        sj = fma(g0, x0.y, sj);
                        ti = fma(g1, x1.x, ti);
                        tj = fma(g1, x1.y, tj);
                        off += 2 * k + 3;
                    }
                    if (k < k1) {
                        const double g0 = cur[off];
                        const double2 x0 = x2[k];
                        si = fma(g0, x0.x, si);
                        sj = fma(g0, x0.y, sj);
                    }
                    bi[q] = si + ti;
                    bj[q] = sj + tj;
                }
            }
        }
        __syncthreads(); // the chunk buffer is free; A-role results visible

        if (c == nch - 1) {
            // column finished: the B-role thread that owns column l adds the A-role result of row l
            if (roleB) {
#pragma unroll
                for (int q = 0; q < QMAX; ++q) {
                    const int l = tb + GB * q;
                    if (l < nmo) {
                        const double2 va = va2[l];
                        p.R1[r * nmo + l] = va.y + bj[q];
                        p.R2[r * nmo + l] = va.x + bi[q];
                    }
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) jw += __shfl_xor_sync(0xffffffffu, jw, o);
            if (lane == 0) p.Jp[r * p.njp + warp] = jw;
            // va2 / x2 are rewritten only behind the next __syncthreads (va2) or at the top of the next column (x2, read
            // for the last time before the trailing barrier above)
        }
        r = rn;
        c = cn;
        ++n;
    }
}

// K[a,k] = sum_{j<=a} R1[(a,j)][k] + sum_{i>a} R2[(i,a)][k];   J[a,k] = sum_w Jp[T(max)+min][w]
// The nmo terms of K[a,k] are dealt to JR_PARTS threads (term t -> part t mod JR_PARTS), each sums its terms in
// order, the parts are added in order: fixed summation order, JR_PARTS times the loads in flight.
constexpr int JR_KX = 32, JR_PARTS = 8;
__global__ void __launch_bounds__(JR_KX *JR_PARTS) jk_reduce_kernel(JkParams p, double *__restrict__ J, double *__restrict__ K, int64_t ldo)
{
    __shared__ double part_s[JR_PARTS][JR_KX];
    const int kx = threadIdx.x % JR_KX, part = threadIdx.x / JR_KX;
    const int k = blockIdx.x * JR_KX + kx;
    const int a = blockIdx.y;
    const int nmo = p.nmo;
    double s = 0.0;
    if (k < nmo) {
        for (int t = part; t < nmo; t += JR_PARTS)
            s += t <= a ? p.R1[(tri(a) + t) * nmo + k] : p.R2[(tri(t) + a) * nmo + k];
    }
    part_s[part][kx] = s;
    __syncthreads();
    if (part == 0 && k < nmo) {
        double t = part_s[0][kx];
#pragma unroll
        for (int q = 1; q < JR_PARTS; ++q) t += part_s[q][kx];
        K[a + ldo * k] = t;
        const int hi = a > k ? a : k, lo = a > k ? k : a;
        const double *jp = p.Jp + (tri(hi) + lo) * p.njp;
        double u = 0.0;
        for (int w = 0; w < p.njp; ++w) u += jp[w];
        J[a + ldo * k] = u;
    }
}

int jk_warps() { return JK_MAXWARPS; }
int jk_max_nmo() { return JK_NMO_MAX; }
int jk_chunk_doubles() { return JK_CHUNK; }

namespace {
template <int QMAX>
cudaError_t launch_rows(const JkParams &p, int num_sms, cudaStream_t s)
{
    const int smem = 2 * JK_BUF * 8 + 2 * JK_NMO_MAX * 16;
    cudaError_t e = cudaFuncSetAttribute(jk_rows_kernel<QMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    const int grid = p.npair < num_sms ? (int)p.npair : num_sms;
    jk_rows_kernel<QMAX><<<grid, 512, smem, s>>>(p);
    return cudaGetLastError();
}
} // namespace

cudaError_t launch_jk(const JkParams &p, double *J, double *K, int64_t ldo, int num_sms, cudaStream_t s)
{
    if (p.nmo <= 0) return cudaSuccess;
    if (p.nmo > JK_NMO_MAX) return cudaErrorInvalidValue;
    cudaError_t e;
    if (p.nmo <= 256)
        e = launch_rows<1>(p, num_sms, s);
    else if (p.nmo <= 512)
        e = launch_rows<2>(p, num_sms, s);
    else
        e = launch_rows<3>(p, num_sms, s);
    if (e != cudaSuccess) return e;
    return launch_jk_reduce(p, J, K, ldo, s);
}

cudaError_t launch_jk_reduce(const JkParams &p, double *J, double *K, int64_t ldo, cudaStream_t s)
{
    dim3 g2((p.nmo + JR_KX - 1) / JR_KX, p.nmo);
    jk_reduce_kernel<<<g2, JR_KX * JR_PARTS, 0, s>>>(p, J, K, ldo);
    return cudaGetLastError();
}

} // namespace qemb
