// jk_stream.cu -- J/K from the device-resident (ij|kl), single-pass streaming kernel (nmo <= 256).
//
// Same contraction and outputs as jk.cu (reference consumers: get_veff, src/quemb/molbe/helper.py:28-69, and the
// two-electron part of update_ebe_hf, src/quemb/molbe/pfrag.py:327-390):
//   J[ij] = sum_{kl} (ij|kl) D[k,l]        K[i,k] = sum_{jl} (ij|kl) D[j,l]
// HBM bound: column r = (i,j) of G is read once (8 npair bytes) and every element P[T(k)+l] of its packed
// triangle is used for five FMAs from ONE shared-memory read:
//   pa[k]  += g x[l]   (x = D[:,i] and D[:,j]: row sums, reduced over the lanes once per row pair)
//   vb[l]  += g x[k]   (column sums, kept in the lane's registers for the whole column, l = lane + 32 q)
//   J      += g w[k,l] (packed weights D[k,l] (2 - delta_kl), coalesced loads issued one row pair ahead)
// jk.cu reads every element three times from shared memory, in three differently shaped loops, behind two CTA
// barriers per chunk (27 % of the HBM peak, 38 instructions per 32 elements); here a warp owns whole rows (lanes
// along l, conflict free), one row pair costs ~12 instructions per 64 elements plus one 4-value butterfly, and
// there is no CTA barrier inside a column:
//   * producer warp: one 1-D bulk copy (cp.async.bulk, SASS UBLKCP) per stage = a block of whole packed rows,
//     completion counted on the stage's "full" mbarrier; it runs ahead over column boundaries by the ring depth
//   * 8 consumer warps: wait "full", process their row pairs of the stage, release the PREVIOUS stage after the
//     next wait (same deferred release as k3_gram.cu: an arrive placed directly behind the loads is scheduled
//     ahead of the FMAs that consume them)
//   * per column two named barriers among the consumers: column sums of the 8 warps meet in shared memory and
//     are added in a fixed order (deterministic), next column's density columns are staged meanwhile
#include "common.cuh"
#include "kernels.cuh"
#include <type_traits>

namespace qemb {

namespace {
template <int I, int N, class F>
__device__ __forceinline__ void static_for(F &&f)
{
    if constexpr (I < N) {
        f(std::integral_constant<int, I>{});
        static_for<I + 1, N>(f);
    }
}
constexpr int JS_NW = 8;                      // consumer warps
constexpr int JS_THREADS = 32 * (JS_NW + 1);  // + producer warp
constexpr int JS_MAX_RING = 6;
constexpr int JS_MAX_STAGES = 40;             // stages per column (<= ceil(256 / 8))

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
__device__ __forceinline__ void consumer_barrier()
{
    asm volatile("bar.sync 1, %0;" ::"n"(32 * JS_NW) : "memory");
}
__device__ __forceinline__ void pair_of(int64_t r, int &i, int &j)
{
    int a = (int)((sqrt(8.0 * (double)r + 1.0) - 1.0) * 0.5);
    while ((int64_t)a * (a + 1) / 2 + a + 1 <= r) ++a; // tri(a + 1) <= r
    while ((int64_t)a * (a + 1) / 2 > r) --a;
    i = a;
    j = (int)(r - (int64_t)a * (a + 1) / 2);
}
} // namespace

// shared memory: ring * stage_doubles doubles | slab[JS_NW][nmoP] double2 | x2[2][nmoP] double2 | va2[nmoP] double2 |
//                jws[JS_NW] | srow[JS_MAX_STAGES + 1] | full[ring], empty[ring]
template <int QL>
__global__ void __launch_bounds__(JS_THREADS, 1) jk_stream_kernel(JkParams p)
{
    constexpr int NMOP = 32 * QL;
    extern __shared__ __align__(16) double js_smem[];
    double *ringbuf = js_smem;
    double2 *slab = reinterpret_cast<double2 *>(ringbuf + (size_t)p.ring * p.stage_doubles);
    double2 *x2 = slab + JS_NW * NMOP;
    double2 *va2 = x2 + 2 * NMOP;
    double *jws = reinterpret_cast<double *>(va2 + NMOP);
    int *srow = reinterpret_cast<int *>(jws + JS_NW);
    uint64_t *full = reinterpret_cast<uint64_t *>(srow + JS_MAX_STAGES + 2);
    uint64_t *empty = full + JS_MAX_RING;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nmo = p.nmo, nst = p.nstages, ring = p.ring;

    for (int q = tid; q <= nst; q += JS_THREADS) srow[q] = p.srow[q]; // producer's stage table
    if (tid == 0) {
        for (int s = 0; s < ring; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], JS_NW);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == JS_NW) {
        // ---------------------------------------------------------------- producer
        if (lane != 0) return;
        const double *Gend = p.G + p.ldg * (p.npair - 1) + p.npair;
        int s = 0, phase = 0;
        for (int64_t r = blockIdx.x; r < p.npair; r += gridDim.x) {
            const double *col = p.G + p.ldg * r;
            for (int st = 0; st < nst; ++st) {
                mbar_wait(&empty[s], phase ^ 1);
                const int k0 = srow[st], k1 = srow[st + 1];
                const int n = tri32(k1) - tri32(k0);
                const double *src = col + tri32(k0);
                const int sh = (int)((reinterpret_cast<uintptr_t>(src) >> 3) & 1);
                const double *srcA = src - sh; // 16-byte aligned
                int cnt = (n + sh + 1) & ~1;
                double *dst = ringbuf + (size_t)s * p.stage_doubles;
                if (srcA + cnt > Gend) { // the pad element would lie behind the matrix: copy the last one by hand
                    cnt -= 2;
                    dst[cnt] = srcA[cnt];
                }
                mbar_arrive_expect_tx(&full[s], cnt * 8);
                if (cnt > 0) bulk_load(dst, srcA, cnt * 8, &full[s]);
                if (++s == ring) {
                    s = 0;
                    phase ^= 1;
                }
            }
        }
        return;
    }

    // -------------------------------------------------------------------- consumers
    int64_t r = blockIdx.x;
    int par = 0;
    {
        int i, j;
        pair_of(r, i, j);
        if (tid < NMOP)
            x2[tid] = tid < nmo ? make_double2(p.D[tid + (int64_t)nmo * i], p.D[tid + (int64_t)nmo * j]) : make_double2(0.0, 0.0);
    }
    consumer_barrier();

    int s = 0, phase = 0, held = -1;
    while (r < p.npair) {
        const double2 *xc = x2 + par * NMOP;
        double xi[QL], xj[QL], bi[QL], bj[QL];
#pragma unroll
        for (int q = 0; q < QL; ++q) {
            const double2 v = xc[lane + 32 * q];
            xi[q] = v.x;
            xj[q] = v.y;
            bi[q] = bj[q] = 0.0;
        }
        double jw = 0.0;
        const int colpar = (int)((p.ldg * r) & 1); // G is 16-byte aligned (checked by the launcher)

        // Stage ST holds the packed rows [32 ST, min(32 ST + 32, nmo)): every row of it has ST blocks of 32 columns
        // strictly left of the diagonal and ONE block that contains it, so the block structure is a compile-time
        // property of the stage and the loop below unrolls without a data-dependent branch per block.
        // Row pairs of this warp in a stage: A = (32 ST + 2 warp, +1), B = A + 16.
        // weights of a row pair: w[q] = dpk[T(k) + lane + 32 q] for lane + 32 q <= k, else 0
        auto load_w = [&](auto nf_c, auto full_c, int ka, bool hasb_, double (&wa)[QL], double (&wb)[QL]) {
            constexpr int NF = decltype(nf_c)::value;
            const bool hasb = decltype(full_c)::value ? true : hasb_; // every stage but the last holds 32 rows
            const double *da = p.dpk + tri32(ka) + lane;
            const double *db = da + ka + 1;
            const int l = lane + 32 * NF;
#pragma unroll
            for (int q = 0; q < NF; ++q) wa[q] = __ldg(da + 32 * q);
            wa[NF] = (l <= ka) ? __ldg(da + 32 * NF) : 0.0;
            if (hasb) {
#pragma unroll
                for (int q = 0; q < NF; ++q) wb[q] = __ldg(db + 32 * q);
                wb[NF] = (l <= ka + 1) ? __ldg(db + 32 * NF) : 0.0;
            } else {
#pragma unroll
                for (int q = 0; q <= NF; ++q) wb[q] = 0.0;
            }
        };
        // one row pair (ka, ka + 1) of a stage with NF full blocks; `base` + T(k) + l addresses P[T(k) + l]
        auto pair = [&](auto nf_c, auto full_c, int ka, bool hasb_, const double *base, const double (&wa)[QL],
                        const double (&wb)[QL]) {
            constexpr int NF = decltype(nf_c)::value;
            const bool hasb = decltype(full_c)::value ? true : hasb_;
            const double *rowa = base + tri32(ka) + lane;
            const double *rowb = hasb ? rowa + ka + 1 : rowa; // a missing row b re-reads row a against zero weights
            const int kb = hasb ? ka + 1 : ka;
            const double2 xka = xc[ka];
            const double2 xkb = hasb ? xc[kb] : make_double2(0.0, 0.0);
            double pia = 0.0, pja = 0.0, pib = 0.0, pjb = 0.0, jw2 = 0.0;
#pragma unroll
            for (int q = 0; q < NF; ++q) {
                const double ga = rowa[32 * q], gb = rowb[32 * q];
                pia = fma(ga, xi[q], pia);
                pja = fma(ga, xj[q], pja);
                pib = fma(gb, xi[q], pib);
                pjb = fma(gb, xj[q], pjb);
                jw = fma(ga, wa[q], jw);
                jw2 = fma(gb, wb[q], jw2);
                bi[q] = fma(ga, xka.x, bi[q]);
                bj[q] = fma(ga, xka.y, bj[q]);
                bi[q] = fma(gb, xkb.x, bi[q]);
                bj[q] = fma(gb, xkb.y, bj[q]);
            }
            {
                const int l = lane + 32 * NF;
                const double ga = (l <= ka) ? rowa[32 * NF] : 0.0;
                const double gb = (l <= kb) ? rowb[32 * NF] : 0.0;
                const double gat = (l < ka) ? ga : 0.0; // the diagonal element belongs to the row sum only
                const double gbt = (l < kb) ? gb : 0.0;
                pia = fma(ga, xi[NF], pia);
                pja = fma(ga, xj[NF], pja);
                pib = fma(gb, xi[NF], pib);
                pjb = fma(gb, xj[NF], pjb);
                jw = fma(ga, wa[NF], jw);
                jw2 = fma(gb, wb[NF], jw2);
                bi[NF] = fma(gat, xka.x, bi[NF]);
                bj[NF] = fma(gat, xka.y, bj[NF]);
                bi[NF] = fma(gbt, xkb.x, bi[NF]);
                bj[NF] = fma(gbt, xkb.y, bj[NF]);
            }
            jw += jw2;
            // 4-value butterfly: lanes 0-7 end with sum pia, 8-15 pja, 16-23 pib, 24-31 pjb
            const bool hi = lane & 16;
            double k0v = hi ? pib : pia, k1v = hi ? pjb : pja;
            const double s0 = hi ? pia : pib, s1 = hi ? pja : pjb;
            k0v += __shfl_xor_sync(0xffffffffu, s0, 16);
            k1v += __shfl_xor_sync(0xffffffffu, s1, 16);
            const bool b8 = lane & 8;
            double kv = b8 ? k1v : k0v;
            const double sv = b8 ? k0v : k1v;
            kv += __shfl_xor_sync(0xffffffffu, sv, 8);
            kv += __shfl_xor_sync(0xffffffffu, kv, 4);
            kv += __shfl_xor_sync(0xffffffffu, kv, 2);
            kv += __shfl_xor_sync(0xffffffffu, kv, 1);
            if ((lane & 7) == 0 && (hasb || !hi)) {
                double *dst = reinterpret_cast<double *>(va2 + (hi ? kb : ka)) + (b8 ? 1 : 0);
                *dst = kv;
            }
        };
        double w0a[QL], w0b[QL], w1a[QL], w1b[QL];
        {
            constexpr bool F0 = QL > 1;
            if (F0 || 2 * warp < nmo)
                load_w(std::integral_constant<int, 0>{}, std::bool_constant<F0>{}, 2 * warp, 2 * warp + 1 < nmo, w0a, w0b);
        }
        static_for<0, QL>([&](auto st_c) {
            constexpr int ST = decltype(st_c)::value;
            constexpr bool FULL = ST + 1 < QL; // all 32 rows exist
            std::bool_constant<FULL> full_c;
            const int kend = FULL ? 32 * ST + 32 : nmo;
            const int kaA = 32 * ST + 2 * warp, kaB = kaA + 16;
            if (FULL || kaB < kend) load_w(st_c, full_c, kaB, kaB + 1 < kend, w1a, w1b);
            mbar_wait(&full[s], phase);
            if (held >= 0) {
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[held]);
            }
            held = s;
            const int sh = (colpar + tri32(32 * ST)) & 1;
            const double *base = ringbuf + (size_t)s * p.stage_doubles + sh - tri32(32 * ST);
            if (++s == ring) {
                s = 0;
                phase ^= 1;
            }
            if (FULL || kaA < kend) pair(st_c, full_c, kaA, kaA + 1 < kend, base, w0a, w0b);
            if constexpr (ST + 1 < QL) {
                constexpr bool NFULL = ST + 2 < QL;
                const int kn = 32 * (ST + 1) + 2 * warp;
                if (NFULL || kn < nmo)
                    load_w(std::integral_constant<int, ST + 1>{}, std::bool_constant<NFULL>{}, kn, kn + 1 < nmo, w0a, w0b);
            }
            if (FULL || kaB < kend) pair(st_c, full_c, kaB, kaB + 1 < kend, base, w1a, w1b);
        });

        // ---- end of column
        const int64_t rn = r + gridDim.x;
        double nx0 = 0.0, nx1 = 0.0;
        if (rn < p.npair && tid < nmo) {
            int i, j;
            pair_of(rn, i, j);
            nx0 = __ldg(p.D + tid + (int64_t)nmo * i);
            nx1 = __ldg(p.D + tid + (int64_t)nmo * j);
        }
#pragma unroll
        for (int q = 0; q < QL; ++q) slab[warp * NMOP + lane + 32 * q] = make_double2(bi[q], bj[q]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) jw += __shfl_xor_sync(0xffffffffu, jw, o);
        if (lane == 0) jws[warp] = jw;
        __syncwarp();
        if (held >= 0 && lane == 0) mbar_arrive(&empty[held]); // every read of the column's last stage is behind us
        held = -1;
        consumer_barrier();
        if (tid < nmo) {
            double2 v = va2[tid];
#pragma unroll
            for (int w = 0; w < JS_NW; ++w) {
                const double2 t = slab[w * NMOP + tid];
                v.x += t.x;
                v.y += t.y;
            }
            p.R1[r * nmo + tid] = v.y;
            p.R2[r * nmo + tid] = v.x;
        }
        if (tid == 0) {
            double t = 0.0;
#pragma unroll
            for (int w = 0; w < JS_NW; ++w) t += jws[w];
            p.Jp[r * p.njp] = t;
        }
        if (tid < NMOP) x2[(par ^ 1) * NMOP + tid] = make_double2(nx0, nx1);
        consumer_barrier();
        par ^= 1;
        r = rn;
    }
}

namespace {
template <int QL>
cudaError_t launch_stream_q(const JkParams &p, int smem, int num_sms, cudaStream_t s)
{
    cudaError_t e = cudaFuncSetAttribute(jk_stream_kernel<QL>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    const int grid = p.npair < num_sms ? (int)p.npair : num_sms;
    jk_stream_kernel<QL><<<grid, JS_THREADS, smem, s>>>(p);
    return cudaGetLastError();
}
} // namespace

int jk_stream_max_nmo() { return 256; }

// Stage table for one column: blocks of `rows` whole packed rows (rows = 32, or 16 when three stages of 32 do not fit).
// Returns the number of stages; fills srow[0..nst], the ring depth and the slot size in doubles.
int jk_stream_plan(int nmo, int *srow, int *ring, int *stage_doubles, int *smem_bytes)
{
    const int QL = (nmo + 31) / 32, nmoP = 32 * QL;
    const int fixed = (JS_NW + 3) * nmoP * 16 + JS_NW * 8 + (JS_MAX_STAGES + 2) * 4 + 2 * JS_MAX_RING * 8 + 64;
    const int budget = 227 * 1024 - fixed;
    for (int rows = 32; rows >= 32; rows /= 2) { // the consumers' stage structure is compiled in: 32 rows
        int nst = 0, biggest = 0;
        srow[0] = 0;
        for (int k = 0; k < nmo; k += rows) {
            const int k1 = k + rows < nmo ? k + rows : nmo;
            const int n = k1 * (k1 + 1) / 2 - k * (k + 1) / 2;
            biggest = n > biggest ? n : biggest;
            srow[++nst] = k1;
        }
        const int slot = (biggest + 2 + 1) & ~1; // + alignment shift / pad element, even
        int depth = budget / (slot * 8);
        if (depth > JS_MAX_RING) depth = JS_MAX_RING;
        if (depth >= 3 && nst <= JS_MAX_STAGES) {
            *ring = depth;
            *stage_doubles = slot;
            *smem_bytes = depth * slot * 8 + fixed;
            return nst;
        }
    }
    return 0;
}

cudaError_t launch_jk_stream(const JkParams &p, int smem, int num_sms, cudaStream_t s)
{
    switch ((p.nmo + 31) / 32) {
    case 1: return launch_stream_q<1>(p, smem, num_sms, s);
    case 2: return launch_stream_q<2>(p, smem, num_sms, s);
    case 3: return launch_stream_q<3>(p, smem, num_sms, s);
    case 4: return launch_stream_q<4>(p, smem, num_sms, s);
    case 5: return launch_stream_q<5>(p, smem, num_sms, s);
    case 6: return launch_stream_q<6>(p, smem, num_sms, s);
    case 7: return launch_stream_q<7>(p, smem, num_sms, s);
    case 8: return launch_stream_q<8>(p, smem, num_sms, s);
    default: return cudaErrorInvalidValue;
    }
}

} // namespace qemb
