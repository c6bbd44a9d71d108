// jk_stream.cu -- J/K from the device-resident (ij|kl), single-pass streaming kernel (nmo <= 255).
//
// Same contraction and outputs as jk.cu (reference consumers: get_veff, src/quemb/molbe/helper.py:28-69, and the
// two-electron part of update_ebe_hf, src/quemb/molbe/pfrag.py:327-390):
//   J[ij] = sum_{kl} (ij|kl) D[k,l]        K[i,k] = sum_{jl} (ij|kl) D[j,l]
// HBM bound: column r = (i,j) of G is read once (8 npair bytes) and every element P[T(k)+l] of its packed
// triangle is used for five FMAs from ONE shared-memory read:
//   pa[k]  += g x[l]   (x = D[:,i] and D[:,j]: row sums, reduced over the lanes once per row pair)
//   vb[l]  += g x[k]   (column sums, kept in the lane's registers for the whole column, l = lane + 32 q)
//   J      += g w[k,l] (packed weights D[k,l] (2 - delta_kl), staged in shared memory next to the rows)
// jk.cu reads every element three times from shared memory, in three differently shaped loops, behind two CTA
// barriers per chunk (27 % of the HBM peak, 38 instructions per 32 elements); here a warp owns whole rows (lanes
// along l, conflict free), one row pair costs ~14 instructions per 64 elements plus one 4-value butterfly, and
// there is no CTA barrier inside a column:
//   * producer warp: one 1-D bulk copy (cp.async.bulk, SASS UBLKCP) per stage = a block of whole packed rows,
//     and one of the same rows of the J weights, completion counted on the stage's "full" mbarrier; stages are
//     placed back to back in a ring that takes all the shared memory left, so ~100 KB are in flight per SM
//   * 8 consumer warps: wait "full", process their row pair of the stage, release it behind a store whose data
//     depends on every read of the stage (an arrive placed directly behind the loads is issued before they return,
//     the race found in k3_gram.cu)
//   * per column two named barriers among the consumers: column sums of the 8 warps meet in shared memory and
//     are added in a fixed order (deterministic), next column's density columns are staged meanwhile
#include "jk_pipe.cuh"
#include "kernels.cuh"

namespace qemb {

namespace {
using namespace jkpipe;
#ifndef JS_COMPUTE
#define JS_COMPUTE 1 // 0: consumers only wait for and release the stages (data-path ceiling of the ring, wrong results)
#endif
constexpr int JS_NW = 8;                      // consumer warps: one row pair each per 16 packed rows of a stage
constexpr int JS_THREADS = 32 * (JS_NW + 1);  // + producer warp
constexpr int JS_NB = 16;                     // mbarrier pairs = stages in flight at most
constexpr int JS_MAX_STAGES = 16;             // stages per column
// Stage boundaries in packed rows (multiples of 16, identical for every fragment): tall stages where the rows are
// short -- a warp then has 4 / 2 independent row pairs between two waits -- and 16 rows from row 128 up.
__host__ __device__ constexpr int js_row0(int S) { return S == 0 ? 0 : S == 1 ? 64 : S == 2 ? 96 : 128 + 16 * (S - 3); }
// Stages below JS_WSM carry their J weights in the ring; from JS_WSM up the ring would hold G only (more bytes of G
// in flight per SM) and the consumers fetch the weights of their row pair from L2 into registers before they wait
// for the stage.  Measured at nmo 201: JS_WSM = 3 / 5 / 6 / 7 / 8 -> 1.01 / 1.16 / 1.13 / 1.10 / 1.05 ms, all stages
// with staged weights (the default, 16 = more than there are stages) 0.92 ms: the L2 latency of the weights costs
// more than the ring space they take.
#ifndef JS_WSM_DEF
#define JS_WSM_DEF 16
#endif
constexpr int JS_WSM = JS_WSM_DEF;
__host__ __device__ constexpr int js_stages(int rows) // stages that cover the packed rows [0, rows)
{
    int S = 0;
    while (js_row0(S) < rows) ++S;
    return S;
}

__device__ __forceinline__ void consumer_barrier()
{
    asm volatile("bar.sync 1, %0;" ::"n"(32 * JS_NW) : "memory");
}
} // namespace

// shared memory: ring of ring_doubles doubles, a stage = [G rows | weight rows] placed back to back (wrapping to 0
//                when the next one does not fit) | slab[JS_NW][nmoP] double2 | x2[2][nmoP] double2 | va2[nmoP] double2 |
//                jws[JS_NW] | full[JS_NB], empty[JS_NB]
// Stage S = packed rows [js_row0(S), min(js_row0(S + 1), nmo)); in every 16 rows of it warp w owns the row pair
// (first row + 2 w, + 1).  A row k has NF = k / 32 blocks of 32 columns strictly left of the diagonal and one block that
// contains it, the same NF for all the rows of a 16-row group, so the block
// structure is a compile-time property of the stage and the stage loop unrolls without data-dependent branches.
template <int QL>
__global__ void __launch_bounds__(JS_THREADS, 1) jk_stream_kernel(JkParams p)
{
    constexpr int NMOP = 32 * QL;
    constexpr int NSTAGE = js_stages(32 * QL); // stages that start at or below row 32 (QL - 1) always exist
    extern __shared__ __align__(16) double js_smem[];
    const int cap = p.ring_doubles; // ring capacity in doubles (even); stages are placed back to back
    double *ringbuf = js_smem;
    double2 *slab = reinterpret_cast<double2 *>(ringbuf + cap);
    double2 *x2 = slab + JS_NW * NMOP;
    double2 *va2 = x2 + 2 * NMOP;
    double *jws = reinterpret_cast<double *>(va2 + NMOP);
    int *prf = reinterpret_cast<int *>(jws + JS_NW); // per-warp sink of the release proofs
    int *stab = prf + 2 * JS_NW; // per stage: elements, T(first row)
    uint64_t *full = reinterpret_cast<uint64_t *>(stab + 2 * JS_MAX_STAGES);
    uint64_t *empty = full + JS_NB;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nmo = p.nmo, nst = p.nstages;

    for (int st = tid; st < nst; st += JS_THREADS) {
        const int k0 = js_row0(st), k1 = js_row0(st + 1) < nmo ? js_row0(st + 1) : nmo;
        stab[2 * st] = tri32(k1) - tri32(k0);
        stab[2 * st + 1] = tri32(k0);
    }
    if (tid == 0) {
        for (int s = 0; s < JS_NB; ++s) {
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
        // Ring bookkeeping in registers: `v` counts doubles handed out since the start (including the gap skipped at
        // a wrap), `rel` is v at the end of the last stage known to be released.  Stages are released in order, so a
        // region is free as soon as (its end in v) - rel <= cap.  The released stages' ends are replayed with the same
        // placement rule instead of being stored.
        // (A stage starts at a multiple of 16 rows, T of it is even: the alignment shift is the parity of the column start.)
        int off = 0, v = 0;
        unsigned n = 0, tail = 0;  // stage counters: next to issue, oldest not known to be released
        int t_off = 0, t_v = 0, t_st = 0; // replay of the placement for stage `tail`
        int rel = 0;
        const int dpar = (int)((p.ldg * gridDim.x) & 1); // parity step from one of my columns to the next
        int colpar = (int)((p.ldg * blockIdx.x) & 1), t_colpar = colpar;
        const int64_t colstep = p.ldg * gridDim.x;
        const double *col = p.G + p.ldg * blockIdx.x;
        for (int64_t r = blockIdx.x; r < p.npair; r += gridDim.x, col += colstep, colpar ^= dpar) {
            for (int st = 0; st < nst; ++st) {
                const int nel = stab[2 * st], tr = stab[2 * st + 1];
                const int sh = colpar;
                const int cnt = (nel + sh + 1) & ~1;
                const int cntw = st < JS_WSM ? (nel + 1) & ~1 : 0; // weight rows: dpk is aligned and padded
                const int sz = cnt + cntw;
                if (off + sz > cap) { // the consumers take the same decision
                    v += cap - off;
                    off = 0;
                }
                const int beg = off;
                off += sz;
                v += sz;
                while (v - rel > cap || n - tail >= JS_NB) {
                    mbar_wait(&empty[tail % JS_NB], (tail / JS_NB) & 1);
                    const int nel2 = stab[2 * t_st];
                    const int sz2 = ((nel2 + t_colpar + 1) & ~1) + (t_st < JS_WSM ? (nel2 + 1) & ~1 : 0);
                    if (t_off + sz2 > cap) {
                        t_v += cap - t_off;
                        t_off = 0;
                    }
                    t_off += sz2;
                    t_v += sz2;
                    rel = t_v;
                    if (++t_st == nst) {
                        t_st = 0;
                        t_colpar ^= dpar;
                    }
                    ++tail;
                }
                const int b = n % JS_NB;
                const double *srcA = col + tr - sh; // 16-byte aligned
                double *dst = ringbuf + beg;
                int cg = cnt;
                if (srcA + cg > Gend) { // the pad element would lie behind the matrix: copy the last one by hand
                    cg -= 2;
                    dst[cg] = srcA[cg];
                }
                mbar_arrive_expect_tx(&full[b], (cg + cntw) * 8);
                if (cg > 0) bulk_load(dst, srcA, cg * 8, &full[b]);
                if (cntw > 0) bulk_load(dst + cnt, p.dpk + tr, cntw * 8, &full[b]);
                ++n;
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

    int off = 0;
    unsigned nctr = 0;
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

        // row sums of the previous pair, reduced over the lanes while the next pair's FMAs are in flight
        double dpia = 0.0, dpja = 0.0, dpib = 0.0, dpjb = 0.0;
        int dka = -1;
        bool dhasb = false;
        // 4-value butterfly: lanes 0-7 end with sum pia, 8-15 pja, 16-23 pib, 24-31 pjb
        auto flush = [&]() {
            const bool hi = lane & 16;
            double k0v = hi ? dpib : dpia, k1v = hi ? dpjb : dpja;
            const double s0 = hi ? dpia : dpib, s1 = hi ? dpja : dpjb;
            k0v += __shfl_xor_sync(0xffffffffu, s0, 16);
            k1v += __shfl_xor_sync(0xffffffffu, s1, 16);
            const bool b8 = lane & 8;
            double kv = b8 ? k1v : k0v;
            const double sv = b8 ? k0v : k1v;
            kv += __shfl_xor_sync(0xffffffffu, sv, 8);
            kv += __shfl_xor_sync(0xffffffffu, kv, 4);
            kv += __shfl_xor_sync(0xffffffffu, kv, 2);
            kv += __shfl_xor_sync(0xffffffffu, kv, 1);
            if (dka >= 0 && (lane & 7) == 0 && (dhasb || !hi)) {
                double *dst = reinterpret_cast<double *>(va2 + (hi ? dka + 1 : dka)) + (b8 ? 1 : 0);
                *dst = kv;
            }
        };

        static_for<0, NSTAGE>([&](auto st_c) {
            constexpr int S = decltype(st_c)::value;
            constexpr int R0 = js_row0(S), R1 = js_row0(S + 1);
            constexpr bool FULL = R1 <= 32 * (QL - 1);   // all its rows exist for every nmo of this QL
            constexpr bool EXISTS = R0 <= 32 * (QL - 1); // the stage exists for every nmo of this QL
            if (EXISTS || S < nst) {
                const int kend = (FULL || R1 < nmo) ? R1 : nmo;
                const int nel = FULL ? (R1 * (R1 + 1) - R0 * (R0 + 1)) / 2 : tri32(kend) - (R0 * (R0 + 1)) / 2;
                const int sh = colpar;
                constexpr bool WSM = S < JS_WSM; // weights in the ring; else (16-row stage, one pair) from L2
                const int cnt = (nel + sh + 1) & ~1, cntw = WSM ? (nel + 1) & ~1 : 0;
                if (off + cnt + cntw > cap) off = 0; // same placement rule as the producer
                const double *gbase = ringbuf + off - (R0 * (R0 + 1)) / 2;
                off += cnt + cntw;
                double wga[QL], wgb[QL];
                if constexpr (!WSM) {
                    static_assert(WSM || R1 - R0 == 16, "one row pair per warp in the stages without staged weights");
                    constexpr int NF0 = R0 / 32;
                    const int ka = R0 + 2 * warp;
                    if (FULL || ka < kend) {
                        const bool hasb = FULL ? true : ka + 1 < kend;
                        const double *da = p.dpk + tri32(ka) + lane, *db = da + ka + 1;
                        const int l = lane + 32 * NF0;
#pragma unroll
                        for (int q = 0; q < NF0; ++q) {
                            wga[q] = __ldg(da + 32 * q);
                            wgb[q] = hasb ? __ldg(db + 32 * q) : 0.0;
                        }
                        wga[NF0] = (l <= ka) ? __ldg(da + 32 * NF0) : 0.0;
                        wgb[NF0] = (hasb && l <= ka + 1) ? __ldg(db + 32 * NF0) : 0.0;
                    }
                }
                const int b = nctr % JS_NB;
                mbar_wait(&full[b], (nctr / JS_NB) & 1);
                ++nctr;
                int proof = 0;
                static_for<0, (R1 - R0) / 16>([&](auto m_c) {
                    constexpr int M = decltype(m_c)::value;
                    constexpr int NF = (R0 + 16 * M) / 32;
                    if constexpr (NF < QL) {
                        const int ka = R0 + 16 * M + 2 * warp;
                        if (JS_COMPUTE && (FULL || ka < kend)) {
                            const bool hasb = FULL ? true : ka + 1 < kend;
                            const int kb = hasb ? ka + 1 : ka;
                            const double *rowa = gbase + sh + tri32(ka) + lane;
                            const double *rowb = hasb ? rowa + ka + 1 : rowa; // a missing row b re-reads row a against zeros
                            [[maybe_unused]] const double *wra = gbase + cnt + tri32(ka) + lane;
                            [[maybe_unused]] const double *wrb = hasb ? wra + ka + 1 : wra;
                            const double2 xka = xc[ka];
                            const double2 xkb = hasb ? xc[kb] : make_double2(0.0, 0.0);
                            const double bscale = hasb ? 1.0 : 0.0;
                            double pia[2] = {0.0, 0.0}, pja[2] = {0.0, 0.0}, pib[2] = {0.0, 0.0}, pjb[2] = {0.0, 0.0};
                            double jwa[2] = {0.0, 0.0}, jwb[2] = {0.0, 0.0};
#pragma unroll
                            for (int q = 0; q < NF; ++q) {
                                const double ga = rowa[32 * q], gb = rowb[32 * q];
                                double wa, wb;
                                if constexpr (WSM) {
                                    wa = wra[32 * q];
                                    wb = wrb[32 * q];
                                } else {
                                    wa = wga[q];
                                    wb = wgb[q];
                                }
                                pia[q & 1] = fma(ga, xi[q], pia[q & 1]);
                                pja[q & 1] = fma(ga, xj[q], pja[q & 1]);
                                pib[q & 1] = fma(gb, xi[q], pib[q & 1]);
                                pjb[q & 1] = fma(gb, xj[q], pjb[q & 1]);
                                jwa[q & 1] = fma(ga, wa, jwa[q & 1]);
                                jwb[q & 1] = fma(gb, wb, jwb[q & 1]);
                                bi[q] = fma(ga, xka.x, bi[q]);
                                bj[q] = fma(ga, xka.y, bj[q]);
                                bi[q] = fma(gb, xkb.x, bi[q]);
                                bj[q] = fma(gb, xkb.y, bj[q]);
                            }
                            {
                                const int l = lane + 32 * NF;
                                const double ga = (l <= ka) ? rowa[32 * NF] : 0.0;
                                const double gb = (l <= kb) ? rowb[32 * NF] : 0.0;
                                double wa, wb;
                                if constexpr (WSM) {
                                    wa = (l <= ka) ? wra[32 * NF] : 0.0;
                                    wb = (l <= kb) ? wrb[32 * NF] : 0.0;
                                } else {
                                    wa = wga[NF];
                                    wb = wgb[NF];
                                }
                                const double gat = (l < ka) ? ga : 0.0; // the diagonal element belongs to the row sum only
                                const double gbt = (l < kb) ? gb : 0.0;
                                pia[NF & 1] = fma(ga, xi[NF], pia[NF & 1]);
                                pja[NF & 1] = fma(ga, xj[NF], pja[NF & 1]);
                                pib[NF & 1] = fma(gb, xi[NF], pib[NF & 1]);
                                pjb[NF & 1] = fma(gb, xj[NF], pjb[NF & 1]);
                                jwa[NF & 1] = fma(ga, wa, jwa[NF & 1]);
                                jwb[NF & 1] = fma(gb, wb, jwb[NF & 1]);
                                bi[NF] = fma(gat, xka.x, bi[NF]);
                                bj[NF] = fma(gat, xka.y, bj[NF]);
                                bi[NF] = fma(gbt, xkb.x, bi[NF]);
                                bj[NF] = fma(gbt, xkb.y, bj[NF]);
                            }
                            flush(); // previous pair
                            jw += fma(jwb[0] + jwb[1], bscale, jwa[0] + jwa[1]);
                            dpia = pia[0] + pia[1];
                            dpja = pja[0] + pja[1];
                            dpib = pib[0] + pib[1];
                            dpjb = pjb[0] + pjb[1];
                            dka = ka;
                            dhasb = hasb;
                            proof ^= (__double2loint(pia[0]) ^ __double2loint(pia[1]) ^ __double2loint(pib[0])) ^
                                     (__double2loint(pib[1]) ^ __double2loint(jwa[0]) ^ __double2loint(jwa[1])) ^
                                     (__double2loint(jwb[0]) ^ __double2loint(jwb[1]));
                        }
                    }
                });
                // Release the stage: lane 0 stores a value that depends on every shared-memory read of the stage (a
                // few bits of the row sums <- all G loads and of the J sums <- all weight loads), so the store cannot
                // issue before the last read has returned, and arrives behind it.  (An arrive placed directly behind
                // the loads is issued before they complete, see k3_gram.cu; deferring it to the next stage keeps two
                // stages per SM out of flight.)
                if (lane == 0) {
                    prf[warp] = proof;
                    mbar_arrive(&empty[b]);
                }
            }
        });
        flush();

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

// Stages of one column: blocks of whole packed rows (js_row0) and their J weights, placed back to back in a ring that
// takes all the shared memory left (the bytes in flight per SM bound the HBM throughput).  Returns the number of stages
// (0: does not fit); fills the ring capacity in doubles and the dynamic shared-memory size.  Three of the largest stage
// must fit: with two, a stage that wraps to the start of the ring can overlap the one still being read (deadlock).
int jk_stream_plan(int nmo, int *ring_doubles, int *smem_bytes)
{
    const int QL = (nmo + 31) / 32, nmoP = 32 * QL;
    const int fixed = (JS_NW + 3) * nmoP * 16 + 2 * JS_NW * 8 + 2 * JS_MAX_STAGES * 4 + 2 * JS_NB * 8 + 64;
    const int cap = ((227 * 1024 - fixed) / 8) & ~1;
    int nst = 0, biggest = 0;
    for (int st = 0; js_row0(st) < nmo; ++st) { // the consumers' stage structure is compiled in (js_row0)
        const int k = js_row0(st), k1 = js_row0(st + 1) < nmo ? js_row0(st + 1) : nmo;
        const int n = k1 * (k1 + 1) / 2 - k * (k + 1) / 2;
        const int need = st < JS_WSM ? 2 * ((n + 3) & ~1) : (n + 3) & ~1; // G rows (+ shift / pad) | weight rows
        biggest = need > biggest ? need : biggest;
        ++nst;
    }
    if (cap / biggest < 3 || nst > JS_MAX_STAGES) return 0;
    *ring_doubles = cap;
    *smem_bytes = cap * 8 + fixed;
    return nst;
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
