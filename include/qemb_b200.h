/*
 * qemb_b200.h -- C ABI of libqemb_b200.so: the B200-native (sm_100a, FP64) sparse-DF fragment ERI
 * transform that replaces QuEmb's Eigen extension `quemb.molbe._cpp.eri_sparse_DF`.
 *
 * Every entry point names the reference interface it stands in for
 * ("cpp:" = src/quemb/molbe/_cpp/eri_sparse_DF.cpp, "hpp:" = .../indexers.hpp,
 *  "pyi:" = .../eri_sparse_DF.pyi, "py:" = src/quemb/molbe/eri_sparse_DF.py of troyvvgroup/quemb).
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ / torch types cross this boundary.
 *   - every function returns 0 on success, a negative QEMB_E* code on failure; the message of the
 *     last failure on the calling thread is returned by qemb_last_error().  Nothing throws.
 *   - all host matrices are column-major FP64 (the reference's Eigen default, cpp:29-33) with an
 *     explicit leading dimension; index lists are CSR pairs (ptr[n+1], idx[nnz]) of int64
 *     (hpp:16-17).
 *   - "dev" pointers are CUDA device pointers of the handle's device; "host" pointers may be pageable
 *     or pinned (pinned makes the copies asynchronous).
 *   - handles are immutable after upload and may be shared by concurrent qemb_fragment_* calls that
 *     use different qemb_workspace objects (the reference releases the GIL around every call,
 *     cpp:889-938).
 *   - there is NO CPU fallback: without a usable CUDA device every compute entry point fails with
 *     QEMB_ENODEV.
 */
#ifndef QEMB_B200_H
#define QEMB_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QEMB_OK 0
#define QEMB_EINVAL (-1)  /* bad argument / shape mismatch  (std::runtime_error, cpp:228-238)      */
#define QEMB_ENOMEM (-2)  /* device or host allocation failed (std::bad_alloc, cpp:56-62)          */
#define QEMB_ECUDA (-3)   /* CUDA runtime error               (CUDA_CHECK_THROW, cpp:40-46)        */
#define QEMB_ENODEV (-4)  /* no CUDA device                                                         */
#define QEMB_EINDEX (-5)  /* (mu,nu) not in the screened set  (std::out_of_range, cpp:210)         */

typedef struct qemb_ctx qemb_ctx;             /* one CUDA device + stream                           */
typedef struct qemb_tensor qemb_tensor;       /* SemiSparseSym3DTensor on the device, cpp:110-298   */
typedef struct qemb_metric qemb_metric;       /* GPU_MatrixHandle, cpp:64-107                       */
typedef struct qemb_screen qemb_screen;       /* S_abs on the device (input of cpp:443-465)         */
typedef struct qemb_workspace qemb_workspace; /* per-caller scratch: plan, B, X, G buffers          */

const char *qemb_last_error(void);
const char *qemb_version(void);
/* get_log_level / set_log_level, hpp:34-46 / pyi:139-157: process-global, Python logging levels (default 40 = Error).
 * At <= 20 (Info) qemb_fragment_transform prints the reference's "[TIMER] <step> finished in <t> s" lines (hpp:170-204)
 * from CUDA-event times of screening, K1, K2a, K2b and K3; reading them waits for the fragment. */
int qemb_get_log_level(void);
int qemb_set_log_level(int level);

/* -- device -------------------------------------------------------------------------------------- */
int qemb_device_count(int *n);
int qemb_ctx_create(int device, qemb_ctx **out);
void qemb_ctx_destroy(qemb_ctx *);
int qemb_ctx_synchronize(qemb_ctx *);
int qemb_ctx_device(const qemb_ctx *, int *device);

/* -- index maps (host, exact integer arithmetic) ------------------------------------------------- */
/* hpp:75-79 */
int64_t qemb_ravel_symmetric(int64_t a, int64_t b);
/* hpp:89-96; returns (min,max) like the C++ reference (exact integer arithmetic, no double sqrt) */
void qemb_unravel_symmetric(int64_t i, int64_t *lo, int64_t *hi);
/* extract_unique, hpp:149-162 / pyi:132: prefix nu<=mu of every sorted list.
 * uniq_ptr has nao+1 entries; uniq_idx may be NULL to count only. Returns n_unique (or <0). */
int64_t qemb_extract_unique(int64_t nao, const int64_t *exch_ptr, const int64_t *exch_idx,
                            int64_t *uniq_ptr, int64_t *uniq_idx);

/* -- SemiSparseSym3DTensor: screened (P|mu nu) ---------------------------------------------------- */
/* ctor #3, cpp:143-164 (+ compute_offsets_and_unique cpp:248-258,
 * initialize_exch_reachable_with_offsets cpp:260-279).  Allocates naux x n_unique on the device
 * (NaN-filled like cpp:160) and builds the column maps. exch lists must be sorted ascending. */
int qemb_tensor_create(qemb_ctx *, int64_t naux, int64_t nao, const int64_t *exch_ptr,
                       const int64_t *exch_idx, qemb_tensor **out);
void qemb_tensor_destroy(qemb_tensor *);
int qemb_tensor_n_unique(const qemb_tensor *, int64_t *n_unique);
/* nonzero_size / size, cpp:196-201 */
int qemb_tensor_shape(const qemb_tensor *, int64_t *naux, int64_t *nao, int64_t *n_unique);
/* column of the pair (mu,nu) (= offsets[ravel_symmetric(mu,nu)], cpp:210); QEMB_EINDEX if unscreened */
int qemb_tensor_column(const qemb_tensor *, int64_t mu, int64_t nu, int64_t *col);
/* exch_reachable_with_offsets, cpp:260-279: col[k] for every entry k of the full exch lists
 * (a pair whose key was never assigned reads column 0, the reference's operator[] default). */
int qemb_tensor_reachable_columns(const qemb_tensor *, int64_t *col /* nnz(exch) */);
/* fills columns [col0, col0+ncols) from host (replaces writes through mut_unique_dense_data,
 * cpp:810-812, py:486-488).  A whitened tensor accepts only a full replacement (which clears the flag): a partial
 * refresh would mix raw and L^-1-scaled columns -- QEMB_EINVAL; clear the flag with qemb_tensor_set_whitened first. */
int qemb_tensor_upload(qemb_tensor *, const double *host, int64_t ld_host, int64_t col0, int64_t ncols);
/* Staging of one libcint block (replaces the Python triple loop py:459-488):
 * block[(i * n_j + j) * naux + P] = (P | p0+i, q0+j), C order as `_aux_e2(...)` returns it; every screened pair
 * with q0+j <= p0+i receives its column on the device, all other entries of the block are ignored. */
int qemb_tensor_scatter_block(qemb_tensor *, const double *block, int64_t n_i, int64_t n_j, int64_t p0, int64_t q0);
/* (asynchronous: `block` is copied into one of two pinned staging slots before the call returns, the H2D copy and the
 * scatter kernel of block b overlap the caller's evaluation of block b+1; any later read of the tensor is ordered
 * behind them on the compute stream) */
/* number of NaN entries left in the tensor: a column nobody filled keeps the NaN of qemb_tensor_create (cpp:160); the
 * reference asserts `not np.isnan(result.unique_dense_data).any()` at the end of get_sparse_P_mu_nu (py:492) */
int qemb_tensor_count_nan(qemb_tensor *, int64_t *n);
/* reads columns back (unique_dense_data / __getitem__, cpp:203-215, 806-826) */
int qemb_tensor_download(const qemb_tensor *, double *host, int64_t ld_host, int64_t col0, int64_t ncols);

/* -- Coulomb metric ------------------------------------------------------------------------------- */
/* GPU_MatrixHandle(L_host), cpp:64-107 / pyi:29-36: lower Cholesky factor of (P|Q), column-major.
 * Also inverts the diagonal blocks once (used by the blocked DMMA triangular solve). */
int qemb_metric_create(qemb_ctx *, const double *L_host, int64_t ld, int64_t naux, qemb_metric **out);
void qemb_metric_destroy(qemb_metric *);

/* Optional (design option O1): W = L^-1 (P|mu nu) once for the whole molecule, in place on the
 * device copy; afterwards qemb_fragment_transform skips the per-fragment solve (cpp:615). */
int qemb_tensor_whiten(qemb_tensor *, const qemb_metric *);
int qemb_tensor_is_whitened(const qemb_tensor *, int *flag);
/* The same solve on the column range [col0, col0+ncols) only (asynchronous): ranks that each hold a replica of
 * (P|mu nu) split the columns, exchange the results on the device (qemb_tensor_data_dev gives the address for the
 * collective) and then declare the tensor whitened. */
int qemb_tensor_whiten_range(qemb_tensor *, const qemb_metric *, int64_t col0, int64_t ncols);
int qemb_tensor_set_whitened(qemb_tensor *, int flag);
int qemb_tensor_data_dev(const qemb_tensor *, void **dev, int64_t *ld);
/* Aux-sharded staging (oversize fragments): `slab` is a tensor created with naux = the number of
 * auxiliary rows this device owns; the call uploads columns [col0, col0+ncols) of the FULL-naux host matrix
 * (rows >= p_begin + slab naux never travel), solves them against L in a scratch buffer and keeps rows
 * [p_begin, p_begin + slab naux) -- the slab of L^-1 (P|mu nu) -- so no device ever holds all of (P|mu nu).
 * Marks the slab whitened; qemb_fragment_transform on it (metric NULL) returns the slab's PARTIAL (ij|kl),
 * to be summed over the devices (NCCL reduce).  p_begin must be a multiple of 16. */
int qemb_tensor_upload_whitened_rows(qemb_tensor *slab, const qemb_metric *, const double *host, int64_t ld_host,
                                     int64_t col0, int64_t ncols, int64_t p_begin);
int qemb_tensor_release_scratch(qemb_tensor *);
/* device -> device copy of the data (and whitened flag) between two tensors built from the same lists
 * (keeps a raw copy around when the same molecule is transformed repeatedly) */
int qemb_tensor_copy_data(qemb_tensor *dst, const qemb_tensor *src);

/* -- screening input ------------------------------------------------------------------------------ */
int qemb_screen_create(qemb_ctx *, const double *S_abs_host, int64_t ld, int64_t nao, qemb_screen **out);
void qemb_screen_destroy(qemb_screen *);

/* -- per-caller workspace -------------------------------------------------------------------------- */
int qemb_workspace_create(qemb_ctx *, qemb_workspace **out);
void qemb_workspace_destroy(qemb_workspace *);
/* bytes currently held on the device by this workspace */
int qemb_workspace_bytes(const qemb_workspace *, int64_t *bytes);

/* -- the fragment path ----------------------------------------------------------------------------- */
/* get_AO_per_MO, cpp:443-465 / pyi:137: X = |S_abs * TA| on the device (plain left-to-right FP64
 * sums, no FMA contraction), A(i) = {mu : X(mu,i) >= eps}.  Stores the lists in the workspace
 * ("the plan") and returns the total count.  TA is nao x nmo column-major on the host. */
int qemb_fragment_screen(qemb_workspace *, const qemb_screen *, const double *TA_host, int64_t ld,
                         int64_t nao, int64_t nmo, double eps, int64_t *n_mu_i);
/* number of (mu, i) of the last qemb_fragment_screen with | |X(mu,i)| - eps | < 1e-12: entries whose membership may
 * differ from a producer that sums in another order (the reference is compiled -ffast-math) */
int qemb_fragment_borderline(const qemb_workspace *, int64_t *n);
/* plan from caller-supplied lists (the AO_by_MO argument of contract_with_TA_1st, pyi:126) */
int qemb_fragment_set_lists(qemb_workspace *, const double *TA_host, int64_t ld, int64_t nao,
                            int64_t nmo, const int64_t *A_ptr, const int64_t *A_idx);
/* read the plan's lists back: A_ptr[nmo+1], A_idx[n_mu_i] */
int qemb_fragment_get_lists(const qemb_workspace *, int64_t *A_ptr, int64_t *A_idx);

/* K1, contract_with_TA_1st cpp:484-532: B[:, off(i,mu)] = sum_{nu in R(mu)} TA(nu,i) M[:, col(mu,nu)],
 * column offsets = running counter over i then mu in A(i) (cpp:467-482).  Result stays on the device
 * inside the workspace. */
int qemb_fragment_contract_1st(qemb_workspace *, const qemb_tensor *);
/* (P | mu i) from the caller instead of from K1: the public constructors of SemiSparse3DTensor (pyi:41-58,
 * cpp:308-352) build the K1 result from host data, and contract_with_TA_2nd_to_sym_dense accepts such a tensor.
 * B_host is naux x n_mu_i column-major (columns in the order of qemb_fragment_set_lists' lists, cpp:467-482), copied
 * before the call returns; qemb_fragment_contract_2nd / _solve / _gram then run on it. */
int qemb_fragment_set_B(qemb_workspace *, int64_t naux, const double *B_host, int64_t ld);
/* K2a, contract_with_TA_2nd_to_sym_dense cpp:560-605: sym[:, T(hi)+lo] = sum_{mu in A(lo)} TA(mu,hi) B[:, off(lo,mu)] */
int qemb_fragment_contract_2nd(qemb_workspace *);
/* K2b, eval_via_cholesky step 1 cpp:615: X = L^-1 sym (in place in the workspace) */
int qemb_fragment_solve(qemb_workspace *, const qemb_metric *);
/* K3, eval_via_cholesky step 2 cpp:620: G = X^T X, full npair x npair square, symmetric, s4 pair
 * order ij = T(max)+min.  Output stays in the workspace. */
int qemb_fragment_gram(qemb_workspace *);

/* debugging / parity taps: copy intermediates of the last fragment to the host (column-major) */
int qemb_fragment_get_B(const qemb_workspace *, double *host, int64_t ld);   /* naux x n_mu_i   */
int qemb_fragment_get_sym(const qemb_workspace *, double *host, int64_t ld); /* naux x npair    */
int qemb_fragment_get_G(const qemb_workspace *, double *host, int64_t ld);   /* npair x npair   */
/* device address of G (npair x npair, ld = npair) for zero-copy consumers (e.g. NCCL reduce) */
int qemb_fragment_G_dev(const qemb_workspace *, void **dev, int64_t *npair);

/* transform_integral / transform_integral_cuda, cpp:724-751 / pyi:158-181: screen -> K1 -> K2a ->
 * (K2b unless the tensor is whitened) -> K3 for one fragment.
 *   out         npair x npair FP64, ld_out >= npair.  Host pointer unless out_is_device != 0.
 *   p_begin/end restrict the auxiliary index to [p_begin, p_end) (aux-sharded fragments; requires a
 *               whitened tensor; pass 0, -1 for the full range).  The result is then the PARTIAL
 *               Gram matrix of that slab, to be summed over ranks.
 * The call is asynchronous when `out` is pinned or device memory; qemb_fragment_wait_output(ticket)
 * or qemb_ctx_synchronize() completes it.  `ticket` may be NULL. */
int qemb_fragment_transform(qemb_workspace *, const qemb_tensor *, const qemb_metric *,
                            const qemb_screen *, const double *TA_host, int64_t ld, int64_t nao,
                            int64_t nmo, double eps, int64_t p_begin, int64_t p_end, double *out,
                            int64_t ld_out, int out_is_device, int64_t *ticket);
/* out_is_device == 2: the lower 128x128 tiles of this call's (partial) result are ADDED into `out` with system-scope
 * red.add.f64 from the K3 epilogue -- `out` may be memory of a PEER GPU opened with qemb_ipc_open, so the reduce-sum of
 * an aux-sharded fragment rides on NVLink tile by tile while the remaining tiles are still being computed.
 * The owner zero-fills the matrix first (qemb_device_zero) and rebuilds the upper triangle afterwards
 * (qemb_device_mirror_lower); ordering between ranks is the caller's (barriers). */
/* `out` == NULL: (ij|kl) stays resident in the workspace (no device->host transfer); read it with
 * qemb_fragment_get_G / qemb_fragment_G_dev or consume it in place with qemb_fragment_jk. */

/* Consumer of the path, SURVEY 8f rank 1 -- scf.hf.dot_eri_dm(eri, dm, hermi=1) as called by get_veff
 * (src/quemb/molbe/helper.py:28-69) and, through E2_i = sum_j D_ij (2J - K)_ij, the two-electron sum of
 * update_ebe_hf (src/quemb/molbe/pfrag.py:368-389):  vj[i,j] = sum_kl (ij|kl) D[k,l],
 * vk[i,k] = sum_jl (ij|kl) D[j,l] from the (ij|kl) of the last transform, read once where it lives in HBM.
 * D, vj, vk: nmo x nmo column-major host matrices; D must be symmetric.  Synchronous. */
int qemb_fragment_jk(qemb_workspace *, const double *D_host, int64_t ldd, double *vj_host, double *vk_host,
                     int64_t ld_out);
/* Same contraction, enqueued on the compute stream: returns as soon as D is staged; vj / vk (pinned host memory for
 * real asynchrony) are complete after qemb_ctx_synchronize.  The all-fragment consumer pipeline (mbe.py:1146-1170 done
 * where (ij|kl) lives) issues it right behind every resident transform, so only nmo x nmo matrices leave the device. */
int qemb_fragment_jk_enqueue(qemb_workspace *, const double *D_host, int64_t ldd, double *vj_host, double *vk_host,
                             int64_t ld_out);
/* device time (ms) of the last qemb_fragment_jk and its algorithmic bytes (8 npair^2) */
int qemb_fragment_jk_time(const qemb_workspace *, float *ms, double *bytes);
/* blocks until the host output of the transform that returned `ticket` is complete */
int qemb_fragment_wait_output(qemb_workspace *, int64_t ticket);
/* how many host-output transforms may be pending per workspace before the oldest ticket's completion state is
 * recycled (a caller keeps at most this many output buffers in flight) */
int qemb_output_ring_depth(void);

/* Algorithmic work of the last planned fragment (SURVEY 8d formulas): flops and bytes of K1, K2a,
 * K2b, K3 in that order (8 doubles). */
int qemb_fragment_work(const qemb_workspace *, const qemb_tensor *, double *flops_bytes8);
/* device time in ms of the kernels of the last fragment: screen, K1, K2a, K2b, K3 (CUDA events on
 * the ctx stream; only recorded when enabled) */
int qemb_ctx_set_timing(qemb_ctx *, int enabled);
/* How a host-bound (ij|kl) matrix travels.  1 (default): only the lower triangle crosses PCIe, in column
 * groups, and the library's host threads rebuild the upper triangle while later groups and the next
 * fragment are in flight (the reference's GPU variant copies the full syrk result and mirrors it on one
 * host thread, cpp:684-708).  0: the full square crosses PCIe. */
int qemb_ctx_set_output_mode(qemb_ctx *, int mode);
/* Output mode 1 only: the last `frac` of the columns of a host-bound (ij|kl) cross PCIe whole and are not mirrored by
 * the host threads (default: QEMB_FULL_COLUMN_FRACTION or 0).  Shifts work from the host cores (the bottleneck of the
 * pure triangle transfer) to the copy engine, which has slack while the kernels run. */
int qemb_ctx_set_full_column_fraction(qemb_ctx *, double frac);
/* Output mode 1 only: bytes per column group of the triangle transfer (default 256 MB).  The host threads start
 * mirroring a group when its copy has completed; smaller groups start earlier and may still find the data in the
 * last-level cache, larger ones amortise the per-copy overhead. */
int qemb_ctx_set_output_group_bytes(qemb_ctx *, int64_t bytes);

/* Self-tuning of the two knobs above.  Measures on this context -- call it on all ranks of a node at the same moment,
 * e.g. right after a barrier -- the device -> host copy rate with idle host threads (P0), the same rate while the host
 * threads mirror a test matrix (P) and that mirror rate (M), and picks the cheaper of: full squares over PCIe
 * (2U / P0), or the fraction f* = max(0, (P - M) / (P + M)) of the upper triangle over PCIe and the rest mirrored
 * (max((1 + f) U / P, (1 - f) U / M)).  Reports P, M (GB/s) and the chosen fraction (1.0 = full squares, output mode
 * 0); qemb_ctx_autotune_copy_alone returns P0. */
int qemb_ctx_autotune_output(qemb_ctx *, double *d2h_gbs, double *mirror_gbs, double *upper_fraction_pcie);
int qemb_ctx_autotune_copy_alone(const qemb_ctx *, double *d2h_alone_gbs);
/* The model only proposes: qemb_ctx_autotune_output then RUNS the candidates -- full squares, pure triangle, the model's
 * split -- as four back-to-back deliveries of a 0.5 GB test matrix each through the real output path and keeps the
 * fastest (the model's choice stands if the trials cannot run).  Seconds per candidate, -1 = not run: */
int qemb_ctx_autotune_trials(const qemb_ctx *, double *secs3);
/* Trial times within 5 % of the fastest are a tie, broken by the rate model (2 / P0 for full squares,
 * max((1 + f) / P, (1 - f) / M) for a split f): the trial matrix is a tenth of a fragment and one box tied two modes whose
 * builds differ by 20 %.  qemb_autotune_pick is that rule as a host-only test hook: candidates 0 full squares, 1 pure
 * triangle, 2 split with upper3[2] of the upper triangle over PCIe; secs <= 0 = not run; *pick = -1 if none ran. */
int qemb_autotune_pick(const double *secs3, const double *upper3, double d2h_alone_gbs, double d2h_gbs, double mirror_gbs,
                       int *pick);

/* OPT-IN variant of K3 (G = X^T X, cpp:620): 0 = FP64 tensor path (default, the parity reference), 5 .. 8 = error-free
 * split of X into that many 8-bit integer slices (balanced base-256 digits of each column rounded below its exponent)
 * multiplied on the int8 tensor cores with exact int32 accumulation and recombined in FP64 (k3_int8.cu).  Deviation
 * from the exact product on K3-shaped operands, as a fraction of max|G|: about 1e-11 (5), 2e-13 (6), FP64 rounding
 * level (7, 8) (profiles/k3_int8_parity_study_r1.txt); cost ~ slices (slices + 1) / 2 int8 products.  Default:
 * QEMB_K3_INT8 or 0.  Applies to the accumulating transform (out_is_device == 2) as well: every slab is sliced with its
 * own exponents and its lower triangle added.  Slabs of more than 2^17 / slices rows stay on the FP64 path (int32). */
int qemb_ctx_set_k3_int8_slices(qemb_ctx *ctx, int slices);
/* Guaranteed-tolerance mode of the same opt-in variant (abs_tol > 0; 0 switches it off; overrides the fixed slice
 * count): per fragment the largest column exponent e_max of X is reduced on the device first, and K3 runs with the
 * smallest slice count S in 5 .. 8 whose WORST-CASE bound  qemb_k3_int8_bound(naux, S) * 2^(2 e_max)  is <= abs_tol for
 * every (ij|kl) -- with the FP64 kernel when no S qualifies.  Costs one pass over X and a stream synchronisation per
 * fragment.  qemb_fragment_k3_choice reports what the last K3 of a workspace ran with (slices 0 = FP64 kernel). */
int qemb_ctx_set_k3_int8_tolerance(qemb_ctx *ctx, double abs_tol);
/* worst case of |G_int8 - X^T X| per unit of 2^(e_i + e_j) for `slices` slices over `naux` auxiliary rows */
int qemb_k3_int8_bound(int64_t naux, int slices, double *bound);
int qemb_fragment_k3_choice(const qemb_workspace *, int *slices, int *max_expo);
/* Host twin of the device slicer of the int8 K3 (same source: csrc/k3_int8_digits.h): exponent per column
 * (expo[c]; a column holding NaN / Inf gets the poison value -2^30 and zero digits) and the digits
 * D[(c * slices + a) * nrows + p].  Test hook for the CPU suite; no reference counterpart, not on the product path. */
int qemb_k3_int8_slice_host(const double *X, int64_t ldx, int64_t nrows, int64_t ncols, int slices, int8_t *D, int *expo);
/* Host-only test hooks (no reference counterpart, not on the product path) for the schedules the persistent kernels
 * walk: the lower-triangular tile order of K3 (2 ints per tile into ti_tj, nt (nt + 1) / 2 tiles; returns the count),
 * the k splits of a small K3 (launch_gram's nsplit and 16-row k tiles per split), and the order in which the
 * gather-GEMM tiles of given costs are dealt to `num_sms` persistent CTAs (order = a permutation of 0 .. n-1). */
int64_t qemb_test_gram_tiles(int nt, int32_t *ti_tj);
int qemb_test_gram_splits(int ntiles, int ktiles, int num_sms, int *nsplit, int *kchunk);
int qemb_test_balance_tiles(int64_t n, const float *cost, int num_sms, int32_t *order);
/* host DRAM bandwidth seen by the library's host threads: non-temporal fill of `bytes` (write_gbs) and parallel memcpy
 * of bytes/2 (copy_gbs = bytes copied per second); best of 3.  Denominator of bench.py's e2e.host_roofline_frac. */
int qemb_host_bandwidth(int64_t bytes, double *write_gbs, double *copy_gbs);
/* orders an external CUDA stream (cudaStream_t passed as void*, e.g. torch's current stream that NCCL collectives run
 * on) against the context's compute stream without a host synchronisation: direction 0 = the external stream waits
 * for everything enqueued on the compute stream so far, 1 = the compute stream waits for the external stream */
int qemb_ctx_stream_handoff(qemb_ctx *, void *external_stream, int direction);
/* number of host worker threads of the library (QEMB_HOST_THREADS, default: cores in the affinity mask - 1) */
int qemb_host_threads(int *n);
/* G[c + ld*r] = G[r + ld*c] for all r > c of an n x n host matrix (the host half of output mode 1; also the
 * replacement of the reference's serial mirror loop, cpp:700-708). */
int qemb_host_mirror_lower(double *G, int64_t ld, int64_t n);
int qemb_fragment_times(const qemb_workspace *, float *ms5);
/* CUDA-event marks for whole-step timing: record mark `id` (0..15) on stream 0 = compute, 1 = prep,
 * 2 = copy; qemb_ctx_mark_elapsed waits for both marks and returns the device time between them. */
int qemb_ctx_mark(qemb_ctx *, int id, int stream);
int qemb_ctx_mark_elapsed(qemb_ctx *, int id0, int id1, float *ms);
/* number of kernel launches issued by this library on this ctx since creation */
int qemb_ctx_launch_count(const qemb_ctx *, int64_t *n);
/* bytes this context has handed to the copy engines so far: host->device (tensor, metric, S_abs, TA, plan
 * tables) and device->host ((ij|kl) transfers; the screening mask travels through mapped memory) */
int qemb_ctx_transfer_bytes(const qemb_ctx *, int64_t *h2d, int64_t *d2h);

/* device buffers that other processes of the node can map (cudaIpc*): alloc + 64-byte handle on the owner, open / close
 * on the peers, free on the owner */
int qemb_ipc_alloc(qemb_ctx *, int64_t bytes, void **dev, unsigned char *handle64);
int qemb_ipc_open(qemb_ctx *, const unsigned char *handle64, void **dev);
int qemb_ipc_close(qemb_ctx *, void *dev);
int qemb_ipc_free(qemb_ctx *, void *dev);
/* asynchronous zero fill / upper-from-lower mirror on the context's compute stream; synchronous device -> host copy */
int qemb_device_zero(qemb_ctx *, void *dev, int64_t bytes);
int qemb_device_mirror_lower(qemb_ctx *, double *G, int64_t ld, int64_t n);
int qemb_device_to_host(qemb_ctx *, void *host, const void *dev, int64_t bytes);

/* pinned host memory helpers (so that Python callers can hand out asynchronous buffers) */
int qemb_host_alloc(void **p, int64_t bytes);
int qemb_host_free(void *p);

/* FP64 DGEMM-shaped peak probe: runs the library's own DMMA mainloop on an n x n x k problem that
 * lives entirely in L2/smem and returns TFLOP/s (used to report the tensor-pipe ceiling next to
 * the cuBLAS figure). */
int qemb_probe_dmma_tflops(qemb_ctx *, int64_t n, int64_t k, int iters, double *tflops);

/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * K3 for one small fragment, G = X^T X (what qemb_fragment_gram computes, cpp:620), with NO allocation of any kind:
 * the split-K partial matrices live in a file-scope `static __device__` array of 32 MiB (csrc/gram_noalloc.cu; every
 * GPU that loads that module holds it until the process exits), whose device address is taken once per device with
 * cudaGetSymbolAddress.  The caller passes input and output:
 *   X_dev   naux x npair FP64, column-major, leading dimension ldx >= naux, ldx even, 16-byte aligned (TMA); rows
 *           [naux, ldx) are never read, so no zero padding is needed
 *   G_dev   npair x npair FP64 (both triangles written), leading dimension ldg >= npair
 *   stream  cudaStream_t as void* (NULL = default stream) of the CURRENT device, which must own X_dev and G_dev;
 *           asynchronous: two kernel launches are enqueued, nothing is synchronised.
 * Limit: npair <= 2048 (qemb_gram_noalloc_limits reports it and the array size); larger sizes fail with QEMB_EINVAL
 * and a message naming the limit -- use qemb_fragment_gram / qemb_fragment_transform there.
 * Every call shares the one array and finds the last call's contents in it (nothing needs to start at zero: every
 * entry the reduction reads is stored earlier in the same call).  Calls that OVERLAP on the GPU -- different streams,
 * threads -- race on the array: serialise them.  The first call on a device does the one-off set-up (symbol address,
 * shared-memory attribute, driver entry point, lazy module load): make it before a graph capture, not inside one. */
int qemb_gram_noalloc(const double *X_dev, int64_t ldx, int64_t naux, int64_t npair, double *G_dev, int64_t ldg,
                      void *stream);
int qemb_gram_noalloc_limits(int64_t *max_npair, int64_t *scratch_bytes);

#ifdef __cplusplus
}
#endif
#endif /* QEMB_B200_H */
