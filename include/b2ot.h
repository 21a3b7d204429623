/* b2ot — C ABI of the B200-native entropic Sinkhorn hot path (libb2ot.so).
 *
 * Drop-in boundary for PythonOT/POT's Sinkhorn path.  The reference has no FFI for this path (it
 * is pure Python over the `nx` array backend, ot/backend.py:245-262), so the entry points below
 * are what a binding of that path has to call; each one cites the reference lines it replaces
 * (paths relative to the reference checkout).  Conventions:
 *
 *   - every pointer named *_dev / K / M / a / b / u / v / workspace is a DEVICE pointer owned by
 *     the caller (the Python host allocates with torch); the library never allocates, frees or
 *     retains them beyond the lifetime of the handle they were given to;
 *   - matrices are row-major with an explicit leading dimension in ELEMENTS;
 *   - dtype is B2OT_F32 or B2OT_F64; `stream` is a cudaStream_t passed as void*;
 *   - work is enqueued on the caller's stream and is asynchronous unless stated otherwise;
 *   - return value 0 = ok, < 0 = error (message in b2ot_last_error(), thread-local);
 *   - solver status words: 0 running, 1 converged, 2 numerical guard fired, 3 max-iter.
 *
 * There is no CPU fallback anywhere in this library.
 */
#ifndef B2OT_H_
#define B2OT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2OT_F32 0
#define B2OT_F64 1

#define B2OT_MODE_KNOPP 0 /* scaling domain, ot/bregman/_sinkhorn.py:487-687  */
#define B2OT_MODE_LOG 1   /* log domain,     ot/bregman/_sinkhorn.py:690-918  */

#define B2OT_STATUS_RUNNING 0
#define B2OT_STATUS_CONVERGED 1
#define B2OT_STATUS_NUMERR 2
#define B2OT_STATUS_MAXITER 3

/* sweep implementation selector for b2ot_dense_create (`variant`) */
#define B2OT_VARIANT_AUTO 0    /* resident when K fits the grid's shared memory, else fused cluster sweep when the layout allows it, else split */
#define B2OT_VARIANT_FUSED 1   /* one pass over K per iteration: TMA-bulk staged row tiles, cluster/DSMEM row reduce */
#define B2OT_VARIANT_SPLIT 2   /* two passes (row kernel + column kernel); any shape / alignment */
#define B2OT_VARIANT_RESIDENT 3 /* small problems, both domains: K resident in the grid's shared memory, one cooperative
                                   launch per chunk of iterations (AUTO picks it when K fits and m <= 4096) */

/* ---- library ---- */
int b2ot_version(void);
/* "src=<first 16 hex digits of sha256 over pot_b200/csrc sources + headers, Makefile order> built=<date time>" */
const char* b2ot_build_info(void);
const char* b2ot_last_error(void);
int b2ot_device_info(int device, int* sm_count, int* cc_major, int* cc_minor);
/* number of kernels this library has launched in this process (bench.py's "gpu_launches") */
unsigned long long b2ot_launch_count(void);

/* ---- Gibbs kernel / scaled cost (once per solve) ----
 * K = exp(M / (-reg))   replaces nx.exp(M / (-reg)),  ot/bregman/_sinkhorn.py:621
 * Mr = -M / reg         replaces -M / reg,            ot/bregman/_sinkhorn.py:860
 * In-place (K == M, ldK == ldM) is allowed. */
int b2ot_kexp(const void* M, void* K, int64_t n, int64_t m, int64_t ldM, int64_t ldK, double reg, int dtype,
              void* stream);
int b2ot_negscale(const void* M, void* Mr, int64_t n, int64_t m, int64_t ldM, int64_t ldMr, double reg, int dtype,
                  void* stream);

/* ---- dense Sinkhorn solver (scaling and log domain) ----
 * One handle = one solve on one device (or one row block of a row-sharded solve).
 * Replaces the loops ot/bregman/_sinkhorn.py:626-670 (knopp) and :880-906 (log).
 *
 *   mode KNOPP: K = exp(-M/reg) (n x m), a (n), b (m); u, v are scalings.
 *   mode LOG  : K = -M/reg      (n x m), a = log(a) (n), b (m), logb = log(b) (m); u, v are log-scalings.
 *
 * Row sharding: `n` is the number of LOCAL rows; b, v are replicated.  When an all-reduce hook is
 * installed the per-iteration column sums (m+1 values of `dtype`, the last one carrying the
 * numerical-guard flag) are summed across ranks through it, on `stream`, once per iteration.
 */
typedef struct b2ot_dense b2ot_dense;
#define B2OT_OP_SUM 0
#define B2OT_OP_MAX 1
/* in-place all-reduce of `count` elements of `dtype` at buf_dev across the row-shard ranks, enqueued on `stream` */
typedef int (*b2ot_allreduce_fn)(void* buf_dev, int64_t count, int dtype, int op, void* stream, void* ctx);

size_t b2ot_dense_workspace_bytes(int64_t n, int64_t m, int dtype, int max_iter, int check_every);

int b2ot_dense_create(b2ot_dense** handle, int mode, int dtype, int64_t n, int64_t m, int64_t ld, const void* K,
                      const void* a, const void* b, const void* logb, void* workspace, size_t workspace_bytes,
                      int max_iter, int check_every, double stop_thr, int variant, void* stream);
/* switch a KNOPP handle (created with check_every = 1) to the UNBALANCED scaling iteration of
 * ot.unbalanced.sinkhorn_knopp_unbalanced (ot/unbalanced/_sinkhorn.py:749-787): per iteration
 * u = (a / K v)^fi_1 (inside the one fused sweep), v = (b / K^T u)^fi_2, fi = reg_m / (reg_m + reg) (:743-744);
 * K is the caller's exp(-M/reg) * c (:736-740).  Guard: K^T u == 0 or nan / inf in u, v -> previous (u, v), status
 * NUMERR (:760-772).  Error every iteration: (|u-uprev|_inf / max(|u|_inf, |uprev|_inf, 1) + same for v) / 2, stop
 * when err < stop_thr (:774-787).  b2ot_dense_init then only loads (u0, v0).
 * Row-sharded runs (b2ot_dense_set_allreduce / b2ot_dense_set_comm, before or after this call): per iteration one SUM
 * all-reduce of the m column sums + the guard flag and one MAX all-reduce of the three u statistics of the error (the
 * error couples all rows); every rank takes the v step and the stop decision on identical data. */
int b2ot_dense_set_unbalanced(b2ot_dense* handle, double fi_1, double fi_2);
/* switch a KNOPP handle to the STABILIZED scaling iteration of ot.bregman.sinkhorn_stabilized
 * (ot/bregman/_sinkhorn.py:1105-1382).  The iteration itself is the balanced one (v = b / K^T u, u = a / K v, :1270-1271,
 * one fused sweep) on K = exp(-(M - alpha - beta)/reg) (:1256-1258), which the caller builds before b2ot_dense_create.
 * In addition, every iteration, on the device and without a host round trip:
 *   - the finalize kernel tests max|u|, max|v| against tau (:1274);
 *   - when exceeded, gated kernels fold the scalings into the potentials (alpha_r += log u, beta_r += log v, i.e.
 *     alpha += reg log u, :1281), reset u = 1/n, v = 1/m (:1283-1287), rewrite K = exp(Mr + alpha_r + beta_r) over the
 *     handle's K buffer (:1288) and re-seed the column pass, all in the enqueued iteration stream;
 *   - error every `check_every` (= print_period) iterations = || sum_i Gamma_ij - b ||_2 (:1303-1304), on the reset
 *     scalings when the iteration absorbed, like the reference; stop test `err <= stop_thr` (:1316).
 * Mr       : -M/reg (n x m, leading dimension ldMr), read by every rebuild
 * alpha_r  : n DOUBLES alpha/reg, beta_r: m DOUBLES beta/reg (device, caller-owned, updated in place; double for both
 *            dtypes: they cancel against Mr at magnitude 1/reg and grow by one log per absorption)
 * The K buffer given to b2ot_dense_create is OVERWRITTEN by the rebuilds.  After the solve the exported (u, v) and
 * (alpha_r, beta_r) give logu = alpha_r + log u, logv = beta_r + log v (:1335-1336).  Single device only. */
int b2ot_dense_set_stabilized(b2ot_dense* handle, double tau, const void* Mr, int64_t ldMr, void* alpha_r,
                              void* beta_r);
/* number of absorptions so far (synchronises the stream) */
int b2ot_dense_absorb_count(b2ot_dense* handle, int* count);
/* Several target histograms on ONE kernel matrix (2-D b: u is n x H, v is m x H, row-major, H = b.shape[1]): one
 * iteration = two passes over K whatever H.  b2ot_multi_scaling_step enqueues step `iter` (out of place; a: n values,
 * B: m x H) and the 16 doubles of `stats_dev`:
 *     [0] max|u_new| [1] max|v_new| [2] max|u_new - u_prev| [3] max|u_prev| [4] max|v_new - v_prev| [5] max|v_prev|
 *     [6] u_new contains nan [7] v_new contains nan [8] sum (v_prev * K^T u_prev - B)^2 (KNOPP) [9] some K^T u == 0
 *     [10] inf in u_new [11] inf in v_new        (the maxima skip nan entries)
 *     [12] status 0 running / 1 converged / 2 numerical guard  [13] iteration at which the solve ended
 *     [14] number of errors written to err_hist_dev  [15] last error
 * mode B2OT_MULTI_STATS      v_new = B / (K^T u_prev), u_new = a / (K v_new); statistics only, the host decides - the loop
 *                            body of sinkhorn_stabilized with a 2-D b (ot/bregman/_sinkhorn.py:1230-1234, :1270-1271; the
 *                            tests at :1274, :1297-1301, :1319 read [0..7]; stats_dev may be null);
 * mode B2OT_MULTI_KNOPP      same update = sinkhorn_knopp with a 2-D b (:626-663).  Decisions on the device: guard
 *                            ([9] or nan / inf -> status 2 at `iter`, the iterate the step started from stays valid,
 *                            :633-645); when (iter - 1) % check_every == 0 the error of the PREVIOUS iterate
 *                            sqrt([8]) = || v * K^T u - b ||_F (:649-653, all histograms jointly) is recorded and
 *                            `err < stop_thr` sets status 1 at iter - 1 (that iterate = (u_prev, v_prev) of this step);
 * mode B2OT_MULTI_UNBALANCED u_new = (a / (K v_prev))^fi_1, v_new = (B / (K^T u_new))^fi_2 = sinkhorn_knopp_unbalanced with
 *                            a 2-D b (ot/unbalanced/_sinkhorn.py:749-787): guard -> status 2 at `iter` (previous iterate
 *                            valid, :760-772), else err = ([2] / max([0], [3], 1) + [4] / max([1], [5], 1)) / 2 over all
 *                            histograms (:774-781) is recorded every step, `err < stop_thr` sets status 1 at `iter`
 *                            (the NEW iterate is the result).
 * Once the status is non-zero every later step is a no-op, so the host enqueues steps in chunks and polls [12]; the
 * caller zeroes stats_dev before step 0.  fi_1 = fi_2 = 1 except for UNBALANCED.
 * scratch_dev >= b2ot_multi_scratch_bytes(n, m, H); deterministic.
 * b2ot_multi_absorb enqueues the joint absorption of sinkhorn_stabilized (:1275-1279) on alpha_r = alpha/reg (n doubles),
 * beta_r = beta/reg (m doubles): alpha_r[i] += max_h log u[i, h], beta_r[j] += max over ALL entries of log v (a scalar,
 * as written in the reference); u, v are kept (that branch has no reset).  The caller then rebuilds
 * K = exp(Mr + alpha_r + beta_r) with b2ot_dense_plan (mode LOG). */
#define B2OT_MULTI_STATS 0
#define B2OT_MULTI_KNOPP 1
#define B2OT_MULTI_UNBALANCED 2
size_t b2ot_multi_scratch_bytes(int64_t n, int64_t m, int64_t H);
int b2ot_multi_scaling_step(int dtype, int mode, const void* K, int64_t n, int64_t m, int64_t ld, int64_t H,
                            const void* a, const void* B, const void* u_prev, const void* v_prev, void* u_new,
                            void* v_new, int iter, int check_every, double stop_thr, double fi_1, double fi_2,
                            double* stats_dev, double* err_hist_dev, int err_capacity, void* scratch_dev,
                            void* stream);
int b2ot_multi_absorb(int dtype, int64_t n, int64_t m, int64_t H, const void* u, const void* v, double* alpha_r,
                      double* beta_r, void* stream);
int b2ot_dense_set_allreduce(b2ot_dense* handle, b2ot_allreduce_fn fn, void* ctx);
/* enqueue: load the initial scalings (u0: n, v0: m; log-scalings in LOG mode) and run the
 * column-only pass that seeds the first v update (:629 / :881 of the reference loop) */
int b2ot_dense_init(b2ot_dense* handle, const void* u0, const void* v0);
/* enqueue n_iters iterations (each: one sweep over K + one finalize); no host sync */
int b2ot_dense_iterate(b2ot_dense* handle, int n_iters);
/* measurement aid: run n_iters iterations with CUDA events (recorded on the solve's stream) around every
 * sweep launch; returns the average duration of the sweep alone and of a whole iteration (milliseconds) */
int b2ot_dense_profile(b2ot_dense* handle, int n_iters, float* sweep_ms_avg, float* iter_ms_avg);
/* synchronise the stream and read the device status word; also services the log-domain
 * exact-column fallback when the device asked for it */
int b2ot_dense_poll(b2ot_dense* handle, int* status, int* iter);
/* init + iterate until status != running, polling every `poll_every` iterations (<= 0: only at the end) */
int b2ot_dense_solve(b2ot_dense* handle, const void* u0, const void* v0, int poll_every, int* status, int* niter,
                     int* n_err);
/* copy the current scalings to u_out (n) / v_out (m) (device) and the error history to err_host (host doubles);
 * synchronises the stream */
int b2ot_dense_export(b2ot_dense* handle, void* u_out, void* v_out, double* err_host, int err_capacity, int* n_err);
/* what the handle resolved to: variant, cluster size, clusters in the grid, pipeline stages, vectors per thread */
int b2ot_dense_info(const b2ot_dense* handle, int* variant, int* cluster_size, int* n_clusters, int* stages,
                    int* vecs_per_thread);
int b2ot_dense_destroy(b2ot_dense* handle);

/* plan: KNOPP out = u_i K_ij v_j (:685-687), LOG out = exp(K_ij + u_i + v_j) (:915-918) */
int b2ot_dense_plan(int mode, int dtype, const void* K, const void* u, const void* v, void* out, int64_t n,
                    int64_t m, int64_t ldK, int64_t ldOut, void* stream);
/* loss <M, plan> without materialising the plan (ot/bregman/_sinkhorn.py:435-438); result: one double at
 * loss_dev; scratch_dev >= b2ot_dense_loss_scratch_bytes(n, m) */
size_t b2ot_dense_loss_scratch_bytes(int64_t n, int64_t m);
int b2ot_dense_loss(int mode, int dtype, const void* K, const void* M, const void* u, const void* v, int64_t n,
                    int64_t m, int64_t ldK, int64_t ldM, double* loss_dev, void* scratch_dev, void* stream);

/* ---- plan / objective epilogue: plan, value_linear, entropy / KL term and row marginal in ONE pass ----
 * Replaces the post-solve temporaries of ot.solve (ot/solvers.py:476-483: value_linear = sum M * plan,
 * value = value_linear + reg * sum plan log(plan + 1e-16)  or  + reg * kl_div(plan, a b^T), ot/backend.py:1542-1546)
 * and of sinkhorn_knopp_unbalanced's log (ot/unbalanced/_sinkhorn.py:785-797: cost, kl_div(plan, c), plan 1).
 *
 *   K (n x m) : Gibbs kernel (KNOPP) / scaled cost (LOG) the solve ran on, or NULL to rebuild it from the cost:
 *               LOG K = -M/reg (ot/bregman/_sinkhorn.py:860); KNOPP K = exp(M/(-reg)) (:621), times q_ij when
 *               B2OT_VALUE_KMULQ is set (ot/unbalanced/_sinkhorn.py:736-740) -- then M is read once, plan written once;
 *   Cref      : optional n x m reference matrix q of the KL term; NULL: q_ij = a_i b_j (needs a, b);
 *   plan_out  : n x m or NULL (values only);  rowsum_out : n values (plan 1) or NULL;
 *   terms     : OR of B2OT_VALUE_*;  values_dev: 3 doubles {sum M*P, sum P log(P + 1e-16), sum P log(P/q + 1e-16)}
 *               (terms not requested are written as 0);  scratch_dev >= b2ot_dense_plan_value_scratch_bytes(n, m). */
#define B2OT_VALUE_LINEAR 1
#define B2OT_VALUE_ENTROPY 2
#define B2OT_VALUE_KL 4
#define B2OT_VALUE_KMULQ 8
size_t b2ot_dense_plan_value_scratch_bytes(int64_t n, int64_t m);
int b2ot_dense_plan_value(int mode, int dtype, const void* K, const void* M, const void* Cref, double reg,
                          const void* u, const void* v, const void* a, const void* b, void* plan_out,
                          void* rowsum_out, int64_t n, int64_t m, int64_t ldK, int64_t ldM, int64_t ldC,
                          int64_t ldOut, int terms, double* values_dev, void* scratch_dev, void* stream);
/* column marginal out_j = sum_i P_ij (nx.sum(plan, 0), ot/unbalanced/_sinkhorn.py:795); deterministic;
 * scratch_dev >= b2ot_colsum_scratch_bytes(n, m) */
size_t b2ot_colsum_scratch_bytes(int64_t n, int64_t m);
int b2ot_colsum(int dtype, const void* P, int64_t n, int64_t m, int64_t ld, void* out, void* scratch_dev,
                void* stream);
/* b2ot_colsum that allocates nothing and takes no scratch pointer: the partials live in a file-scope device array of
 * 32 x B2OT_COLSUM_STATIC_MAX_COLS doubles (4 MiB per GPU that loads the library, held until exit).  No counterpart in
 * the reference: added for callers that cannot call cudaMalloc (CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  m > B2OT_COLSUM_STATIC_MAX_COLS is rejected (-1, message in
 * b2ot_last_error()).  Same kernels and summation order as b2ot_colsum: identical results.  All calls share the array:
 * calls that overlap on the GPU race; order them on one stream (or by events). */
#define B2OT_COLSUM_STATIC_MAX_COLS 16384
int b2ot_colsum_static(int dtype, const void* P, int64_t n, int64_t m, int64_t ld, void* out, void* stream);

/* ---- batched Sinkhorn (ot.batch): B independent n x m problems, each solved on chip ----
 * Replaces bregman_projection_batch (ot/batch/_utils.py:46-162; mode KNOPP, potentials = scalings f, g),
 * bregman_log_projection_batch (ot/batch/_utils.py:165-290; mode LOG, potentials = log-scalings u, v) and
 * the value epilogue of solve_batch (ot/batch/_linear.py:317-345, entropy_batch ot/batch/_utils.py:4-22).
 *
 *   X (B, n, m) contiguous: `transform` 0 = X is the kernel K (KNOPP) / log-kernel (LOG) itself,
 *                           1 = X is the cost M, K = exp(-M/reg) (_linear.py:323), 2 = X is M, K = -M/reg (:318);
 *   a (B, n), b (B, m); pa (B, n), pb (B, m) receive the potentials;
 *   T_out / logT_out (B, n, m), value_linear / value_reg (B doubles, device) may be null;
 *   value_linear = sum M * T needs Mval (B, n, m); value_reg = sum T log(T + 1e-16) (reg_type 1, "entropy")
 *   or sum T log(T / (a b^T) + 1e-16) (reg_type 2, "kl"): the caller forms value = value_linear + reg * value_reg;
 *   update order f then g; every 10th iteration (it % 10 == 0) the maximum over the batch of the L2 row-marginal
 *   violation is tested against tol with `<` on the device (:144-149, :273-278);
 *   *n_iters (host, may be null) = last iteration index, as the reference reports it.  Reading it synchronises
 *   the stream; with tol <= 0 the whole solve is a single launch and nothing synchronises.
 * Returns -4 when the shape does not fit the on-chip solver (see b2ot_batch_config). */
#define B2OT_BATCH_FORCE_EXACT 1 /* flags: always use the exact column pass of the log domain (test hook) */
size_t b2ot_batch_workspace_bytes(int64_t B);
int b2ot_batch_config(int mode, int dtype, int64_t n, int64_t m, int* cluster_size, int* smem_bytes,
                      int* cols_per_lane);
int b2ot_batch_solve(int mode, int dtype, int64_t B, int64_t n, int64_t m, const void* X, int transform, double reg,
                     const void* Mval, const void* a, const void* b, void* pa, void* pb, void* T_out, void* logT_out,
                     double* value_linear, double* value_reg, int reg_type, int max_iter, double tol, void* workspace,
                     size_t workspace_bytes, int* n_iters, int flags, void* stream);

/* ---- batched cost matrices and per-problem reductions of ot.batch ----
 * b2ot_batch_dist replaces ot.batch.dist_batch (ot/batch/_linear.py:178-233): out[b] = dist(X[b], Y[b]) for
 *   X (B, n, d), Y (B, m, d), out (B, n, m), all contiguous, in ONE launch (bound by the HBM write of out):
 *   SQEUCLIDEAN |x|^2 + |y|^2 - 2 x.y and EUCLIDEAN = its square root, both WITHOUT the clamp at 0 that ot.dist
 *   applies (dist_euclidean_batch, :54-85 -- tiny negative values / nan under the root are the reference's behaviour);
 *   MINKOWSKI (sum_d w_d |x - y|^p)^(1/p) (dist_lp_batch with q = 1, :19-51; w may be null);
 *   KL sum_d y log(y + eps) - sum_d y log(x + eps) (dist_kl_batch, logits_X = False, :88-120, eps 1e-10 there).
 * b2ot_batch_reduce: out[b] = sign * sum over the (n, m) plane of A*Bm (DOT: loss_linear_batch, :123-147) or of
 *   A*log(A + eps) (ENTROPY with sign -1: entropy_batch, ot/batch/_utils.py:4-22); `per` = n*m elements per problem,
 *   out = B values of `dtype`; fp64 accumulation in a fixed order (deterministic); one HBM read of the operands. */
#define B2OT_BMETRIC_SQEUCLIDEAN 0
#define B2OT_BMETRIC_EUCLIDEAN 1
#define B2OT_BMETRIC_MINKOWSKI 2
#define B2OT_BMETRIC_KL 3
#define B2OT_BREDUCE_DOT 0
#define B2OT_BREDUCE_ENTROPY 1
int b2ot_batch_dist(int dtype, const void* X, const void* Y, int64_t B, int64_t n, int64_t m, int d, int metric,
                    double p, const void* w, double eps, void* out, void* stream);
size_t b2ot_batch_reduce_scratch_bytes(int64_t B, int64_t per);
int b2ot_batch_reduce(int dtype, int op, const void* A, const void* Bm, int64_t B, int64_t per, double eps,
                      double sign, void* out, void* scratch_dev, void* stream);

/* ---- Gromov-Wasserstein contraction: strided, batched matrix product ----
 * C[b, f](i, j) = beta * C0[b](i, j) + alpha * sum_{s < Sf} sum_k A[b, f + s](i, k) * B[b, f + s](k, j)
 * for b < Zb, f < Zf; every operand addressed by ELEMENT strides (batch, feature, row, column; C0 and C have unit
 * column stride; C0 may be null, and may alias C), so transposes and the interleaved feature axis of (B, n, n, d)
 * structure tensors need no copies.  Replaces `-nx.dot(nx.dot(hC1, T), hC2.T)` of tensor_product
 * (ot/gromov/_utils.py:244-280) and the two einsums + feature sum of tensor_product_batch
 * (ot/batch/_quadratic.py:535-572).  Own register-tiled FMA kernel (gemm_kernels.cuh), deterministic. */
int b2ot_strided_gemm(int dtype, int64_t Zb, int64_t Zf, int64_t Sf, int64_t M, int64_t N, int64_t K, const void* A,
                      int64_t sAb, int64_t sAf, int64_t sAi, int64_t sAk, const void* B, int64_t sBb, int64_t sBf,
                      int64_t sBk, int64_t sBj, const void* C0, int64_t sC0b, int64_t sC0i, void* C, int64_t sCb,
                      int64_t sCf, int64_t sCi, double alpha, double beta, void* stream);

/* ---- online-cost (empirical) Sinkhorn: cost rebuilt from the point clouds, never materialised ----
 * Replaces the lazy branch of ot.bregman.empirical_sinkhorn (ot/bregman/_empirical.py:193-271): log-domain
 * updates f = log a - LSE_j(g_j - C_ij/reg), g = log b - LSE_i(f_i - C_ij/reg) (f first, :218-233) with
 * C = max(|x|^2 + |y|^2 - 2 x.y, 0) (ot/utils.py:292-299; metric EUCLIDEAN takes the square root, :300-301);
 * every 10th iteration ((it + 1) % 10 == 0) the L1 violation of the ROW marginal is tested with `<=` (:235-256),
 * on the device.  (f, g) are log-scalings, as the reference returns them.
 *
 *   Xs (n x d, leading dimension ldXs), Xt (m x d), a / loga (n), logb (m): device pointers kept by the handle;
 *   row sharding: n = LOCAL source points, Xt / g / logb replicated; with an all-reduce hook installed the column
 *   log-sum-exp partials are combined across ranks (one max + one sum all-reduce of m values per iteration) and
 *   the marginal check is summed (one double, every 10th iteration). */
#define B2OT_METRIC_SQEUCLIDEAN 0
#define B2OT_METRIC_EUCLIDEAN 1
#define B2OT_EMP_AUTO 0   /* tensor-core sweep (tcgen05, 3xTF32) when dtype/shape allow it, else the register-tiled sweep */
#define B2OT_EMP_SIMT 1   /* register-tiled FMA sweep: any d, fp32 and fp64 */
#define B2OT_EMP_TC 2     /* tcgen05 sweep: fp32, sqeuclidean */
typedef struct b2ot_emp b2ot_emp;
size_t b2ot_emp_workspace_bytes(int64_t n, int64_t m, int d, int dtype, int metric, int variant, int max_iter);
int b2ot_emp_create(b2ot_emp** handle, int dtype, int64_t n, int64_t m, int d, const void* Xs, int64_t ldXs,
                    const void* Xt, int64_t ldXt, const void* a, const void* loga, const void* logb, double reg,
                    int metric, void* workspace, size_t workspace_bytes, int max_iter, double stop_thr, int variant,
                    void* stream);
/* operand split of the tcgen05 sweep's cross term -2 x.y^T (call before b2ot_emp_init; no effect on the SIMT sweep):
 *   F16  (default) each fp32 coordinate = hi + lo fp16 halves under one power-of-two scale per cloud, 3 MMAs
 *        (hi*hi + hi*lo + lo*hi), fp32 accumulate: 22 significant bits, half the operand bytes of TF32;
 *   TF32 the 3xTF32 split (8-bit exponent: no dependence on the dynamic range inside a cloud). */
#define B2OT_EMP_OPERAND_F16 0
#define B2OT_EMP_OPERAND_TF32 1
int b2ot_emp_set_operand(b2ot_emp* handle, int operand);
int b2ot_emp_set_allreduce(b2ot_emp* handle, b2ot_allreduce_fn fn, void* ctx);
/* enqueue: load the initial potentials f0 (n), g0 (m) and compute the squared norms */
int b2ot_emp_init(b2ot_emp* handle, const void* f0, const void* g0);
/* enqueue n_iters iterations (row sweep, f update + pending check, column sweep, g update); no host sync */
int b2ot_emp_iterate(b2ot_emp* handle, int n_iters);
/* measurement aid (as b2ot_dense_profile): CUDA events around the two sweep launches of each iteration; average
 * duration of ONE sweep kernel and of a whole iteration, in milliseconds */
int b2ot_emp_profile(b2ot_emp* handle, int n_iters, float* sweep_ms_avg, float* iter_ms_avg);
int b2ot_emp_poll(b2ot_emp* handle, int* status, int* iter);
/* what the handle resolved to: B2OT_EMP_SIMT or B2OT_EMP_TC, and the reduction splits of the two sweeps */
int b2ot_emp_info(const b2ot_emp* handle, int* variant, int* nsplit_rows, int* nsplit_cols);
/* init + iterate until status != running, polling every `poll_every` iterations (<= 0: only at the end) */
int b2ot_emp_solve(b2ot_emp* handle, const void* f0, const void* g0, int poll_every, int* status, int* niter,
                   int* n_err);
int b2ot_emp_export(b2ot_emp* handle, void* f_out, void* g_out, double* err_host, int err_capacity, int* n_err);
int b2ot_emp_destroy(b2ot_emp* handle);

/* cost matrix out = dist(X, Y) (ot.dist / euclidean_distances, ot/utils.py:270-307, :400-526), metrics above;
 * scratch_dev >= b2ot_dist_scratch_bytes(n, m, dtype) */
size_t b2ot_dist_scratch_bytes(int64_t n, int64_t m, int dtype);
int b2ot_dist(int dtype, const void* X, const void* Y, int64_t n, int64_t m, int d, int64_t ldX, int64_t ldY, int metric,
              void* out, int64_t ldOut, void* scratch_dev, void* stream);
/* block of the lazy plan: out_ij = exp(f_i + g_j - C_ij/reg) for the n x m block whose points / potentials start at
 * the given pointers (get_sinkhorn_lazytensor, ot/bregman/_empirical.py:26-71); scratch as for b2ot_dist */
int b2ot_emp_plan_block(int dtype, const void* Xs, const void* Xt, const void* f, const void* g, int64_t n, int64_t m,
                        int d, int64_t ldXs, int64_t ldXt, double reg, int metric, void* out, int64_t ldOut,
                        void* scratch_dev, void* stream);
/* loss <C, plan> = sum_ij C_ij exp(f_i + g_j - C_ij/reg) reduced on the fly (lazy branch of empirical_sinkhorn2,
 * ot/bregman/_empirical.py:466-483); result: one double at loss_dev; scratch >= b2ot_emp_loss_scratch_bytes */
size_t b2ot_emp_loss_scratch_bytes(int64_t n, int64_t m, int dtype);
int b2ot_emp_loss(int dtype, const void* Xs, const void* Xt, const void* f, const void* g, int64_t n, int64_t m, int d,
                  int64_t ldXs, int64_t ldXt, double reg, int metric, double* loss_dev, void* scratch_dev, void* stream);

/* ---- row-sharded runs: one-shot all-reduce over NVLink peer memory (one process per GPU, one node) ----
 * The reference has no multi-device path (no NCCL / MPI call site anywhere in /root/reference/ot); this is the
 * exchange step of the row-sharded solvers: per iteration each rank stores its reduced column vector into a slot
 * of every peer's exchange buffer (peer-mapped with CUDA IPC) and publishes a flag; the solver's own finalize
 * kernel waits on the flags and sums the slots in rank order (bit-identical on every rank).  See csrc/comm.cuh.
 *
 *   b2ot_comm_alloc   cudaMalloc + zero an exchange buffer on the current device, return its IPC handle
 *   b2ot_comm_open    map a peer's buffer from its handle (exchanged by the host, e.g. torch.distributed)
 *   b2ot_comm_create  bufs[q] = rank q's buffer as seen from this process (own pointer at q == rank);
 *                     slot size = (bytes - 4096) / (2 * nranks), at most 8 ranks
 * The library allocates nothing else; these buffers are the one exception to "the caller owns all memory". */
#define B2OT_IPC_HANDLE_BYTES 64
typedef struct b2ot_comm b2ot_comm;
int b2ot_comm_alloc(size_t bytes, void** buf_dev, unsigned char* ipc_handle);
int b2ot_comm_free(void* buf_dev);
int b2ot_comm_open(const unsigned char* ipc_handle, void** peer_ptr);
int b2ot_comm_close(void* peer_ptr);
int b2ot_comm_create(b2ot_comm** comm, int nranks, int rank, void* const* bufs, size_t bytes);
int b2ot_comm_destroy(b2ot_comm* comm);
size_t b2ot_comm_slot_bytes(const b2ot_comm* comm);
/* in-place all-reduce (B2OT_OP_SUM / B2OT_OP_MAX) of count elements through the exchange buffers, on `stream` */
int b2ot_comm_allreduce(b2ot_comm* comm, void* buf_dev, int64_t count, int dtype, int op, void* stream);
/* attach to a solver handle (replaces the all-reduce hook); the communicator must outlive the handle */
int b2ot_dense_set_comm(b2ot_dense* handle, b2ot_comm* comm);
/* on = 1: the per-iteration column sums travel as tagged 8-byte words {value bits, sequence} stored straight into the
 * peers' slots and polled inside the finalize kernel (no put kernel, no flags, no system-scope fences); needs
 * slot_bytes >= roundup64(8 (m + 1)) + 2 (m + 1) sizeof(dtype).  Default off (put kernel + sequence flags). */
int b2ot_dense_set_comm_ll(b2ot_dense* h, int on);
int b2ot_emp_set_comm(b2ot_emp* handle, b2ot_comm* comm);

#ifdef __cplusplus
}
#endif
#endif /* B2OT_H_ */
