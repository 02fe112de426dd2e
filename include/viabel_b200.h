/*
 * viabel_b200 — C ABI of the B200-native Monte Carlo variational hot path.
 *
 * This header is the drop-in boundary.  The reference (kunalghosh/viabel) has no FFI: its
 * boundary is a set of Python callables.  Each entry point below sits directly underneath
 * one of those callables and cites the reference lines it replaces (paths are relative to
 * the reference checkout).  INTEGRATION.md shows the ctypes stub a maintainer of the
 * reference would add.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch types.  Every pointer whose name starts with
 *     d_ is DEVICE memory owned by the caller; h_ is host memory.
 *   - the library never allocates device memory: every op that needs scratch has a
 *     *_workspace_bytes query and takes (d_ws, ws_bytes).
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); nothing blocks.
 *   - return value: 0 ok, > 0 a cudaError_t, < 0 an argument error (VB_ERR_*).
 *   - scalars are returned through device pointers.
 *   - float64 everywhere unless a `dtype` argument says otherwise; accumulators are always
 *     float64.
 */
#ifndef VIABEL_B200_H
#define VIABEL_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VB_OK               0
#define VB_ERR_ARG         (-1)
#define VB_ERR_UNSUPPORTED (-2)
#define VB_ERR_WORKSPACE   (-3)
#define VB_ERR_ALIGN       (-4)
#define VB_ERR_TOO_LARGE   (-5)   /* problem size beyond the stated maximum of a fixed-scratch entry point */

enum { VB_F32 = 0, VB_F64 = 1 };
enum { VB_FAMILY_MF_GAUSS = 0, VB_FAMILY_MF_T = 1, VB_FAMILY_FR_T = 2 };
/* VB_OBJ_KLVI_PD: black_box_klvi_pd / _pd2 (viabel/vb.py:268-296): value -(mean log p - mean log q(x; lambda)) on the
 * drawn samples instead of the analytic entropy; same gradient as VB_OBJ_KLVI. */
enum { VB_OBJ_KLVI = 0, VB_OBJ_CHIVI = 1, VB_OBJ_KLVI_PD = 2 };
enum { VB_MODEL_FUNNEL = 0, VB_MODEL_MIXTURE = 1, VB_MODEL_EIGHT_SCHOOLS = 2,
       VB_MODEL_REGRESSION = 3, VB_MODEL_GAUSS = 4 };

/* Library/ABI version and the SM architecture the kernels were compiled for (100). */
int vb_abi_version(void);
int vb_compiled_sm(void);
const char* vb_error_string(int status);
/* Measured fp64 tensor-core peak of the current device in TFLOP/s (back-to-back DMMA.8x8x4 on register operands, every
 * SM, best of 4 launches of ~2 ms): the roofline denominator bench.py quotes the full-rank chain kernels against.  No
 * reference counterpart (measurement aid). */
size_t vb_measure_dmma_peak_workspace_bytes(void);
int vb_measure_dmma_peak(double* h_out_tflops, void* d_ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * Model plugins: log p(x) and grad log p(x) on the unconstrained space.
 * Replaces the `logdensity` argument of black_box_klvi / black_box_chivi
 * (viabel/vb.py:236,248) and the Stan bridge make_stan_log_density (viabel/vb.py:314-321).
 *   funnel            notebooks/funnel-distribution.ipynb:48-54
 *   mixture           notebooks/normal-mixture.ipynb:43
 *   eight schools ncp notebooks/eight_schools_ncp.stan:1-23 (+log tau Jacobian)
 *   regression        notebooks/robust-regression.ipynb:73-86
 *   gauss             BASELINE.json config 3 (not in the reference)
 * Handles are small host objects; device buffers they reference stay owned by the caller.
 * ---------------------------------------------------------------------------------- */
typedef struct vb_model vb_model;

int  vb_model_create_funnel(double log_sigma_stdev, vb_model** out);
int  vb_model_create_mixture(double loc, vb_model** out);
int  vb_model_create_eight_schools(const double* h_y, const double* h_sigma, int n_schools,
                                   vb_model** out);
/* d_X is (n_rows, dim) row-major, d_y is (n_rows); both of `dtype`. */
int  vb_model_create_regression(const void* d_X, const void* d_y, long long n_rows, int dim,
                                double df, double prior_sd, int dtype, vb_model** out);
/* The design matrix is model data: attach its tf32 (hi, lo) planes, made ONCE with
 * vb_split_f32(d_X, n_rows, dim, dim, d_X_planes) into vb_split_bytes(n_rows, dim) bytes that outlive the
 * model, so that the fp32 / 3xTF32 path (vb_model_logp_grad, dtype VB_F32) does not re-split X on every call
 * and its workspace shrinks by that amount.  NULL detaches.  fp32 data only. */
int  vb_model_regression_attach_split(vb_model* model, const float* d_X_planes);
/* Fused regression kernel (one persistent tcgen05 launch, "fp16 x 3" arithmetic: every fp32 operand as an fp16
 * (hi, lo) pair after an exact power-of-two scaling, fp32 accumulation; same 22 significant bits as 3xTF32).
 * vb_split_f16 writes the scaled planes of a (rows, cols) fp32 matrix plus a 256-byte header (the scale) into
 * vb_split_f16_bytes(rows, cols) bytes (256-byte aligned, must outlive the model); a regression model of
 * dim <= 128 with such planes attached evaluates vb_model_logp_grad(dtype f32) through the fused kernel.
 * Replaces the per-sample Stan log_prob / grad_log_prob loop of viabel/vb.py:301-321. */
size_t vb_split_f16_bytes(long long rows, long long cols);
int  vb_split_f16(const float* d_src, long long rows, long long cols, long long ld, void* d_dst_planes,
                  void* stream);
int  vb_model_regression_attach_split_f16(vb_model* model, const void* d_X_planes16);
/* d_mean (dim) and d_precision (dim, dim) symmetric, float64. */
int  vb_model_create_gauss(const double* d_mean, const double* d_precision, int dim,
                           vb_model** out);
/* Constants of the correlated-Gaussian plugin's fp32 tensor path (tf32 planes of float(P), bias = mean' P), made
 * once into vb_model_gauss_cache_bytes(dim) bytes (256-byte aligned, must outlive the model) instead of per call. */
size_t vb_model_gauss_cache_bytes(int dim);
int  vb_model_gauss_attach_cache(vb_model* model, float* d_cache, void* stream);
int  vb_model_kind(const vb_model* m);
int  vb_model_dim(const vb_model* m);
void vb_model_destroy(vb_model* m);

/* log p (n) [float64] and optionally grad (n, dim) [dtype] for x (n, dim) [dtype].
 * Plugin contract of viabel/vb.py:314-321. */
size_t vb_model_logp_grad_workspace_bytes(const vb_model* m, long long n, int dtype);
int vb_model_logp_grad(const vb_model* m, int dtype, const void* d_x, long long n,
                       double* d_out_logp, void* d_out_grad_or_null,
                       void* d_ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * Base draws.  Philox4x32-10, key = seed, counter = (global sample index, lane, call
 * offset): results do not depend on the launch geometry or on how samples are split
 * across GPUs (rank r passes sample_index0 = first global index it owns).
 *   mean-field Gaussian  eps (n, dim) ~ N(0,1)              viabel/vb.py:57
 *   mean-field t         eps (n, dim) ~ standard_t(df)      viabel/vb.py:151
 *   full-rank t          chi2 (n) ~ chisquare(df) first, then z (n, dim) ~ N(0,1)
 *                                                           viabel/vb.py:204,206
 * d_out_chi2 is only written for VB_FAMILY_FR_T.
 * ---------------------------------------------------------------------------------- */
int vb_base_draws(int family, double df, int dtype, long long n, int dim,
                  unsigned long long seed, unsigned int call_offset,
                  unsigned long long sample_index0,
                  void* d_out_eps, void* d_out_chi2_or_null, void* stream);
/* vb_base_draws for the full-rank family in fp32 storage that also writes the 3xTF32 operand planes of the normals
 * (viabel/vb.py:203-206 draws; the planes feed vb_fr_weighted_split). */
int vb_base_draws_fr_split(double df, long long n, int dim, unsigned long long seed, unsigned int call_offset,
                           unsigned long long sample_index0, float* d_out_z, float* d_out_chi2,
                           float* d_out_z_planes, void* stream);

/* ------------------------------------------------------------------------------------
 * Fused mean-field step: draw/load eps -> x = m + exp(l) eps -> log p, grad log p ->
 * (log q) -> weighted reductions -> (value, grad).  One launch.
 * Replaces the objective_and_grad closures of black_box_klvi (viabel/vb.py:236-245) and
 * black_box_chivi (viabel/vb.py:248-266) for the mean-field families
 * (viabel/vb.py:48-82, 140-182) and the elementwise model plugins.
 *
 *   d_var_param   [2*dim] float64: (mean, log scale)
 *   n             samples this call (this rank) draws;  n_total = samples over all ranks
 *   d_base        (n, dim) of `dtype`, or NULL -> Philox(seed, call_offset, sample_index0)
 *   d_out_partials[VB_MF_NPART(dim)] float64, always written:
 *                 [0] running max M of the log weights (CHIVI; 0 for KLVI)
 *                 [1] sum_i w_i            (w = 1 for KLVI, exp(alpha (lw_i - M)) for CHIVI)
 *                 [2] sum_i log p(x_i)     (KLVI)  /  sum_i lw_i (CHIVI)
 *                 [3] number of samples
 *                 [4 .. 4+dim)        sum_i w_i g_ij
 *                 [4+dim .. 4+2dim)   sum_i w_i g_ij eps_ij
 *   d_out_value / d_out_grad: if non-NULL the kernel also finalises for n_total == n
 *                 (single rank); multi-rank callers pass NULL and call vb_mf_finalize
 *                 on the gathered partials.
 * ---------------------------------------------------------------------------------- */
#define VB_MF_NPART(dim) (2 * (dim) + 4)

size_t vb_mf_step_workspace_bytes(int dim);
int vb_mf_step(int family, double df, const vb_model* model, int objective, double alpha,
               int dtype, const double* d_var_param, long long n,
               const void* d_base_or_null,
               unsigned long long seed, unsigned int call_offset,
               unsigned long long sample_index0,
               double* d_out_value_or_null, double* d_out_grad_or_null,
               double* d_out_partials, void* d_ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * Device-resident AdaGrad loop over the fused mean-field step.
 * Replaces the Python loop of adagrad_optimize (viabel/vb.py:345-389, has_log_norm=False):
 * windowed sum of squared gradients (vb.py:368-373), the update
 * `param - rate * grad / sqrt(epsilon + accum)` (vb.py:374), the value history (vb.py:364) and
 * the iterates kept from iteration 3*n_iters/4 on (vb.py:375-376).  The iterate, the
 * gradient window and the histories stay in HBM; groups of `graph_iters` iterations are
 * replayed as one CUDA graph.  The update is evaluated with individually rounded fp64
 * operations in numpy's order, so for equal gradients it is bit-identical to vb.py:371-374.
 *
 *   d_var_param      [2*dim] float64, in: init_param, updated in place every iteration
 *   d_base_or_null   NULL -> Philox stream (d_seeds[i], d_offsets[i]) for iteration i;
 *                    else (n_iters, n_samples, dim) external base draws of `dtype`
 *   d_rates          [n_iters] float64 learning-rate schedule (viabel/vb.py:324-343)
 *   d_seeds/d_offsets[n_iters] per-iteration Philox (seed, call offset)
 *   d_value_history  [n_iters] float64;  d_param_history [n_iters - 3*n_iters/4][2*dim]
 *   d_state          vb_mf_adagrad_state_bytes(dim, window) bytes, 16-byte aligned, owned by
 *                    the plan until vb_mf_adagrad_destroy
 *   d_ws             vb_mf_step_workspace_bytes(dim)
 *   graph_iters      0: plain launches; k > 0: k iterations per CUDA graph replay
 *   fused            1: the last CTA of each step launch applies the update (one launch per iteration);
 *                    0: separate update launch (two launches per iteration)
 * vb_mf_adagrad_run enqueues the next `n_steps` iterations on `stream` (no host sync).
 * ---------------------------------------------------------------------------------- */
typedef struct vb_adagrad_plan vb_adagrad_plan;
size_t vb_mf_adagrad_state_bytes(int dim, int window);
int vb_mf_adagrad_create(int family, double df, const vb_model* model, int objective, double alpha,
                         int dtype, double* d_var_param, long long n_samples,
                         const void* d_base_or_null, int n_iters, int window, double epsilon,
                         const double* d_rates, const unsigned long long* d_seeds,
                         const unsigned int* d_offsets, double* d_value_history,
                         double* d_param_history, void* d_state, size_t state_bytes,
                         void* d_ws, size_t ws_bytes, int graph_iters, int fused, void* stream,
                         vb_adagrad_plan** out_plan);
/* The fork's RMSProp / Adam update rules (viabel/vb.py:448-462, 616-629) on the same device-resident loop:
 * rule = VB_OPT_RMSPROP or VB_OPT_ADAM.  d_param_history (n_iters, 2*dim) receives the PRE-update iterate of every
 * iteration (vb.py:464), d_value_history the objective values; d_bias1 / d_bias2 (Adam) = 1 - beta^(i+2) for
 * i = 0 .. n_iters-1 with beta1 = 0.9, beta2 = 0.999, made on the host.  d_state: vb_mf_adagrad_state_bytes(dim, 2)
 * bytes, zero-filled.  Run and destroy with vb_mf_adagrad_run / vb_mf_adagrad_destroy. */
enum { VB_OPT_ADAGRAD = 0, VB_OPT_RMSPROP = 1, VB_OPT_ADAM = 2 };
int vb_mf_optimizer_create(int rule, int family, double df, const vb_model* model, int objective,
                           double alpha, int dtype, double* d_var_param, long long n_samples,
                           const void* d_base_or_null, int n_iters, double epsilon,
                           const double* d_rates, const unsigned long long* d_seeds,
                           const unsigned int* d_offsets, const double* d_bias1_or_null,
                           const double* d_bias2_or_null, double* d_value_history,
                           double* d_param_history, void* d_state, size_t state_bytes,
                           void* d_ws, size_t ws_bytes, int graph_iters, int fused, void* stream,
                           vb_adagrad_plan** out_plan);
int vb_mf_adagrad_run(vb_adagrad_plan* plan, int n_steps, void* stream);
int vb_mf_adagrad_destroy(vb_adagrad_plan* plan);

/* Deterministic fixed-order merge of `n_ranks` partial vectors (laid out
 * [n_ranks][VB_MF_NPART(dim)]) followed by the (value, grad) formulas. */
int vb_mf_finalize(int family, int objective, double alpha, const double* d_var_param,
                   int dim, const double* d_partials, int n_ranks,
                   double* d_out_value, double* d_out_grad, void* stream);

/* Generic (unfused) mean-field reductions for models whose log p / grad are produced by
 * vb_model_logp_grad (regression, gauss): same partial layout as vb_mf_step.
 *   d_eps (n, dim) dtype; d_logp (n) float64; d_grad (n, dim) dtype. */
size_t vb_mf_reduce_workspace_bytes(long long n, int dim);
int vb_mf_reduce(int family, double df, int objective, double alpha, int dtype,
                 const double* d_var_param, long long n, int dim,
                 const void* d_eps, const double* d_logp, const void* d_grad,
                 double* d_out_value_or_null, double* d_out_grad_or_null,
                 double* d_out_partials, void* d_ws, size_t ws_bytes, void* stream);

/* get_samples_and_log_weights (notebooks/experiments.py:60-63) in one pass for the mean-field families and
 * the elementwise plugins: samples (n, dim) [dtype], log p (n) (optional) and
 * lw = log p(x) - log q(x; var_param) (n) float64, all left on the device for all_bounds / psislw. */
int vb_mf_samples_logweights(int family, double df, const vb_model* model, int dtype,
                             const double* d_var_param, long long n, const void* d_base_or_null,
                             unsigned long long seed, unsigned int call_offset,
                             unsigned long long sample_index0, void* d_out_samples,
                             double* d_out_logp_or_null, double* d_out_lw, void* stream);

/* Importance-weighted mean and covariance (notebooks/experiments.py:81-86, improve_with_psis):
 * w = exp(lw - max lw); mean = sum w x / sum w; cov = sum w (x - mean)(x - mean)' / sum w
 * (== np.cov(samples.T, aweights=w, ddof=0)).  d_x (n, dim) of `dtype`, d_log_weights (n) float64. */
size_t vb_weighted_moments_workspace_bytes(long long n, int dim);
int vb_weighted_moments(int dtype, const void* d_x, const double* d_log_weights, long long n, int dim,
                        double* d_out_mean, double* d_out_cov_or_null,
                        void* d_ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * Variational-family closures (viabel/vb.py:54-79, 148-179, 202-230).
 *   sample:      x (n, dim) = m + exp(l) * eps                      (mean field)
 *   logdensity:  log q(x; var_param) (n), all normalising constants kept
 * entropy / mean_and_cov / pth_moment are O(dim) host arithmetic on var_param and live in
 * the Python layer.
 * ---------------------------------------------------------------------------------- */
int vb_mf_sample(int dtype, const double* d_var_param, long long n, int dim,
                 const void* d_eps, void* d_out_x, void* stream);
int vb_mf_logdensity(int family, double df, int dtype, const double* d_var_param,
                     long long n, int dim, const void* d_x, double* d_out_logq, void* stream);

/* ------------------------------------------------------------------------------------
 * Full-rank Student-t family (viabel/vb.py:192-233, viabel/_distributions.py:8-38).
 * The O(dim^3) parameter chain (fold, eigh, Sylvester) is driven from the Python layer with
 * the dense kernels below; the per-sample work is:
 *   vb_fr_sample     x_i = mu + (z_i S) / sqrt(chi2_i/df)          viabel/vb.py:204-208
 *   vb_fr_rows       lw_i = log p_i - log q_i with q evaluated through the base draw
 *                    (maha_i = |z_i|^2 / (chi2_i/df)), running max
 *   vb_fr_weighted   g_mu = sum_i w_i g_i ;  G_S = Ztilde' diag(w) G  (dim x dim)
 * ---------------------------------------------------------------------------------- */
size_t vb_fr_sample_workspace_bytes(int dtype, long long n, int dim);
int vb_fr_sample(int dtype, const double* d_mu, const void* d_scale /* (dim,dim) dtype */,
                 double df, long long n, int dim, const void* d_z, const void* d_chi2,
                 void* d_out_x, void* d_ws, size_t ws_bytes, void* stream);
/* log q(x; mu, Sigma) for arbitrary x (n, dim): d_U (dim, dim) float64 = u * sqrt(pinv(s)) from
 * the eigendecomposition Sigma = u diag(s) u', log_pdet = sum log s.  viabel/_distributions.py:8-38. */
size_t vb_fr_logdensity_workspace_bytes(long long n, int dim);
int vb_fr_logdensity(int dtype, const double* d_mu, const double* d_U, double log_pdet, double df,
                     long long n, int dim, const void* d_x, double* d_out_logq,
                     void* d_ws, size_t ws_bytes, void* stream);
/* d_out_rows[4]: max lw, sum exp(alpha (lw - max)) (CHIVI) or n (KLVI), sum log p, sum lw.
 * d_out_lw (n) float64; d_out_w (n) float64 weights (1 or exp(alpha (lw - max))). */
size_t vb_fr_rows_workspace_bytes(long long n);
int vb_fr_rows(int objective, double alpha, double df, int dtype, long long n, int dim,
               const void* d_z, const void* d_chi2, const double* d_logp,
               double logq_const_minus_entropy,
               double* d_out_lw, double* d_out_w, double* d_out_rows,
               void* d_ws, size_t ws_bytes, void* stream);
/* G_S (dim, dim) float64 = sum_i (w_i / sqrt(chi2_i/df)) z_i' g_i ; g_mu (dim) = sum_i w_i g_i */
size_t vb_fr_weighted_workspace_bytes(long long n, int dim);
int vb_fr_weighted(int dtype, double df, long long n, int dim, const void* d_z,
                   const void* d_chi2, const double* d_w, const void* d_grad,
                   double* d_out_gmu, double* d_out_GS, void* d_ws, size_t ws_bytes,
                   void* stream);
/* The same with the 3xTF32 operand planes of z (vb_split_bytes(n, dim) bytes, layout of vb_split_f32) supplied by the
 * caller -- vb_base_draws_fr_split writes them together with the draws -- so that no split pass over z is made.
 * fp32 storage only. */
int vb_fr_weighted_split(double df, long long n, int dim, const float* d_z, const float* d_z_planes,
                         const float* d_chi2, const double* d_w, const float* d_grad,
                         double* d_out_gmu, double* d_out_GS, void* d_ws, size_t ws_bytes, void* stream);

/* Dense helper: C (m, n) = alpha * op(A) op(B) + beta * C, row-major, float64 (SIMT) or
 * float32 storage with 3xTF32 tcgen05 arithmetic.  transa/transb: 0 = as stored, 1 = transposed. */
size_t vb_gemm_workspace_bytes(int dtype, long long m, long long n, long long k);
int vb_gemm(int dtype, int transa, int transb, long long m, long long n, long long k,
            double alpha, const void* d_A, long long lda, const void* d_B, long long ldb,
            double beta, void* d_C, long long ldc, void* d_ws, size_t ws_bytes, void* stream);

/* Operands of the 3xTF32 tensor path are "split" matrices: two fp32 planes (tf32 hi part, tf32 lo
 * part) of shape (rows, pad4(cols)), plane 1 following plane 0.  vb_split_f32 produces them;
 * vb_gemm_tf32x3_split multiplies pre-split operands: C (m, n) fp32 = A B with
 *   A planes stored (m, k) [a_mn_major = 0] or (k, m) [a_mn_major = 1],
 *   B planes stored (n, k) [b_mn_major = 0] or (k, n) [b_mn_major = 1].
 * Serves the full-rank scale product z S (viabel/vb.py:208) and the design-matrix products. */
size_t vb_split_bytes(long long rows, long long cols);
int vb_split_f32(const float* d_src, long long rows, long long cols, long long ld,
                 float* d_dst_planes, void* stream);
int vb_gemm_tf32x3_split(int a_mn_major, int b_mn_major, long long m, long long n, long long k,
                         const float* d_A_planes, const float* d_B_planes,
                         float* d_C, long long ldc, void* stream);

/* ------------------------------------------------------------------------------------
 * O(d^3) parameter chain of the full-rank Student-t family as two persistent cooperative kernels (csrc/fr_chain.cu):
 * no eigendecomposition, no library call.  Replaces paragami's fold + scipy sqrtm (viabel/vb.py:185-189, 207) and
 * autograd's sqrtm VJP (solve_sylvester) + the assembly of value_and_grad's result (vb.py:236-266).
 *   vb_fr_chain_forward   var_param -> L (d, d), S (d, d) = sqrtm(L L') [scale_mode 0] or L' [1: x = mu + L ztilde],
 *                         optional fp32 copy of S, d_out_scal[4] = (sum of the log-diagonal = entropy, |L|_F^2,
 *                         an upper bound of sigma_max(L) that the backward's inverse iteration starts from, reserved)
 *   vb_fr_chain_backward  G_S = Ztilde' diag(w) G, g_mu = sum w g, rows = [max lw, sum w, sum log p, sum lw]; with
 *                         n_total = 0 the global sample count is read from rows[4] on the device (sharded callers
 *                         whose count travelled with the exchange of the sums)
 *                         -> d_out[1 + P] = (value, gradient) of KLVI (objective 0), CHIVI (1), path-derivative KLVI (2)
 *   vb_fr_chain_info      h_out8 = [polar, inverse, Smith iterations, status (0 = every iteration met its tolerance),
 *                         GEMMs executed by the forward, by the backward, symmetric (half-work) GEMMs among them:
 *                         forward, backward] of the last calls on a workspace (synchronises)
 * ---------------------------------------------------------------------------------- */
size_t vb_fr_chain_workspace_bytes(int dim);
int vb_fr_chain_forward(const double* d_var_param, int dim, int scale_mode, double* d_out_L, double* d_out_S,
                        float* d_out_S32_or_null, double* d_out_scal, void* d_ws, size_t ws_bytes, void* stream);
int vb_fr_chain_backward(int dim, int scale_mode, int objective, double alpha, double n_total,
                         const double* d_L, const double* d_S, const double* d_GS, const double* d_gmu,
                         const double* d_rows, const double* d_scal, double* d_out, void* d_ws, size_t ws_bytes,
                         void* stream);
int vb_fr_chain_info(const void* d_ws, int dim, int* h_out8, void* stream);

/* ------------------------------------------------------------------------------------
 * Bounds (viabel/bounds.py).
 *   vb_bounds_divergence  one streaming pass over lw (n):  max, sum exp(alpha (lw - max)),
 *                         its second moment, mean/variance of lw.   bounds.py:142-192
 *     d_out[8]: [0] max  [1] mean w  [2] std(w)/sqrt(n) (ddof 0)  [3] mean lw
 *               [4] std(lw)/sqrt(n)  [5] cubo  [6] n  [7] reserved
 *   vb_bounds_moments     two passes over samples (n, dim): column means, then
 *                         C2 = mean_i sum_j c_ij^2, C4 = mean_i sum_j c_ij^4 and optionally
 *                         the covariance (ddof 1).              bounds.py:130-135, 55-56
 * Multi-rank: *_partials variants return mergeable partials, see vb_bounds_merge_*.
 * ---------------------------------------------------------------------------------- */
size_t vb_bounds_divergence_workspace_bytes(long long n);
int vb_bounds_divergence(const double* d_lw, long long n, double alpha, double* d_out,
                         void* d_ws, size_t ws_bytes, void* stream);
/* partials[6]: max, sum w, sum w^2 (both relative to max), count, mean lw, M2 lw */
int vb_bounds_divergence_partials(const double* d_lw, long long n, double alpha,
                                  double* d_out_partials, void* d_ws, size_t ws_bytes,
                                  void* stream);
int vb_bounds_divergence_merge(const double* d_partials, int n_ranks, double alpha,
                               double* d_out, void* stream);

/* vb_bounds_divergence_static: NO COUNTERPART IN THE REFERENCE -- added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  Same result as vb_bounds_divergence (same two kernels, same d_out[8]); the scratch
 * (per-CTA partials, 1184 x 6 doubles = 56 832 bytes) is a file-scope `static __device__` array of the library
 * instead of a caller workspace, so the call allocates nothing (no cudaMalloc / cudaFree / allocating library
 * routine) and takes no workspace arguments.  Every GPU that loads the library holds those 56 832 bytes until the
 * process exits.  Limit: n <= VB_BOUNDS_DIVERGENCE_STATIC_MAX_N (2^31 weights), larger n returns
 * VB_ERR_TOO_LARGE ("problem larger than the fixed scratch of this entry point allows").  All calls on one
 * device share the array and it keeps the last call's contents (nothing has to start at zero: each call
 * overwrites every slot it reads).  Calls that OVERLAP on the GPU (two streams, two host threads) race on it:
 * serialise them or use vb_bounds_divergence with separate workspaces. */
#define VB_BOUNDS_DIVERGENCE_STATIC_MAX_N 2147483648LL
long long vb_bounds_divergence_static_max_n(void);
int vb_bounds_divergence_static(const double* d_lw, long long n, double alpha, double* d_out,
                                void* stream);

size_t vb_bounds_moments_workspace_bytes(long long n, int dim);
int vb_bounds_colsum(int dtype, const void* d_x, long long n, int dim, double* d_out_sum,
                     void* d_ws, size_t ws_bytes, void* stream);
/* d_mean (dim): the GLOBAL column means.  d_out[2] = sum_i sum_j c^2, sum_i sum_j c^4 (raw sums);
 * d_out_scatter_or_null (dim, dim) = sum_i c_i c_i' (raw). */
int vb_bounds_central(int dtype, const void* d_x, long long n, int dim, const double* d_mean,
                      double* d_out, double* d_out_scatter_or_null,
                      void* d_ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * PSIS (notebooks/psis.py).
 *   vb_psis_smooth  psislw for one column (psis.py:112-208): max, exact order-statistic
 *                   select, tail compaction in index order, tail sort by (value, index),
 *                   GPD fit (psis.py:211-331), gpinv + log + scatter + clamp
 *                   (psis.py:187-198), log-sum-exp normalise (psis.py:200).
 *     d_out_diag[8]: [0] k  [1] sigma  [2] xcutoff  [3] n2 (tail size)  [4] smoothed (0/1)
 *                    [5] max(lw)  [6] logsumexp after smoothing  [7] candidates gathered
 *     d_out_tail_idx_or_null: int64, capacity vb_psis_tail_capacity(n, Reff); entry r is the
 *                    original index of the r-th smallest tail element, i.e. `tailinds[x2si]`
 *                    of psis.py:174,182,196 (ties broken by ascending index).
 *   The call synchronises `stream` internally (it reads back candidate counts).
 *   vb_gpd_fit      gpdfitnew on an ascending-sorted array (psis.py:211-331) -> (k, sigma)
 *   vb_gpd_inv      gpinv (psis.py:334-376)
 *   vb_logsumexp    sumlogs (psis.py:379-395)
 * ---------------------------------------------------------------------------------- */
long long vb_psis_tail_capacity(long long n, double reff);
size_t vb_psis_smooth_workspace_bytes(long long n, double reff);
int vb_psis_smooth(const double* d_lw_in, double* d_lw_out, long long n, double reff,
                   double* d_out_diag, long long* d_out_tail_idx_or_null,
                   void* d_ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------
 * Batched columns: psislw on an (n, m) matrix (notebooks/psis.py:136-137,163 -- one independent
 * smoothing per column) and the leave-one-out reduction of psisloo (notebooks/psis.py:68-109).
 * One CTA per column, the whole column in shared memory (n <= vb_psis_columns_max_n() = 8192;
 * longer columns: call vb_psis_smooth per column).
 *   d_lw_cols / d_out_cols   (m, n) float64: column j of the reference's (n, m) array is row j here
 *   d_out_k[m]               k-hat per column (inf where the tail has <= 4 points, psis.py:177-179)
 *   d_out_loo_or_null[m]     log sum_i exp(out_ij - lw_ij): with lw = -log_lik this is `loos` of
 *                            psisloo (psis.py:104-107)
 *   d_status_or_null[m]      scratch ints; when given, columns longer than 2048 draws first try the selection
 *                            mode (sample pivot -> <= 2048 candidates sorted instead of the whole column) and
 *                            only the columns it flags are redone with the full in-CTA sort
 * ---------------------------------------------------------------------------------- */
long long vb_psis_columns_max_n(void);
int vb_psis_smooth_columns(const double* d_lw_cols, double* d_out_cols, long long n, long long m,
                           double reff, double* d_out_k, double* d_out_loo_or_null, int* d_status_or_null,
                           void* stream);

/* Staged PSIS for log weights sharded by contiguous index ranges (rank r owns global indices
 * [index0, index0 + n_local)); vb_psis_smooth is the same sequence on one GPU.  Candidates are
 * 16-byte (key, global index) pairs; between the stages the caller exchanges a few small buffers:
 *   S1 vb_psis_sample   VB_PSIS_SAMPLE strided (ordered key of x, index) pairs of this shard
 *      -> all-gather the samples
 *   S2 vb_psis_pivot    sort them (capacity: power of two >= 4096, >= count); pivot_x = the value
 *                       with ceil(frac * count) sampled values above it (-inf when frac >= 1)
 *   S3 vb_psis_filter   ONE pass over the shard: x >= pivot_x -> candidate (raw bits of x, index),
 *                       otherwise body += exp(x - pivot_x); d_out_body_max = [body, max candidate]
 *      -> read d_out_count (host), all-reduce(max) the local maxima
 *   S4 vb_psis_prepare  keys := ~bits(M - x) (the reference sorts the shifted weights,
 *                       psis.py:165-167); d_out_term0 = body * exp(pivot_x - M)
 *   S4b vb_psis_refine  keep about `keep` largest local candidates; the rest add exp(-(M - x)) to
 *                       d_out_term   -> all-gather candidates and the per-rank [term0, term]
 *   S5 vb_psis_tail     replicated: sort, exact cutoff (psis.py:157,169-176), GPD fit, smoothing,
 *                       log-sum-exp; fills d_out_diag like vb_psis_smooth
 *   S6 vb_psis_write    out = (x - M) - lse over the shard + the smoothed tail entries it owns
 */
#define VB_PSIS_SAMPLE 4096
int vb_psis_sample(const double* d_lw, long long n_local, long long index0, void* d_out_sample,
                   void* stream);
int vb_psis_pivot(void* d_sample, long long count, long long capacity, double frac,
                  double* d_out_pivot_x, void* stream);
size_t vb_psis_filter_workspace_bytes(long long n_local);
int vb_psis_filter(const double* d_lw, long long n_local, long long index0,
                   const double* d_pivot_x, void* d_out_cand, long long cap,
                   unsigned long long* d_out_count, double* d_out_body_max,
                   void* d_ws, size_t ws_bytes, void* stream);
int vb_psis_prepare(void* d_cand, long long count, const double* d_M, const double* d_pivot_x,
                    const double* d_body_pre, double* d_out_term0, void* stream);
size_t vb_psis_refine_workspace_bytes(void);
int vb_psis_refine(const void* d_cand, long long count, const double* d_M,
                   const double* d_pivot_x, long long keep, void* d_out_cand,
                   unsigned long long* d_out_count, double* d_out_term,
                   void* d_ws, size_t ws_bytes, void* stream);
/* Global refinement for world > 1 (the bins depend only on the all-reduced maximum and the common pivot):
 * vb_psis_refine_hist counts this rank's candidates into d_hist[vb_psis_hist_bins()], the caller all-reduces
 * (sum) the counters, vb_psis_refine_select keeps the candidates above the ONE threshold that leaves
 * >= keep_global of them over all ranks; the dropped ones add exp(-u) to *d_out_term. */
int vb_psis_hist_bins(void);
int vb_psis_refine_hist(const void* d_cand, long long count, const double* d_M, const double* d_pivot_x,
                        unsigned int* d_hist, void* d_ws, size_t ws_bytes, void* stream);
int vb_psis_refine_select(const void* d_cand, long long count, const double* d_M, const double* d_pivot_x,
                          const unsigned int* d_hist_global, long long keep_global, void* d_out_cand,
                          unsigned long long* d_out_count, double* d_out_term, void* d_ws, size_t ws_bytes,
                          void* stream);
size_t vb_psis_tail_workspace_bytes(long long count, long long K1);
int vb_psis_tail(const void* d_cands, long long count, long long K1, const double* d_M,
                 const double* d_pivot_x, const double* d_terms, int n_terms,
                 double* d_out_diag, long long* d_out_tail_idx_or_null,
                 void* d_ws, size_t ws_bytes, void* stream);
int vb_psis_write(const double* d_lw, double* d_out, long long n_local, long long index0,
                  const double* d_M, long long count, long long K1, const void* d_tail_ws,
                  void* stream);
/* S6 / the single-GPU composite with the one-pass statistics of divergence_bound(alpha = 2) (bounds.py:169-176) of the
 * RAW weights taken in the same final pass (the pass is HBM-bound and knows the maximum, so all_bounds after psislw
 * costs no third read of the weights): d_out_div_partials[6] has the layout of vb_bounds_divergence_partials. */
/* S5 with the candidate count in device memory (at most `cap` entries in d_cands): no host read between the
 * stages.  d_count1 / cap1 (optional): count and capacity of the filter pass that produced the candidates, checked
 * on the device too.  When a precondition fails (fewer than K1 + 1 candidates, a buffer overflow, a refinement bin
 * too large for the per-warp ranking) the remaining kernels are inert and d_out_diag[7] = -(status code): run the
 * host-driven vb_psis_tail then.  Workspace: vb_psis_tail_workspace_bytes(cap, K1). */
int vb_psis_tail_dev(const void* d_cands, long long cap, const unsigned long long* d_count, long long K1,
                     const double* d_M, const double* d_pivot_x, const double* d_terms, int n_terms,
                     const unsigned long long* d_count1, long long cap1, const int* d_status_in_or_null,
                     double* d_out_diag, long long* d_out_tail_idx_or_null,
                     void* d_ws, size_t ws_bytes, void* stream);
/* Device-driven stages of the SHARDED protocol (no rank waits on a host read; replaces the count exchanges around
 * notebooks/psis.py:165-200).  vb_psis_prepare_hist_dev = S4 + this rank's refinement histogram with the candidate
 * count in device memory; vb_psis_refine_select_dev cuts at the global threshold straight into a fixed-capacity
 * exchange slot of capx + 3 rows of 16 bytes: candidates | (term0, term1) | (kept count, flags) | (shard size,
 * filter count); vb_psis_gather_compact turns the all-gathered slots (world, capx + 3) into the dense candidate
 * list in rank order, the body terms in rank order, the total and a status (any rank's flag, a changed shard size,
 * a total beyond capx) that vb_psis_tail_dev takes as d_status_in. */
int vb_psis_prepare_hist_dev(void* d_cand, long long cap, const unsigned long long* d_count, const double* d_M,
                             const double* d_pivot_x, const double* d_body_pre, double* d_out_term0,
                             unsigned int* d_hist, void* d_ws, size_t ws_bytes, void* stream);
int vb_psis_refine_select_dev(const void* d_cand, long long cap, const unsigned long long* d_count,
                              const double* d_M, const double* d_pivot_x, const unsigned int* d_hist_global,
                              long long keep_global, void* d_send, long long capx, double* d_terms,
                              long long n_local, void* d_ws, size_t ws_bytes, void* stream);
int vb_psis_gather_compact(const void* d_recv, int world, long long capx, const long long* d_sizes_or_null,
                           void* d_out_cands, unsigned long long* d_out_total, double* d_out_terms,
                           int* d_out_status, void* stream);
size_t vb_psis_div_workspace_bytes(void);
int vb_psis_write_div(const double* d_lw, double* d_out, long long n_local, long long index0,
                      const double* d_M, long long count, long long K1, const void* d_tail_ws,
                      double* d_out_div_partials, void* d_ws, size_t ws_bytes, void* stream);
int vb_psis_smooth_div(const double* d_lw_in, double* d_lw_out, long long n, double reff,
                       double* d_out_diag, long long* d_out_tail_idx_or_null,
                       double* d_out_div_partials_or_null, void* d_ws, size_t ws_bytes, void* stream);
/* vb_psis_smooth_div whose small results also reach the HOST in the one synchronisation the call has anyway:
 * h_out18 = [diag (8) | the alpha = 2 statistics as vb_bounds_divergence writes them (8; zeros when no partials were
 * requested) | path taken: 0 device-driven, 1 host-driven | reserved].  Serves psislw followed by all_bounds on the
 * same weights (notebooks/experiments.py:69,159-166) without a second device read. */
int vb_psis_smooth_host(const double* d_lw_in, double* d_lw_out, long long n, double reff,
                        double* d_out_diag, long long* d_out_tail_idx_or_null,
                        double* d_out_div_partials_or_null, double* h_out18, void* d_ws, size_t ws_bytes, void* stream);

size_t vb_gpd_fit_workspace_bytes(long long n);
int vb_gpd_fit(const double* d_x_sorted, long long n, double* d_out_k_sigma,
               void* d_ws, size_t ws_bytes, void* stream);
/* ascending sort of a float64 array (gpdfitnew(sort=True), psis.py:254-265) */
size_t vb_sort_f64_workspace_bytes(long long n);
int vb_sort_f64(const double* d_x, long long n, double* d_out, void* d_ws, size_t ws_bytes, void* stream);
/* gpdfitnew(return_quadrature=True) (psis.py:314-331): also the m = 30 + floor(sqrt(n)) quadrature points ks (weak
 * prior applied) and their unnormalised weights; the caller drops weights below 10 eps and normalises (psis.py:300-305). */
int vb_gpd_fit_quadrature(const double* d_x_sorted, long long n, double* d_out_k_sigma, double* d_out_ks,
                          double* d_out_w, void* d_ws, size_t ws_bytes, void* stream);
int vb_gpd_inv(const double* d_p, long long n, double k, double sigma, double* d_out,
               void* stream);
/* d_dst (cols, rows) = scale * transpose of d_src (rows, cols): the layout change of psislw / psisloo for 2-D input
 * (psis.py:101 negates: scale = -1). */
int vb_transpose_scale_f64(const double* d_src, long long rows, long long cols, double scale, double* d_dst, void* stream);
size_t vb_logsumexp_workspace_bytes(long long n);
int vb_logsumexp(const double* d_x, long long n, double* d_out, void* d_ws, size_t ws_bytes,
                 void* stream);

/* lw_out -= lse (psis.py:200) for a shard whose log-sum-exp was merged across ranks. */
int vb_psis_normalise(double* d_lw_out, long long n_local, double lse, void* stream);

/* max of a float64 array (psis.py:165, bounds.py:169) */
size_t vb_max_workspace_bytes(long long n);
int vb_max(const double* d_x, long long n, double* d_out, void* d_ws, size_t ws_bytes,
           void* stream);

/* ------------------------------------------------------------------------------------
 * Convergence diagnostics of the R-hat-monitored optimisers (viabel/functions.py, called from
 * rmsprop_IA_optimize_with_rhat / adam_IA_optimize_with_rhat, viabel/vb.py:486-522).
 *   vb_rhat_windows     compute_R_hat (functions.py:8-31) on n_windows iteration windows of a
 *                       (n_chains, n_iters, n_params) fp64 history at once: window w covers the
 *                       iterations [d_windows[2w], d_windows[2w] + d_windows[2w+1]) of every chain
 *                       (the caller guarantees even lengths >= 2 inside the history), each chain
 *                       cut in two halves.  One call serves compute_R_hat (1 window),
 *                       compute_R_hat_adaptive_numpy (functions.py:44-52: consecutive windows) and
 *                       compute_R_hat_halfway (functions.py:54-65: second halves of growing
 *                       prefixes).  Outputs (n_windows, n_params): var_hat and R-hat = sqrt(var_hat).
 *   vb_iterate_average  stochastic_iterate_averaging (functions.py:68-77): out[t - start] =
 *                       mean(est[start..t]) for t in [start, n_iters); est has row stride ld, so a
 *                       column block of the history (the means, or the log scales) is passed in
 *                       place.  Summation order is that of numpy's cumsum.
 * ---------------------------------------------------------------------------------- */
size_t vb_rhat_workspace_bytes(int n_chains, long long n_params, int n_windows);
int vb_rhat_windows(const double* d_chains, int n_chains, long long n_iters, long long n_params,
                    const long long* d_windows, int n_windows, double* d_out_var_hat,
                    double* d_out_rhat, void* d_ws, size_t ws_bytes, void* stream);
int vb_iterate_average(const double* d_est, long long n_iters, long long n_params, long long ld,
                       long long start, double* d_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VIABEL_B200_H */
