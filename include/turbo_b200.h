/* turbo_b200 -- C ABI of the B200-native GP surrogate + acquisition hot path.
 *
 * Drop-in boundary for mbway/turbo's surrogate / acquisition plugins.  The reference has no
 * native code and no FFI: the arithmetic behind its plugin interface runs inside scikit-learn /
 * SciPy.  Each entry point below therefore cites the reference (or sklearn) call it replaces;
 * INTEGRATION.md shows the ctypes binding and the plugin classes that sit on top.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host; the caller owns all
 *     DATA buffers (PyTorch tensors in the shipped host layer): the library allocates no device
 *     memory.  It does own, per device it is used on, a handful of CUDA streams and events (two
 *     internal streams of the acquisition pipeline, two of the batched fit; created on first use
 *     under a mutex, released by tb200_shutdown()) and sets per-device function attributes once;
 *   - calls may come from any host thread (turbo's Qt GUI runs the optimiser on a QThread,
 *     turbo/gui/qt.py:139-147); a call is not re-entrant on the same workspace;
 *   - all reals are IEEE FP64, indices int64, matrices C-contiguous row-major;
 *   - NP = tb200_padded_n(n) = n rounded up to 128.  Padded rows/cols of a factor are the
 *     identity, padded vector entries are 0, so padding never changes a result;
 *   - theta is scikit-learn's log-space hyper-parameter vector
 *       [log const (if has_const), log length_scale (d if ard else 1), log noise (if has_white)];
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *   - return value: 0 ok; <0 argument / CUDA error (text in tb200_last_error()).  Numerical
 *     failure of a Cholesky is reported per matrix in `info` (0 ok, k>0 = first non-positive
 *     pivot, 1-based, as LAPACK dpotrf / sklearn:_gpr.py:590-593 & :351-362).
 */
#ifndef TURBO_B200_H
#define TURBO_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { TB200_RBF = 0, TB200_MATERN12 = 1, TB200_MATERN32 = 2, TB200_MATERN52 = 3 };
enum { TB200_EI = 0, TB200_PI = 1, TB200_UCB = 2 };
enum { TB200_CAND_GIVEN = 0, TB200_CAND_UNIFORM = 1, TB200_CAND_LHS = 2 };
#define TB200_MAX_D 64

/* kernel family + layout of theta (sklearn:kernels.py Sum/Product/ConstantKernel/WhiteKernel/RBF/Matern) */
typedef struct {
    int kind;      /* TB200_RBF .. TB200_MATERN52 */
    int d;         /* input dimensions (<= TB200_MAX_D) */
    int ard;       /* 1: one length scale per dimension */
    int has_const; /* ConstantKernel factor present */
    int has_white; /* WhiteKernel term present */
} tb200_kernel_spec;

/* acquisition function of turbo/modules/acquisition_functions.py */
typedef struct {
    int acq;          /* TB200_EI (:336-358), TB200_PI (:225-247), TB200_UCB (:147-158) */
    double sign;      /* scale_factor: +1 'max', -1 'min' */
    double incumbent; /* incumbent_cost (EI/PI) */
    double param;     /* xi (EI/PI) or beta (UCB; +inf -> sigma alone) */
    double y_mean;    /* GaussianProcessRegressor._y_train_mean */
    double y_std;     /* GaussianProcessRegressor._y_train_std  */
} tb200_acq_spec;

/* where candidates come from (turbo/modules/naive_selectors.py:41-46 uniform, :65-83 LHS) */
typedef struct {
    int mode;            /* TB200_CAND_GIVEN: read Xc; else generate on device from (seed, global index) */
    uint64_t seed;
    int64_t m_total;     /* LHS: number of strata = size of the whole (unsharded) candidate set */
    double lo[TB200_MAX_D];
    double hi[TB200_MAX_D];
} tb200_cand_spec;

/* points no selected candidate may come too close to: the reference discards a Bayesian trial whose
 * point x has np.sum((X - x)**2, axis=1) <= tol for some row of X (turbo/utils.py:111-130 via
 * turbo/modules/fallback.py:49-54, turbo/optimiser.py:349-352) and draws a random one instead; with an exclusion set
 * the argmax skips such candidates in the same pass.  tol < 0 or n == 0 disables it (fallback.py:51-52). */
typedef struct {
    const double* X;  /* device, (n, d) row-major, UNSCALED coordinates (the space the candidates live in) */
    int n;
    double tol;       /* close_tolerance: largest SQUARED Euclidean distance considered too close */
    double* min_out;  /* optional device (m,) : every candidate's smallest squared distance; NULL = scratch */
} tb200_exclusion;

int tb200_version(void);
const char* tb200_last_error(void);
int tb200_padded_n(int n);
/* destroy the per-device streams and events the library created (they are re-created on the next use);
 * synchronises them first.  Call before unloading the library or resetting a device. */
int tb200_shutdown(void);
/* the library's single debug hook is the environment variable TB200_DEBUG ("key=value,key=value": fit_engine=dmma,
 * fit_gmax=7|8|9, fit_gmax_trtri, fit_gmax_kinv; csrc/common.cuh), read ONCE on first use -- never on the
 * call path.  tb200_debug_reload re-reads it (for tests that change it inside one process). */
int tb200_debug_reload(void);

/* bytes of scratch the calls below need */
size_t tb200_fit_workspace_bytes(int n, int d, int R, int with_grad);
size_t tb200_acquire_workspace_bytes(int n, int d, int64_t batch);
/* candidate batch size the library would pick for this n (multiple of 128) */
int64_t tb200_default_batch(int n);

/* log-marginal likelihood (and gradient w.r.t. theta) for R hyper-parameter vectors at once.
 * Replaces GaussianProcessRegressor.log_marginal_likelihood(theta, eval_gradient=True)
 * (sklearn:_gpr.py:541-656), called R x n_iter times by fit (:302-333) behind
 * turbo/modules/surrogates.py:315-318.  y_n is the normalised target (:275-280).
 * Non-PD member r: info[r] > 0, lml[r] = -inf, grad[r,:] = 0. */
int tb200_lml_batched(const double* X, const double* y_n, int n, const tb200_kernel_spec* spec_host,
                      const double* theta, int R, double jitter, int with_grad,
                      double* lml, double* grad, int* info, void* work, size_t work_bytes, void* stream);

/* final factorisation at one theta: what fit() keeps for predict (sklearn:_gpr.py:349-367).
 * Outputs: XsT (d x NP) = training inputs / length_scale, transposed; Linv (NP x NP) = inverse of
 * the Cholesky factor of K + jitter I (lower); alpha (NP) = K^-1 y_n; lml (1); Ldiag (NP) = diag(L). */
int tb200_factor(const double* X, const double* y_n, int n, const tb200_kernel_spec* spec_host,
                 const double* theta, double jitter, double* XsT, double* Linv, double* alpha,
                 double* Ldiag, double* lml, int* info, void* work, size_t work_bytes, void* stream);

/* fused predict + acquisition + argmax over candidates [index_base, index_base + m).
 * Replaces acq(random_x) + argsort of turbo/modules/auxiliary_optimisers.py:60-66, i.e.
 * EI/PI/UCB.FunctionInstance.__call__ -> ModelInstance.predict -> sklearn predict
 * (sklearn:_gpr.py:446-500).  Optional outputs (NULL to skip; indexed by local candidate 0..m-1):
 * mu_out, sigma_out, acq_out.  best_val/best_idx (1 each) are READ-MODIFY-WRITE: initialise to
 * (-inf, -1) with tb200_best_reset; ties keep the lowest global index; NaN never wins. */
int tb200_acquire(const double* Xc, int64_t m, int64_t index_base, const tb200_cand_spec* cand_host,
                  const double* XsT, const double* Linv, const double* alpha, int n,
                  const tb200_kernel_spec* spec_host, const double* theta,
                  const tb200_acq_spec* acq_host, double* mu_out, double* sigma_out, double* acq_out,
                  double* best_val, int64_t* best_idx, int64_t batch, void* work, size_t work_bytes,
                  void* stream);
int tb200_best_reset(double* best_val, int64_t* best_idx, void* stream);
/* tb200_acquire with an exclusion set (NULL excl = tb200_acquire).  acq_out still holds the value of an excluded
 * candidate; only the argmax skips it. */
int tb200_acquire_excl(const tb200_exclusion* excl_host, const double* Xc, int64_t m, int64_t index_base,
                       const tb200_cand_spec* cand_host, const double* XsT, const double* Linv, const double* alpha,
                       int n, const tb200_kernel_spec* spec_host, const double* theta, const tb200_acq_spec* acq_host,
                       double* mu_out, double* sigma_out, double* acq_out, double* best_val, int64_t* best_idx,
                       int64_t batch, void* work, size_t work_bytes, void* stream);
/* smallest squared Euclidean distance (and its row, optional) from each of m query rows to the n rows of X: the
 * batched form of turbo/utils.py:111-130 `close_to_any` (x close <=> min_out <= tol), summed in NumPy's order so the
 * comparison agrees bit for bit.  n == 0: +inf / -1 (close_to_any of an empty set is False). */
int tb200_min_sqdist(const double* Xq, int64_t m, const double* X, int n, int d, double* min_out, int64_t* arg_out,
                     void* stream);

/* INT8-tensor-core engine for the predictive variance (tcgen05.mma kind::i8 + TMEM, csrc/ozaki_i8.cuh).
 * Same contract and results as tb200_acquire (same sklearn lines); the O(n^2)-per-candidate contraction
 * V = L^-1 k* is assembled exactly from 7 byte slices per operand with INT32 accumulation.
 *   tb200_i8_prepare: slice the inverse factor once per model.  linvS: tb200_i8_factor_bytes(n) bytes;
 *     rowscale: 2*NP doubles (scale sf2*2^e_r per row, then the exponents e_r).
 *   gmax = 7: 21 slice products of 6 digit planes (truncation ~2^-46 of row-max*sf2 per term);
 *   gmax = 8: 28 slice products (truncation ~2^-54 of row-max*sf2 per term, >= 40x inside the parity floor);
 *   gmax = 9: 34 products, indistinguishable from an FP64 GEMM.  INT32 accumulation is exact for 8192 training
 *   points per pass (balanced s8 digits of L^-1); rows longer than that are accumulated in several passes that are
 *   summed in FP64 before squaring.  n <= TB200_I8_MAX_N (workspace sizing only). */
#define TB200_I8_MAX_N 32768
size_t tb200_i8_factor_bytes(int n);
int tb200_i8_prepare(const double* Linv, int n, const tb200_kernel_spec* spec_host, const double* theta,
                     uint8_t* linvS, double* rowscale, void* stream);
size_t tb200_acquire_i8_workspace_bytes(int n, int d, int64_t batch);
int tb200_acquire_i8(const double* Xc, int64_t m, int64_t index_base, const tb200_cand_spec* cand_host,
                     const double* XsT, const uint8_t* linvS, const double* rowscale, int gmax,
                     const double* alpha, int n, const tb200_kernel_spec* spec_host, const double* theta,
                     const tb200_acq_spec* acq_host, double* mu_out, double* sigma_out, double* acq_out,
                     double* best_val, int64_t* best_idx, int64_t batch, void* work, size_t work_bytes,
                     void* stream);
int tb200_acquire_i8_excl(const tb200_exclusion* excl_host, const double* Xc, int64_t m, int64_t index_base,
                          const tb200_cand_spec* cand_host, const double* XsT, const uint8_t* linvS,
                          const double* rowscale, int gmax, const double* alpha, int n,
                          const tb200_kernel_spec* spec_host, const double* theta, const tb200_acq_spec* acq_host,
                          double* mu_out, double* sigma_out, double* acq_out, double* best_val, int64_t* best_idx,
                          int64_t batch, void* work, size_t work_bytes, void* stream);

/* Choosing the slice cut-off per model (the host layer's variance engine 'auto'): tb200_i8_probe_points writes 2P probe
 * points -- P training points evenly spaced through X (n x d, unscaled), where k** - v.v cancels completely, then the P
 * midpoints between each of them and its successor -- into out (2P x d); tb200_max_sqdiff returns max |a_i^2 - b_i^2| of
 * two standard-deviation vectors (one double; +inf if either holds a NaN).  The host evaluates the probe with gmax 7 and
 * gmax 9 (tb200_acquire_i8 with sigma_out) and uses 21 products only when the two variances agree to 1/8 of the parity
 * floor (8 eps n prior_var); profiles/r2_predict_precision.md. */
int tb200_i8_probe_points(const double* X, int n, int d, int P, double* out, void* stream);
int tb200_max_sqdiff(const double* a, const double* b, int64_t m, double* out_max, void* stream);

/* write candidates [index_base, index_base+m) of a generated set into out (m x d) */
int tb200_generate_candidates(const tb200_cand_spec* cand_host, int d, int64_t index_base, int64_t m,
                              double* out, void* stream);

/* measured FP64 peaks of this device (roofline denominators for the DMMA kernels):
 * runs register-resident DMMA.8x8x4 / DFMA loops on every SM for ~`ms` milliseconds each. */
int tb200_fp64_peak(double ms, double* dmma_tflops_host, double* dfma_tflops_host);
/* measured INT8 tensor-core peak of this device in TOP/s (2 operations per multiply-accumulate): the roofline
 * denominator of the tcgen05.mma kind::i8 kernels (predictive variance, batched fit products).  One CTA per SM issues
 * M128 x N256 x K32 MMAs on shared-memory-resident operands for ~`ms` milliseconds; best of 5 repetitions.
 * random_digits = 0: zero operands (the rate at the maximum clock: nothing toggles, the power limit is never reached);
 * random_digits = 1: pseudo-random digits (the rate at the clock the power limit allows for real data). */
int tb200_i8_peak(double ms, int random_digits, double* int8_tops_host);
/* the same loop with MMAs of N = n_cols columns (a multiple of 16, <= 256): the rate as a function of N, which decides
 * how the kernels cut their stacked operands into instructions (csrc/ozaki_i8.cuh issue_range) */
int tb200_i8_mma_rate(double ms, int random_digits, int n_cols, double* int8_tops_host);
/* diagnostic: SM cycles per k-block (two K = 32 steps) of one of the compile-time MMA lists of csrc/ozaki_i8.cuh
 * (MmaList<variant>: the variance kernel's own instruction list and re-ordered / re-chunked / re-aligned variants of
 * it).  fill_planes = -1: issued back to back on resident shared-memory operands; fill_planes = 0..7: inside the
 * kernels' two-stage pipeline, a producer warp refilling the idle stage with that many digit planes of both operands
 * per k-block (0 = stage hand-over only).  Separates what the instruction mix costs on the tensor pipe from what the
 * fill traffic into shared memory costs it (tools/i8_mma_list.py). */
int tb200_i8_mma_list(int variant, int kblocks, int fill_planes, double* cycles_per_kblock_host, int* mmas_per_step_host,
                      int* columns_per_step_host);

/* Incremental refit at fixed hyper-parameters (the reference's `training_iterations` callable returning 0,
 * turbo/modules/surrogates.py:250-269, :306-310): append ONE training point to the factor in O(n^2).
 * tb200_factor_append overwrites the padded row n_old of Linv, column n_old of XsT and Ldiag[n_old]; it needs
 * n_old < tb200_padded_n(n_old) (otherwise rebuild with tb200_factor).  info: 0, or n_old+1 if the new pivot is not
 * positive (nothing is modified then).  tb200_alpha_lml recomputes alpha = K^-1 y and the log-marginal likelihood
 * (sklearn:_gpr.py:613-617) for a (re-normalised) target vector from Linv and diag(L). */
size_t tb200_append_workspace_bytes(int n);
int tb200_factor_append(const double* x_new, double* XsT, double* Linv, double* Ldiag, int n_old,
                        const tb200_kernel_spec* spec_host, const double* theta, double jitter, int* info, void* work,
                        size_t work_bytes, void* stream);
int tb200_alpha_lml(const double* y_n, int n, const double* Linv, const double* Ldiag, double* alpha, double* lml,
                    void* work, size_t work_bytes, void* stream);

/* tb200_alpha_lml without a workspace and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses
 * that forbid it).  The scratch vectors live in a file-scope device array of the library (512.5 KB per GPU that loads
 * it, held until exit); n <= 16384, larger n is rejected with an error (use tb200_alpha_lml).  Same results as
 * tb200_alpha_lml.  Every call on a device shares that array: calls that overlap on the GPU (different streams) race;
 * calls on one stream are safe back to back. */
int tb200_alpha_lml_static(const double* y_n, int n, const double* Linv, const double* Ldiag, double* alpha, double* lml,
                           void* stream);

/* Per-epilogue-CTA running bests (value, global index; index -1 = unused) of the last tb200_acquire(_i8) call that
 * used `work`: tb200_acquire_local_best_count(batch, n) distinct good candidates, the starting points of the gradient
 * polish (the reference's `start_from_best` best random samples, auxiliary_optimisers.py:73-75). */
int tb200_acquire_local_best_count(int64_t batch, int n);
int tb200_acquire_local_bests(void* work, int n, int d, int64_t batch, int use_i8, double* vals_out, int64_t* idx_out,
                              void* stream);

/* Acquisition value and its gradient with respect to the point, for P <= 128 points per call: the
 * analytic counterpart of the numerical gradients SciPy's L-BFGS-B takes of `-acq(x)` in
 * turbo/modules/auxiliary_optimisers.py:84-93.  val_out[P], grad_out[P][d] (d a / d x in the caller's coordinates);
 * mu_out / sigma_out optional.  FP64 DMMA path (V = Linv k*, w = Linv^T V). */
size_t tb200_acquire_grad_workspace_bytes(int n, int d);
int tb200_acquire_grad(const double* Xp, int P, const double* XsT, const double* Linv, const double* alpha, int n,
                       const tb200_kernel_spec* spec_host, const double* theta, const tb200_acq_spec* acq_host,
                       double* val_out, double* grad_out, double* mu_out, double* sigma_out, void* work,
                       size_t work_bytes, void* stream);

/* CUDA-event timing of the dominant kernel (the variance GEMM inside tb200_acquire), recorded on the
 * stream the kernel is launched on.  enable(1) starts collecting; read() synchronises, returns the
 * summed duration and the number of timed launches, and clears the record. */
int tb200_timing_enable(int on);
/* Diagnostic trace: one CUDA event before/after every kernel of tb200_acquire(_i8) on the stream it runs on,
 * keeping the two-stream schedule.  tag = slot*16 + {0: before K*, 4: after K*, 1: before variance kernel,
 * 2: after it, 3: after the epilogue}.  tb200_trace_read returns the number of records (times in ms from the first). */
int tb200_trace_enable(int on);
int tb200_trace_read(double* t_ms_host, int* tags_host, int cap);
int tb200_timing_read(double* gemm_ms_host, int64_t* gemm_launches_host);

/* plain C = alpha*op(A)*op(B) + beta*C on the library's DMMA GEMM (testing / roofline runs);
 * M,N multiples of 128, K multiple of 16; ta/tb: 0 = [m][k] / [k][n] storage, 1 = transposed. */
int tb200_dgemm(int ta, int tb, int M, int N, int K, double alpha, const double* A, int lda,
                const double* B, int ldb, double beta, double* C, int ldc, void* stream);

/* number of kernel launches issued by this library since load (bench.py's gpu_launches) */
int64_t tb200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif
