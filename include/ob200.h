/*
 * ob200.h -- C ABI of libob200.so: the B200 (sm_100a) implementation of Orion's
 * candidate-scoring hot path (sample -> transform -> normalise -> GP posterior ->
 * acquisition -> argmax).
 *
 * The reference (bouthilx/metaopt = Orion v0.1.6) is pure Python and has no FFI; each entry
 * point below names the reference (or third-party, where the reference delegates) function
 * whose work it replaces.  Paths are relative to /root/reference; `sklearn:` means
 * scikit-learn 1.9.0 `sklearn/gaussian_process/`.
 *
 * Conventions
 *  - every function returns 0 on success or a negative OB200_E* code; nothing throws or exits.
 *    `ob200_last_error(ctx)` gives the message of the last failure on that context.
 *  - one context = one CUDA device (one process per GPU; multi-GPU is done above this ABI with
 *    torch.distributed/NCCL).  Calls on one context are not re-entrant.
 *  - pointers named *_dev are device pointers on the context's device, *_host are host pointers;
 *    all matrices are dense row-major float64.  Caller owns every buffer it passes in; the
 *    context owns everything else and frees it in ob200_ctx_destroy.
 *  - work is enqueued on the context's stream (ob200_set_stream); functions that return values
 *    to host memory synchronise that stream before returning, the others are asynchronous.
 */
#ifndef OB200_H
#define OB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ob200_ctx ob200_ctx;

enum {
  OB200_OK = 0,
  OB200_EINVAL = -1,   /* bad argument */
  OB200_ECUDA = -2,    /* CUDA runtime error */
  OB200_ENOTSPD = -3,  /* Gram matrix not positive definite (numpy LinAlgError in the reference path) */
  OB200_ENOMEM = -4,   /* device allocation failed */
  OB200_ESTATE = -5,   /* call order (e.g. score before fit) */
  OB200_EBOUNDS = -6   /* "Improbable bounds": 4 rejected draws (ValueError in space.py:387-389) */
};

/* slot kinds / priors (one slot = one scalar source variate of an Orion dimension) */
enum { OB200_KIND_REAL = 0, OB200_KIND_INT = 1, OB200_KIND_CAT = 2 };
enum { OB200_PRIOR_UNIFORM = 0, OB200_PRIOR_LOGUNIFORM = 1, OB200_PRIOR_NORMAL = 2, OB200_PRIOR_CHOICE = 3 };
enum { OB200_KERNEL_MATERN52 = 0, OB200_KERNEL_RBF = 1 };
enum { OB200_ACQ_EI = 0, OB200_ACQ_PI = 1, OB200_ACQ_LCB = 2 };

/* Flat dimension-descriptor table row.  Replaces the per-dimension Python objects
 * Real/Integer/Categorical (src/orion/algo/space.py:291-647) + the transformer chain built by
 * build_required_space(requires='real') (src/orion/core/worker/transformer.py:21-77). */
typedef struct ob200_slot {
  int32_t kind;     /* OB200_KIND_*                                                     */
  int32_t prior;    /* OB200_PRIOR_*                                                    */
  int32_t out_col;  /* first output column of this slot                                 */
  int32_t width;    /* number of output columns: 1, or k for a one-hot of k>2 categories */
  int32_t ncat;     /* k for categorical slots, else 0                                  */
  int32_t cdf_off;  /* offset of this slot's k cumulative probabilities in the cdf table */
  double p0, p1;    /* uniform: loc, scale; loguniform: log a, log b - log a; normal: loc, scale */
  double low, high; /* acceptance interval [low, high) (Real/Integer.interval(), space.py:358-369,425-448) */
  double nlo, nhi;  /* normalisation bounds of the output columns (TransformedDimension.interval(), transformer.py:384-392) */
} ob200_slot;

/* GP hyper-parameters: amplitude * Matern52|RBF(length_scale[D]); replaces the sklearn kernel
 * object ConstantKernel*Matern / ConstantKernel*RBF (sklearn:kernels.py:1685-1745,1530-1573). */
typedef struct ob200_kernel {
  int32_t kind;             /* OB200_KERNEL_*                       */
  int32_t n_length_scale;   /* 1 (isotropic) or D (anisotropic)     */
  double amplitude;         /* ConstantKernel value (= kernel diag) */
  const double* length_scale_host;
} ob200_kernel;

/* winner of one acquisition: utility value (EI, PI or -LCB), global candidate index, posterior */
typedef struct ob200_best {
  double value;
  int64_t index;
  double mu, sigma;
} ob200_best;

/* ---- context ------------------------------------------------------------------------- */
/* No reference counterpart (Orion has no device state); lifetime rules follow producer.py:161
 * (the algorithm object is deep-copied freely, so device state lives here, not in the object). */
int ob200_ctx_create(int device, ob200_ctx** out);
void ob200_ctx_destroy(ob200_ctx* ctx);
const char* ob200_last_error(const ob200_ctx* ctx);
/* cudaStream_t to enqueue on (0 = legacy default stream). */
int ob200_set_stream(ob200_ctx* ctx, void* cuda_stream);
int ob200_sync(ob200_ctx* ctx);
/* number of kernels this library has launched on this context since creation */
int64_t ob200_launch_count(const ob200_ctx* ctx);
/* ABI version / build info */
int ob200_version(void);

/* ---- a1-a12: Space.sample + TransformedSpace.transform + normalise -------------------- */
/* Uploads the slot table; replaces Space/Dimension construction (space.py:732-858) and
 * build_required_space (transformer.py:21-77).  cdf_host: concatenated, normalised cumulative
 * probabilities of the categorical slots (numpy rng.choice semantics, space.py:542-552). */
int ob200_space_set(ob200_ctx* ctx, const ob200_slot* slots_host, int n_slots,
                    const double* cdf_host, int n_cdf);

/* Replaces Space.sample(n, seed) (space.py:745-773) -> Real/_Discrete/Categorical.sample
 * (:371-391,410-423,542-552) -> TransformedSpace.transform (transformer.py:473-475) -> skopt
 * Normalize.  Point i (global index first_index+i) is a pure function of (seed, index): Philox
 * 4x32-10 counter-based draws, so shards reproduce the same candidates.
 *   out_dev       [count x ld] row-major, ld >= D_out (count = 0: nothing is written, out_dev may be NULL)
 *   normalise     0: transformed values; 1: (x-nlo)/(nhi-nlo)
 *   variates_dev  NULL, or [count x n_slots] host-chosen U (uniform/loguniform/choice) / Z (normal)
 *                 variates: single try per slot (parity mode against the reference's own draws)
 * Asynchronous.  Rejection failures are latched; read them with ob200_sample_status. */
int ob200_sample_transform(ob200_ctx* ctx, uint64_t seed, uint64_t first_index, int64_t count,
                           int normalise, double* out_dev, int64_t ld, const double* variates_dev);
/* Synchronises; returns OB200_EBOUNDS if any draw since the last call exhausted its 4 tries
 * (the reference raises ValueError("Improbable bounds"), space.py:387-389), and clears the latch. */
int ob200_sample_status(ob200_ctx* ctx);
/* Host-buffer convenience: same as ob200_sample_transform but writes to host memory (D2H inside). */
int ob200_sample_transform_host(ob200_ctx* ctx, uint64_t seed, uint64_t first_index, int64_t count,
                                int normalise, double* out_host, int64_t ld,
                                const double* variates_host);

/* ---- a13: GP fit ------------------------------------------------------------------------ */
/* Replaces GaussianProcessRegressor.fit with fixed hyper-parameters (sklearn:_gpr.py:233-368):
 * y normalisation (:275-280), K = kernel(X) (:349), K[diag] += alpha (:350), L = cholesky(K)
 * (:352), alpha_ = cho_solve(L, y) (:363-367).  Additionally forms inv(L) (tile-packed) so the
 * posterior variance is a triangular GEMM.  X_dev [n x D], y_dev [n]; 1 <= n <= 8192.
 * Returns OB200_ENOTSPD like numpy.linalg.LinAlgError (:351-361).  Synchronous. */
int ob200_gp_fit(ob200_ctx* ctx, const double* X_dev, const double* y_dev, int n, int D,
                 const ob200_kernel* kernel, double alpha, int normalize_y);
int ob200_gp_fit_host(ob200_ctx* ctx, const double* X_host, const double* y_host, int n, int D,
                      const ob200_kernel* kernel, double alpha, int normalize_y);
/* ob200_gp_fit for ONE evaluation of a hyper-parameter search (what sklearn's fit does inside
 * `_constrained_optimization`, _gpr.py:299-340: log_marginal_likelihood(theta, eval_gradient=True) per step): same
 * arguments; produces everything ob200_gp_lml and ob200_gp_lml_grad read (L, inv(L), alpha, the LML), skips the
 * Newton-Schulz refinement of inv(L) and the tile packing the score kernel needs, and leaves the context NOT
 * scoreable (ob200_score* return OB200_ESTATE) until the next ob200_gp_fit / ob200_gp_import. */
int ob200_gp_fit_search(ob200_ctx* ctx, const double* X_dev, const double* y_dev, int n, int D,
                        const ob200_kernel* kernel, double alpha, int normalize_y);
/* Append ONE observation to the current fit in O(n^2) instead of refitting in O(n^3): new row of L from
 * inv(L) k, new row of inv(L), alpha_ recomputed in full from inv(L) (the y statistics move with every observation
 * when normalize_y is on), same hyper-parameters and diagonal term as the fit.  Counterpart of what
 * GaussianProcessRegressor.fit (sklearn _gpr.py:233-368) recomputes from scratch after every skopt tell(); used for
 * constant-liar batches and fixed-hyper-parameter runs.  x_new_dev [D], y_dev [n_total] = ALL
 * targets, n_total = current n + 1.  OB200_ESTATE when the padded factor has no free row (caller refits) or the fit
 * was imported; OB200_ENOTSPD like the fit.  Synchronous. */
int ob200_gp_append(ob200_ctx* ctx, const double* x_new_dev, const double* y_dev, int n_total, int normalize_y);
/* log marginal likelihood of the current fit (sklearn:_gpr.py:541-656, value only). */
int ob200_gp_lml(ob200_ctx* ctx, double* lml_host);
/* Gradient of the log marginal likelihood of the current fit (sklearn:_gpr.py:541-656,
 * log_marginal_likelihood(theta, eval_gradient=True)): grad_host[0] = d/dlog(amplitude),
 * grad_host[1..D] = d/dlog(length_scale_d) (sum them for an isotropic kernel), grad_host[D+1] =
 * d/d(diagonal term) (multiply by the noise level for d/dlog(noise)).  n_grad must be D+2.  Needs a
 * fit made on this context (not an imported one).  Synchronous. */
int ob200_gp_lml_grad(ob200_ctx* ctx, double* lml_host, double* grad_host, int n_grad);
/* Posterior mean / standard deviation and their gradients with respect to the (normalised) input at m <= 64
 * points: what skopt's acquisition optimiser "lbfgs" evaluates per step through its GaussianProcessRegressor
 * subclass, predict(X, return_std=True, return_mean_grad=True, return_std_grad=True) (scikit-optimize, no version pinned
 * by the reference -- SURVEY.md row 19; skopt/learning/gaussian_process/gpr.py, third-party and absent from the
 * reference tree; reached from the fmin_l_bfgs_b loop of Optimizer._tell, skopt/optimizer/optimizer.py).  x_host [m x ldx] row-major,
 * out_host [m x (2 + 2 D)]: mu, sigma, dmu/dx[D], dsigma/dx[D] per point, same units as ob200_score.
 * sigma == 0 gives a zero dsigma.  Needs a fit made on this context (not an imported one).  Synchronous. */
int ob200_gp_predict_grad_host(ob200_ctx* ctx, const double* x_host, int64_t ldx, int m, double* out_host);
/* Parity/debug read-back of L [n x n] (lower, row-major, upper = 0) and alpha_ [n] (either may be NULL). */
int ob200_gp_get_factor(ob200_ctx* ctx, double* L_host, double* alpha_host);
/* y statistics used by the fit: y_mean, y_std (after _handle_zeros_in_scale), min(y). */
int ob200_gp_get_ystats(ob200_ctx* ctx, double* y_mean, double* y_std, double* y_min);
/* Size in bytes of the opaque fit state, and export/import of it as one device blob, so that
 * rank 0 can fit and the other ranks of a multi-GPU job receive it by NCCL broadcast. */
int ob200_gp_state_bytes(ob200_ctx* ctx, int n, int D, int64_t* bytes);
int ob200_gp_export(ob200_ctx* ctx, void* blob_dev, int64_t bytes);
int ob200_gp_import(ob200_ctx* ctx, const void* blob_dev, int64_t bytes);

/* ---- a14-a17: cross-kernel + posterior + acquisition + argmax ---------------------------- */
/* Replaces, for M candidates, GaussianProcessRegressor.predict(X, return_std=True)
 * (sklearn:_gpr.py:444-500: K* = kernel(X, X_train) :446, mu = K* alpha_ :447-450,
 * V = solve_triangular(L, K*^T) :460-462, var = diag - sum V^2 :480-481, clamp :485-491,
 * *y_std^2 :494, sqrt :500), skopt gaussian_ei / gaussian_pi / gaussian_lcb and
 * X[np.argmin(values)] (first occurrence wins ties).
 *   cand_dev   [M x ldc] normalised candidates (columns 0..D-1 used); or NULL to generate
 *              candidate i from (seed, first_index+i) with the space set by ob200_space_set
 *   acq_mask   bit OB200_ACQ_* set = compute that acquisition's winner into best_host[that]
 *   mu_dev / sigma_dev / acq_dev  optional dense outputs [M] (acq_dev: [3 x M], rows by OB200_ACQ_*,
 *              only rows in acq_mask written); NULL to skip
 * Index reported = first_index + row.  Synchronous (copies the winners back). */
int ob200_score(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, uint64_t seed,
                uint64_t first_index, int acq_mask, double y_opt, double xi, double kappa,
                ob200_best* best_host, double* mu_dev, double* sigma_dev, double* acq_dev);
/* Same with candidates (and optional dense outputs) in HOST memory: H2D / D2H inside the call. */
int ob200_score_host(ob200_ctx* ctx, const double* cand_host, int64_t ldc, int64_t M,
                     int acq_mask, double y_opt, double xi, double kappa, ob200_best* best_host,
                     double* mu_host, double* sigma_host, double* acq_host);
/* Asynchronous form: winners stay on the device ([3] ob200_best at best_dev), no sync. */
int ob200_score_async(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, uint64_t seed,
                      uint64_t first_index, int acq_mask, double y_opt, double xi, double kappa,
                      ob200_best* best_dev, double* mu_dev, double* sigma_dev, double* acq_dev);

/* Exact pruned form of ob200_score for ONE acquisition (acq = OB200_ACQ_*): same winner, bit for bit,
 * as the dense call (same reference functions: sklearn:_gpr.py:444-500 + skopt gaussian_ei/pi/lcb +
 * np.argmin), but the O(n^2) variance is computed only for candidates whose utility upper bound
 * U(mu, sigma_ub) -- from the O(nD) cross-kernel alone: the mean, and the variance the nearest observation
 * leaves, sigma_ub^2 = (amplitude - k_max^2 / (amplitude + diag)) * y_std^2 -- reaches the best utility found so far.  best_host[0] receives the winner; n_scored_host (may be NULL) the number of
 * candidates that went through the full posterior.  Synchronous. */
int ob200_score_pruned(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, uint64_t seed,
                       uint64_t first_index, int acq, double y_opt, double xi, double kappa,
                       ob200_best* best_host, int64_t* n_scored_host);
/* The same for several acquisitions at once (acq_mask bits 1 << OB200_ACQ_*; the gp_hedge portfolio): one bound
 * pass, one selection (the union over the acquisitions), every selected candidate scored for all of them.
 * best_host [3] and n_scored_host [3] (may be NULL) are indexed by OB200_ACQ_*; entries outside the mask are zeroed.
 * The bound pass runs in FP32 on the TF32 tensor cores with a rigorous error budget (mu_lb <= mu, sigma_ub >= sigma;
 * csrc/ob200_bound.cu); candidates it cannot exclude get exact FP64 bounds before the O(n^2) scoring, so the winner
 * stays bit-identical to the dense call.  The FP32 pass is used only for fits whose error budget leaves a usable
 * bound (decided once per fit from ||alpha||_1); OB200_BOUND_F32=0 in the environment selects the FP64 bound pass for
 * everything.  When many candidates are still standing after that, a GP on 192 evenly spaced observations bounds
 * their variance from above once more (conditioning on a subset can only raise it) before the O(n^2) scoring. 
 * Calls that take the dense kernel instead: M up to a few waves of the dense kernel (one tile of 32 candidates per SM
 * each; one wave from n_obs ~ 1000, more for smaller fits, where a dense wave is cheaper than the bound pass's fixed cost);
 * -LCB with kappa < 0, which is not increasing in sigma.  y_opt, xi, kappa must
 * be finite (OB200_EINVAL otherwise, in every score entry point).  One host synchronisation per call. */
int ob200_score_pruned_multi(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, uint64_t seed,
                             uint64_t first_index, int acq_mask, double y_opt, double xi, double kappa,
                             ob200_best* best_host, int64_t* n_scored_host);

/* Pass 1 of the pruned path on its own (no reference counterpart; used by the tests and tools to check the
 * bounds): mu_lb_dev[i] <= posterior mean and sigma_ub_dev[i] >= posterior std of candidate i, same units as
 * ob200_score.  *used_f32 (may be NULL) = 1 when the FP32 bound kernel produced them, 0 when the FP64 mean-only
 * kernel did (mu_lb is then the mean itself).  Asynchronous on the context's stream. */
int ob200_score_bounds(ob200_ctx* ctx, const double* cand_dev, int64_t ldc, int64_t M, double* mu_lb_dev,
                       double* sigma_ub_dev, int* used_f32);
/* Counters of the last ob200_score_pruned[_multi] call that took the pruned path: out[0] = FP32 bound pass used,
 * out[1] = candidates in the first batch, out[2] = candidates the pass-1 bounds left for the second batch,
 * out[3] = of those, the ones still standing after their bounds were recomputed in FP64 (0 when out[0] == 0),
 * out[4] = the ones still standing after the subset-GP variance bound (-1 when that step did not run: it is taken
 * only when more than four first batches' worth of candidates are left). */
int ob200_prune_stats(const ob200_ctx* ctx, int64_t out[5]);

/* a16-a17 alone: acquisition + first-occurrence argmax over given posterior arrays
 * (skopt.acquisition.gaussian_ei/pi/lcb + numpy.argmin).  mu_dev, sigma_dev [M]. */
int ob200_acq_argmax(ob200_ctx* ctx, const double* mu_dev, const double* sigma_dev, int64_t M,
                     uint64_t first_index, int acq_mask, double y_opt, double xi, double kappa,
                     ob200_best* best_host, double* acq_dev);

/* Allocation-free variant of ob200_acq_argmax.  It has no counterpart in the reference; it exists
 * for callers that must not allocate device memory when they call in -- services that reserve all
 * GPU memory up front and treat a later cudaMalloc as a fault, or code running after its host
 * framework has frozen allocation -- and so can neither create a context nor let one grow.
 *   - no context: works on the calling thread's CURRENT device; mu_dev, sigma_dev [M], best_dev [3]
 *     and the optional acq_dev [3 x M] are caller-owned device buffers on that device; the result
 *     is written to best_dev asynchronously on cuda_stream (a cudaStream_t, 0 = default stream).
 *   - the per-block scratch is a fixed 384 KiB array inside the library's device image (held by
 *     every GPU that loads the library until exit); it limits one call to 2^24 = 16,777,216
 *     candidates, larger M is refused with OB200_EINVAL and a message in errbuf (may be NULL).
 *   - it calls no cudaMalloc / cudaFree or any other allocating routine.  The first call on each
 *     device looks the scratch array's device address up once; make that first call outside any
 *     stream capture.
 *   - all calls share the one scratch array (it keeps the previous call's records, none of which
 *     a later call reads before overwriting): two calls that overlap on the GPU -- different
 *     streams or host threads -- race on it; serialise them.
 *   - the looked-up address is cached per device for the life of the process; a host that calls
 *     cudaDeviceReset() invalidates it and must not call this function on that device again. */
int ob200_acq_argmax_static(const double* mu_dev, const double* sigma_dev, int64_t M, uint64_t first_index,
                            int acq_mask, double y_opt, double xi, double kappa, ob200_best* best_dev,
                            double* acq_dev, void* cuda_stream, char* errbuf, int errbuf_len);

#ifdef __cplusplus
}
#endif
#endif /* OB200_H */
