/*
 * gpb200.h -- C ABI of libgpb200.so: the B200 (sm_100a) implementation of the GP inference hot
 * path of jonnylaw/gaussian-processes (covariance build -> Cholesky -> triangular solves ->
 * log marginal likelihood per Metropolis-Hastings proposal, + posterior predictive mean/variance).
 *
 * The reference (Scala) has no FFI of its own; the seam is a set of Scala object methods whose
 * signatures stay unchanged.  Each entry point below names the reference method
 * (file:line under core/src/main/scala/gp/) it stands in for.  INTEGRATION.md shows the Scala
 * (Panama / JNI) binding a maintainer adds on the reference side.
 *
 * Conventions
 *   - every call returns an int status: 0 ok, <0 usage / CUDA error (gpb_strerror); numerical
 *     failure (matrix not positive definite) is NOT a status -- it is reported in `info`
 *     (LAPACK convention: k>0 = leading minor k not positive), never by abort;
 *   - plain pointers and sizes only; caller owns every host buffer; the library owns device
 *     memory behind the opaque handles; no exception crosses the boundary;
 *   - all matrices are column-major (Breeze DenseMatrix layout), locations are SoA [dim][n];
 *   - one gpb_ctx = one device + its streams + workspace; not shared between threads (one ctx per
 *     JVM thread / chain group); calls are synchronous on return;
 *   - there is NO CPU fallback: without a CUDA device gpb_ctx_create fails with GPB_E_NODEVICE.
 */
#ifndef GPB200_H
#define GPB200_H

#ifdef __cplusplus
extern "C" {
#endif

#define GPB_OK 0
#define GPB_E_INVALID (-1)     /* bad argument */
#define GPB_E_NODEVICE (-2)    /* no CUDA device / wrong architecture */
#define GPB_E_CUDA (-3)        /* CUDA runtime error (see gpb_last_cuda_error) */
#define GPB_E_NOMEM (-4)       /* device or host allocation failed */
#define GPB_E_UNSUPPORTED (-5) /* e.g. Matern term, dim > 2 */
#define GPB_E_WATCHDOG (-6)    /* scheduler watchdog fired (internal error, never a hang) */
#define GPB_E_NOTPD (-7)       /* used only by calls that have no info out-parameter */

#define GPB_KIND_SE 0    /* SquaredExp(h, sigma)   KernelParameters.scala:21-35 */
#define GPB_KIND_WHITE 1 /* White(sigma)           KernelParameters.scala:53-67 */
#define GPB_KIND_MATERN 2 /* reserved -> GPB_E_UNSUPPORTED (out of scope) */
#define GPB_MAX_TERMS 8

#define GPB_FORMULA_AS_PARSED 0 /* what GaussianProcess.scala:126 executes */
#define GPB_FORMULA_GPML 1      /* what its docs intend: -Q/2 - S - n/2 log 2pi */

typedef struct gpb_ctx gpb_ctx;
typedef struct gpb_dataset gpb_dataset;

/* One covariance term, in Parameters.toList order: SE -> (p0=h, p1=sigma), White -> (p0=sigma). */
typedef struct {
  int kind;
  double p0, p1, p2;
} gpb_term;

/* The three reductions the likelihood needs: Q = u.u, U1 = sum u, S = sum log L_ii
 * with u = L^-1 (y - m(x)), K + nugget I = L L^T. */
typedef struct {
  double Q, U1, S;
  int n;
  int info;
} gpb_ll;

/* ---- context / dataset ------------------------------------------------------------------- */
int gpb_ctx_create(int device, gpb_ctx **out);
void gpb_ctx_destroy(gpb_ctx *ctx);

/* Upload the observations once (GaussianProcess.loglikelihood's outer application,
 * GaussianProcess.scala:109-113: `observed.map(_.x)`; Data/Location types GaussianProcess.scala:36,
 * Location.scala:5-13).  x: SoA [dim][n], dim in {1,2}; y: [n].  Copies; caller keeps ownership. */
int gpb_dataset_create(gpb_ctx *ctx, int n, int dim, const double *x, const double *y,
                       gpb_dataset **out);
void gpb_dataset_destroy(gpb_dataset *ds);

/* ---- log marginal likelihood ------------------------------------------------------------- */
/* GaussianProcess.loglikelihood(observed, dist)(p)  GaussianProcess.scala:109-127
 * (buildCov KernelFunction.scala:84-89, distanceMatrix GaussianProcess.scala:50-61, euclidean
 * Location.scala:27-33, kernels KernelFunction.scala:13-15,33-35,47-55, nugget :117-118, residual
 * :120-122 with MeanFunction.scala:7-17, cholesky :123, forwardSolve :124 / Predict.scala:40-45).
 * beta: plane mean coefficients, nbeta = 0 (Zero) | dim+1 (Plane).  nugget: 1e-3 in the reference. */
int gpb_loglik(gpb_ctx *ctx, const gpb_dataset *ds, const gpb_term *terms, int nterms,
               const double *beta, int nbeta, double nugget, gpb_ll *out);

/* `batch` independent proposals on one dataset in one pass (the per-proposal call of
 * KernelParameters.sample, KernelParameters.scala:242-243, for many chains at once).
 * terms: [batch][nterms]; beta: [batch][nbeta]; out: [batch]. */
int gpb_loglik_batch(gpb_ctx *ctx, const gpb_dataset *ds, int batch, const gpb_term *terms,
                     int nterms, const double *beta, int nbeta, double nugget, gpb_ll *out);

/* Host helper: turn the reductions into the scalar the reference returns.
 * formula GPB_FORMULA_AS_PARSED: -Q/2 + (S + C) U1 / 2 ; GPB_FORMULA_GPML: -Q/2 - S - C ;
 * C = n/2 log(2 pi).  NaN when info != 0. */
double gpb_ll_value(const gpb_ll *r, int formula);

/* ---- posterior predictive ---------------------------------------------------------------- */
/* Predict.fit(newXs, observed, dist, p)  Predict.scala:57-94 (buildDistVec :21-29).  Mirrors the
 * reference: raw y (no mean subtraction, :80), var = k(0) - u.u with k(0) including the white
 * term (:78,:92); nugget 1e-6 in the reference (:67).  xstar: SoA [dim][m]; mean, var: [m].
 * The caller forms Gaussian(mean, sqrt(var)). */
int gpb_predict(gpb_ctx *ctx, const gpb_dataset *ds, const gpb_term *terms, int nterms,
                double nugget, int m, const double *xstar, double *mean, double *var, int *info);

/* ---- matrix-level entry points (host column-major in / out) ------------------------------- */
/* GaussianProcess.distanceMatrix  GaussianProcess.scala:50-61 (Euclidean only).  D: [n*n]. */
int gpb_distance_matrix(gpb_ctx *ctx, int n, int dim, const double *x, double *D);
/* KernelFunction.buildCov(xs, covFn, dist) (+ nugget on the diagonal)  KernelFunction.scala:84-89.
 * K: [n*n], full symmetric matrix. */
int gpb_build_cov(gpb_ctx *ctx, int n, int dim, const double *x, const gpb_term *terms, int nterms,
                  double nugget, double *K);
/* breeze.linalg.cholesky call sites GaussianProcess.scala:123, Predict.scala:69.  A: [n*n] in,
 * lower factor with zeroed upper triangle out. */
int gpb_potrf(gpb_ctx *ctx, int n, double *A, int *info);
/* Predict.forwardSolve(A, y)  Predict.scala:40-45  (dtrsv "L","N","N").  nrhs columns in B [n*nrhs]. */
int gpb_trsm_lower(gpb_ctx *ctx, int n, int nrhs, const double *L, double *B);
/* Mcmc.backSolve(A, y) / (A, Y)  Mcmc.scala:35-41, 52-58  (dtrsv "U" / dtrsm "L","U","N","N"). */
int gpb_trsm_upper(gpb_ctx *ctx, int n, int nrhs, const double *U, double *B);

/* ---- Metropolis-Hastings over the kernel hyper-parameters, many chains -------------------- */
/* The batched form of KernelParameters.sample (KernelParameters.scala:231-246) driven by Breeze's
 * MarkovChain.Kernels.metropolis, with the shipped example's proposal and prior
 * (examples/.../SimulatedGp.scala:115-151): log-normal random walk with step `delta`, independent
 * InverseGamma(shape, scale) priors per parameter.  Explicit RNG streams:
 *   z  [iters][chains][P] standard normals, uu [iters][chains] uniforms.
 * init: [chains][nterms]; ig: [P][2].  The whole loop is device resident: one proposal batch per
 * iteration, ll(t) cached (1 factorisation / iteration / chain; the reference does 2).
 * out_states [iters][chains][P], out_ll [iters][chains] (prior + loglik after the step),
 * out_accept [iters][chains].  info_out (may be NULL): first non-PD info seen (the reference throws
 * there; the chain then rejects that proposal when on_notpd == 1, or the call stops with
 * GPB_E_NOTPD when on_notpd == 0, the default that mirrors the reference). */
typedef struct {
  int formula;  /* GPB_FORMULA_* */
  int on_notpd; /* 0 = stop with GPB_E_NOTPD (mirror), 1 = treat as reject */
  double delta;
  double nugget;
} gpb_mh_opts;

int gpb_mh_run(gpb_ctx *ctx, const gpb_dataset *ds, int chains, int iters, const gpb_term *init,
               int nterms, const double *beta, int nbeta, const double *ig, const gpb_mh_opts *opts,
               const double *z, const double *uu, double *out_states, double *out_ll,
               int *out_accept, int *info_out);

/* Counter-based streams (Philox4x32-10, key = (seed, chain), counter = (iter, slot)) for callers
 * that do not bring their own: fills z [iters][chains][P] and uu [iters][chains] on the host. */
int gpb_philox_streams(unsigned long long seed, int chain0, int chains, int iters, int P, double *z,
                       double *uu);

/* ---- "next" rows of the scope table ------------------------------------------ */
/* MeanParameters.samplePlane(prior, obs, xs, dist, p)  MeanParameters.scala:74-103, the Gibbs step
 * of the plane mean in Mcmc.sample (Mcmc.scala:73).  X: the d = dim+1 design-matrix columns
 * [d][n] exactly as makeDesignMatrix lays them out (MeanParameters.scala:55-65; the caller keeps its
 * layout quirk).  They ride the factorisation as appended rows, so X^T K^-1 X = V V^T needs no
 * back-solve.  gpb_plane_stats returns G = X^T (K + nugget I)^-1 X [d*d] and w = X^T L^-1 y [d];
 * gpb_sample_plane finishes the d x d algebra on the host: prec = G + I/var, mean = prec \
 * (mean0/var + w), beta = mean + chol(prec) \ z with z the d standard normals the reference draws.
 * nugget is 1e-6 in the reference (:84). */
int gpb_plane_stats(gpb_ctx *ctx, const gpb_dataset *ds, const gpb_term *terms, int nterms,
                    double nugget, int d, const double *X, double *G, double *w, int *info);
int gpb_sample_plane(gpb_ctx *ctx, const gpb_dataset *ds, const gpb_term *terms, int nterms,
                     double nugget, int d, const double *X, double prior_mean, double prior_var,
                     const double *z, double *beta, int *info);
/* GaussianProcess.draw(xs, dist, p)  GaussianProcess.scala:67-83: mu + root z with nugget 1e-3.
 * The reference takes the symmetric eigen root (eigSym); this uses the Cholesky root of the same
 * covariance -- equal in distribution, not bitwise.  mu may be NULL (Zero mean). */
int gpb_draw(gpb_ctx *ctx, int n, int dim, const double *x, const gpb_term *terms, int nterms,
             double nugget, const double *mu, const double *z, double *out, int *info);

/* GaussianProcess.mllGradient(observed, dist)(p)  GaussianProcess.scala:166-187, with the tangent
 * matrices of KernelParameters.tangentMatrix / gradient (KernelParameters.scala:270-300): the input
 * of the HMC samplers (KernelParameters.scala:121-198).  `terms` hold the parameters as the reference
 * passes them to mllGradient, i.e. UNCONSTRAINED (log scale; sampleHmc :141-144).  grad: [batch][P] in
 * Parameters.toList order of the kernel terms; info: [batch].  K^-1 is never materialised: L^-T rides
 * the factorisation as an appended identity block, and a fused DMMA contraction + epilogue forms
 * 0.5 sum_ij (alpha_i alpha_j - Kinv_ij) G_p(d_ij).
 *   GPB_GRAD_AS_EXECUTED  what the Scala executes: kxx is built from the unconstrained values
 *       themselves (:171-176), the tangent entries are exp(-d^2/s^2) + 1, -(2h/s^3) exp(-d^2/s^2) + 1
 *       and 2 (white), with h, s the exponentials (:272-281).  The reference solves with LU; this
 *       path needs kxx positive definite and reports info > 0 otherwise (grad = NaN).
 *   GPB_GRAD_ANALYTIC     the gradient of the GPML log marginal likelihood with respect to the log
 *       parameters: kxx from the exponentials, dK/dlog h = h E, dK/dlog s = h E 2 d^2/s^2,
 *       dK/dlog s_w = s_w [d == 0].
 * nugget is 1e-3 in the reference (:175). */
#define GPB_GRAD_AS_EXECUTED 0
#define GPB_GRAD_ANALYTIC 1
int gpb_mll_gradient(gpb_ctx *ctx, const gpb_dataset *ds, int batch, const gpb_term *terms, int nterms,
                     const double *beta, int nbeta, double nugget, int mode, double *grad, int *info);

/* ---- misc -------------------------------------------------------------------------------- */
const char *gpb_strerror(int status);
const char *gpb_last_cuda_error(const gpb_ctx *ctx);
/* Number of kernels this context has launched since creation (bench.py's gpu_launches). */
long long gpb_launch_count(const gpb_ctx *ctx);
/* Device time (ms, CUDA events on the context's stream) of the last call's kernels, by phase:
 * t[0] build, t[1] factor+solve, t[2] epilogue.  Used by bench.py for the roofline object. */
int gpb_last_timing(const gpb_ctx *ctx, float *t3);
/* Scheduler selection: 0 = persistent task-graph kernel (default), 1 = one launch per block
 * column (debug / fallback on the same device code).  */
int gpb_set_sched(gpb_ctx *ctx, int mode);
/* Schedule analysis (profiling aid): when enabled, the factorisation kernel records 14 u64 per task
 * (timestamps of receipt / first data / contraction done / dependency satisfied / factor-or-solve
 * done / release, and the SM id).  gpb_get_trace copies up to max_tasks records of the last call
 * and returns the number of tasks of that call. */
int gpb_set_trace(gpb_ctx *ctx, int enable);
long long gpb_get_trace(gpb_ctx *ctx, unsigned long long *out, long long max_tasks);
/* Schedule introspection (host only, no device needed): the task that ticket `ticket` of the
 * factorisation kernel stands for, in a system of n training rows + `appended` extra rows and `batch`
 * proposals.  out4 = {proposal, tile row, tile column, kind}; kind 0 = plain tile (contract, solve),
 * 1 = first diagonal block, 2 = pre-accumulation of diagonal block `tile row` over all but its last
 * block column, 3 = fused: solve tile (j+1, j), then update and factor diagonal block j+1.  Returns the
 * number of tickets (out4 untouched when the ticket is out of range), or a negative status.  Tickets
 * are drawn in increasing order and every task only depends on lower tickets (tests/). */
long long gpb_schedule_task(int n, int appended, int batch, long long ticket, int *out4);
const char *gpb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GPB200_H */
