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
#define GPB_E_NOCOMM (-8)      /* NCCL unavailable, or the ctx has no communicator (gpb_comm_*) */

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

/* gpb_loglik for callers that cannot allocate.  No counterpart in the reference: added
 * for code under CUDA graph capture, tight serving loops and harnesses that
 * forbid cudaMalloc.  Same kernels and the same result (bit for bit) as gpb_loglik on the same inputs,
 * but the call makes no cudaMalloc / cudaFree / allocating library call and needs no gpb_ctx: the
 * factorisation's workspace is a static __device__ array inside the library, sized for
 * n <= GPB_NOALLOC_MAX_N (2.7 MB, held by every GPU that loads the library until the process exits).
 * A larger n is refused with GPB_E_INVALID and a message on stderr.
 *   device      CUDA device ordinal (made current);  stream: the caller's cudaStream_t (NULL = the
 *               default stream); the call is synchronous on return.
 *   d_x, d_y    DEVICE pointers owned by the caller: locations SoA [dim][n], observations [n].
 *   terms, beta, out   HOST pointers, as in gpb_loglik.
 * All calls share the one array, which keeps the last call's contents (the call rewrites everything
 * it reads).  Calls that overlap on the GPU race on it: the library serialises the calls a process
 * makes through this entry point (a mutex; the call is synchronous on return), and anything that
 * replays these kernels outside that mutex (a captured graph) must not overlap another call. */
#define GPB_NOALLOC_MAX_N 512
int gpb_loglik_noalloc(int device, void *stream, int n, int dim, const double *d_x, const double *d_y,
                       const gpb_term *terms, int nterms, const double *beta, int nbeta, double nugget,
                       gpb_ll *out);

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
/* The same matrix in the form every factorising entry point consumes (k_build's hot-path mode:
 * h exp(d^2 (-1/s^2)), no sqrt / re-squaring / division per entry): lower triangle (rows >= columns)
 * of K + nugget I, strict upper triangle zero.  For entrywise parity tests of the hot path. */
int gpb_build_cov_hot(gpb_ctx *ctx, int n, int dim, const double *x, const gpb_term *terms, int nterms,
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
 * init: [chains][nterms]; beta: [nbeta] plane coefficients shared by all chains (fixed); ig: [P][2].
 * The whole loop is device resident: one proposal batch per iteration, ll(t) cached
 * (1 factorisation / iteration / chain; the reference does 2).
 * out_states [iters][chains][P], out_ll [iters][chains] (prior + loglik after the step),
 * out_accept [iters][chains].
 * Accept rule, exactly Breeze's: u < min(1, exp(new - old)) with JVM NaN semantics -- a NaN on
 * either side rejects (scala.math.min propagates NaN and `u < NaN` is false).
 * Proposal whose covariance is not positive definite (the reference throws from `cholesky` on that
 * chain's thread and the chain ends there; the other chains are separate threads and carry on):
 *   on_notpd == 0 (mirror, default): that chain STOPS.  Its rows from the failing iteration on hold
 *       NaN states / NaN ll / accept 0; the other chains run to the end; the call then returns
 *       GPB_E_NOTPD and *info_out = the LAPACK info of the first failure (gpb_mh_run_ex also returns
 *       the iteration at which each chain stopped).
 *   on_notpd == 1: the proposal is rejected and the chain continues; the call returns GPB_OK and
 *       *info_out = the first info seen.
 * This entry point is gpb_mh_run_ex with host streams, a fixed mean and no gather. */
typedef struct {
  int formula;  /* GPB_FORMULA_* */
  int on_notpd; /* 0 = the failing chain stops, GPB_E_NOTPD (mirror); 1 = treat as reject */
  double delta;
  double nugget;
} gpb_mh_opts;

int gpb_mh_run(gpb_ctx *ctx, const gpb_dataset *ds, int chains, int iters, const gpb_term *init,
               int nterms, const double *beta, int nbeta, const double *ig, const gpb_mh_opts *opts,
               const double *z, const double *uu, double *out_states, double *out_ll,
               int *out_accept, int *info_out);

/* The full Mcmc.sample iteration (Mcmc.scala:63-76) for many chains, device resident:
 *   step = KernelParameters.sample compose MeanParameters.sample   (mean Gibbs step FIRST, :73)
 * mean_step == 1 (Plane mean, MeanParameters.sample / samplePlane, MeanParameters.scala:74-121): every
 *   iteration first draws beta from its full conditional -- kxx = K(current kernel parameters) +
 *   mean_nugget I (1e-6 in the reference, :84), raw y (:91), Gaussian(mean_prior_mean,
 *   mean_prior_var) prior on every coefficient, `design` = the d = nbeta design-matrix columns [d][n]
 *   exactly as makeDesignMatrix lays them out (:55-65, its layout quirk included); the design rows
 *   ride the factorisation (X^T K^-1 X = V V^T) and the d x d algebra runs in a kernel -- then runs
 *   the Metropolis step on the kernel parameters with BOTH log-likelihoods evaluated at the new beta
 *   (Breeze's metropolis re-evaluates ll(t); with a moving mean the cached value would be stale):
 *   three factorisations per chain and iteration, like the reference.
 *   Streams per (iteration, chain): zdim = P + d normals -- [0, P) the proposal, [P, P + d) the Gibbs
 *   draw -- and one uniform.  States carry W = P + d values: kernel parameters then beta, the column
 *   order of the examples' chain files (SimulatedGp.scala:159-160).
 * mean_step == 0: fixed beta per chain (beta0 [chains][nbeta], or NULL / nbeta = 0 for the Zero mean),
 *   zdim = W = P, one factorisation per chain and iteration.
 * rng: GPB_RNG_HOST  z [iters][chains][zdim], uu [iters][chains] supplied by the caller;
 *      GPB_RNG_PHILOX generated on the device, Philox4x32-10, key = seed, counter = (iteration,
 *      slot, chain0 + chain, stream) -- the layout of gpb_philox_streams; z and uu are ignored.  The
 *      uniforms are bit-identical to gpb_philox_streams; the normals go through the device's log /
 *      sincos (<= 2 ulp from the host generator): gpb_philox_streams_device returns exactly what the
 *      loop consumes, for parity runs against a host implementation.
 * Output ring: chain outputs are written to a device ring and flushed every `flush_every`
 *   iterations (default 32; device memory does not grow with `iters`) on a second stream, overlapped
 *   with the following iterations.
 * gather == 1 (needs gpb_comm_init_rank / gpb_comm_init_all on this ctx): this ctx runs `chains` of
 *   `total_chains` chains (block partition: rank r owns chains [first_r, first_r + count_r), see
 *   gpb_comm_shard); every flush all-gathers the ring segments device to device with NCCL and every
 *   rank receives the outputs of ALL chains: out_* are [iters][total_chains][...].  chain0 must be
 *   this rank's first_r.
 * small_n: GPB_SMALL_AUTO runs systems with n <= 64 (no mean step) as ONE launch -- one CTA per
 *   chain walks all iterations, matrix in shared memory (the shipped example thins to n = 20,
 *   SimulatedGp.scala:112); GPB_SMALL_OFF forces the general path. */
#define GPB_RNG_HOST 0
#define GPB_RNG_PHILOX 1
#define GPB_SMALL_AUTO 0
#define GPB_SMALL_OFF 1
typedef struct {
  int struct_size; /* sizeof(gpb_mh_config): lets the struct grow without breaking callers */
  int formula;     /* GPB_FORMULA_* */
  int on_notpd;    /* see gpb_mh_run */
  int rng;         /* GPB_RNG_* */
  double delta;    /* proposal step (SimulatedGp.scala:115-130) */
  double nugget;   /* 1e-3 (GaussianProcess.scala:117) */
  unsigned long long seed; /* GPB_RNG_PHILOX */
  int chain0;      /* global index of this ctx's first chain (Philox counter; gather offset) */
  int mean_step;   /* 0 | 1 */
  double mean_prior_mean, mean_prior_var; /* priorMean: Gaussian (Mcmc.scala:66) */
  double mean_nugget;                     /* 1e-6 (MeanParameters.scala:84) */
  int flush_every; /* iterations per ring segment; <= 0: default */
  int gather;      /* 0 | 1 */
  int total_chains;/* gather: chains over all ranks */
  int small_n;     /* GPB_SMALL_* */
} gpb_mh_config;

typedef struct {
  int info;            /* first LAPACK info > 0 seen on any chain (0: none) */
  int stopped_chains;  /* chains that stopped on a non-PD proposal (on_notpd == 0) */
  int path;            /* 0 general (k_build / k_chol_tasks per iteration), 1 one-launch small-n kernel */
  int flushes;         /* ring flushes (and all-gathers when gather == 1) */
  float device_ms;     /* device time of the loop (CUDA events on the ctx stream) */
} gpb_mh_result;

/* out_states [iters][C][W], out_ll [iters][C], out_accept [iters][C], out_stop [C] (may be NULL):
 * iteration at which the chain stopped, -1 = ran to the end; C = gather ? total_chains : chains. */
int gpb_mh_run_ex(gpb_ctx *ctx, const gpb_dataset *ds, const gpb_mh_config *cfg, int chains, int iters,
                  const gpb_term *init, int nterms, const double *beta0, int nbeta, const double *ig,
                  const double *z, const double *uu, const double *design, double *out_states,
                  double *out_ll, int *out_accept, int *out_stop, gpb_mh_result *res);

/* Generic prior / proposal (KernelParameters.sample takes arbitrary closures,
 * KernelParameters.scala:231-235): the host supplies them as callbacks, the device evaluates the
 * likelihoods.  Per iteration: propose(user, iter, chains, cur [chains][P], next [chains][P]) fills the
 * proposals; log_prior(user, chains, params [chains][P], out [chains]) their log prior densities; one
 * gpb_loglik_batch over all chains; accept with uu [iters][chains] (Breeze's rule).  One ABI-internal
 * round trip per iteration -- the price of host closures.  Same outputs as gpb_mh_run. */
typedef void (*gpb_propose_fn)(void *user, int iter, int chains, const double *cur, double *next);
typedef void (*gpb_log_prior_fn)(void *user, int chains, const double *params, double *out);
int gpb_mh_run_host(gpb_ctx *ctx, const gpb_dataset *ds, int chains, int iters, const gpb_term *init,
                    int nterms, const double *beta, int nbeta, const gpb_mh_opts *opts,
                    gpb_propose_fn propose, gpb_log_prior_fn log_prior, void *user, const double *uu,
                    double *out_states, double *out_ll, int *out_accept, int *info_out);

/* Counter-based streams (Philox4x32-10, key = seed, counter = (iter, slot, chain, stream)) for
 * callers that do not bring their own: fills z [iters][chains][P] and uu [iters][chains] on the host. */
int gpb_philox_streams(unsigned long long seed, int chain0, int chains, int iters, int P, double *z,
                       double *uu);
/* The same streams as the device generator of gpb_mh_run_ex (GPB_RNG_PHILOX) produces them. */
int gpb_philox_streams_device(gpb_ctx *ctx, unsigned long long seed, int chain0, int chains, int iters,
                              int P, double *z, double *uu);

/* ---- multi-GPU: chains shard across devices, one gather ------------------------ */
/* The reference runs its chains in parallel with Streaming.writeParallelChain (call sites
 * examples/.../SimulatedGp.scala:163-165, TemperatureDlm.scala:353-358).  Here chains shard across the
 * GPUs of one box, one gpb_ctx per device -- one per process under torchrun, or one per JVM thread --
 * and the only exchange is the all-gather of accepted states, log-likelihoods and accept flags
 * (NCCL over NVLink), device to device, inside gpb_mh_run_ex.
 *   gpb_comm_unique_id : rank 0 creates the 128-byte id and hands it to the other ranks by its own
 *                        means (torch.distributed store, a JVM field, ...).
 *   gpb_comm_init_rank : every rank, concurrently, with its own ctx.
 *   gpb_comm_init_all  : single process, one thread: communicators for ctxs[0..n) at once
 *                        (ncclCommInitAll); the ctxs are then driven from n threads.
 *   gpb_comm_shard     : the block partition gpb_mh_run_ex assumes: first chain and count of `rank`.
 *   gpb_gather         : all-gather of `count` doubles per rank through the same communicator (host
 *                        buffers; e.g. the log-likelihoods of sharded proposal batches).  recv: [nranks][count].
 * NCCL is loaded at run time (dlopen libnccl.so.2): without it these calls return GPB_E_NOCOMM and
 * everything single-GPU keeps working. */
#define GPB_COMM_ID_BYTES 128
int gpb_comm_unique_id(void *id128);
int gpb_comm_init_rank(gpb_ctx *ctx, int nranks, int rank, const void *id128);
int gpb_comm_init_all(gpb_ctx **ctxs, int n);
int gpb_comm_destroy(gpb_ctx *ctx);
int gpb_comm_info(const gpb_ctx *ctx, int *rank, int *nranks);
int gpb_comm_shard(int total, int nranks, int rank, int *first, int *count);
int gpb_gather(gpb_ctx *ctx, const double *send, int count, double *recv);

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
/* Schedule analysis (profiling aid): when enabled, the factorisation kernel records 16 u64 per task
 * (timestamps of receipt / first data / contraction done / dependency satisfied / factor-or-solve
 * done / release, the producer lane's issue times, start / end of a diagonal factorisation).  gpb_get_trace copies up to max_tasks records of the last call
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
