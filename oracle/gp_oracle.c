/*
 * gp_oracle.c -- CPU restatement of the reference's GP inference hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under gaussian-processes_b200/ may link, import or
 * execute this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, and only as the checker or the timed CPU baseline.
 *
 * PARITY UNPINNED: the reference (jonnylaw/gaussian-processes, Scala 2.12 + Breeze 0.13.2 +
 * netlib-java F2J) ships no golden vectors, no fixtures and no live test for this path
 *  and cannot be compiled here (no JVM).  This file follows the cited
 * Scala source line by line; its authority comes from (i) that fidelity, (ii) agreement of
 * three evaluations (double in LAPACK order, long double, scipy LAPACK -- tests/test_oracle.py)
 * and (iii) hand-checkable closed forms for n = 1, 2.
 *
 * Third-party arithmetic restated here (not vendored under /root/reference):
 *   - org.scalanlp:breeze_2.12:0.13.2  `cholesky`  -> LAPACK dpotrf("L")   (call sites
 *     GaussianProcess.scala:123, Predict.scala:69)
 *   - com.github.fommil.netlib:core:1.1.2 F2J  dpotrf / dpotf2 / dtrsv / dtrsm (reference
 *     LAPACK 3.1 loop structure: blocked left-looking, NB = 64, dpotf2 on the diagonal block)
 *   - breeze MarkovChain.Kernels.metropolis (Breeze 0.13.2 MarkovChain.scala): per step
 *     next <- proposal(t); a = min(1, exp(ll(next) - ll(t))); u ~ U(0,1); u < a ? next : t,
 *     re-evaluating ll(t) every step.
 *
 * All matrices are column-major with leading dimension n (Breeze DenseMatrix layout).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

#define ORC_SE 0
#define ORC_WHITE 1

typedef struct {
  int kind;      /* 0 = SquaredExp(h, sigma), 1 = White(sigma) */
  double p0, p1; /* SE: h, sigma; White: sigma */
  double p2;     /* unused (Matern reserved) */
} orc_term;

/* Location.scala:27-33  euclidean: sqrt(pow(dx,2) [+ pow(dy,2)]).  Math.pow(v, 2) == v*v exactly. */
double orc_euclidean(int dim, const double *a, const double *b) {
  if (dim == 1) {
    double dx = a[0] - b[0];
    return sqrt(dx * dx);
  }
  double dx = a[0] - b[0], dy = a[1] - b[1];
  return sqrt(dx * dx + dy * dy);
}

/* KernelFunction.scala:13-15  squaredExponential: h * exp(-(d*d) / (sigma*sigma)) */
static double se(double h, double sigma, double d) { return h * exp(-(d * d) / (sigma * sigma)); }

/* KernelFunction.scala:33-35  white: if (dist == 0) sigma else 0.0 */
static double white(double sigma, double d) { return d == 0 ? sigma : 0.0; }

/* KernelFunction.scala:47-55  apply: ps.map(k_i(x)).sum  -- left fold from 0.0 in list order */
double orc_cov_fn(const orc_term *terms, int nterms, double d) {
  double s = 0.0;
  for (int t = 0; t < nterms; ++t) {
    if (terms[t].kind == ORC_SE)
      s = s + se(terms[t].p0, terms[t].p1, d);
    else
      s = s + white(terms[t].p0, d);
  }
  return s;
}

/* x is SoA [dim][n] (the C-ABI layout); fetch location i into loc[2]. */
static void loc_of(int n, int dim, const double *x, int i, double *loc) {
  loc[0] = x[i];
  loc[1] = dim == 2 ? x[n + i] : 0.0;
}

/* GaussianProcess.scala:50-61 distanceMatrix: res = for (x <- xs; y <- xs) yield dist(x, y);
 * new DenseMatrix(n, n, res)  => column-major, element (row=j, col=i) = dist(xs(i), xs(j)). */
void orc_distance_matrix(int n, int dim, const double *x, double *D) {
  double a[2], b[2];
  for (int i = 0; i < n; ++i) {
    loc_of(n, dim, x, i, a);
    for (int j = 0; j < n; ++j) {
      loc_of(n, dim, x, j, b);
      D[(size_t)i * n + j] = orc_euclidean(dim, a, b);
    }
  }
}

/* KernelFunction.scala:84-89 buildCov = distanceMatrix(xs, dist).map(covFn), then
 * `+ diag(fill(n)(nugget))` (GaussianProcess.scala:117-118, Predict.scala:67-68). Full n x n. */
void orc_build_cov(int n, int dim, const double *x, const orc_term *terms, int nterms,
                   double nugget, double *K) {
  orc_distance_matrix(n, dim, x, K);
  for (size_t e = 0; e < (size_t)n * n; ++e) K[e] = orc_cov_fn(terms, nterms, K[e]);
  for (int i = 0; i < n; ++i) K[(size_t)i * n + i] += nugget;
}

/* MeanFunction.scala:7-17: zero, or plane beta0 + x*beta1 (+ y*beta2). nbeta = 0 | 2 | 3. */
double orc_mean_fn(int dim, const double *beta, int nbeta, const double *loc) {
  if (nbeta == 0) return 0.0;
  if (dim == 1) return beta[0] + loc[0] * beta[1];
  return beta[0] + loc[0] * beta[1] + loc[1] * beta[2];
}

/* ---------------------------------------------------------------------------------------------
 * Breeze `cholesky` -> F2J dpotrf("L").  Reference-LAPACK 3.1 loop structure, restated.
 * A(i,j) is a[i + j*lda].  Returns LAPACK info (0 ok; k>0: leading minor k not positive).
 * ------------------------------------------------------------------------------------------- */
#define A_(i, j) a[(size_t)(i) + (size_t)(j) * lda]

/* dpotf2 'L': per column j: ajj = A(j,j) - dot(row j); sqrt; dgemv update of the column; dscal */
static int potf2_lower(int n, double *a, int lda) {
  for (int j = 0; j < n; ++j) {
    double dot = 0.0;
    for (int k = 0; k < j; ++k) dot += A_(j, k) * A_(j, k);
    double ajj = A_(j, j) - dot;
    if (!(ajj > 0.0)) { /* also catches NaN, like LAPACK's disnan test */
      A_(j, j) = ajj;
      return j + 1;
    }
    ajj = sqrt(ajj);
    A_(j, j) = ajj;
    if (j < n - 1) {
      /* dgemv('N', n-j-1, j, -1, A(j+1,0), lda, A(j,0), lda, 1, A(j+1,j), 1) */
      for (int k = 0; k < j; ++k) {
        double temp = -A_(j, k);
        if (temp != 0.0)
          for (int i = j + 1; i < n; ++i) A_(i, j) += temp * A_(i, k);
      }
      double r = 1.0 / ajj; /* dscal(n-j-1, 1/ajj, ...) */
      for (int i = j + 1; i < n; ++i) A_(i, j) *= r;
    }
  }
  return 0;
}


/* y[0..m) += sum_{l<k} (-w[l*ldw]) * a[l*lda + 0..m), applied in ascending l per element. */
static void axpy_cols(int m, int k, const double *w, int ldw, const double *a, int lda, double *y) {
  int l = 0;
  for (; l + 4 <= k; l += 4) {
    const double t0 = -w[(size_t)l * ldw], t1 = -w[(size_t)(l + 1) * ldw];
    const double t2 = -w[(size_t)(l + 2) * ldw], t3 = -w[(size_t)(l + 3) * ldw];
    const double *a0 = a + (size_t)l * lda, *a1 = a0 + lda, *a2 = a1 + lda, *a3 = a2 + lda;
    for (int i = 0; i < m; ++i) y[i] = (((y[i] + t0 * a0[i]) + t1 * a1[i]) + t2 * a2[i]) + t3 * a3[i];
  }
  for (; l < k; ++l) {
    const double t0 = -w[(size_t)l * ldw];
    const double *a0 = a + (size_t)l * lda;
    for (int i = 0; i < m; ++i) y[i] += t0 * a0[i];
  }
}

/* dpotrf 'L', NB = 64 (ilaenv default for DPOTRF). */
int orc_potrf_lower(int n, double *a, int lda) {
  const int nb = 64;
  if (nb >= n) return potf2_lower(n, a, lda);
  for (int j = 0; j < n; j += nb) {
    int jb = n - j < nb ? n - j : nb;
    /* dsyrk('L','N', jb, j, -1, A(j,0), lda, 1, A(j,j), lda): C(i,c) += -A(c,l) * A(i,l), l ascending.
     * The l loop is unrolled-and-jammed by 4 (fewer passes over the C column); every element
     * still receives its updates in ascending l, so results equal the reference BLAS loop nest. */
    for (int c = 0; c < jb; ++c)
      axpy_cols(jb - c, j, &A_(j + c, 0), lda, &A_(j + c, 0), lda, &A_(j + c, j + c));
    int info = potf2_lower(jb, &A_(j, j), lda);
    if (info != 0) return info + j;
    if (j + jb < n) {
      int m = n - j - jb;
      double *b = &A_(j + jb, j);
      /* dgemm('N','T', m, jb, j, -1, A(j+jb,0), lda, A(j,0), lda, 1, A(j+jb,j), lda) */
      for (int c = 0; c < jb; ++c)
        axpy_cols(m, j, &A_(j + c, 0), lda, &A_(j + jb, 0), lda, b + (size_t)c * lda);
      /* dtrsm('R','L','T','N', m, jb, 1, A(j,j), lda, A(j+jb,j), lda) */
      for (int k = 0; k < jb; ++k) {
        double temp = 1.0 / A_(j + k, j + k);
        double *bk = b + (size_t)k * lda;
        for (int i = 0; i < m; ++i) bk[i] *= temp;
        for (int c = k + 1; c < jb; ++c) {
          double v = A_(j + c, j + k);
          if (v != 0.0) {
            double *bc = b + (size_t)c * lda;
            for (int i = 0; i < m; ++i) bc[i] -= v * bk[i];
          }
        }
      }
    }
  }
  return 0;
}

/* Breeze cholesky returns L with the strict upper triangle zeroed. */
void orc_zero_upper(int n, double *a, int lda) {
  for (int j = 1; j < n; ++j)
    for (int i = 0; i < j; ++i) A_(i, j) = 0.0;
}

/* Predict.scala:40-45 forwardSolve -> dtrsv("L","N","N", n, A, n, x, 1) (reference BLAS loop). */
void orc_trsv_lower(int n, const double *a, int lda, double *x) {
  for (int j = 0; j < n; ++j) {
    if (x[j] != 0.0) {
      x[j] /= A_(j, j);
      double temp = x[j];
      for (int i = j + 1; i < n; ++i) x[i] -= temp * A_(i, j);
    }
  }
}

/* Mcmc.scala:35-41 backSolve(A, y) -> dtrsv("U","N","N") */
void orc_trsv_upper(int n, const double *a, int lda, double *x) {
  for (int j = n - 1; j >= 0; --j) {
    if (x[j] != 0.0) {
      x[j] /= A_(j, j);
      double temp = x[j];
      for (int i = j - 1; i >= 0; --i) x[i] -= temp * A_(i, j);
    }
  }
}

/* Mcmc.scala:52-58 backSolve(A, Y) -> dtrsm("L","U","N","N", m, nrhs, 1.0, A, m, Y, m) */
void orc_trsm_upper(int m, int nrhs, const double *a, int lda, double *b, int ldb) {
  for (int j = 0; j < nrhs; ++j) {
    double *bj = b + (size_t)j * ldb;
    for (int k = m - 1; k >= 0; --k) {
      if (bj[k] != 0.0) {
        bj[k] /= A_(k, k);
        for (int i = 0; i < k; ++i) bj[i] -= bj[k] * A_(i, k);
      }
    }
  }
}
#undef A_

/* ---------------------------------------------------------------------------------------------
 * GaussianProcess.loglikelihood  (GaussianProcess.scala:109-127)
 * ------------------------------------------------------------------------------------------- */
typedef struct {
  double Q;       /* u . u                     */
  double U1;      /* sum_i u_i                 */
  double S;       /* sum_i log L_ii            */
  double ll_ref;  /* as-parsed Scala: (-0.5*u) dot (u - S - C)    -- parity target */
  double ll_gpml; /* intended: -0.5 Q - S - C  (docs parameter_inference.md:64-66)            */
  int n;
  int info;       /* LAPACK info from the factorisation */
} orc_ll;

/* work: n*n + n doubles. */
int orc_loglik(int n, int dim, const double *x, const double *y, const orc_term *terms, int nterms,
               const double *beta, int nbeta, double nugget, double *work, orc_ll *out) {
  double *K = work, *u = work + (size_t)n * n;
  orc_build_cov(n, dim, x, terms, nterms, nugget, K);             /* :117-118 */
  double loc[2];
  for (int i = 0; i < n; ++i) {                                   /* :120-122 ys - mean(xs) */
    loc_of(n, dim, x, i, loc);
    u[i] = y[i] - orc_mean_fn(dim, beta, nbeta, loc);
  }
  out->n = n;
  out->info = orc_potrf_lower(n, K, n);                           /* :123 cholesky(kxx) */
  if (out->info != 0) {
    out->Q = out->U1 = out->S = out->ll_ref = out->ll_gpml = NAN;
    return 0;
  }
  orc_trsv_lower(n, K, n, u);                                     /* :124 forwardSolve(l, ys) */
  double S = 0.0, Q = 0.0, U1 = 0.0;
  for (int i = 0; i < n; ++i) S += log(K[(size_t)i * n + i]);     /* sum(log(diag(l))) */
  double C = n * 0.5 * log(2 * M_PI);
  double ref = 0.0;
  for (int i = 0; i < n; ++i) {
    Q += u[i] * u[i];
    U1 += u[i];
    ref += (-0.5 * u[i]) * ((u[i] - S) - C);                      /* :126 as parsed */
  }
  out->Q = Q; out->U1 = U1; out->S = S;
  out->ll_ref = ref;
  out->ll_gpml = -0.5 * Q - S - C;
  return 0;
}

/* The same evaluation carried in long double (x87 80-bit) with an unblocked Cholesky: used only
 * to measure the double-precision noise floor of the quantities above (tests/test_oracle.py). */
int orc_loglik_ld(int n, int dim, const double *x, const double *y, const orc_term *terms,
                  int nterms, const double *beta, int nbeta, double nugget, orc_ll *out) {
  long double *K = (long double *)malloc(sizeof(long double) * ((size_t)n * n + n));
  if (!K) return -1;
  long double *u = K + (size_t)n * n;
  double a[2], b[2];
  for (int i = 0; i < n; ++i) {
    loc_of(n, dim, x, i, a);
    for (int j = 0; j < n; ++j) {
      loc_of(n, dim, x, j, b);
      long double dx = (long double)a[0] - b[0], dy = (long double)a[1] - b[1];
      long double d = sqrtl(dx * dx + dy * dy), s = 0.0L;
      for (int t = 0; t < nterms; ++t) {
        if (terms[t].kind == ORC_SE)
          s += (long double)terms[t].p0 * expl(-(d * d) / ((long double)terms[t].p1 * terms[t].p1));
        else
          s += d == 0 ? (long double)terms[t].p0 : 0.0L;
      }
      K[(size_t)i * n + j] = s + (i == j ? (long double)nugget : 0.0L);
    }
    u[i] = (long double)y[i] - orc_mean_fn(dim, beta, nbeta, a);
  }
  out->n = n; out->info = 0;
  for (int j = 0; j < n && out->info == 0; ++j) {
    long double *cj = K + (size_t)j * n;
    for (int k = 0; k < j; ++k) {
      long double ljk = K[(size_t)k * n + j];
      const long double *ck = K + (size_t)k * n;
      for (int i = j; i < n; ++i) cj[i] -= ljk * ck[i];
    }
    if (!(cj[j] > 0)) { out->info = j + 1; break; }
    long double d = sqrtl(cj[j]);
    cj[j] = d;
    for (int i = j + 1; i < n; ++i) cj[i] /= d;
  }
  if (out->info == 0) {
    long double S = 0, Q = 0, U1 = 0;
    for (int j = 0; j < n; ++j) {
      u[j] /= K[(size_t)j * n + j];
      for (int i = j + 1; i < n; ++i) u[i] -= u[j] * K[(size_t)j * n + i];
      S += logl(K[(size_t)j * n + j]);
      Q += u[j] * u[j];
      U1 += u[j];
    }
    long double C = n * 0.5L * logl(2 * M_PIl);
    out->Q = (double)Q; out->U1 = (double)U1; out->S = (double)S;
    out->ll_ref = (double)(-0.5L * Q + 0.5L * (S + C) * U1);
    out->ll_gpml = (double)(-0.5L * Q - S - C);
  } else {
    out->Q = out->U1 = out->S = out->ll_ref = out->ll_gpml = NAN;
  }
  free(K);
  return 0;
}

/* ---------------------------------------------------------------------------------------------
 * Predict.fit  (Predict.scala:57-94).  xstar SoA [dim][m].  mean/var length m.
 * Mirrors the reference exactly: nugget passed by caller (1e-6 at :67), raw y with NO mean
 * subtraction (:80), kyy = covFn(dist(xs,xs)) = covFn(0) including the white term (:78),
 * u = L^-1 k*, v = L^-1 y, mean = u.v, var = kyy - u.u (:89-92).  v is loop-invariant, computed
 * once here (the reference recomputes it per point with the identical result).
 * work: n*n + 2n doubles.
 * ------------------------------------------------------------------------------------------- */
int orc_predict(int n, int dim, const double *x, const double *y, const orc_term *terms, int nterms,
                double nugget, int m, const double *xstar, double *mean, double *var, double *work,
                int *info) {
  double *K = work, *v = work + (size_t)n * n, *u = v + n;
  orc_build_cov(n, dim, x, terms, nterms, nugget, K);
  *info = orc_potrf_lower(n, K, n);
  if (*info != 0) return 0;
  memcpy(v, y, sizeof(double) * n);
  orc_trsv_lower(n, K, n, v);
  double a[2], b[2];
  for (int j = 0; j < m; ++j) {
    loc_of(m, dim, xstar, j, b);
    for (int i = 0; i < n; ++i) { /* buildDistVec :21-29: covFn(dist(d, newx)) */
      loc_of(n, dim, x, i, a);
      u[i] = orc_cov_fn(terms, nterms, orc_euclidean(dim, a, b));
    }
    double kyy = orc_cov_fn(terms, nterms, orc_euclidean(dim, b, b));
    orc_trsv_lower(n, K, n, u);
    double mu = 0.0, uu = 0.0;
    for (int i = 0; i < n; ++i) { mu += u[i] * v[i]; uu += u[i] * u[i]; }
    mean[j] = mu;
    var[j] = kyy - uu;
  }
  return 0;
}

/* ---------------------------------------------------------------------------------------------
 * Metropolis-Hastings over the kernel hyper-parameters.
 *   KernelParameters.sample (KernelParameters.scala:231-246): ll = prior(kp) + loglikelihood
 *   Breeze metropolis (see header): evaluates ll(next) AND ll(t) every step.
 *   Proposal = the shipped example's log-normal random walk (SimulatedGp.scala:115-130):
 *     SE(h,s) -> SE(h*exp(z1), s*exp(z2)), White(s) -> White(s*exp(z)), z ~ Gaussian(0, delta)
 *     drawn left to right over the term list (cats traverse).
 *   Prior = sum over terms of InverseGamma(shape, scale).logPdf per parameter
 *     (SimulatedGp.scala:132-151; InverseGamma from bayesian_dlms 0.4.0, unvendored:
 *      shape*ln(scale) - lnGamma(shape) - (shape+1)*ln(x) - scale/x).
 * RNG: explicit streams.  z[it*P + p] standard normals (scaled by delta here), uu[it] uniforms.
 * Flattened parameter order = Parameters.toList order (GaussianProcess.scala:24-25,
 * KernelParameters.scala:26,58): per term SE -> (h, sigma), White -> (sigma).
 * ------------------------------------------------------------------------------------------- */
int orc_nparams(const orc_term *terms, int nterms) {
  int p = 0;
  for (int t = 0; t < nterms; ++t) p += terms[t].kind == ORC_SE ? 2 : 1;
  return p;
}

double orc_invgamma_logpdf(double shape, double scale, double xv) {
  return shape * log(scale) - lgamma(shape) - (shape + 1.0) * log(xv) - scale / xv;
}

/* priorKernel: ps.map(...).sum ; ig = [P][2] (shape, scale) per flattened parameter */
double orc_prior(const orc_term *terms, int nterms, const double *ig) {
  double s = 0.0;
  int p = 0;
  for (int t = 0; t < nterms; ++t) {
    if (terms[t].kind == ORC_SE) {
      double v = orc_invgamma_logpdf(ig[2 * p], ig[2 * p + 1], terms[t].p0) +
                 orc_invgamma_logpdf(ig[2 * p + 2], ig[2 * p + 3], terms[t].p1);
      s = s + v;
      p += 2;
    } else {
      s = s + orc_invgamma_logpdf(ig[2 * p], ig[2 * p + 1], terms[t].p0);
      p += 1;
    }
  }
  return s;
}

/* formula: 0 = as-parsed ll_ref (what the reference executes), 1 = intended ll_gpml.
 * evals_per_iter: 2 = faithful to Breeze (re-evaluate ll(t)); 1 = cache it (same numbers).
 * out_states [iters][P], out_ll [iters] (prior + loglik of the state after the step),
 * out_accept [iters].  Returns 0, or LAPACK info > 0 if a proposal's K was not PD (the
 * reference throws there: SURVEY 0.2l). */
int orc_mh_run_a(int n, int dim, const double *x, const double *y, const orc_term *init, int nterms,
                 const double *beta, int nbeta, double nugget, const double *ig, double delta,
                 int iters, const double *z, const double *uu, int formula, int evals_per_iter,
                 double *out_states, double *out_ll, int *out_accept, double *out_a, int *out_done) {
  int P = orc_nparams(init, nterms);
  double *work = (double *)malloc(sizeof(double) * ((size_t)n * n + n));
  orc_term *cur = (orc_term *)malloc(sizeof(orc_term) * nterms * 2);
  if (!work || !cur) return -1;
  orc_term *next = cur + nterms;
  memcpy(cur, init, sizeof(orc_term) * nterms);
  orc_ll r;
  orc_loglik(n, dim, x, y, cur, nterms, beta, nbeta, nugget, work, &r);
  if (out_done) *out_done = 0;
  if (r.info) { free(work); free(cur); return r.info; }
  double old_ll = orc_prior(cur, nterms, ig) + (formula ? r.ll_gpml : r.ll_ref);
  for (int it = 0; it < iters; ++it) {
    if (out_done) *out_done = it;   /* iterations completed before a throw */
    memcpy(next, cur, sizeof(orc_term) * nterms);
    int p = 0;
    for (int t = 0; t < nterms; ++t) {
      if (next[t].kind == ORC_SE) {
        next[t].p0 = cur[t].p0 * exp(delta * z[(size_t)it * P + p]);
        next[t].p1 = cur[t].p1 * exp(delta * z[(size_t)it * P + p + 1]);
        p += 2;
      } else {
        next[t].p0 = cur[t].p0 * exp(delta * z[(size_t)it * P + p]);
        p += 1;
      }
    }
    orc_loglik(n, dim, x, y, next, nterms, beta, nbeta, nugget, work, &r);
    if (r.info) { free(work); free(cur); return r.info; }
    double new_ll = orc_prior(next, nterms, ig) + (formula ? r.ll_gpml : r.ll_ref);
    if (evals_per_iter == 2) {
      orc_loglik(n, dim, x, y, cur, nterms, beta, nbeta, nugget, work, &r);
      old_ll = orc_prior(cur, nterms, ig) + (formula ? r.ll_gpml : r.ll_ref);
    }
    /* JVM semantics: scala.math.min(1.0, NaN) is NaN and `u < NaN` is false, so a NaN on either side
     * rejects; C's fmin(1.0, NaN) returns 1.0 and would always accept */
    double dl = new_ll - old_ll;
    int acc = (dl == dl) && uu[it] < fmin(1.0, exp(dl));
    if (out_a) out_a[it] = (dl == dl) ? fmin(1.0, exp(dl)) : NAN;   /* the acceptance probability itself */
    if (acc) {
      memcpy(cur, next, sizeof(orc_term) * nterms);
      old_ll = new_ll;
    }
    p = 0;
    for (int t = 0; t < nterms; ++t) {
      out_states[(size_t)it * P + p++] = cur[t].p0;
      if (cur[t].kind == ORC_SE) out_states[(size_t)it * P + p++] = cur[t].p1;
    }
    out_ll[it] = old_ll;
    out_accept[it] = acc;
  }
  if (out_done) *out_done = iters;
  free(work);
  free(cur);
  return 0;
}

int orc_mh_run(int n, int dim, const double *x, const double *y, const orc_term *init, int nterms,
               const double *beta, int nbeta, double nugget, const double *ig, double delta,
               int iters, const double *z, const double *uu, int formula, int evals_per_iter,
               double *out_states, double *out_ll, int *out_accept) {
  return orc_mh_run_a(n, dim, x, y, init, nterms, beta, nbeta, nugget, ig, delta, iters, z, uu, formula,
                      evals_per_iter, out_states, out_ll, out_accept, NULL, NULL);
}
