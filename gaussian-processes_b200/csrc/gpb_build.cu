// gpb_build.cu -- k_build: tiled covariance assembly, fused distance + sum-of-kernels + white
// noise at exact zero distance + nugget + identity padding + appended K* / residual rows.
//
// Replaces, per proposal (reference file:line under core/src/main/scala/gp/):
//   GaussianProcess.distanceMatrix   GaussianProcess.scala:50-61   (n^2 closure calls, boxed)
//   Location.euclidean               Location.scala:27-33          sqrt(dx^2 [+ dy^2])
//   KernelFunction.apply / SE / white KernelFunction.scala:13-15,33-35,47-55 (left fold from 0.0)
//   KernelFunction.buildCov          KernelFunction.scala:84-89
//   `+ diag(fill(n)(nugget))`        GaussianProcess.scala:117-118, Predict.scala:67-68
//   residual y - m(x)                GaussianProcess.scala:120-122, MeanFunction.scala:7-17
//   Predict.buildDistVec             Predict.scala:21-29           (K* rows)
//
// HBM-bound by design: one 8-byte store per lower-triangle entry, coalesced along rows
// (column-major), inputs (x, y: KBs) staged in shared memory.  Algorithmic bytes: 4 n^2 per
// proposal for the likelihood (lower triangle incl. diagonal tiles), 8 n m for K*.
#include "gpb_internal.h"

namespace gpb {

__device__ __forceinline__ double cov_of_dist(const Hyper& h, double d) {
  // KernelFunction.scala:47-55: ps.map(k_i(d)).sum == ((0.0 + k0) + k1) + ...
  double s = 0.0;
#pragma unroll 1
  for (int t = 0; t < h.nterms; ++t) {
    double v;
    if (h.kind[t] == GPB_KIND_SE) {
      double sg = h.p1[t];
      // KernelFunction.scala:14; explicit _rn ops: the JVM never contracts a*b+c into an FMA
      v = __dmul_rn(h.p0[t], exp(-__ddiv_rn(__dmul_rn(d, d), __dmul_rn(sg, sg))));
    } else {
      v = (d == 0.0) ? h.p0[t] : 0.0;  // KernelFunction.scala:34
    }
    s = __dadd_rn(s, v);
  }
  return s;
}

// Hot-path form of the same sum: exp(d2 * (-1/s^2)) with -1/s^2 formed once per CTA -- no square
// root, no re-squaring, no IEEE division per entry (k_build is FP64-ALU bound on exactly those).  d2 is
// within an ulp of sqrt(d2)^2, so an entry moves by at most a few ulps of the term's scale h:
// |dk| <= h x e^-x 2^-52 with x = d^2/s^2, i.e. < 0.4 ulp(h) -- five orders below the 1e-10 tolerance
// of the log-likelihood.  The matrix-level gpb_build_cov keeps the reference's exact operation order.
__device__ __forceinline__ double cov_of_dist2_fast(const Hyper& h, const double* __restrict__ ninv2, double d2) {
  double s = 0.0;
#pragma unroll 1
  for (int t = 0; t < h.nterms; ++t) {
    const double v = (h.kind[t] == GPB_KIND_SE) ? h.p0[t] * exp(d2 * ninv2[t]) : ((d2 == 0.0) ? h.p0[t] : 0.0);
    s = __dadd_rn(s, v);
  }
  return s;
}

// mode 0: covariance system, the reference's exact operation order (lower trapezoid only unless
// full != 0); mode 1: distance matrix; mode 2: covariance system, hot-path form (cov_of_dist2_fast).
__global__ void __launch_bounds__(256)
k_build(Geom g, const double* __restrict__ x, const double* __restrict__ y,
        const double* __restrict__ xstar, const Hyper* __restrict__ hyper, double nugget,
        double* __restrict__ mats, int mode, int full, const double* __restrict__ rows) {
  int ti, tj;
  if (full) {
    ti = blockIdx.x / g.nt; tj = blockIdx.x % g.nt;
  } else {
    // strictly-upper tiles are never read: the grid enumerates the lower trapezoid only
    const int tri = g.nt * (g.nt + 1) / 2;
    const int idx = blockIdx.x;
    if (idx < tri) {
      ti = (int)((sqrtf(8.0f * (float)idx + 1.0f) - 1.0f) * 0.5f);
      while (ti * (ti + 1) / 2 > idx) --ti;
      while ((ti + 1) * (ti + 2) / 2 <= idx) ++ti;
      tj = idx - ti * (ti + 1) / 2;
    } else {
      ti = g.nt + (idx - tri) / g.nt; tj = (idx - tri) % g.nt;
    }
  }
  const int b = blockIdx.y;
  __shared__ double rx[2][TILE], cx[2][TILE], cy[TILE], cm[TILE];
  __shared__ Hyper h;
  __shared__ double ninv2[MAXT];
  const int tid = threadIdx.x;
  if (tid == 0) h = hyper[b];
  __syncthreads();
  if (tid < MAXT) ninv2[tid] = (tid < h.nterms && h.kind[tid] == GPB_KIND_SE) ? -1.0 / (h.p1[tid] * h.p1[tid]) : 0.0;
  if (tid < TILE) {
    // column entities are always training points (or padding)
    int c = tj * TILE + tid;
    double c0 = 0.0, c1 = 0.0, yy = 0.0, mu = 0.0;
    if (c < g.n) {
      c0 = x[c];
      if (g.dim == 2) c1 = x[g.n + c];
      if (y != nullptr) yy = y[c];
      if (h.nbeta > 0) {  // MeanFunction.scala:7-10 plane
        mu = __dadd_rn(h.beta[0], __dmul_rn(c0, h.beta[1]));
        if (g.dim == 2) mu = __dadd_rn(mu, __dmul_rn(c1, h.beta[2]));
      }
    }
    cx[0][tid] = c0; cx[1][tid] = c1; cy[tid] = yy; cm[tid] = mu;
  } else if (tid < 2 * TILE) {
    int r = ti * TILE + (tid - TILE);
    double r0 = 0.0, r1 = 0.0;
    if (r < g.n) {
      r0 = x[r];
      if (g.dim == 2) r1 = x[g.n + r];
    } else if (r >= g.npad && r - g.npad < g.m) {
      int a = r - g.npad;
      r0 = xstar[a];
      if (g.dim == 2) r1 = xstar[g.m + a];
    }
    rx[0][tid - TILE] = r0; rx[1][tid - TILE] = r1;
  }
  __syncthreads();

  double* M = mats + (size_t)b * g.mat_stride;
  const int rl = tid & (TILE - 1);
  const int R = ti * TILE + rl;
  const double r0 = rx[0][rl], r1 = rx[1][rl];
  // row class: 0 training, 1 padding (identity), 2 test point, 3 rhs row, 4 zero, 5 explicit row
  int rclass;
  if (R < g.n) rclass = 0;
  else if (R < g.npad) rclass = 1;
  else if (R - g.npad < g.m) rclass = 2;
  else if (R - g.npad - g.m < g.design) rclass = 5;
  else if (R == g.rhs_row) rclass = 3;
  else rclass = 4;

#pragma unroll 4
  for (int q = 0; q < TILE / 4; ++q) {
    const int cl = (tid >> 6) + 4 * q;
    const int Cc = tj * TILE + cl;
    double v;
    if (Cc >= g.n) {
      v = (rclass == 1 && R == Cc) ? 1.0 : 0.0;
    } else if (rclass == 0 || rclass == 2) {
      double dx = r0 - cx[0][cl];
      double d2 = __dmul_rn(dx, dx);
      if (g.dim == 2) {
        double dy = r1 - cx[1][cl];
        d2 = __dadd_rn(d2, __dmul_rn(dy, dy));  // no FMA contraction: mirror Location.scala:30-31
      }
      if (mode == 2) {
        v = cov_of_dist2_fast(h, ninv2, d2);
        if (rclass == 0 && R == Cc) v = __dadd_rn(v, nugget);
      } else {
        double d = sqrt(d2);  // Location.scala:28-31
        if (mode == 1) {
          v = d;
        } else {
          v = cov_of_dist(h, d);
          if (rclass == 0 && R == Cc) v = __dadd_rn(v, nugget);  // GaussianProcess.scala:117-118
        }
      }
    } else if (rclass == 5) {
      // design-matrix column (MeanParameters.scala:55-65), or a row of the identity block
      v = (g.ident0 >= 0) ? ((R - g.npad == Cc) ? 1.0 : 0.0) : rows[(size_t)(R - g.npad - g.m) * g.n + Cc];
    } else if (rclass == 3) {
      v = cy[cl] - cm[cl];  // GaussianProcess.scala:120-122
    } else {
      v = 0.0;
    }
    M[(size_t)Cc * g.ld + R] = v;
  }
}

void launch_build(cudaStream_t st, const Geom& g, const double* x, const double* y,
                  const double* xstar, const Hyper* hyper, double nugget, double* mats, int mode,
                  int full, const double* rows) {
  const int tiles = full ? g.rt * g.nt : g.nt * (g.nt + 1) / 2 + (g.rt - g.nt) * g.nt;
  dim3 grid(tiles, g.batch);
  k_build<<<grid, 256, 0, st>>>(g, x, y, xstar, hyper, nugget, mats, mode, full, rows);
}

}  // namespace gpb
