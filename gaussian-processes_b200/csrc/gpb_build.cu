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
#include "gpb_cov.cuh"

namespace gpb {

#ifndef GPB_BUILD_MINB
#define GPB_BUILD_MINB 4
#endif

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

// exp(x) for x <= 0 on the hot path, four independent arguments at a time: the textbook scheme of
// every FP64 libm (k = rint(x log2 e) by the magic-number add, r = x - k ln2 in two FMAs, degree-11
// minimax polynomial, scale by 2^k through the exponent field) with the rare cases moved out of line:
// no two-step scaling and no range tests per argument.  Max error 0.84 ulp over [-700, 0] (checked in
// exact rational arithmetic, tools/exp_check.py) -- the same <= 1 ulp bound as CUDA's exp and
// java.lang.Math.exp.  If any of the four is below -700 (results under 1e-304), -inf or NaN, the group
// takes the library routine.  15 FP64 + 1 integer instruction per argument, coefficients straight
// from the constant bank, the four Horner chains interleaved.
__constant__ double EXP_C[10] = {2.5052097064908941e-08, 2.7626262793835868e-07, 2.7557414788000726e-06,
                                 2.4801504602132958e-05, 1.9841269707468915e-04, 1.3888888932258898e-03,
                                 8.3333333333978320e-03, 4.1666666666573905e-02, 1.6666666666666563e-01,
                                 5.0000000000000056e-01};
__device__ __noinline__ double exp_library(double x) { return exp(x); }
__device__ __forceinline__ void exp_nonpos4(const double (&x)[4], double (&e)[4]) {
  const double MAGIC = 6755399441055744.0;   // 1.5 * 2^52
  double t[4], r[4], p[4];
  int k[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    t[u] = fma(x[u], 1.4426950408889634, MAGIC);
    k[u] = __double2loint(t[u]);
    t[u] -= MAGIC;
    r[u] = fma(t[u], -6.93147180559945286e-01, x[u]);
    r[u] = fma(t[u], -2.31904681384629956e-17, r[u]);
    p[u] = EXP_C[0];
  }
#pragma unroll
  for (int c = 1; c < 10; ++c)
#pragma unroll
    for (int u = 0; u < 4; ++u) p[u] = fma(p[u], r[u], EXP_C[c]);
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    p[u] = fma(p[u], r[u], 1.0);
    p[u] = fma(p[u], r[u], 1.0);
    // p in [0.70, 1.42], k >= -1010: p * 2^k is normal, the exponent add is exact
    e[u] = __hiloint2double(__double2hiint(p[u]) + (k[u] << 20), __double2loint(p[u]));
  }
  // negative doubles order like their high words read as unsigned: hi(x) > hi(-700.0) <=> x < -700, -inf or -NaN
  const unsigned h01 = max((unsigned)__double2hiint(x[0]), (unsigned)__double2hiint(x[1]));
  const unsigned h23 = max((unsigned)__double2hiint(x[2]), (unsigned)__double2hiint(x[3]));
  if (max(h01, h23) > 0xC085E000u) {
#pragma unroll
    for (int u = 0; u < 4; ++u) e[u] = exp_library(x[u]);
  }
}

// mode 0: covariance system, the reference's exact operation order (lower trapezoid only unless
// full != 0); mode 1: distance matrix; mode 2: covariance system, hot-path form.
// A CTA builds `tpc` consecutive tiles of one proposal: the hyper-parameters are fetched once (all
// words in parallel), and the coordinates of the next tile are fetched into registers while the
// current tile is computed.
__device__ __forceinline__ void tile_of(const Geom& g, int full, int idx, int& ti, int& tj) {
  if (full) { ti = idx / g.nt; tj = idx % g.nt; return; }
  // strictly-upper tiles are never read: the grid enumerates the lower trapezoid only
  const int tri = g.nt * (g.nt + 1) / 2;
  if (idx < tri) {
    ti = (int)((sqrtf(8.0f * (float)idx + 1.0f) - 1.0f) * 0.5f);
    while (ti * (ti + 1) / 2 > idx) --ti;
    while ((ti + 1) * (ti + 2) / 2 <= idx) ++ti;
    tj = idx - ti * (ti + 1) / 2;
  } else {
    ti = g.nt + (idx - tri) / g.nt; tj = (idx - tri) % g.nt;
  }
}

__global__ void __launch_bounds__(256, GPB_BUILD_MINB)
k_build(Geom g, const double* __restrict__ x, const double* __restrict__ y,
        const double* __restrict__ xstar, const Hyper* __restrict__ hyper, double nugget,
        double* __restrict__ mats, int mode, int full, const double* __restrict__ rows, int ntiles, int tpc) {
  const int b = blockIdx.y;
  __shared__ double rx[2][TILE], cx[2][TILE], cy[TILE];
  __shared__ Hyper h;
  __shared__ double ninv2[MAXT];
  const int tid = threadIdx.x;
  static_assert(sizeof(Hyper) % 4 == 0 && sizeof(Hyper) / 4 <= 256, "Hyper is copied one word per thread");
  if (tid < (int)(sizeof(Hyper) / 4))
    reinterpret_cast<int*>(&h)[tid] = reinterpret_cast<const int*>(hyper + b)[tid];

  // coordinates of one tile: threads 0..63 the column entities (always training points or padding),
  // threads 64..127 the row entities (training, padding, or test points)
  auto fetch = [&](int ti, int tj, double& v0, double& v1, double& v2) {
    v0 = v1 = v2 = 0.0;
    if (tid < TILE) {
      const int c = tj * TILE + tid;
      if (c < g.n) {
        v0 = x[c];
        if (g.dim == 2) v1 = x[g.n + c];
        if (y != nullptr) v2 = y[c];
      }
    } else if (tid < 2 * TILE) {
      const int r = ti * TILE + (tid - TILE);
      if (r < g.n) {
        v0 = x[r];
        if (g.dim == 2) v1 = x[g.n + r];
      } else if (r >= g.npad && r - g.npad < g.m) {
        const int a = r - g.npad;
        v0 = xstar[a];
        if (g.dim == 2) v1 = xstar[g.m + a];
      }
    }
  };
  const int idx0 = blockIdx.x * tpc, idx1 = min(idx0 + tpc, ntiles);
  int ti, tj;
  tile_of(g, full, idx0, ti, tj);
  double v0, v1, v2;
  fetch(ti, tj, v0, v1, v2);
  __syncthreads();
  if (tid < MAXT) ninv2[tid] = (tid < h.nterms && h.kind[tid] == GPB_KIND_SE) ? -1.0 / (h.p1[tid] * h.p1[tid]) : 0.0;

  double* M = mats + (size_t)b * g.mat_stride;
  const int rl = tid & (TILE - 1);
  for (int idx = idx0; idx < idx1; ++idx) {
    if (idx > idx0) __syncthreads();          // every thread is done with the previous tile's coordinates
    if (tid < TILE) { cx[0][tid] = v0; cx[1][tid] = v1; cy[tid] = v2; }
    else if (tid < 2 * TILE) { rx[0][tid - TILE] = v0; rx[1][tid - TILE] = v1; }
    __syncthreads();
    const int cti = ti, ctj = tj;
    if (idx + 1 < idx1) { tile_of(g, full, idx + 1, ti, tj); fetch(ti, tj, v0, v1, v2); }

    const int R = cti * TILE + rl;
    const double r0 = rx[0][rl], r1 = rx[1][rl];

    // Hot path: a tile whose 64 rows are all training points (or all test points) and whose 64 columns
    // are all training points -- every tile of K and K* when n is a multiple of 64.  No per-entry
    // classification; the term loop runs once per FOUR entries; exp_nonpos4 instead of the library exp.
    if (mode == 2 && ctj * TILE + TILE <= g.n &&
        (cti * TILE + TILE <= g.n || (cti * TILE >= g.npad && cti * TILE + TILE <= g.npad + g.m))) {
      const int c0 = tid >> 6;
      const int nterms_reg = h.nterms;
      const int tk0 = h.kind[0], tk1 = h.kind[1], tk2 = h.kind[2];
      const double tp0 = h.p0[0], tp1 = h.p0[1], tp2 = h.p0[2];
      const double tn0 = ninv2[0], tn1 = ninv2[1], tn2 = ninv2[2];
      const size_t step = (size_t)4 * g.ld;                       // this thread's columns are c0 + 4 e, e = 0..15
      double* pc = M + (size_t)(ctj * TILE + c0) * g.ld + R;
      // the nugget lands on this thread's entry e = (rl - c0) / 4 of a diagonal tile, if that is an integer
      const int ediag = (cti == ctj && ((rl - c0) & 3) == 0) ? (rl - c0) >> 2 : -1;
#pragma unroll 1
      for (int q = 0; q < TILE / 16; ++q, pc += 4 * step) {
        double d2[4], s[4];
        unsigned nz = 0xffffffffu;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int cl = c0 + 4 * (4 * q + u);
          const double dx = r0 - cx[0][cl];
          d2[u] = __dmul_rn(dx, dx);
          if (g.dim == 2) {
            const double dy = r1 - cx[1][cl];
            d2[u] = __dadd_rn(d2[u], __dmul_rn(dy, dy));
          }
          s[u] = 0.0;
          // d2 >= 0 (or NaN): it is zero iff all its bits are (a NaN never equals zero either way)
          nz = min(nz, (unsigned)__double2hiint(d2[u]) | (unsigned)__double2loint(d2[u]));
        }
        // one term: the same operations in the same order whichever way the terms are walked (left fold)
        auto add_term = [&](int kind, double p0, double ni) {
          if (kind == GPB_KIND_SE) {
            double xa[4], ea[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) xa[u] = __dmul_rn(d2[u], ni);
            exp_nonpos4(xa, ea);
#pragma unroll
            for (int u = 0; u < 4; ++u) s[u] = __dadd_rn(s[u], __dmul_rn(p0, ea[u]));
          } else if (nz == 0u) {
            // white noise: exact zero distance only (the diagonal and duplicated points); s + 0.0 == s otherwise
#pragma unroll
            for (int u = 0; u < 4; ++u) s[u] = __dadd_rn(s[u], (d2[u] == 0.0) ? p0 : 0.0);
          }
        };
        if (nterms_reg <= 3) {
          // the usual kernels (SE + white, SE + SE + white): parameters in registers, no loop over shared memory
          add_term(tk0, tp0, tn0);
          if (nterms_reg > 1) add_term(tk1, tp1, tn1);
          if (nterms_reg > 2) add_term(tk2, tp2, tn2);
        } else {
#pragma unroll 1
          for (int t = 0; t < h.nterms; ++t) add_term(h.kind[t], h.p0[t], ninv2[t]);
        }
        if ((ediag >> 2) == q) {
#pragma unroll
          for (int u = 0; u < 4; ++u) s[u] = __dadd_rn(s[u], ((ediag & 3) == u) ? nugget : 0.0);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) pc[u * step] = s[u];
      }
      continue;
    }

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
      const int Cc = ctj * TILE + cl;
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
        double mu = 0.0;
        if (h.nbeta > 0) {  // MeanFunction.scala:7-10 plane
          mu = __dadd_rn(h.beta[0], __dmul_rn(cx[0][cl], h.beta[1]));
          if (g.dim == 2) mu = __dadd_rn(mu, __dmul_rn(cx[1][cl], h.beta[2]));
        }
        v = cy[cl] - mu;  // GaussianProcess.scala:120-122
      } else {
        v = 0.0;
      }
      M[(size_t)Cc * g.ld + R] = v;
    }
  }
}

void launch_build(cudaStream_t st, const Geom& g, const double* x, const double* y,
                  const double* xstar, const Hyper* hyper, double nugget, double* mats, int mode,
                  int full, const double* rows) {
  static const int num_sms = [] {
    int dev = 0, v = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    return v > 0 ? v : 148;
  }();
  const int tiles = full ? g.rt * g.nt : g.nt * (g.nt + 1) / 2 + (g.rt - g.nt) * g.nt;
  // several tiles per CTA once there are enough CTAs to fill the machine a few times over
  int tpc = 1;
  while (tpc < 8 && (long long)((tiles + 2 * tpc - 1) / (2 * tpc)) * g.batch >= num_sms * 4 * 4) tpc *= 2;
  dim3 grid((tiles + tpc - 1) / tpc, g.batch);
  k_build<<<grid, 256, 0, st>>>(g, x, y, xstar, hyper, nugget, mats, mode, full, rows, tiles, tpc);
}

}  // namespace gpb
