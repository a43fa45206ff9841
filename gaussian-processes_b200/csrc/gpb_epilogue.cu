// gpb_epilogue.cu -- warp-reduced epilogues of the factorised tall system.
//
//   k_ll_epilogue   : Q = u.u, U1 = sum u, S = sum log L_ii   (GaussianProcess.scala:126; the three
//                     reductions both the as-parsed and the intended formula need)
//   k_pred_epilogue : mean_j = V_j . u, var_j = k(0) - V_j . V_j  (Predict.scala:78,89-93)
//   k_zero_upper    : Breeze `cholesky` returns L with a zeroed strict upper triangle
//
// u is the solved rhs row (tile row of appended rows), V the solved K* rows.  HBM-bound, tiny.
#include "gpb_internal.h"

namespace gpb {

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(256)
k_ll_epilogue(Geom g, const double* __restrict__ mats, const int* __restrict__ info, gpb_ll* out) {
  const int b = blockIdx.x;
  const double* M = mats + (size_t)b * g.mat_stride;
  double q = 0.0, u1 = 0.0, s = 0.0;
  for (int c = threadIdx.x; c < g.n; c += blockDim.x) {
    const double u = M[(size_t)c * g.ld + g.rhs_row];
    const double l = M[(size_t)c * g.ld + c];
    q = fma(u, u, q);
    u1 += u;
    s += log(l);
  }
  __shared__ double red[3][8];
  q = warp_sum(q); u1 = warp_sum(u1); s = warp_sum(s);
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) { red[0][w] = q; red[1][w] = u1; red[2][w] = s; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double Q = 0, U1 = 0, S = 0;
    for (int i = 0; i < 8; ++i) { Q += red[0][i]; U1 += red[1][i]; S += red[2][i]; }
    gpb_ll r;
    r.Q = Q; r.U1 = U1; r.S = S; r.n = g.n; r.info = info[b];
    out[b] = r;
  }
}

// one thread per test point j; consecutive threads read consecutive rows of a column (coalesced).
// The n columns are cut into gridDim.y contiguous slices (partial sums in `part`, then a fixed-order
// second pass): enough CTAs to fill the machine at m ~ 10^4, and still bit-reproducible.
__global__ void __launch_bounds__(128)
k_pred_partial(Geom g, const double* __restrict__ mats, double* __restrict__ part, int cols_per_slice) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const double* M = mats;  // predictive is single-proposal
  __shared__ double us[128];
  double mu = 0.0, vv = 0.0;
  const size_t row = (size_t)g.npad + j;
  const int cbeg = blockIdx.y * cols_per_slice, cstop = min(g.n, cbeg + cols_per_slice);
  for (int c0 = cbeg; c0 < cstop; c0 += 128) {
    __syncthreads();
    if (c0 + threadIdx.x < cstop) us[threadIdx.x] = M[(size_t)(c0 + threadIdx.x) * g.ld + g.rhs_row];
    __syncthreads();
    if (j < g.m) {
      const int cend = min(128, cstop - c0);
      for (int c = 0; c < cend; ++c) {
        const double v = M[(size_t)(c0 + c) * g.ld + row];
        mu = fma(v, us[c], mu);
        vv = fma(v, v, vv);
      }
    }
  }
  if (j < g.m) {
    part[((size_t)blockIdx.y * 2 + 0) * g.m + j] = mu;
    part[((size_t)blockIdx.y * 2 + 1) * g.m + j] = vv;
  }
}

__global__ void __launch_bounds__(128)
k_pred_finalize(Geom g, const double* __restrict__ part, int slices, const Hyper* __restrict__ hyper,
                double* __restrict__ mean, double* __restrict__ var) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= g.m) return;
  double mu = 0.0, vv = 0.0;
  for (int s = 0; s < slices; ++s) {
    mu += part[((size_t)s * 2 + 0) * g.m + j];
    vv += part[((size_t)s * 2 + 1) * g.m + j];
  }
  // kyy = covFn(dist(xs, xs)) = covFn(0): every SE contributes h, every White its sigma
  const Hyper& h = hyper[0];
  double k0 = 0.0;
  for (int t = 0; t < h.nterms; ++t) k0 = k0 + h.p0[t];
  mean[j] = mu;
  var[j] = k0 - vv;
}

// Gibbs step of the plane mean (MeanParameters.samplePlane, MeanParameters.scala:74-103): with the
// design-matrix columns X_a carried through the panel solves as appended rows V_a = X_a^T L^-T,
//   X^T K^-1 X = V V^T   and   X^T u  (u = L^-1 y, the reference multiplies X^T by u, :94)
// are d*(d+1)/2 + d reductions over n.  out: G [d*d] (symmetric, row-major) then w [d].
__global__ void __launch_bounds__(256)
k_plane_epilogue(Geom g, const double* __restrict__ mats, const double* __restrict__ rows,
                 double* __restrict__ out) {
  constexpr int MAXD = 4;
  const int d = g.design;
  double G[MAXD][MAXD], w[MAXD];
#pragma unroll
  for (int a = 0; a < MAXD; ++a) { w[a] = 0.0;
#pragma unroll
    for (int b = 0; b < MAXD; ++b) G[a][b] = 0.0; }
  const size_t r0 = (size_t)g.npad + g.m;
  mats += (size_t)blockIdx.x * g.mat_stride;          // one CTA per proposal (the MH loop's Gibbs step is batched)
  out += (size_t)blockIdx.x * (d * d + d);
  for (int c = threadIdx.x; c < g.n; c += blockDim.x) {
    const double* col = mats + (size_t)c * g.ld;
    const double u = col[g.rhs_row];
    double v[MAXD];
#pragma unroll
    for (int a = 0; a < MAXD; ++a) v[a] = (a < d) ? col[r0 + a] : 0.0;
#pragma unroll
    for (int a = 0; a < MAXD; ++a) {
      if (a < d) w[a] = fma(rows[(size_t)a * g.n + c], u, w[a]);
#pragma unroll
      for (int b = a; b < MAXD; ++b) G[a][b] = fma(v[a], v[b], G[a][b]);
    }
  }
  __shared__ double red[8][MAXD * MAXD + MAXD];
  const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int a = 0; a < MAXD; ++a) {
    const double ws = warp_sum(w[a]);
    if (lane == 0) red[wid][MAXD * MAXD + a] = ws;
#pragma unroll
    for (int b = a; b < MAXD; ++b) {
      const double gs = warp_sum(G[a][b]);
      if (lane == 0) red[wid][a * MAXD + b] = gs;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int a = 0; a < d; ++a) {
      for (int b = a; b < d; ++b) {
        double t = 0.0;
        for (int i = 0; i < 8; ++i) t += red[i][a * MAXD + b];
        out[a * d + b] = t; out[b * d + a] = t;
      }
      double t = 0.0;
      for (int i = 0; i < 8; ++i) t += red[i][MAXD * MAXD + a];
      out[d * d + a] = t;
    }
  }
}
void launch_plane_epilogue(cudaStream_t st, const Geom& g, const double* mats, const double* rows,
                           double* out) {
  k_plane_epilogue<<<g.batch, 256, 0, st>>>(g, mats, rows, out);
}

// GaussianProcess.draw (GaussianProcess.scala:67-83) with the Cholesky root: out = mu + L z, one
// thread per row, columns ascending (coalesced along rows of the column-major factor).
__global__ void __launch_bounds__(128)
k_draw(Geom g, const double* __restrict__ mats, const double* __restrict__ z,
       const double* __restrict__ mu, double* __restrict__ out) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  __shared__ double zs[128];
  double acc = 0.0;
  const int cmax = min(g.n, (int)(blockIdx.x + 1) * (int)blockDim.x);
  for (int c0 = 0; c0 < cmax; c0 += 128) {
    __syncthreads();
    if (c0 + threadIdx.x < g.n) zs[threadIdx.x] = z[c0 + threadIdx.x];
    __syncthreads();
    if (r < g.n) {
      const int cend = min(128, r + 1 - c0);
      for (int c = 0; c < cend; ++c) acc = fma(mats[(size_t)(c0 + c) * g.ld + r], zs[c], acc);
    }
  }
  if (r < g.n) out[r] = mu[r] + acc;
}
void launch_draw(cudaStream_t st, const Geom& g, const double* mats, const double* z, const double* mu,
                 double* out) {
  k_draw<<<(g.n + 127) / 128, 128, 0, st>>>(g, mats, z, mu, out);
}

__global__ void k_zero_upper(Geom g, double* __restrict__ mats) {
  const int c = blockIdx.x;
  double* col = mats + (size_t)blockIdx.y * g.mat_stride + (size_t)c * g.ld;
  for (int r = threadIdx.x; r < c && r < g.n; r += blockDim.x) col[r] = 0.0;
}

__global__ void k_pad_identity(Geom g, double* __restrict__ mats) {
  const int r = g.n + blockIdx.x * blockDim.x + threadIdx.x;
  if (r < g.npad) mats[(size_t)blockIdx.y * g.mat_stride + (size_t)r * g.ld + r] = 1.0;
}
void launch_pad_identity(cudaStream_t st, const Geom& g, double* mats) {
  if (g.npad == g.n) return;
  dim3 grid((g.npad - g.n + 63) / 64, g.batch);
  k_pad_identity<<<grid, 64, 0, st>>>(g, mats);
}

void launch_ll_epilogue(cudaStream_t st, const Geom& g, const double* mats, const int* info,
                        gpb_ll* out) {
  k_ll_epilogue<<<g.batch, 256, 0, st>>>(g, mats, info, out);
}
int pred_slices(const Geom& g) {
  int s = (g.n + 2047) / 2048;
  return s < 1 ? 1 : (s > 32 ? 32 : s);
}
void launch_pred_epilogue(cudaStream_t st, const Geom& g, const double* mats, const Hyper* hyper,
                          double* part, double* mean, double* var) {
  const int slices = pred_slices(g);
  const int cols = ((g.n + slices - 1) / slices + 127) / 128 * 128;
  dim3 grid((g.m + 127) / 128, slices);
  k_pred_partial<<<grid, 128, 0, st>>>(g, mats, part, cols);
  k_pred_finalize<<<(g.m + 127) / 128, 128, 0, st>>>(g, part, slices, hyper, mean, var);
}
void launch_zero_upper(cudaStream_t st, const Geom& g, double* mats) {
  dim3 grid(g.n, g.batch);
  k_zero_upper<<<grid, 128, 0, st>>>(g, mats);
}

}  // namespace gpb
